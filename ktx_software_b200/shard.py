"""Work partitioning across the GPUs of one box (north_star): by image (level x layer x face/slice), then
by block-row slabs.  Pure host logic: no compute, no collective -- blocks are independent, the host gathers the slabs.

A *unit* is (image index, first block row, number of block rows).  A slab of whole block rows is a contiguous byte
range both in the RGBA source rows and in the UASTC output, so a rank can encode it with one
b200_uastc_encode / _encode_image_dev call on a sub-image and write the result at `out_offset`.
"""
from dataclasses import dataclass
from typing import List, Sequence, Tuple


@dataclass(frozen=True)
class Unit:
    image: int        # index into the caller's image list
    row0: int         # first block row
    rows: int         # number of block rows
    blocks_x: int     # blocks per row of that image

    @property
    def n_blocks(self) -> int:
        return self.rows * self.blocks_x

    @property
    def out_offset(self) -> int:  # byte offset of the slab inside the image's UASTC output
        return self.row0 * self.blocks_x * 16


def image_blocks(width: int, height: int) -> Tuple[int, int]:
    return (max(1, width) + 3) // 4, (max(1, height) + 3) // 4


def make_units(sizes: Sequence[Tuple[int, int]], n_gpus: int) -> List[Unit]:
    """sizes: (width, height) per image.  Any image larger than total_blocks / (4 * n_gpus) is split into block-row
    slabs of about that size; smaller images stay whole."""
    dims = [image_blocks(w, h) for w, h in sizes]
    total = sum(bx * by for bx, by in dims)
    limit = max(1, total // (4 * max(1, n_gpus)))
    units = []
    for i, (bx, by) in enumerate(dims):
        if bx * by <= limit or by == 1:
            units.append(Unit(i, 0, by, bx))
            continue
        rows_per = max(1, limit // bx)
        r = 0
        while r < by:
            n = min(rows_per, by - r)
            units.append(Unit(i, r, n, bx))
            r += n
    return units


def assign(units: Sequence[Unit], n_gpus: int) -> List[List[Unit]]:
    """Greedy longest-processing-time assignment: units by block count descending, each to the least loaded GPU.
    Deterministic (ties by unit order, then by GPU index)."""
    order = sorted(range(len(units)), key=lambda k: (-units[k].n_blocks, k))
    load = [0] * n_gpus
    out: List[List[Unit]] = [[] for _ in range(n_gpus)]
    for k in order:
        g = min(range(n_gpus), key=lambda j: (load[j], j))
        out[g].append(units[k])
        load[g] += units[k].n_blocks
    for lst in out:
        lst.sort(key=lambda u: (u.image, u.row0))
    return out


def plan(sizes: Sequence[Tuple[int, int]], n_gpus: int) -> List[List[Unit]]:
    return assign(make_units(sizes, n_gpus), n_gpus)


def slab_view(img, unit: Unit):
    """The pixel rows of `img` ([h,w,4] array) that `unit` covers (edge replication is the encoder's business: the
    last slab simply carries the remaining, possibly non-multiple-of-4, rows)."""
    y0 = unit.row0 * 4
    y1 = min(img.shape[0], (unit.row0 + unit.rows) * 4)
    return img[y0:y1]
