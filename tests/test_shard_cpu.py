"""Host-side sharding logic (no GPU): the partition covers every block exactly once, is balanced, and a world_size-2
gloo run in which each rank handles its shard (with the CPU port standing in for the device, as the checker) gathers
to the same bytes as the unsharded run."""
import os, sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import refutil as R
from ktx_software_b200 import shard as S


def _chain(size):
    out = []
    while True:
        out.append((size, size))
        if size == 1:
            return out
        size //= 2


@pytest.mark.parametrize("n_gpus", [1, 2, 4, 8])
def test_plan_covers_everything_once_and_balances(n_gpus):
    sizes = _chain(8192)                       # C2: level 0 holds 75 % of the work and must be slab-split
    shards = S.plan(sizes, n_gpus)
    seen = {}
    for g, lst in enumerate(shards):
        for u in lst:
            for r in range(u.row0, u.row0 + u.rows):
                assert (u.image, r) not in seen
                seen[(u.image, r)] = g
    want = sum(S.image_blocks(w, h)[1] for w, h in sizes)
    assert len(seen) == want
    loads = [sum(u.n_blocks for u in lst) for lst in shards]
    assert sum(loads) == sum(bx * by for bx, by in (S.image_blocks(w, h) for w, h in sizes)) == 5592407
    if n_gpus > 1:
        assert max(loads) <= 1.10 * (sum(loads) / n_gpus)


def test_plan_c4_whole_faces():
    shards = S.plan([(2048, 2048)] * 384, 8)   # C4: 48 whole faces per GPU, nothing split
    assert all(len(lst) == 48 and all(u.row0 == 0 and u.rows == 512 for u in lst) for lst in shards)


def test_ragged_and_tiny_images():
    sizes = [(70, 37), (2, 2), (1, 1), (5, 1000)]
    for n in (1, 2, 3):
        shards = S.plan(sizes, n)
        rows = sorted((u.image, r) for lst in shards for u in lst for r in range(u.row0, u.row0 + u.rows))
        assert rows == sorted((i, r) for i, (w, h) in enumerate(sizes) for r in range(S.image_blocks(w, h)[1]))


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    imgs = [R.img_mix(72, 50, 3), R.img_mix(32, 32, 4), R.img_mix(2, 2, 5)]
    shards = S.plan([(im.shape[1], im.shape[0]) for im in imgs], world)
    mine = []
    for u in shards[rank]:
        sub = S.slab_view(imgs[u.image], u)
        h, w = sub.shape[:2]
        pad = np.pad(sub, ((0, (-h) % 4), (0, (-w) % 4), (0, 0)), mode="edge")
        px = pad.reshape(pad.shape[0] // 4, 4, pad.shape[1] // 4, 4, 4).transpose(0, 2, 1, 3, 4).reshape(-1, 64)
        mine.append((u.image, u.out_offset, R.port_encode_blocks(px, 1).tobytes()))
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)          # the "host gathers the slabs" step; no data-path collective
    if rank == 0:
        outs = [bytearray(S.image_blocks(im.shape[1], im.shape[0])[0] * S.image_blocks(im.shape[1], im.shape[0])[1] * 16) for im in imgs]
        for lst in gathered:
            for image, off, data in lst:
                outs[image][off:off + len(data)] = data
        np.save(tmp, np.frombuffer(b"".join(bytes(o) for o in outs), np.uint8))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_gather_matches_unsharded(tmp_path):
    tmp = str(tmp_path / "out.npy")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, tmp), nprocs=2, join=True)
    got = np.load(tmp)
    imgs = [R.img_mix(72, 50, 3), R.img_mix(32, 32, 4), R.img_mix(2, 2, 5)]
    want = []
    for im in imgs:
        h, w = im.shape[:2]
        pad = np.pad(im, ((0, (-h) % 4), (0, (-w) % 4), (0, 0)), mode="edge")
        px = pad.reshape(pad.shape[0] // 4, 4, pad.shape[1] // 4, 4, 4).transpose(0, 2, 1, 3, 4).reshape(-1, 64)
        want.append(R.port_encode_blocks(px, 1).reshape(-1))
    assert (got == np.concatenate(want)).all()
