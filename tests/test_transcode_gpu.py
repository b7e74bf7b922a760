"""GPU parity tests for the transcode path, through the C ABI (include/ktx_b200_uastc.h)."""
import json, os
import numpy as np
import pytest
import refutil as R

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def K():
    import ktx_software_b200 as K
    K.init()
    return K


def _hex(h):
    return np.frombuffer(bytes.fromhex(h), np.uint8).reshape(-1, 16)


def test_cts_goldens(K):
    cts = json.load(open(os.path.join(GOLD, "cts_uastc.json")))["transcode"]
    for sub, c in cts.items():
        u = _hex(c["uastc"])
        assert (K.transcode(u, K.TTF_BC7_RGBA).reshape(-1, 16) == _hex(c["bc7"])).all(), sub
        assert (K.transcode(u, K.TTF_ASTC_4x4_RGBA).reshape(-1, 16) == _hex(c["astc"])).all(), sub


def test_golden_corpus_bit_exact(K):
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    assert (K.transcode(g["uastc"], K.TTF_BC7_RGBA).reshape(-1, 16) == g["bc7"]).all()
    assert (K.transcode(g["uastc"], K.TTF_ASTC_4x4_RGBA).reshape(-1, 16) == g["astc"]).all()


def test_invalid_block_fails_like_reference(K):
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    good, inv = g["uastc"][:64], g["invalid"][:1]
    # three spans; the bad block sits in span 1
    buf = np.concatenate([good, good[:10], inv, good[:5], good]).reshape(-1)
    spans = [(0, 0, 64), (64 * 16, 64 * 16, 16), (80 * 16, 80 * 16, 64)]
    with pytest.raises(K.B200Error) as ei:
        K.transcode(buf, K.TTF_BC7_RGBA, spans=spans)
    assert ei.value.code == -5 and ei.value.bad_span == 1


def test_empty_and_ragged_spans(K):
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    u = g["uastc"][:1000]
    assert K.transcode(np.zeros(0, np.uint8), K.TTF_BC7_RGBA, spans=[]).size == 0
    # non-contiguous, reordered destination; zero-length span in the middle
    spans = [(0, 500 * 16, 300), (300 * 16, 0, 0), (300 * 16, 0, 500), (800 * 16, 800 * 16, 200)]
    out = K.transcode(u, K.TTF_ASTC_4x4_RGBA, spans=spans).reshape(-1, 16)
    assert (out[500:800] == g["astc"][:300]).all()
    assert (out[:500] == g["astc"][300:800]).all()
    assert (out[800:1000] == g["astc"][800:1000]).all()


@pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not shipped")
def test_large_random_vs_reference(K):
    """2M blocks: reference-encoded image blocks tiled + random payloads; multi-slab pipeline path."""
    px = R.ref_extract_blocks(R.img_mix(512, 512, 11))
    enc = np.concatenate([R.ref_encode_blocks(px, l) for l in (0, 1, 2, 3)])
    rp = R.random_payload_blocks(5, 4096, include_invalid=False)
    ok = np.array([R.ref_transcode(rp[k:k + 1], 0)[1] < 0 for k in range(len(rp))])
    base = np.concatenate([enc, rp[ok]])
    rng = np.random.default_rng(3)
    big = base[rng.integers(0, len(base), 2_200_000)]
    for tgt, fmt in ((K.TTF_BC7_RGBA, 0), (K.TTF_ASTC_4x4_RGBA, 1)):
        want = R.ref_transcode(base, fmt)[0]
        out = K.transcode(big, tgt).reshape(-1, 16)
        # permutation property: transcode is per-block, so out == want[perm]
        rng2 = np.random.default_rng(3)
        perm = rng2.integers(0, len(base), 2_200_000)
        assert (out == want[perm]).all()


def test_device_variant_matches_host_variant(K):
    import torch
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    u = torch.from_numpy(g["uastc"].copy()).cuda()
    out = torch.empty_like(u)
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    K.transcode_dev(K.TTF_BC7_RGBA, u.data_ptr(), out.data_ptr(), u.shape[0], 0, bad.data_ptr())
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == g["bc7"]).all()
    assert bad.item() == -1  # UINT64_MAX
    # a bad block at index 7
    u2 = u.clone()
    u2[7] = torch.from_numpy(g["invalid"][0].copy()).cuda()
    K.transcode_dev(K.TTF_ASTC_4x4_RGBA, u2.data_ptr(), out.data_ptr(), u2.shape[0], 0, bad.data_ptr())
    torch.cuda.synchronize()
    assert bad.item() == 7


@pytest.mark.parametrize("n", [1, 31, 1023, 1024, 1025, 2048 + 17, 300000])
def test_tile_boundaries_and_mode_binning(K, n):
    """The tiled kernel bins each 1024-block tile by mode; sizes around the tile edge and mode-sorted / mode-shuffled
    inputs must give the same bytes as the per-block golden answers."""
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    rng = np.random.default_rng(n)
    pick = rng.integers(0, len(g["uastc"]), n)
    if n == 300000:
        pick = np.sort(pick)  # long single-mode runs (the corpus is grouped by source)
    u = g["uastc"][pick]
    assert (K.transcode(u, K.TTF_BC7_RGBA).reshape(-1, 16) == g["bc7"][pick]).all()
    assert (K.transcode(u, K.TTF_ASTC_4x4_RGBA).reshape(-1, 16) == g["astc"][pick]).all()


def test_many_invalid_blocks_report_first(K):
    import torch
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    u = torch.from_numpy(g["uastc"][:5000].copy()).cuda()
    inv = torch.from_numpy(g["invalid"][:3].copy()).cuda()
    u[4097] = inv[0]; u[1500] = inv[1]; u[4999] = inv[2]
    out = torch.empty_like(u)
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    K.transcode_dev(K.TTF_BC7_RGBA, u.data_ptr(), out.data_ptr(), u.shape[0], 0, bad.data_ptr())
    torch.cuda.synchronize()
    assert bad.item() == 1500
    ok = np.ones(5000, bool); ok[[1500, 4097, 4999]] = False
    assert (out.cpu().numpy()[ok] == g["bc7"][:5000][ok]).all()
    assert (out.cpu().numpy()[~ok] == 0).all()
