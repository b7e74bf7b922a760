"""GPU parity tests for the hint-driven transcode targets (KTX_TTF_ETC1_RGB / ETC2_RGBA / BC1_RGB / BC3_RGBA / BC4_R / BC5_RG / ETC2_EAC_R11 / ETC2_EAC_RG11)
through the C ABI (b200_uastc_transcode_ex, b200_uastc_transcode_dev_ex).  SURVEY 8(f).2; bar: bit-exact with the
reference's transcode_image -> transcode_slice (XC/basisu_transcoder.cpp:10141-10180) -> transcode_uastc_to_etc1 /
_etc2_rgba / _bc1 / _bc3 / _bc4 / _bc5, including the high-quality and alpha-to-opaque decode flags."""
import os
import numpy as np
import pytest
import refutil as R

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
needs_ref = pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not shipped")
NAMES = {0: "etc1", 1: "etc2", 2: "bc1", 3: "bc3", 4: "bc4", 5: "bc5", 20: "r11", 21: "rg11"}


@pytest.fixture(scope="module")
def K():
    import ktx_software_b200 as K
    K.init()
    return K


@pytest.fixture(scope="module")
def gold():
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    h = np.load(os.path.join(GOLD, "hinted_golden.npz"))
    return g["uastc"][h["index"]], h, g["invalid"]


@pytest.mark.parametrize("tf", sorted(NAMES))
def test_golden_corpus_bit_exact(K, gold, tf):
    u, h, _ = gold
    got = K.transcode(u, tf, decode_flags=0).reshape(-1, K.target_block_bytes(tf))
    assert (got == h[NAMES[tf]]).all()


def test_decode_flags_on_golden_corpus(K, gold):
    u, h, _ = gold
    assert (K.transcode(u, K.TTF_BC1_RGB, decode_flags=K.TF_HIGH_QUALITY).reshape(-1, 8) == h["bc1_hq"]).all()
    assert (K.transcode(u, K.TTF_BC3_RGBA, decode_flags=K.TF_HIGH_QUALITY).reshape(-1, 16) == h["bc3_hq"]).all()
    assert (K.transcode(u, K.TTF_ETC2_EAC_R11, decode_flags=K.TF_HIGH_QUALITY).reshape(-1, 8) == h["r11_hq"]).all()
    assert (K.transcode(u, K.TTF_ETC2_EAC_RG11, decode_flags=K.TF_HIGH_QUALITY).reshape(-1, 16) == h["rg11_hq"]).all()
    a2o = K.TF_TRANSCODE_ALPHA_DATA_TO_OPAQUE_FORMATS
    assert (K.transcode(u, K.TTF_ETC2_EAC_R11, decode_flags=a2o, has_alpha=True).reshape(-1, 8) == h["r11_alpha"]).all()
    assert (K.transcode(u, K.TTF_BC4_R, decode_flags=a2o, has_alpha=True).reshape(-1, 8) == h["bc4_alpha"]).all()
    # without alpha content the flag changes nothing (XC:10417)
    assert (K.transcode(u, K.TTF_BC4_R, decode_flags=a2o, has_alpha=False).reshape(-1, 8) == h["bc4"]).all()
    # ETC1 from the alpha channel: transcode_uastc_to_etc1(blk, dst, 3) (XC:10142-10147, :16560) -- no hints, a search of its own
    assert (K.transcode(u, K.TTF_ETC1_RGB, decode_flags=a2o, has_alpha=True).reshape(-1, 8) == h["etc1_alpha"]).all()
    assert (K.transcode(u, K.TTF_ETC1_RGB, decode_flags=a2o, has_alpha=False).reshape(-1, 8) == h["etc1"]).all()


@pytest.mark.parametrize("tf", sorted(NAMES))
def test_spans_ragged_tiles_and_bad_span(K, gold, tf):
    """Several spans (odd block counts, a tile boundary inside a span, an empty span), destination offsets in the
    target's own block size; then an invalid block names its span."""
    u, h, inv = gold
    ob = K.target_block_bytes(tf)
    want = h[NAMES[tf]]
    counts = [1, 4097, 0, 333, 4096, 7]
    spans, so, do = [], 0, 0
    for n in counts:
        spans.append((so * 16, do * ob, n)); so += n; do += n
    got = K.transcode(u[:so], tf, spans=spans, decode_flags=0).reshape(-1, ob)
    assert (got == want[:so]).all()
    buf = u[:so].copy()
    buf[1 + 4097 + 100] = inv[0]   # inside span 3
    with pytest.raises(K.B200Error) as ei:
        K.transcode(buf, tf, spans=spans, decode_flags=0)
    assert ei.value.code == -5 and ei.value.bad_span == 3


def test_device_entry(K, gold):
    import torch
    u, h, inv = gold
    d_src = torch.from_numpy(u.copy()).cuda()
    d_bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    for tf in sorted(NAMES):
        ob = K.target_block_bytes(tf)
        for n in (u.shape[0], 4095, 1):
            d_dst = torch.zeros(n * ob + 16, dtype=torch.uint8, device="cuda")
            K.transcode_dev_ex(tf, 0, True, d_src.data_ptr(), d_dst.data_ptr(), n, 0, d_bad.data_ptr())
            torch.cuda.synchronize()
            out = d_dst.cpu().numpy()
            assert (out[:n * ob].reshape(-1, ob) == h[NAMES[tf]][:n]).all(), (tf, n)
            assert (out[n * ob:] == 0).all(), "bytes past the last block must not be touched"
            assert d_bad.item() == -1
    bad = u[:5000].copy(); bad[4321] = inv[0]
    d_b = torch.from_numpy(bad).cuda()
    d_dst = torch.zeros(5000 * 8, dtype=torch.uint8, device="cuda")
    K.transcode_dev_ex(K.TTF_BC1_RGB, 0, True, d_b.data_ptr(), d_dst.data_ptr(), 5000, 0, d_bad.data_ptr())
    torch.cuda.synchronize()
    assert d_bad.item() == 4321
    assert (d_dst.cpu().numpy().reshape(-1, 8)[4321] == 0).all()


@needs_ref
@pytest.mark.parametrize("level", [0, 2])
def test_image_blocks_vs_reference_all_targets(K, level):
    """Reference-encoded image blocks (real hint bits) through every target and flag combination, live against oracle/_ref."""
    blocks = R.ref_encode_blocks(R.ref_extract_blocks(R.img_mix(512, 256, 40 + level)), level)
    for tf in sorted(NAMES):
        for fl in (0, 32, 4, 36):
            for has_alpha in (True, False):
                if tf == 0 and (fl & 4) and has_alpha:
                    continue
                want, ok = R.ref_transcode_hinted(blocks, tf, fl, has_alpha)
                assert ok
                got = K.transcode(blocks, tf, decode_flags=fl, has_alpha=has_alpha).reshape(want.shape)
                assert (got == want).all(), (tf, fl, has_alpha)


def test_empty(K):
    assert K.transcode(np.zeros(0, np.uint8), K.TTF_BC1_RGB, decode_flags=0).size == 0
