"""GPU parity tests for the uncompressed transcode targets (KTX_TTF_RGBA32 / RGB565 / BGR565 / RGBA4444), through the
C ABI (b200_uastc_decode, b200_uastc_decode_image_dev).  SURVEY 8(f).1; bar: bit-exact with the reference's
transcode_image (XC/basisu_transcoder.cpp:10331 -> transcode_slice :10208-10292 -> unpack_uastc :15378)."""
import os
import numpy as np
import pytest
import refutil as R

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
needs_ref = pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not shipped")


@pytest.fixture(scope="module")
def K():
    import ktx_software_b200 as K
    K.init()
    return K


def _strip(K, blocks, target=13):
    """n blocks as one 4-texel-high image (n*4 x 4): the decode of block i lands at columns 4i..4i+3."""
    n = blocks.shape[0]
    return K.decode(blocks, target, [(0, n * 4, 4)])[0]


def test_golden_corpus_bit_exact(K):
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    img = _strip(K, g["uastc"])
    assert (img == R.texels_to_image(g["rgba"], g["uastc"].shape[0] * 4, 4)).all()


def test_16bit_targets_match_packers_on_golden_corpus(K):
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    want = R.texels_to_image(g["rgba"], g["uastc"].shape[0] * 4, 4)
    for tf in (K.TTF_RGB565, K.TTF_BGR565, K.TTF_RGBA4444):
        assert (_strip(K, g["uastc"], tf) == R.pack_image16(want, tf)).all(), tf


def test_invalid_block_names_the_image(K):
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    good, inv = g["uastc"][:64], g["invalid"][:1]
    buf = np.concatenate([good, good[:15], inv, good])   # image 0: 8x8 blocks, image 1: 4x4 with a bad block, image 2: 8x8
    images = [(0, 32, 32), (64 * 16, 16, 16), (80 * 16, 32, 32)]
    with pytest.raises(K.B200Error) as ei:
        K.decode(buf, K.TTF_RGBA32, images)
    assert ei.value.code == -5 and ei.value.bad_image == 1


def test_empty_inputs(K):
    assert K.decode(np.zeros(0, np.uint8), K.TTF_RGBA32, []) == []
    out = K.decode(np.zeros(16, np.uint8), K.TTF_RGBA32, [(0, 0, 0)])
    assert out[0].size == 0


@needs_ref
@pytest.mark.parametrize("w,h", [(256, 256), (13, 7), (1, 1), (2, 2), (130, 5), (1023, 517)])
def test_images_vs_reference_all_targets(K, w, h):
    img = R.img_mix(w, h, 21)
    blocks = R.ref_encode_blocks(R.ref_extract_blocks(img), 1)
    for tf in (K.TTF_RGBA32, K.TTF_RGB565, K.TTF_BGR565, K.TTF_RGBA4444):
        want, ok = R.ref_transcode_image(blocks, w, h, tf)
        assert ok
        got = K.decode(blocks, tf, [(0, w, h)])[0]
        assert got.shape == want.shape and (got == want).all(), (w, h, tf)


@needs_ref
def test_mip_chain_in_one_call(K):
    """Several images (a mip chain down to 1x1) in one call, offsets as libktx lays out pData."""
    base = R.img_mix(96, 80, 5)
    chain, blocks, images, off = [], [], [], 0
    w, h, im = 96, 80, base
    while True:
        b = R.ref_encode_blocks(R.ref_extract_blocks(im), 0)
        chain.append((w, h, b)); blocks.append(b); images.append((off, w, h)); off += b.shape[0] * 16
        if w == 1 and h == 1:
            break
        w, h = max(1, w // 2), max(1, h // 2)
        im = R.img_mix(w, h, 5 + len(chain))
    outs = K.decode(np.concatenate(blocks), K.TTF_RGBA32, images)
    for (w, h, b), got in zip(chain, outs):
        want, ok = R.ref_transcode_image(b, w, h, 13)
        assert ok and (got == want).all(), (w, h)


def test_device_entry_pitched_and_slabbed(K):
    """Device-pointer entry with a destination pitch wider than the image, unaligned pitch (scalar store path), and an
    image large enough to be cut into several slabs by the host entry."""
    import torch
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    n = 20000
    blocks = g["uastc"][:n]
    bx, by = 200, 100
    w, h = bx * 4 - 1, by * 4 - 2
    want = R.texels_to_image(g["rgba"][:n], w, h)
    d_src = torch.from_numpy(blocks.copy()).cuda()
    for pitch in (w, w + 1, 1024):
        d_dst = torch.zeros(h * pitch * 4, dtype=torch.uint8, device="cuda")
        d_bad = torch.zeros(1, dtype=torch.int64, device="cuda")
        K.decode_image_dev(K.TTF_RGBA32, d_src.data_ptr(), d_dst.data_ptr(), w, h, pitch, 0, d_bad.data_ptr())
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy().reshape(h, pitch, 4)
        assert (got[:, :w] == want).all(), pitch
        assert (got[:, w:] == 0).all(), "texels outside the image must not be touched"
        assert d_bad.item() == -1
    # host entry, many slabs: 2048 x 2048 texels = 262144 blocks from a repeated corpus
    rep = np.tile(g["uastc"][:16384], (16, 1))
    got = K.decode(rep, K.TTF_RGBA32, [(0, 2048, 2048)])[0]
    want = R.texels_to_image(np.tile(g["rgba"][:16384], (16, 1)), 2048, 2048)
    assert (got == want).all()
