"""GPU parity tests for the UASTC RDO post-pass (b200_uastc_rdo_blocks_dev and b200_uastc_encode with rdo_enable).
north_star puts RDO in the PSNR tier (|dPSNR| <= 0.01 dB + mode agreement); the CUDA walk restates the reference's
serial chain exactly, so these tests assert byte equality with oracle/_ref and report PSNR on top."""
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


def _psnr(a, b):
    mse = ((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean()
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def _rdo_dev(K, u, px, flags, **kw):
    import torch
    d_u = torch.from_numpy(np.ascontiguousarray(u)).cuda()
    d_px = torch.from_numpy(np.ascontiguousarray(px)).cuda()
    K.rdo_blocks_dev(d_u.data_ptr(), d_px.data_ptr(), d_u.shape[0], flags, **kw)
    torch.cuda.synchronize()
    return d_u.cpu().numpy()


@needs_ref
@pytest.mark.parametrize("flags,kw,jobs", [(2 | 512, {}, 4), (3, {"rdo_lambda": 3.0}, 0), (0 | 512, {"rdo_lambda": 0.5, "dict_size": 1024}, 3),
                                           (2 | 8, {"rdo_lambda": 2.0, "dict_size": 32768}, 4), (1, {"rdo_lambda": 1.0, "dict_size": 16}, 1)])
def test_rdo_matches_reference(K, flags, kw, jobs):
    px = np.concatenate([R.ref_extract_blocks(R.img_mix(192, 192, 3)), R.adversarial_blocks(5, 40), R.ref_extract_blocks(R.img_smooth(64, 64, 1))])
    u = R.ref_encode_blocks(px, flags)
    want = R.ref_rdo(u, px, flags, 4, jobs, lam=kw.get("rdo_lambda", 1.0), dict_size=kw.get("dict_size", 4096))
    got = _rdo_dev(K, u, px, flags, total_jobs=jobs, **kw)
    assert (want != u).any(1).sum() > (20 if kw.get("dict_size", 4096) >= 1024 else 0), "the case must actually modify blocks"
    bad = np.nonzero((got != want).any(1))[0]
    pa, pb = _psnr(R.ref_unpack_rgba(got)[0], px), _psnr(R.ref_unpack_rgba(want)[0], px)
    assert abs(pa - pb) <= 0.01
    assert len(bad) == 0, "%d/%d blocks differ from the reference, first %s (PSNR %.3f vs %.3f)" % (len(bad), len(px), bad[:8], pa, pb)


@needs_ref
def test_rdo_chunk_boundaries(K):
    """Chains longer than one launch chunk (8192 steps) and a window that reaches back across the chunk edge."""
    img = R.img_mix(512, 384, 9)
    px = R.ref_extract_blocks(img)                      # 12288 blocks -> one chain of 12288, or 4 of 3072
    u = R.ref_encode_blocks(px, 1 | 512)
    for jobs in (0, 4):
        want = R.ref_rdo(u, px, 1 | 512, 4, jobs, lam=2.0)
        got = _rdo_dev(K, u, px, 1 | 512, total_jobs=jobs, rdo_lambda=2.0)
        assert (got == want).all(), jobs


@needs_ref
def test_encode_with_rdo_through_host_api(K):
    """b200_uastc_encode with rdo_enable: several images (job ranges never cross an image), ragged sizes."""
    imgs = [R.img_mix(128, 96, 3), R.img_mix(70, 37, 4), R.img_mix(2, 2, 5), R.img_smooth(64, 64, 2)]
    flags = 2 | 512
    outs, any_alpha = K.encode_images(imgs, flags, rdo=True, rdo_lambda=1.5, total_jobs=4)
    assert any_alpha
    for im, o in zip(imgs, outs):
        px = R.ref_extract_blocks(im)
        want = R.ref_rdo(R.ref_encode_blocks(px, flags), px, flags, 4, 4, lam=1.5)
        assert (o == want).all()


def test_rdo_properties_without_oracle(K):
    """Size-independent properties: output stays valid UASTC (transcodes), solid blocks untouched, idempotent input
    handling (same input -> same bytes), and lambda -> 0+ leaves (almost) everything alone while a large lambda
    rewrites many blocks and never raises the per-block mode."""
    img = R.img_mix(256, 256, 11)
    a, _ = K.encode_images([img], 2)
    r1, _ = K.encode_images([img], 2, rdo=True, rdo_lambda=4.0, total_jobs=4)
    r2, _ = K.encode_images([img], 2, rdo=True, rdo_lambda=4.0, total_jobs=4)
    assert (r1[0] == r2[0]).all()
    K.transcode(r1[0], K.TTF_BC7_RGBA)
    changed = (r1[0] != a[0]).any(1)
    assert changed.sum() > 100
    assert (R.block_modes(r1[0]) == R.block_modes(a[0])).all()      # RDO only rewrites selectors + hints (+ mode-0 endpoints)
    assert not changed[R.block_modes(a[0]) == 8].any()
    tiny, _ = K.encode_images([img], 2, rdo=True, rdo_lambda=1e-6, total_jobs=4)
    assert (tiny[0] != a[0]).any(1).sum() < changed.sum()
