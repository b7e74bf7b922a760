"""b200_uastc_encode_image_dev_noalloc (no counterpart in the reference): the device
image encode with the kernel's scratch in a static device array instead of a per-launch allocation.  Its output must equal
b200_uastc_encode_image_dev's byte for byte (the encode tier is bit-exact), on two calls in a row -- the second call finds the
first call's contents in the shared scratch.  Collected last on purpose (file name)."""
import os
import numpy as np
import pytest
import refutil as R

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    import ktx_software_b200 as K
    K.init()
    return K


def _both(K, img, flags):
    import torch
    h, w = img.shape[:2]
    n = ((w + 3) // 4) * ((h + 3) // 4)
    d_img = torch.from_numpy(np.ascontiguousarray(img)).cuda()
    a = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
    b = torch.full((n, 16), 0xAB, dtype=torch.uint8, device="cuda")
    fa = torch.zeros(1, dtype=torch.int32, device="cuda")
    fb = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    K.encode_image_dev(d_img.data_ptr(), w, h, a.data_ptr(), flags, st, d_any_alpha=fa.data_ptr())
    K.encode_image_dev_noalloc(d_img.data_ptr(), w, h, b.data_ptr(), flags, st, d_any_alpha=fb.data_ptr())
    torch.cuda.synchronize()
    return a.cpu().numpy(), b.cpu().numpy(), int(fa.item()), int(fb.item())


def test_noalloc_matches_existing_entry_point_twice_in_a_row(K):
    # first call: many tiles, every block class; second call: a different, ragged image and another level over the same scratch
    for img, flags in ((R.img_mix(640, 448, 31), 2), (R.img_mix(203, 77, 32), 1), (R.img_noise(64, 64, 3), 3), (R.img_smooth(8, 4, 1), 0)):
        want, got, fa, fb = _both(K, img, flags)
        bad = np.nonzero((want != got).any(1))[0]
        assert len(bad) == 0, (img.shape, flags, bad[:8])
        assert fa == fb
    # the synchronous form (stream = NULL: legacy default stream)
    import torch
    img = R.img_mix(96, 64, 5)
    d_img = torch.from_numpy(img).cuda()
    out = torch.zeros((24 * 16, 16), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    K.encode_image_dev_noalloc(d_img.data_ptr(), 96, 64, out.data_ptr(), 2)
    want, _ = K.encode_images([img], 2)
    assert (out.cpu().numpy() == want[0]).all()
    if R.have_ref():
        assert (want[0] == R.ref_encode_blocks(R.ref_extract_blocks(img), 2)).all()


def test_noalloc_limits_are_rejected(K):
    import torch
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    with pytest.raises(K.B200Error) as e:      # 8196 x 8192 texels = 2049 x 2048 blocks > B200_NOALLOC_MAX_BLOCKS; rejected before any access
        K.encode_image_dev_noalloc(d.data_ptr(), 8196, 8192, d.data_ptr(), 2)
    assert e.value.code == -2 and "B200_NOALLOC_MAX_BLOCKS" in str(e.value)
    with pytest.raises(K.B200Error) as e:      # level 4 does not run the scratch-based kernel
        K.encode_image_dev_noalloc(d.data_ptr(), 4, 4, d.data_ptr(), 4)
    assert e.value.code == -2
    torch.cuda.synchronize()
