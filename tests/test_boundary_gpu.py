"""GPU tests of the C-ABI boundary itself: re-entrancy from several host threads (the reference pins this for
ktxTexture2_TranscodeBasis with tests/threadtests/threadtests.cc:90-128), the multi-device planner (one job spread over the
devices of a mask must produce the bytes of the one-device run), pointer-alignment rejections, shutdown / re-init, and the
RDO scratch sizing with hundreds of small images."""
import ctypes, os, threading
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


def test_concurrent_calls_from_8_threads(K):
    """8 host threads x (transcode BC7 + ASTC, encode level 1) at once; every result must equal the serial one."""
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    u = np.ascontiguousarray(np.tile(g["uastc"], (8, 1)))          # ~170k blocks: several tiles per call
    imgs = [R.img_mix(96 + 8 * t, 64 + 4 * t, 30 + t) for t in range(8)]
    want_b = K.transcode(u, K.TTF_BC7_RGBA).copy()
    want_a = K.transcode(u, K.TTF_ASTC_4x4_RGBA).copy()
    want_e = [K.encode_images([im], 1)[0][0].copy() for im in imgs]
    errs = []

    def work(t):
        try:
            for _ in range(3):
                assert (K.transcode(u, K.TTF_BC7_RGBA) == want_b).all()
                assert (K.encode_images([imgs[t]], 1)[0][0] == want_e[t]).all()
                assert (K.transcode(u, K.TTF_ASTC_4x4_RGBA) == want_a).all()
        except Exception as e:   # noqa: BLE001
            errs.append((t, repr(e)))

    th = [threading.Thread(target=work, args=(t,)) for t in range(8)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    assert not errs, errs


def _mask(K, want):
    n = K.device_count()
    return (1 << min(n, want)) - 1, min(n, want)


def test_one_job_over_the_device_mask_matches_one_device(K):
    """One mip chain (level 0 must be slab-split) and one batch of faces through b200_uastc_encode with a device mask: the
    gathered bytes equal the single-device bytes.  Runs over 2 devices when the box has them (gpurun --gpus 2), else 1."""
    mask, nd = _mask(K, 8)
    base = R.img_mix(1024, 1024, 9)
    chain = [base]
    while chain[-1].shape[0] > 1:
        a = chain[-1].astype(np.uint16)
        chain.append(((a[0::2, 0::2] + a[1::2, 0::2] + a[0::2, 1::2] + a[1::2, 1::2] + 2) >> 2).astype(np.uint8))
    one, alpha1 = K.encode_images(chain, 2)
    many, alpha2 = K.encode_images(chain, 2, device_mask=mask)
    assert alpha1 == alpha2
    for a, b in zip(one, many):
        assert (a == b).all()
    faces = [R.img_mix(256, 256, 100 + i) for i in range(12)]
    one, _ = K.encode_images(faces, 1)
    many, _ = K.encode_images(faces, 1, device_mask=mask)
    assert all((a == b).all() for a, b in zip(one, many))
    # RDO spreads whole images
    one, _ = K.encode_images(faces[:5], 1 | 512, rdo=True, rdo_lambda=2.0, total_jobs=4)
    many, _ = K.encode_images(faces[:5], 1 | 512, rdo=True, rdo_lambda=2.0, total_jobs=4, device_mask=mask)
    assert all((a == b).all() for a, b in zip(one, many))
    # transcode over the process-wide device set
    u = np.concatenate(one + [np.tile(one[0], (40, 1))])
    want = K.transcode(u, K.TTF_BC7_RGBA).copy()
    K.set_devices(mask)
    try:
        got = K.transcode(u, K.TTF_BC7_RGBA)
        got1 = K.transcode(u, K.TTF_BC1_RGB, decode_flags=0)
    finally:
        K.set_devices(0)
    assert (got == want).all()
    assert (got1 == K.transcode(u, K.TTF_BC1_RGB, decode_flags=0)).all()
    print("devices used:", nd)


def test_bad_device_mask_is_rejected(K):
    n = K.device_count()
    assert n >= 1
    with pytest.raises(K.B200Error) as e:
        K.encode_images([R.img_mix(8, 8, 1)], 0, device_mask=1 << 15 if n < 16 else 0xFFFF0000)
    assert e.value.code in (-1, -2)
    with pytest.raises(K.B200Error):
        K.set_devices(1 << 15 if n < 16 else 1 << 20)


def test_misaligned_device_pointers(K):
    """Block pointers must be 16-byte aligned (rejected, not faulted); a cropped, unaligned image pointer is legal and takes
    the narrow load path (a misaligned uint4 load is a sticky context error)."""
    import torch
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    raw = torch.zeros(16 * 1024 + 64, dtype=torch.uint8, device="cuda")
    dst = torch.zeros(16 * 1024 + 64, dtype=torch.uint8, device="cuda")
    raw[8:8 + 16 * 1024] = torch.from_numpy(g["uastc"][:1024].reshape(-1)).cuda()
    for tf in (K.TTF_BC7_RGBA, K.TTF_BC4_R):
        with pytest.raises(K.B200Error) as e:
            K.transcode_dev(tf, raw.data_ptr() + 8, dst.data_ptr(), 1024) if tf == K.TTF_BC7_RGBA else \
                K.transcode_dev_ex(tf, 0, True, raw.data_ptr() + 8, dst.data_ptr(), 1024)
        assert e.value.code == -2
        with pytest.raises(K.B200Error):
            K.transcode_dev_ex(tf, 0, True, raw.data_ptr(), dst.data_ptr() + 4, 1024)
    # a crop of a larger image: base 4-byte aligned only, pitch of the parent
    big = R.img_mix(100, 64, 4)
    d_big = torch.from_numpy(big).cuda()
    x0, y0, w, h = 3, 5, 64, 40
    out = torch.zeros(((w // 4) * (h // 4), 16), dtype=torch.uint8, device="cuda")
    K.encode_image_dev(d_big.data_ptr() + (y0 * 100 + x0) * 4, w, h, out.data_ptr(), 1, row_pitch_bytes=400)
    want, _ = K.encode_images([np.ascontiguousarray(big[y0:y0 + h, x0:x0 + w])], 1)
    assert (out.cpu().numpy() == want[0]).all()
    with pytest.raises(K.B200Error):
        K.encode_image_dev(d_big.data_ptr(), 64, 40, out.data_ptr() + 8, 1, row_pitch_bytes=400)
    torch.cuda.synchronize()   # no sticky error left behind
    assert (K.transcode(g["uastc"][:64], K.TTF_BC7_RGBA).reshape(-1, 16) == g["bc7"][:64]).all()


def test_shutdown_and_reinit(K):
    g = np.load(os.path.join(GOLD, "transcode_golden.npz"))
    K.transcode(g["uastc"][:4096], K.TTF_BC7_RGBA)
    K.shutdown()
    K.shutdown()
    K.init()
    assert (K.transcode(g["uastc"][:4096], K.TTF_ASTC_4x4_RGBA).reshape(-1, 16) == g["astc"][:4096]).all()


@pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not shipped")
def test_rdo_with_hundreds_of_small_images(K):
    """RDO scratch was RDO_CHUNK table slots per chain whatever its length (256 small images x 4 jobs ~ 16 GB)."""
    imgs = [R.img_mix(32 + 4 * (i % 5), 24 + 4 * (i % 3), 400 + i) for i in range(256)]
    flags = 1 | 512
    outs, _ = K.encode_images(imgs, flags, rdo=True, rdo_lambda=2.0, total_jobs=4)
    for i in (0, 1, 7, 100, 255):
        px = R.ref_extract_blocks(imgs[i])
        want = R.ref_rdo(R.ref_encode_blocks(px, flags), px, flags, 4, 4, lam=2.0)
        assert (outs[i] == want).all(), i


def test_mipchain_from_base_level(K):
    """b200_uastc_encode_mipchain: generated levels equal the harness's integer box filter byte for byte, and every level's
    blocks equal b200_uastc_encode of that level (non-square, odd tail sizes included)."""
    for w, h in ((256, 256), (200, 72), (37, 5)):
        base = R.img_mix(w, h, 21)
        n = max(w, h).bit_length()
        outs, levels, alpha = K.encode_mipchain(base, n, 2, want_levels=True)
        prev = base
        for l in range(1, n):
            a = prev.astype(np.uint16)
            sh, sw = a.shape[:2]
            lw, lh = max(1, w >> l), max(1, h >> l)
            xs0 = np.minimum(2 * np.arange(lw), sw - 1); xs1 = np.minimum(2 * np.arange(lw) + 1, sw - 1)
            ys0 = np.minimum(2 * np.arange(lh), sh - 1); ys1 = np.minimum(2 * np.arange(lh) + 1, sh - 1)
            m = ((a[ys0][:, xs0] + a[ys0][:, xs1] + a[ys1][:, xs0] + a[ys1][:, xs1] + 2) >> 2).astype(np.uint8)
            assert (levels[l - 1] == m).all(), (w, h, l)
            prev = m
        want, alpha2 = K.encode_images([base] + levels, 2)
        assert alpha == alpha2
        for l in range(n):
            assert (outs[l] == want[l]).all(), (w, h, l)


def test_stream_callbacks_see_finished_images(K):
    """b200_uastc_encode_stream: one call per image, in submission order on one device, each made when the image's
    bytes are final (a copy taken inside the callback equals the result of the plain call); also with RDO (calls at the end) and
    over the device mask (any order, every image once)."""
    base = R.img_mix(2048, 2048, 13)
    chain = [base]
    while chain[-1].shape[0] > 4:
        a = chain[-1].astype(np.uint16)
        chain.append(((a[0::2, 0::2] + a[1::2, 0::2] + a[0::2, 1::2] + a[1::2, 1::2] + 2) >> 2).astype(np.uint8))
    imgs = chain + [R.img_noise(64, 48, 3), R.img_smooth(1024, 512, 4)]
    want, alpha = K.encode_images(imgs, 2)
    seen, order = {}, []

    def on_done(i, blocks):
        order.append(i)
        seen[i] = blocks.copy()

    got, alpha2 = K.encode_images_stream(imgs, 2, on_done)
    assert alpha == alpha2
    assert order == list(range(len(imgs)))
    for i in range(len(imgs)):
        assert (seen[i] == want[i]).all(), i
        assert (got[i] == want[i]).all(), i
    # RDO: callbacks after the walk
    small = imgs[4:]
    want_r, _ = K.encode_images(small, 1 | 512, rdo=True, rdo_lambda=2.0, total_jobs=4)
    seen.clear(); order.clear()
    K.encode_images_stream(small, 1 | 512, on_done, rdo=True, rdo_lambda=2.0, total_jobs=4)
    assert sorted(order) == list(range(len(small)))
    assert all((seen[i] == want_r[i]).all() for i in range(len(small)))
    # several devices: every image exactly once
    mask, _ = _mask(K, 8)
    seen.clear(); order.clear()
    K.encode_images_stream(imgs, 2, on_done, device_mask=mask)
    assert sorted(order) == list(range(len(imgs)))
    assert all((seen[i] == want[i]).all() for i in range(len(imgs)))
    # an exception in the callback surfaces after the call, the library stays usable
    with pytest.raises(ZeroDivisionError):
        K.encode_images_stream(imgs[:3], 1, lambda i, b: 1 / 0)
    assert (K.encode_images(imgs[5:6], 2)[0][0] == want[5]).all()
