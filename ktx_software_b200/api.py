"""ctypes bindings over include/ktx_b200_uastc.h.  No compute happens in Python."""
import ctypes, os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_UASTC_LIB") or os.path.join(HERE, "libuastc_b200.so")  # env override: development only

TTF_BC7_RGBA = 6        # ktx_transcode_fmt_e KTX_TTF_BC7_RGBA (lib/include/ktx.h)
TTF_ASTC_4x4_RGBA = 10  # KTX_TTF_ASTC_4x4_RGBA
TTF_RGBA32, TTF_RGB565, TTF_BGR565, TTF_RGBA4444 = 13, 14, 15, 16  # the uncompressed targets (ktx.h:1728-1737)
# the hint-driven block targets (ktx.h:1654-1676); 8 bytes per block for ETC1_RGB, BC1_RGB, BC4_R, else 16
TTF_ETC1_RGB, TTF_ETC2_RGBA, TTF_BC1_RGB, TTF_BC3_RGBA, TTF_BC4_R, TTF_BC5_RG = 0, 1, 2, 3, 4, 5
TTF_ETC2_EAC_R11, TTF_ETC2_EAC_RG11 = 20, 21   # ktx.h:1718-1722
TF_TRANSCODE_ALPHA_DATA_TO_OPAQUE_FORMATS, TF_HIGH_QUALITY = 4, 32   # ktx_transcode_flag_bits_e


def target_block_bytes(target):
    return 8 if target in (TTF_ETC1_RGB, TTF_BC1_RGB, TTF_BC4_R, TTF_ETC2_EAC_R11) else 16


EXPORTED_SYMBOLS = ["b200_uastc_init", "b200_uastc_shutdown", "b200_uastc_last_error", "b200_uastc_launch_count",
                    "b200_uastc_transcode", "b200_uastc_transcode_dev", "b200_uastc_encode", "b200_uastc_encode_blocks_dev",
                    "b200_uastc_encode_image_dev", "b200_uastc_rdo_blocks_dev", "b200_uastc_encode_images_dev",
                    "b200_uastc_decode", "b200_uastc_decode_image_dev", "b200_uastc_transcode_ex", "b200_uastc_transcode_dev_ex", "b200_uastc_encode_ex",
                    "b200_uastc_set_devices", "b200_uastc_device_count", "b200_uastc_build_id", "b200_uastc_encode_mipchain",
                    "b200_uastc_encode_image_dev_noalloc", "b200_uastc_encode_stream"]

_u8p = ctypes.POINTER(ctypes.c_uint8)


class B200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("b200_uastc error %d: %s" % (code, msg))
        self.code = code


class XcodeSpan(ctypes.Structure):
    _fields_ = [("src_offset", ctypes.c_uint64), ("dst_offset", ctypes.c_uint64), ("n_blocks", ctypes.c_uint64)]


class DecodeImage(ctypes.Structure):
    _fields_ = [("src_offset", ctypes.c_uint64), ("dst_offset", ctypes.c_uint64), ("width", ctypes.c_uint32), ("height", ctypes.c_uint32)]


class Image(ctypes.Structure):
    _fields_ = [("rgba", ctypes.c_void_p), ("width", ctypes.c_uint32), ("height", ctypes.c_uint32), ("row_pitch_bytes", ctypes.c_uint32)]


class DevImage(ctypes.Structure):
    _fields_ = [("d_rgba", ctypes.c_void_p), ("width", ctypes.c_uint32), ("height", ctypes.c_uint32), ("row_pitch_bytes", ctypes.c_uint32),
                ("d_out", ctypes.c_void_p)]


class SrcFormat(ctypes.Structure):
    _fields_ = [("num_components", ctypes.c_uint32), ("swizzle", ctypes.c_uint8 * 4)]


class UastcParams(ctypes.Structure):
    _fields_ = [("struct_size", ctypes.c_uint32), ("pack_flags", ctypes.c_uint32), ("rdo_enable", ctypes.c_uint32),
                ("rdo_lambda", ctypes.c_float), ("rdo_dict_size", ctypes.c_uint32),
                ("rdo_max_smooth_block_error_scale", ctypes.c_float), ("rdo_max_smooth_block_std_dev", ctypes.c_float),
                ("rdo_total_jobs", ctypes.c_uint32), ("device_mask", ctypes.c_uint32), ("reserved", ctypes.c_uint32 * 3)]


_lib = None


def lib():
    """Load the CUDA library; raise loudly if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error(-1, "CUDA library %s is missing: run `python -m ktx_software_b200.build` (no CPU fallback)" % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        l.b200_uastc_last_error.restype = ctypes.c_char_p
        l.b200_uastc_launch_count.restype = ctypes.c_uint64
        l.b200_uastc_set_devices.argtypes = [ctypes.c_uint64]
        l.b200_uastc_build_id.restype = ctypes.c_char_p
        l.b200_uastc_encode_mipchain.argtypes = [ctypes.POINTER(Image), ctypes.c_uint32, ctypes.POINTER(UastcParams), ctypes.POINTER(ctypes.c_void_p),
                                                 ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint32)]
        l.b200_uastc_transcode.argtypes = [ctypes.c_uint32, ctypes.POINTER(XcodeSpan), ctypes.c_uint32, ctypes.c_void_p,
                                           ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64)]
        l.b200_uastc_transcode_dev.argtypes = [ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                                               ctypes.c_void_p, ctypes.c_void_p]
        l.b200_uastc_transcode_ex.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.POINTER(XcodeSpan), ctypes.c_uint32,
                                              ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64)]
        l.b200_uastc_transcode_dev_ex.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
                                                  ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p]
        l.b200_uastc_encode.argtypes = [ctypes.POINTER(Image), ctypes.c_uint32, ctypes.POINTER(UastcParams),
                                        ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint32)]
        l.b200_uastc_encode_ex.argtypes = [ctypes.POINTER(Image), ctypes.c_uint32, ctypes.POINTER(SrcFormat), ctypes.POINTER(UastcParams),
                                           ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint32)]
        l.b200_uastc_encode_stream.argtypes = [ctypes.POINTER(Image), ctypes.c_uint32, ctypes.POINTER(SrcFormat), ctypes.POINTER(UastcParams),
                                               ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint32), IMAGE_DONE_FN, ctypes.c_void_p]
        l.b200_uastc_encode_blocks_dev.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_void_p]
        l.b200_uastc_encode_image_dev.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p,
                                                  ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
        l.b200_uastc_encode_image_dev_noalloc.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p,
                                                          ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
        l.b200_uastc_rdo_blocks_dev.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.POINTER(UastcParams), ctypes.c_void_p]
        l.b200_uastc_encode_images_dev.argtypes = [ctypes.POINTER(DevImage), ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
        l.b200_uastc_decode.argtypes = [ctypes.c_uint32, ctypes.POINTER(DecodeImage), ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
                                        ctypes.POINTER(ctypes.c_int64)]
        l.b200_uastc_decode_image_dev.argtypes = [ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32,
                                                  ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
        _lib = l
    return _lib


def _check(rc):
    if rc != 0:
        raise B200Error(rc, lib().b200_uastc_last_error().decode())


def init():
    _check(lib().b200_uastc_init())


def launch_count():
    return int(lib().b200_uastc_launch_count())


def build_id():
    return lib().b200_uastc_build_id().decode()


def set_devices(device_mask):
    """Devices the host-buffer calls spread over (bit d = CUDA device d); 0 = the calling thread's current device."""
    _check(lib().b200_uastc_set_devices(device_mask))


def device_count():
    return int(lib().b200_uastc_device_count())


def shutdown():
    lib().b200_uastc_shutdown()


def transcode(src, target, spans=None, dst=None, decode_flags=None, has_alpha=True):
    """Host-buffer transcode through the C ABI.  src: uint8 array holding UASTC blocks; spans: list of
    (src_offset, dst_offset, n_blocks) or None for one span over the whole buffer.  Returns the dst array.
    decode_flags = None: b200_uastc_transcode (BC7 / ASTC); otherwise b200_uastc_transcode_ex with libktx's transcodeFlags and
    has_alpha (any block target).  Raises B200Error(code=-5) with .bad_span set when the input holds an invalid block."""
    src = np.ascontiguousarray(src, dtype=np.uint8).reshape(-1)
    ob = target_block_bytes(target)
    if spans is None:
        spans = [(0, 0, src.size // 16)]
    if dst is None:
        need = max([d + n * ob for _, d, n in spans] + [0])
        dst = np.zeros(need, np.uint8)
    arr = (XcodeSpan * max(1, len(spans)))(*[XcodeSpan(s, d, n) for s, d, n in spans])
    bad = ctypes.c_int64(-1)
    if decode_flags is None and target in (TTF_BC7_RGBA, TTF_ASTC_4x4_RGBA):
        rc = lib().b200_uastc_transcode(target, arr, len(spans), src.ctypes.data, dst.ctypes.data, ctypes.byref(bad))
    else:
        rc = lib().b200_uastc_transcode_ex(target, decode_flags or 0, 1 if has_alpha else 0, arr, len(spans), src.ctypes.data,
                                           dst.ctypes.data, ctypes.byref(bad))
    if rc != 0:
        e = B200Error(rc, lib().b200_uastc_last_error().decode())
        e.bad_span = bad.value
        raise e
    return dst


def transcode_dev(target, d_src, d_dst, n_blocks, stream=0, d_first_bad=0):
    """Device-pointer transcode (ints / torch data_ptr()); asynchronous when a stream handle is given."""
    _check(lib().b200_uastc_transcode_dev(target, d_src, d_dst, n_blocks, stream or None, d_first_bad or None))


def transcode_dev_ex(target, decode_flags, has_alpha, d_src, d_dst, n_blocks, stream=0, d_first_bad=0):
    """Device-pointer transcode with libktx's transcodeFlags / has_alpha; every block target (b200_uastc_transcode_dev_ex)."""
    _check(lib().b200_uastc_transcode_dev_ex(target, decode_flags, 1 if has_alpha else 0, d_src, d_dst, n_blocks, stream or None,
                                             d_first_bad or None))


def decode(src, target, images):
    """Host-buffer decode to texels through the C ABI (b200_uastc_decode).  src: uint8 array of UASTC blocks;
    images: list of (src_offset, width, height), each image's blocks in raster order.  Returns a list of arrays:
    uint8 [h,w,4] for TTF_RGBA32, uint16 [h,w] for the 16-bit targets.  B200Error(code=-5) carries .bad_image."""
    src = np.ascontiguousarray(src, dtype=np.uint8).reshape(-1)
    bpp = 4 if target == TTF_RGBA32 else 2
    offs, total = [], 0
    for _, w, h in images:
        offs.append(total)
        total += (w * h * bpp + 15) & ~15
    dst = np.zeros(max(total, 16), np.uint8)
    arr = (DecodeImage * max(1, len(images)))(*[DecodeImage(so, do, w, h) for (so, w, h), do in zip(images, offs)])
    bad = ctypes.c_int64(-1)
    rc = lib().b200_uastc_decode(target, arr, len(images), src.ctypes.data, dst.ctypes.data, ctypes.byref(bad))
    if rc != 0:
        e = B200Error(rc, lib().b200_uastc_last_error().decode())
        e.bad_image = bad.value
        raise e
    out = []
    for (_, w, h), do in zip(images, offs):
        raw = dst[do:do + w * h * bpp]
        out.append(raw.reshape(h, w, 4) if bpp == 4 else raw.view(np.uint16).reshape(h, w))
    return out


def decode_image_dev(target, d_src, d_dst, width, height, row_pitch_texels=0, stream=0, d_first_bad=0):
    """Device-pointer decode of one image's blocks to width x height texels (b200_uastc_decode_image_dev)."""
    _check(lib().b200_uastc_decode_image_dev(target, d_src, d_dst, width, height, row_pitch_texels, stream or None, d_first_bad or None))


def encode_blocks_dev(d_block_pixels, d_out, n_blocks, pack_flags, stream=0):
    """Device-pointer encode of n_blocks pre-gathered 4x4 RGBA8 blocks (64 B each) -> 16 B UASTC blocks."""
    _check(lib().b200_uastc_encode_blocks_dev(d_block_pixels, d_out, n_blocks, pack_flags, stream or None))


def encode_image_dev(d_rgba, width, height, d_out, pack_flags, stream=0, row_pitch_bytes=0, d_any_alpha=0):
    """Device-pointer encode of one RGBA8 image resident in HBM; the 4x4 gather is fused into the kernel."""
    _check(lib().b200_uastc_encode_image_dev(d_rgba, width, height, row_pitch_bytes, d_out, pack_flags, stream or None, d_any_alpha or None))


def encode_image_dev_noalloc(d_rgba, width, height, d_out, pack_flags, stream=0, row_pitch_bytes=0, d_any_alpha=0):
    """b200_uastc_encode_image_dev_noalloc: the same encode with the kernel's scratch in a static device array (no allocation)."""
    _check(lib().b200_uastc_encode_image_dev_noalloc(d_rgba, width, height, row_pitch_bytes, d_out, pack_flags, stream or None, d_any_alpha or None))


def encode_images_dev(images, pack_flags, stream=0, d_any_alpha=0):
    """One launch over several device-resident images: images = [(d_rgba, width, height, d_out), ...] (ints)."""
    arr = (DevImage * max(1, len(images)))(*[DevImage(p, w, h, 0, o) for p, w, h, o in images])
    _check(lib().b200_uastc_encode_images_dev(arr, len(images), pack_flags, stream or None, d_any_alpha or None))


def make_params(pack_flags, rdo=False, rdo_lambda=0.0, dict_size=0, smooth_scale=0.0, smooth_std_dev=0.0, total_jobs=0, device_mask=0):
    p = UastcParams()
    p.struct_size = ctypes.sizeof(UastcParams)
    p.pack_flags = pack_flags
    p.rdo_enable = 1 if rdo else 0
    p.rdo_lambda = rdo_lambda
    p.rdo_dict_size = dict_size
    p.rdo_max_smooth_block_error_scale = smooth_scale
    p.rdo_max_smooth_block_std_dev = smooth_std_dev
    p.rdo_total_jobs = total_jobs
    p.device_mask = device_mask
    return p


def rdo_blocks_dev(d_blocks, d_block_pixels, n_blocks, pack_flags, stream=0, **rdo):
    """RDO post-pass in place on device-resident blocks of one image (b200_uastc_rdo_blocks_dev)."""
    p = make_params(pack_flags, rdo=True, **rdo)
    _check(lib().b200_uastc_rdo_blocks_dev(d_blocks, d_block_pixels, n_blocks, ctypes.byref(p), stream or None))


def encode_images(images, pack_flags, **rdo):
    """Host-buffer encode through the C ABI (b200_uastc_encode).  images: list of uint8 [h,w,4] arrays.
    Returns (list of uint8 [nblocks,16] arrays in raster block order, any_alpha)."""
    imgs = [np.ascontiguousarray(im, dtype=np.uint8) for im in images]
    arr = (Image * max(1, len(imgs)))()
    outs = []
    optr = (ctypes.c_void_p * max(1, len(imgs)))()
    for i, im in enumerate(imgs):
        h, w = im.shape[:2]
        assert im.shape[2] == 4
        arr[i] = Image(im.ctypes.data, w, h, 0)
        o = np.zeros((((w + 3) // 4) * ((h + 3) // 4), 16), np.uint8)
        outs.append(o)
        optr[i] = o.ctypes.data
    p = make_params(pack_flags, **rdo)
    any_alpha = ctypes.c_uint32(0)
    _check(lib().b200_uastc_encode(arr, len(imgs), ctypes.byref(p), optr, ctypes.byref(any_alpha)))
    return outs, bool(any_alpha.value)


IMAGE_DONE_FN = ctypes.CFUNCTYPE(None, ctypes.c_uint32, ctypes.c_void_p)


def encode_images_stream(images, pack_flags, on_done, **rdo):
    """b200_uastc_encode_stream: as encode_images, and on_done(i, blocks_of_image_i) is called as soon as image i is complete
    in host memory (while later images are still on the GPU).  Returns (outs, any_alpha)."""
    imgs = [np.ascontiguousarray(im, dtype=np.uint8) for im in images]
    arr = (Image * max(1, len(imgs)))()
    outs = []
    optr = (ctypes.c_void_p * max(1, len(imgs)))()
    for i, im in enumerate(imgs):
        h, w = im.shape[:2]
        assert im.shape[2] == 4
        arr[i] = Image(im.ctypes.data, w, h, 0)
        o = np.zeros((((w + 3) // 4) * ((h + 3) // 4), 16), np.uint8)
        outs.append(o)
        optr[i] = o.ctypes.data
    p = make_params(pack_flags, **rdo)
    any_alpha = ctypes.c_uint32(0)
    errs = []

    def _cb(i, _user):
        try:
            on_done(int(i), outs[int(i)])
        except BaseException as e:   # noqa: BLE001 -- must not unwind through the C frames
            errs.append(e)

    cb = IMAGE_DONE_FN(_cb)
    _check(lib().b200_uastc_encode_stream(arr, len(imgs), None, ctypes.byref(p), optr, ctypes.byref(any_alpha), cb, None))
    if errs:
        raise errs[0]
    return outs, bool(any_alpha.value)


def encode_mipchain(base, n_levels, pack_flags, want_levels=False):
    """Base level in, all levels out (b200_uastc_encode_mipchain): base uint8 [h,w,4].  Returns (blocks per level, generated
    RGBA8 levels 1.. or None, any_alpha)."""
    base = np.ascontiguousarray(base, dtype=np.uint8)
    h, w = base.shape[:2]
    img = Image(base.ctypes.data, w, h, 0)
    outs, lv = [], []
    optr = (ctypes.c_void_p * n_levels)()
    lptr = (ctypes.c_void_p * n_levels)()
    for l in range(n_levels):
        lw, lh = max(1, w >> l), max(1, h >> l)
        o = np.zeros((((lw + 3) // 4) * ((lh + 3) // 4), 16), np.uint8)
        outs.append(o)
        optr[l] = o.ctypes.data
        if want_levels and l:
            m = np.zeros((lh, lw, 4), np.uint8)
            lv.append(m)
            lptr[l] = m.ctypes.data
    p = make_params(pack_flags)
    any_alpha = ctypes.c_uint32(0)
    _check(lib().b200_uastc_encode_mipchain(ctypes.byref(img), n_levels, ctypes.byref(p), optr, lptr if want_levels else None, ctypes.byref(any_alpha)))
    return outs, (lv if want_levels else None), bool(any_alpha.value)


SWIZZLE_CODES = {"r": 1, "g": 2, "b": 3, "a": 4, "0": 5, "1": 6}   # swizzle_e (lib/src/basis_encode.cpp:77-84)


def encode_images_ex(images, swizzle, pack_flags, **rdo):
    """Host-buffer encode of images in the texture's own layout (b200_uastc_encode_ex).  images: uint8 [h,w,c] arrays, c = 1..4
    (all the same c); swizzle: 4 characters of "rgba01" as in ktxBasisParams::inputSwizzle.  Returns (blocks per image, any_alpha)."""
    imgs = [np.ascontiguousarray(im, dtype=np.uint8) for im in images]
    comps = imgs[0].shape[2] if imgs else 4
    fmt = SrcFormat(comps, (ctypes.c_uint8 * 4)(*[SWIZZLE_CODES[ch] for ch in swizzle]))
    arr = (Image * max(1, len(imgs)))()
    outs = []
    optr = (ctypes.c_void_p * max(1, len(imgs)))()
    for i, im in enumerate(imgs):
        h, w = im.shape[:2]
        assert im.shape[2] == comps
        arr[i] = Image(im.ctypes.data, w, h, 0)
        o = np.zeros((((w + 3) // 4) * ((h + 3) // 4), 16), np.uint8)
        outs.append(o)
        optr[i] = o.ctypes.data
    p = make_params(pack_flags, **rdo)
    any_alpha = ctypes.c_uint32(0)
    _check(lib().b200_uastc_encode_ex(arr, len(imgs), ctypes.byref(fmt), ctypes.byref(p), optr, ctypes.byref(any_alpha)))
    return outs, bool(any_alpha.value)


def encode_blocks(block_pixels, pack_flags):
    """Convenience for tests: n pre-gathered 4x4 blocks (uint8 [n,64]) -> uint8 [n,16], via the image entry point
    (the blocks are laid out as one 4-texel-high strip image so the C ABI's gather reproduces them)."""
    px = np.ascontiguousarray(block_pixels, dtype=np.uint8).reshape(-1, 4, 4, 4)
    n = px.shape[0]
    if n == 0:
        return np.zeros((0, 16), np.uint8)
    strip = np.ascontiguousarray(px.transpose(1, 0, 2, 3).reshape(4, n * 4, 4))
    outs, _ = encode_images([strip], pack_flags)
    return outs[0]
