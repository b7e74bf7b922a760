"""ctypes access to the checkers: oracle/_ref/libref_uastc.so (the UNMODIFIED reference, built by oracle/Makefile)
and oracle/libport_uastc.so (host build of the product's own algorithm headers).  Test infrastructure only."""
import ctypes, os, subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle/_ref/libref_uastc.so")
PORT_SO = os.path.join(ROOT, "oracle/libport_uastc.so")
_u8p = ctypes.POINTER(ctypes.c_uint8)


def _p(a):
    return a.ctypes.data_as(_u8p)


def have_ref():
    return os.path.exists(REF_SO)


_ref = None
_port = None


def ref():
    global _ref
    if _ref is None:
        _ref = ctypes.CDLL(REF_SO)
        _ref.ref_transcode.restype = ctypes.c_int64
        _ref.ref_transcode.argtypes = [_u8p, ctypes.c_uint64, ctypes.c_uint32, _u8p]
        _ref.ref_transcode_mt.restype = ctypes.c_int64
        _ref.ref_transcode_mt.argtypes = [_u8p, ctypes.c_uint64, ctypes.c_uint32, _u8p, ctypes.c_uint32]
        _ref.ref_encode_blocks.argtypes = [_u8p, ctypes.c_uint32, ctypes.c_uint32, _u8p, ctypes.c_uint32]
        _ref.ref_encode_image.argtypes = [_u8p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, _u8p, ctypes.c_uint32]
        _ref.ref_extract_blocks.argtypes = [_u8p, ctypes.c_uint32, ctypes.c_uint32, _u8p]
        _ref.ref_unpack_rgba.restype = ctypes.c_int64
        _ref.ref_unpack_rgba.argtypes = [_u8p, ctypes.c_uint64, _u8p, ctypes.c_int]
        _ref.ref_unpack_bc7.argtypes = [_u8p, ctypes.c_uint64, _u8p]
        _ref.ref_rdo.argtypes = [_u8p, ctypes.c_uint32, _u8p, ctypes.c_float, ctypes.c_uint32, ctypes.c_float, ctypes.c_float,
                                 ctypes.c_float, ctypes.c_float, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32]
        _ref.ref_init()
    return _ref


def port():
    global _port
    if _port is None:
        if not os.path.exists(PORT_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"])
        _port = ctypes.CDLL(PORT_SO)
        _port.port_transcode.restype = ctypes.c_int64
        _port.port_transcode.argtypes = [_u8p, ctypes.c_uint64, ctypes.c_uint32, _u8p]
    return _port


def ref_transcode(blocks, fmt):
    """blocks: uint8 [n,16]; fmt 0=BC7 1=ASTC -> (out [n,16], first_bad or -1)"""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)
    out = np.zeros_like(blocks)
    bad = ref().ref_transcode(_p(blocks), blocks.shape[0], fmt, _p(out))
    return out, bad


def port_transcode(blocks, fmt):
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)
    out = np.zeros_like(blocks)
    bad = port().port_transcode(_p(blocks), blocks.shape[0], fmt, _p(out))
    return out, bad


def port_decode_rgba(blocks):
    """blocks [n,16] -> (texels [n,64] as the KTX_TTF_RGBA32 decode lays a block out, first_bad or -1)"""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)
    out = np.zeros((blocks.shape[0], 64), np.uint8)
    f = port().port_decode_rgba
    f.restype = ctypes.c_int64
    f.argtypes = [_u8p, ctypes.c_uint64, _u8p]
    return out, f(_p(blocks), blocks.shape[0], _p(out))


TTF_BYTES_PER_PIXEL = {13: 4, 14: 2, 15: 2, 16: 2}


def ref_transcode_image(blocks, w, h, tf):
    """One w x h image (raster UASTC blocks) through the reference's low-level transcoder the way libktx calls it.
    tf 6 / 10 -> blocks [n,16]; tf 13 -> uint8 [h,w,4]; tf 14/15/16 -> uint16 [h,w].  Returns (out, ok)."""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)
    bx, by = (w + 3) // 4, (h + 3) // 4
    assert blocks.shape[0] == bx * by
    f = ref().ref_transcode_image
    f.restype = ctypes.c_int
    f.argtypes = [ctypes.c_uint32, _u8p, ctypes.c_uint32, ctypes.c_uint32, _u8p, ctypes.c_uint32]
    if tf in TTF_BYTES_PER_PIXEL:
        out = np.zeros(w * h * TTF_BYTES_PER_PIXEL[tf], np.uint8)
        rc = f(tf, _p(blocks), w, h, _p(out), w * h)
        out = out.reshape(h, w, 4) if tf == 13 else out.view(np.uint16).reshape(h, w)
    else:
        out = np.zeros_like(blocks)
        rc = f(tf, _p(blocks), w, h, _p(out), bx * by)
    return out, rc == 0


HINTED_TARGET_BYTES = {0: 8, 1: 16, 2: 8, 3: 16, 4: 8, 5: 16, 20: 8, 21: 16}   # ETC1_RGB, ETC2_RGBA, BC1_RGB, BC3_RGBA, BC4_R, BC5_RG, ETC2_EAC_R11, _RG11


def ref_transcode_hinted(blocks, tf, decode_flags=0, has_alpha=True):
    """n blocks as one (4n x 4) image through the reference's transcode_image with the caller's decode flags
    (4 = alpha data to opaque formats, 32 = high quality).  tf in HINTED_TARGET_BYTES -> (out [n, 8|16], ok)."""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)
    n = blocks.shape[0]
    f = ref().ref_transcode_image_flags
    f.restype = ctypes.c_int
    f.argtypes = [ctypes.c_uint32, _u8p, ctypes.c_uint32, ctypes.c_uint32, _u8p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32]
    out = np.zeros((n, HINTED_TARGET_BYTES[tf]), np.uint8)
    rc = f(tf, _p(blocks), n * 4, 4, _p(out), n, decode_flags, 1 if has_alpha else 0)
    return out, rc == 0


def port_transcode_hinted(blocks, tf, decode_flags=0, has_alpha=True):
    """The product headers' host build; channel selection as transcode_image does it (XC:10417-10438)."""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)
    f = port().port_transcode_hinted
    f.restype = ctypes.c_int64
    f.argtypes = [_u8p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, _u8p]
    out = np.zeros((blocks.shape[0], HINTED_TARGET_BYTES[tf]), np.uint8)
    from_alpha = bool(has_alpha and (decode_flags & 4))
    chan0 = 3 if (tf in (0, 4, 20) and from_alpha) else 0
    kind = 8 if (tf == 0 and from_alpha) else {20: 6, 21: 7}.get(tf, tf)   # 8 = ETC1 from one channel (XC:16560)
    bad = f(_p(blocks), blocks.shape[0], kind, 1 if (decode_flags & 32) else 0, chan0, 3, _p(out))
    return out, bad


def etc2_transcodable(blocks):
    """Mask of blocks the reference's ETC2 path accepts without tripping its `multiplier >= 1` assert (random payloads
    behind an alpha mode prefix can carry a zero EAC multiplier; the reference encoder never writes one)."""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)
    out = np.zeros(blocks.shape[0], np.uint16)
    f = port().port_etc2_hints
    f.restype = None
    f.argtypes = [_u8p, ctypes.c_uint64, ctypes.c_void_p]
    f(_p(blocks), blocks.shape[0], out.ctypes.data)
    return (out == 0xFFFF) | ((out >> 4) >= 1)


def texels_to_image(texels, w, h):
    """[n,64] per-block texels (raster inside the block) -> uint8 [h,w,4], clipped like transcode_slice does."""
    bx, by = (w + 3) // 4, (h + 3) // 4
    t = np.asarray(texels, np.uint8).reshape(by, bx, 4, 4, 4).transpose(0, 2, 1, 3, 4).reshape(by * 4, bx * 4, 4)
    return np.ascontiguousarray(t[:h, :w])


def pack_image16(img, tf):
    """uint8 [h,w,4] -> uint16 [h,w] with the product's texel packer (host build)."""
    px = np.ascontiguousarray(img, np.uint8).reshape(-1, 4).view(np.uint32).reshape(-1)
    out = np.zeros(px.shape[0], np.uint16)
    f = port().port_pack_texels16
    f.restype = None
    f.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_void_p]
    f(px.ctypes.data, px.shape[0], tf, out.ctypes.data)
    return out.reshape(img.shape[0], img.shape[1])


def ref_extract_blocks(img):
    """img uint8 [h,w,4] -> [nblocks,64] edge-clamped 4x4 gathers in raster block order"""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w = img.shape[:2]
    n = ((w + 3) // 4) * ((h + 3) // 4)
    out = np.zeros((n, 64), np.uint8)
    ref().ref_extract_blocks(_p(img), w, h, _p(out))
    return out


def ref_encode_blocks(px, flags, threads=8):
    px = np.ascontiguousarray(px, dtype=np.uint8).reshape(-1, 64)
    out = np.zeros((px.shape[0], 16), np.uint8)
    ref().ref_encode_blocks(_p(px), px.shape[0], flags, _p(out), threads)
    return out


def ref_unpack_rgba(blocks, srgb=False):
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)
    out = np.zeros((blocks.shape[0], 64), np.uint8)
    bad = ref().ref_unpack_rgba(_p(blocks), blocks.shape[0], _p(out), int(srgb))
    return out, bad


def ref_unpack_bc7(blocks):
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16)
    out = np.zeros((blocks.shape[0], 64), np.uint8)
    ref().ref_unpack_bc7(_p(blocks), blocks.shape[0], _p(out))
    return out


# ---------------------------------------------------------------------------------------------------------
# Synthetic inputs: deterministic, numpy-vectorised.
def _lcg_noise(n, seed):
    """n values of the 32-bit LCG s = s*1664525 + 1013904223; noise = (s>>8)%32 - 16 (vectorised by jump-ahead)."""
    a, c = np.uint64(1664525), np.uint64(1013904223)
    # sequential in chunks: compute by doubling (A_k, C_k) s.t. s_{i+k} = A_k s_i + C_k mod 2^32
    s = np.empty(n, np.uint64)
    cur = np.uint64(seed & 0xFFFFFFFF)
    m = np.uint64(0xFFFFFFFF)
    # generate the first block sequentially then extend by jump-ahead
    blk = min(n, 1024)
    for i in range(blk):
        cur = (cur * a + c) & m
        s[i] = cur
    filled = blk
    A, C = np.uint64(1), np.uint64(0)
    for _ in range(blk):
        A, C = (A * a) & m, (C * a + c) & m
    while filled < n:
        k = min(filled, n - filled)
        # jump every element ahead by `filled` steps: need A_filled, C_filled
        s[filled:filled + k] = (s[:k] * A + C) & m
        filled += k
        A, C = (A * A) & m, (C * A + C) & m
    return ((s >> np.uint64(8)) % np.uint64(32)).astype(np.int32) - 16


def img_mix(w, h, seed):
    n = _lcg_noise(w * h, seed).reshape(h, w)
    x = np.arange(w, dtype=np.int32)[None, :].repeat(h, 0)
    y = np.arange(h, dtype=np.int32)[:, None].repeat(w, 1)
    img = np.empty((h, w, 4), np.uint8)
    img[..., 0] = np.clip(255 * x // max(w, 1) + n, 0, 255)
    img[..., 1] = np.clip(255 * y // max(h, 1) + n // 2, 0, 255)
    img[..., 2] = np.clip(((x ^ y) & 255) + n, 0, 255)
    img[..., 3] = np.where(x >= w // 2, 255, np.clip((x + y) // 4 + n, 0, 255))
    # every 16th 64x64 tile solid, every 17th grey (r=g=b)
    tx, ty = x // 64, y // 64
    tid = ty * ((w + 63) // 64) + tx
    solid = (tid % 16) == 15
    img[solid] = np.stack([(tid[solid] * 37) & 255, (tid[solid] * 59) & 255, (tid[solid] * 83) & 255,
                           np.where(tid[solid] % 32 == 15, 255, (tid[solid] * 11) & 255)], -1).astype(np.uint8)
    grey = (tid % 17) == 16
    img[grey, 1] = img[grey, 0]
    img[grey, 2] = img[grey, 0]
    return img


def img_noise(w, h, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)


def img_smooth(w, h, seed):
    x = np.arange(w, dtype=np.int32)[None, :].repeat(h, 0)
    y = np.arange(h, dtype=np.int32)[:, None].repeat(w, 1)
    img = np.empty((h, w, 4), np.uint8)
    img[..., 0] = 255 * x // max(w, 1)
    img[..., 1] = 255 * y // max(h, 1)
    img[..., 2] = (x + y + seed) // 8 & 255
    img[..., 3] = 255
    return img


def adversarial_blocks(seed, n_each=256):
    """Block families from SURVEY 8(c): solid, 2-colour, grey ramps, LA, binary alpha, single channel, noise."""
    r = np.random.default_rng(seed)
    fams = []
    solid = r.integers(0, 256, (n_each, 1, 4), dtype=np.uint8).repeat(16, 1)
    fams.append(solid)
    solid_opaque = solid.copy(); solid_opaque[..., 3] = 255
    fams.append(solid_opaque)
    two = r.integers(0, 256, (n_each, 2, 4), dtype=np.uint8)
    pick = r.integers(0, 2, (n_each, 16))
    fams.append(np.take_along_axis(two, pick[..., None].repeat(4, -1), 1))
    t2 = fams[-1].copy(); t2[..., 3] = 255; fams.append(t2)
    ramp = (np.arange(16)[None, :] * r.integers(1, 17, (n_each, 1)) + r.integers(0, 16, (n_each, 1))).clip(0, 255).astype(np.uint8)
    g = np.stack([ramp, ramp, ramp, np.full_like(ramp, 255)], -1); fams.append(g)
    la = np.stack([ramp, ramp, ramp, r.integers(0, 256, (n_each, 16), dtype=np.uint8)], -1); fams.append(la)
    ba = r.integers(0, 256, (n_each, 16, 4), dtype=np.uint8); ba[..., 3] = r.integers(0, 2, (n_each, 16)) * 255; fams.append(ba)
    sc = np.zeros((n_each, 16, 4), np.uint8); sc[..., 3] = 255
    ch = r.integers(0, 3, n_each)
    sc[np.arange(n_each), :, ch] = r.integers(0, 256, (n_each, 16), dtype=np.uint8); fams.append(sc)
    fams.append(r.integers(0, 256, (n_each, 16, 4), dtype=np.uint8))
    op = r.integers(0, 256, (n_each, 16, 4), dtype=np.uint8); op[..., 3] = 255; fams.append(op)
    # smooth gradients + small noise, opaque and translucent
    base = r.integers(0, 256, (n_each, 1, 4)).astype(np.int32)
    dx = r.integers(-12, 13, (n_each, 1, 4)); dy = r.integers(-12, 13, (n_each, 1, 4))
    xs = (np.arange(16) % 4)[None, :, None]; ys = (np.arange(16) // 4)[None, :, None]
    sm = (base + dx * xs + dy * ys + r.integers(-2, 3, (n_each, 16, 4))).clip(0, 255).astype(np.uint8)
    fams.append(sm)
    smo = sm.copy(); smo[..., 3] = 255; fams.append(smo)
    return np.concatenate(fams, 0).reshape(-1, 64)


HUFF = [(0x1, 4), (0x35, 6), (0x1D, 5), (0x3, 5), (0x13, 5), (0xB, 5), (0x1B, 5), (0x7, 5), (0x17, 5), (0xF, 5), (0x2, 3),
        (0x0, 2), (0x6, 3), (0x1F, 5), (0xD, 5), (0x5, 7), (0x15, 6), (0x25, 6), (0x9, 4), (0x45, 7)]  # UASTC mode prefix codes


def block_modes(blocks):
    """UASTC mode index of each block (19 = reserved/invalid)."""
    blocks = np.asarray(blocks, np.uint8).reshape(-1, 16)
    lut = np.full(128, 255, np.uint8)
    for m, (code, ln) in enumerate(HUFF):
        for hi in range(1 << (7 - ln)):
            lut[code | (hi << ln)] = m
    return lut[blocks[:, 0] & 127]


def random_payload_blocks(seed, n_per_mode=512, include_invalid=True):
    """Random 128-bit payloads behind every mode prefix (trit/quint bundles may then hold out-of-range digits and
    pattern fields may be invalid -- exactly the corrupt-input space a transcoder must treat like the reference)."""
    r = np.random.default_rng(seed)
    out = []
    nm = 20 if include_invalid else 19
    for m in range(nm):
        code, ln = HUFF[m]
        b = r.integers(0, 256, (n_per_mode, 16), dtype=np.uint8)
        lo = b[:, 0].astype(np.uint32) | (b[:, 1].astype(np.uint32) << 8)
        lo = (lo & ~np.uint32((1 << ln) - 1)) | np.uint32(code)
        b[:, 0] = lo & 255
        b[:, 1] = (lo >> 8) & 255
        out.append(b)
    return np.concatenate(out)


HINT_BITS = [15, 15, 15, 15, 15, 15, 15, 15, 0, 23, 17, 17, 17, 23, 23, 23, 23, 23, 15]


def hint_mask(blocks):
    """uint8 [n,16] mask with 1-bits where a block stores transcoding hints (mode-dependent position)."""
    blocks = np.asarray(blocks, np.uint8).reshape(-1, 16)
    modes = block_modes(blocks)
    masks = np.zeros((20, 16), np.uint8)
    for m in range(19):
        ln = HUFF[m][1]
        lo, hi = (ln + 32, ln + 32 + 21) if m == 8 else (ln, ln + HINT_BITS[m])
        v = ((1 << hi) - 1) ^ ((1 << lo) - 1)
        masks[m] = np.frombuffer(v.to_bytes(16, "little"), np.uint8)
    return masks[np.minimum(modes, 19)]


def port_encode_blocks(px, flags):
    px = np.ascontiguousarray(px, dtype=np.uint8).reshape(-1, 64)
    out = np.zeros((px.shape[0], 16), np.uint8)
    f = port().port_encode_blocks
    f.argtypes = [_u8p, ctypes.c_uint64, ctypes.c_uint32, _u8p]
    f.restype = None
    f(_p(px), px.shape[0], flags, _p(out))
    return out


RDO_DEFAULTS = dict(lam=1.0, dict_size=4096, max_rms_ratio=10.0, skip_rms=8.0, smooth_scale=10.0, smooth_std_dev=18.0)


def ref_rdo(blocks, px, flags, threads=4, total_jobs=0, **kw):
    """Reference uastc_rdo in place on a copy.  total_jobs as basisu_comp.cpp:1576 (min(4, threads) or 0)."""
    k = dict(RDO_DEFAULTS); k.update(kw)
    b = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16).copy()
    px = np.ascontiguousarray(px, dtype=np.uint8).reshape(-1, 64)
    rc = ref().ref_rdo(_p(b), b.shape[0], _p(px), k["lam"], k["dict_size"], k["max_rms_ratio"], k["skip_rms"], k["smooth_scale"],
                       k["smooth_std_dev"], flags, threads, total_jobs)
    assert rc == 0
    return b


def port_rdo(blocks, px, flags, total_jobs=0, **kw):
    k = dict(RDO_DEFAULTS); k.update(kw)
    b = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1, 16).copy()
    px = np.ascontiguousarray(px, dtype=np.uint8).reshape(-1, 64)
    f = port().port_rdo
    f.restype = None
    f.argtypes = [_u8p, ctypes.c_uint32, _u8p, ctypes.c_float, ctypes.c_uint32, ctypes.c_float, ctypes.c_float, ctypes.c_float,
                  ctypes.c_float, ctypes.c_uint32, ctypes.c_uint32]
    f(_p(b), b.shape[0], _p(px), k["lam"], k["dict_size"], k["max_rms_ratio"], k["skip_rms"], k["smooth_scale"], k["smooth_std_dev"],
      flags, total_jobs)
    return b
