"""The drop-in boundary COMPILED: the reference's own libktx built with integration/basis_encode.patch and
integration/basis_transcode.patch (integration/Makefile -> oracle/_ref/ktx/libktx_b200.so, test infrastructure) and driven
through its public API -- ktxTexture2_CreateFromMemory / ktxTexture2_Create, ktxTexture2_CompressBasisEx,
ktxTexture2_TranscodeBasis (lib/include/ktx.h:1627, :1823) -- against the reference's CTS goldens
(tests/cts/clitests: transcode/transcode_uastc, create/encode_uastc) and the unmodified reference encoder."""
import ctypes, json, os
import numpy as np
import pytest
import refutil as R

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(os.path.dirname(__file__), "golden")
KTX_SO = os.path.join(ROOT, "oracle", "_ref", "ktx", "libktx_b200.so")
needs_ktx = pytest.mark.skipif(not os.path.exists(KTX_SO), reason="patched libktx not built (make -C integration)")
_u8p = ctypes.POINTER(ctypes.c_uint8)
_u64p = ctypes.POINTER(ctypes.c_uint64)


@pytest.fixture(scope="module")
def ktx():
    import ktx_software_b200 as K
    K.init()
    ctypes.CDLL(K.api.LIB_PATH, mode=ctypes.RTLD_GLOBAL)   # the patched libktx links against the product library
    l = ctypes.CDLL(KTX_SO)
    l.drv_transcode.argtypes = [_u8p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32, _u8p, _u64p]
    l.drv_compress_uastc.argtypes = [_u8p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint32,
                                     ctypes.c_int, ctypes.c_float, ctypes.c_uint32, ctypes.c_int, _u8p, _u64p, _u64p]
    l.drv_compress_then_transcode.argtypes = [_u8p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, _u8p, _u64p]
    return l


def _transcode(ktx, file_bytes, fmt, flags=0, cap=1 << 20):
    src = np.frombuffer(file_bytes, np.uint8).copy()
    out = np.zeros(cap, np.uint8)
    n = ctypes.c_uint64(cap)
    rc = ktx.drv_transcode(src.ctypes.data_as(_u8p), len(src), fmt, flags, out.ctypes.data_as(_u8p), ctypes.byref(n))
    return rc, out[:n.value]


def _compress(ktx, img, flags, levels=None, srgb=False, rdo=False, lam=0.0, threads=1, no_mt=False):
    chain = levels if levels is not None else [img]
    h, w, comps = chain[0].shape
    px = np.concatenate([np.ascontiguousarray(l).reshape(-1) for l in chain])
    cap = sum(((l.shape[0] + 3) // 4) * ((l.shape[1] + 3) // 4) * 16 for l in chain) + 64
    out = np.zeros(cap, np.uint8)
    n = ctypes.c_uint64(cap)
    offs = (ctypes.c_uint64 * len(chain))()
    rc = ktx.drv_compress_uastc(px.ctypes.data_as(_u8p), w, h, comps, int(srgb), len(chain), flags, int(rdo), lam, threads, int(no_mt),
                                out.ctypes.data_as(_u8p), ctypes.byref(n), offs)
    return rc, out[:n.value], list(offs)


@needs_ktx
def test_cts_transcode_uastc_through_patched_libktx(ktx):
    """tests/cts/clitests/tests/transcode/transcode_uastc.json: the 8 valid_*_2D_UASTC.ktx2 inputs -> BC7 and ASTC goldens."""
    cts = json.load(open(os.path.join(GOLD, "cts_uastc.json")))["transcode"]
    for sub, c in cts.items():
        f = bytes.fromhex(c["ktx2_file"])
        for fmt, key in ((6, "bc7"), (10, "astc")):
            rc, got = _transcode(ktx, f, fmt)
            assert rc == 0, (sub, key, rc)
            assert got.tobytes().hex() == c[key], (sub, key)
        # the other accelerated targets agree with the product's own C ABI output (itself checked against the reference elsewhere)
        import ktx_software_b200 as K
        u = np.frombuffer(bytes.fromhex(c["uastc"]), np.uint8)
        # alphaContent as ktxTexture2_TranscodeBasis derives it (basis_transcode.cpp:246-251): the DFD's first sample is
        # KHR_DF_CHANNEL_UASTC_RGBA (3) or _RRRG (5); an RGBA8 source whose alpha was all 255 is stored as UASTC_RGB (0)
        dfd_off = int.from_bytes(f[48:52], "little")
        has_alpha = (f[dfd_off + 4 + 24 + 3] & 0xF) in (3, 5)
        for fmt in (K.TTF_BC1_RGB, K.TTF_BC3_RGBA, K.TTF_ETC1_RGB, K.TTF_ETC2_RGBA, K.TTF_BC4_R, K.TTF_BC5_RG):
            rc, got = _transcode(ktx, f, fmt)
            assert rc == 0, (sub, fmt, rc)
        # ETC1 with KTX_TF_TRANSCODE_ALPHA_DATA_TO_OPAQUE_FORMATS: the alpha channel alone for textures that have alpha, plain ETC1 otherwise
        rc, got = _transcode(ktx, f, K.TTF_ETC1_RGB, flags=4)
        assert rc == 0 and (got == K.transcode(u, K.TTF_ETC1_RGB, decode_flags=4, has_alpha=has_alpha)).all(), sub
        if R.have_ref():
            assert (got.reshape(-1, 8) == R.ref_transcode_hinted(u, 0, 4, has_alpha)[0]).all(), sub
        rc, texels = _transcode(ktx, f, K.TTF_RGBA32)
        assert rc == 0 and (texels.reshape(8, 8, 4) == K.decode(u, K.TTF_RGBA32, [(0, 8, 8)])[0]).all()


@needs_ktx
def test_cts_create_encode_uastc_through_patched_libktx(ktx):
    """tests/cts/clitests/tests/create/encode_uastc.json (`ktx create --encode uastc`, level 2, 1 thread): raw 8x8 inputs of 1-4
    components through ktxTexture2_CompressBasisEx -> the golden files' payloads."""
    cts = json.load(open(os.path.join(GOLD, "cts_uastc.json")))["encode"]
    n = 0
    for sub, c in cts.items():
        if not c.get("raw"):
            continue
        img = np.frombuffer(bytes.fromhex(c["raw"]), np.uint8).reshape(c["height"], c["width"], c["comps"])
        rc, got, _ = _compress(ktx, img, 2)
        assert rc == 0, (sub, rc)
        assert got.tobytes().hex() == c["uastc"], sub
        n += 1
    assert n == 4


@needs_ktx
@pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not shipped")
def test_compress_basis_ex_mip_chain_and_rdo(ktx):
    """A mip-mapped RGBA8 texture through ktxTexture2_CompressBasisEx: level layout of the new pData (smallest level first,
    16-byte aligned) and every level's blocks against the unmodified reference encoder; then uastcRDO with the glue's
    parameter mapping (threadCount -> job ranges, the FavorSimplerModes quirk of basis_encode.cpp:900-903)."""
    base = R.img_mix(64, 40, 5)
    chain = [base]
    for l in range(1, 7):
        a = chain[-1].astype(np.uint16)
        sh, sw = a.shape[:2]
        lw, lh = max(1, 64 >> l), max(1, 40 >> l)
        xs0 = np.minimum(2 * np.arange(lw), sw - 1); xs1 = np.minimum(2 * np.arange(lw) + 1, sw - 1)
        ys0 = np.minimum(2 * np.arange(lh), sh - 1); ys1 = np.minimum(2 * np.arange(lh) + 1, sh - 1)
        chain.append(((a[ys0][:, xs0] + a[ys0][:, xs1] + a[ys1][:, xs0] + a[ys1][:, xs1] + 2) >> 2).astype(np.uint8))
    rc, data, offs = _compress(ktx, None, 2, levels=chain, threads=4)
    assert rc == 0
    sizes = [((l.shape[0] + 3) // 4) * ((l.shape[1] + 3) // 4) * 16 for l in chain]
    assert offs[-1] == 0 and all(offs[i] == offs[i + 1] + sizes[i + 1] for i in range(len(chain) - 1))   # smallest first, no gaps (multiples of 16)
    for l, img in enumerate(chain):
        want = R.ref_encode_blocks(R.ref_extract_blocks(img), 2)
        assert (data[offs[l]:offs[l] + sizes[l]].reshape(-1, 16) == want).all(), l
    # RDO: flags 1 | FavorSimplerModes (set by the glue because uastcRDONoMultithreading is false), 4 job ranges
    img = R.img_mix(128, 96, 3)
    rc, data, _ = _compress(ktx, img, 1, rdo=True, lam=1.5, threads=8)
    assert rc == 0
    px = R.ref_extract_blocks(img)
    want = R.ref_rdo(R.ref_encode_blocks(px, 1 | 512), px, 1 | 512, 4, 4, lam=1.5)
    assert (data.reshape(-1, 16) == want).all()
    rc, data, _ = _compress(ktx, img, 1, rdo=True, lam=1.5, threads=1, no_mt=True)    # one serial chain, no FavorSimplerModes
    want = R.ref_rdo(R.ref_encode_blocks(px, 1), px, 1, 1, 1, lam=1.5)
    assert rc == 0 and (data.reshape(-1, 16) == want).all()


@needs_ktx
@pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not shipped")
def test_round_trip_in_one_texture_object(ktx):
    """CompressBasisEx then TranscodeBasis on the same ktxTexture2 (the prototype swap of basis_transcode.cpp:448-478 runs)."""
    img = R.img_mix(96, 64, 8)
    out = np.zeros(1 << 16, np.uint8)
    n = ctypes.c_uint64(out.size)
    rc = ktx.drv_compress_then_transcode(np.ascontiguousarray(img).ctypes.data_as(_u8p), 96, 64, 2, 6, out.ctypes.data_as(_u8p), ctypes.byref(n))
    assert rc == 0
    u = R.ref_encode_blocks(R.ref_extract_blocks(img), 2)
    assert (out[:n.value].reshape(-1, 16) == R.ref_transcode(u, 0)[0]).all()
