"""ctypes bindings over include/ktx_b200_uastc.h.  No compute happens in Python."""
import ctypes, os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libuastc_b200.so")

TTF_BC7_RGBA = 6        # ktx_transcode_fmt_e KTX_TTF_BC7_RGBA (lib/include/ktx.h)
TTF_ASTC_4x4_RGBA = 10  # KTX_TTF_ASTC_4x4_RGBA

EXPORTED_SYMBOLS = ["b200_uastc_init", "b200_uastc_shutdown", "b200_uastc_last_error", "b200_uastc_launch_count",
                    "b200_uastc_transcode", "b200_uastc_transcode_dev", "b200_uastc_encode", "b200_uastc_encode_blocks_dev"]

_u8p = ctypes.POINTER(ctypes.c_uint8)


class B200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("b200_uastc error %d: %s" % (code, msg))
        self.code = code


class XcodeSpan(ctypes.Structure):
    _fields_ = [("src_offset", ctypes.c_uint64), ("dst_offset", ctypes.c_uint64), ("n_blocks", ctypes.c_uint64)]


class Image(ctypes.Structure):
    _fields_ = [("rgba", ctypes.c_void_p), ("width", ctypes.c_uint32), ("height", ctypes.c_uint32), ("row_pitch_bytes", ctypes.c_uint32)]


class UastcParams(ctypes.Structure):
    _fields_ = [("struct_size", ctypes.c_uint32), ("pack_flags", ctypes.c_uint32), ("rdo_enable", ctypes.c_uint32),
                ("rdo_lambda", ctypes.c_float), ("rdo_dict_size", ctypes.c_uint32),
                ("rdo_max_smooth_block_error_scale", ctypes.c_float), ("rdo_max_smooth_block_std_dev", ctypes.c_float),
                ("rdo_total_jobs", ctypes.c_uint32), ("reserved", ctypes.c_uint32 * 4)]


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
        l.b200_uastc_transcode.argtypes = [ctypes.c_uint32, ctypes.POINTER(XcodeSpan), ctypes.c_uint32, ctypes.c_void_p,
                                           ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64)]
        l.b200_uastc_transcode_dev.argtypes = [ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                                               ctypes.c_void_p, ctypes.c_void_p]
        l.b200_uastc_encode.argtypes = [ctypes.POINTER(Image), ctypes.c_uint32, ctypes.POINTER(UastcParams),
                                        ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint32)]
        l.b200_uastc_encode_blocks_dev.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_void_p]
        _lib = l
    return _lib


def _check(rc):
    if rc != 0:
        raise B200Error(rc, lib().b200_uastc_last_error().decode())


def init():
    _check(lib().b200_uastc_init())


def launch_count():
    return int(lib().b200_uastc_launch_count())


def transcode(src, target, spans=None, dst=None):
    """Host-buffer transcode through the C ABI.  src: uint8 array holding UASTC blocks; spans: list of
    (src_offset, dst_offset, n_blocks) or None for one span over the whole buffer.  Returns the dst array.
    Raises B200Error(code=-5) with .bad_span set when the input holds an invalid block."""
    src = np.ascontiguousarray(src, dtype=np.uint8).reshape(-1)
    if spans is None:
        spans = [(0, 0, src.size // 16)]
    if dst is None:
        need = max([d + n * 16 for _, d, n in spans] + [0])
        dst = np.zeros(need, np.uint8)
    arr = (XcodeSpan * max(1, len(spans)))(*[XcodeSpan(s, d, n) for s, d, n in spans])
    bad = ctypes.c_int64(-1)
    rc = lib().b200_uastc_transcode(target, arr, len(spans), src.ctypes.data, dst.ctypes.data, ctypes.byref(bad))
    if rc != 0:
        e = B200Error(rc, lib().b200_uastc_last_error().decode())
        e.bad_span = bad.value
        raise e
    return dst


def transcode_dev(target, d_src, d_dst, n_blocks, stream=0, d_first_bad=0):
    """Device-pointer transcode (ints / torch data_ptr()); asynchronous when a stream handle is given."""
    _check(lib().b200_uastc_transcode_dev(target, d_src, d_dst, n_blocks, stream or None, d_first_bad or None))
