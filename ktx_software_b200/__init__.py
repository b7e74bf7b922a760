"""ktx_software_b200 -- B200-native UASTC LDR 4x4 encode / transcode behind libktx's C API.

The product is ktx_software_b200/libuastc_b200.so (CUDA, sm_100a; C ABI in include/ktx_b200_uastc.h).  This
package is the thin host-side mirror used by the tests and the bench: ctypes bindings (api.py) and shard.py, the
Python mirror of the library's multi-device planner (test aid).  There is no CPU fallback anywhere in this package:
if the CUDA library is missing, or no B200 is visible, calls raise.
"""
from .api import (B200Error, lib, init, transcode, transcode_dev, launch_count, encode_images, encode_images_stream, encode_blocks, encode_blocks_dev, encode_image_dev, encode_images_dev, rdo_blocks_dev, TTF_BC7_RGBA, TTF_ASTC_4x4_RGBA,  # noqa: F401
                  decode, decode_image_dev, TTF_RGBA32, TTF_RGB565, TTF_BGR565, TTF_RGBA4444,
                  transcode_dev_ex, target_block_bytes, encode_images_ex, TTF_ETC1_RGB, TTF_ETC2_RGBA, TTF_BC1_RGB, TTF_BC3_RGBA, TTF_BC4_R, TTF_BC5_RG, TTF_ETC2_EAC_R11, TTF_ETC2_EAC_RG11,
                  TF_TRANSCODE_ALPHA_DATA_TO_OPAQUE_FORMATS, TF_HIGH_QUALITY, EXPORTED_SYMBOLS, set_devices, device_count, shutdown, build_id, encode_mipchain, encode_image_dev_noalloc)
