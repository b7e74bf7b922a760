/* Entry points of libktx that the UASTC boundary test never reaches but that the object files of this subset reference
 * (ASTC codec, ETC unpack, GL / Vulkan upload): each fails with KTX_UNSUPPORTED_FEATURE.  Test infrastructure only. */
#include "ktx.h"
struct ktxAstcParams;
KTX_error_code ktxTexture2_CompressAstcEx(ktxTexture2* This, struct ktxAstcParams* params) { (void)This; (void)params; return KTX_UNSUPPORTED_FEATURE; }
KTX_error_code ktxTexture2_CompressAstc(ktxTexture2* This, ktx_uint32_t quality) { (void)This; (void)quality; return KTX_UNSUPPORTED_FEATURE; }
KTX_error_code ktxTexture2_DecodeAstc(ktxTexture2* This) { (void)This; return KTX_UNSUPPORTED_FEATURE; }
