/* Plain-C entry points over the PATCHED libktx for tests/test_libktx_boundary_gpu.py: they call the reference's public API
 * (ktxTexture2_Create / _CreateFromMemory / _CompressBasisEx / _TranscodeBasis, lib/include/ktx.h) the way an application does
 * and hand the resulting payload back.  Test infrastructure: linked into oracle/_ref/ktx/libktx_b200.so. */
#include <stdlib.h>
#include <string.h>
#include "ktx.h"

/* pixels: `levels` mip levels, level 0 first, tightly packed, comps bytes per texel, each level max(1, w >> l) x max(1, h >> l).
 * Encodes to UASTC through ktxTexture2_CompressBasisEx; copies the texture's pData to out (capacity *out_size) and stores its
 * size and the level offsets (level_offsets[levels]).  Returns the KTX_error_code. */
int drv_compress_uastc(const unsigned char* pixels, unsigned w, unsigned h, unsigned comps, int srgb, unsigned levels,
                       unsigned uastc_flags, int rdo, float rdo_lambda, unsigned thread_count, int no_multithreading,
                       unsigned char* out, unsigned long long* out_size, unsigned long long* level_offsets) {
  static const ktx_uint32_t fmt_unorm[5] = {0, 9 /*R8_UNORM*/, 16 /*R8G8_UNORM*/, 23 /*R8G8B8_UNORM*/, 37 /*R8G8B8A8_UNORM*/};
  static const ktx_uint32_t fmt_srgb[5] = {0, 15, 22, 29, 43};
  ktxTextureCreateInfo ci;
  memset(&ci, 0, sizeof(ci));
  ci.vkFormat = srgb ? fmt_srgb[comps] : fmt_unorm[comps];
  ci.baseWidth = w; ci.baseHeight = h; ci.baseDepth = 1; ci.numDimensions = 2;
  ci.numLevels = levels; ci.numLayers = 1; ci.numFaces = 1; ci.isArray = KTX_FALSE; ci.generateMipmaps = KTX_FALSE;
  ktxTexture2* tex = NULL;
  KTX_error_code rc = ktxTexture2_Create(&ci, KTX_TEXTURE_CREATE_ALLOC_STORAGE, &tex);
  if (rc != KTX_SUCCESS) return (int)rc;
  const unsigned char* p = pixels;
  for (unsigned l = 0; l < levels && rc == KTX_SUCCESS; l++) {
    const unsigned lw = (w >> l) ? (w >> l) : 1, lh = (h >> l) ? (h >> l) : 1;
    rc = ktxTexture_SetImageFromMemory(ktxTexture(tex), l, 0, 0, p, (ktx_size_t)lw * lh * comps);
    p += (size_t)lw * lh * comps;
  }
  if (rc == KTX_SUCCESS) {
    ktxBasisParams bp;
    memset(&bp, 0, sizeof(bp));
    bp.structSize = sizeof(bp);
    bp.codec = KTX_BASIS_CODEC_UASTC_LDR_4x4;
    bp.threadCount = thread_count;
    bp.uastcFlags = uastc_flags;
    bp.uastcRDO = rdo ? KTX_TRUE : KTX_FALSE;
    bp.uastcRDOQualityScalar = rdo_lambda;
    bp.uastcRDONoMultithreading = no_multithreading ? KTX_TRUE : KTX_FALSE;
    rc = ktxTexture2_CompressBasisEx(tex, &bp);
  }
  if (rc == KTX_SUCCESS) {
    if (tex->dataSize > *out_size) rc = KTX_INVALID_VALUE;
    else {
      memcpy(out, tex->pData, tex->dataSize);
      for (unsigned l = 0; l < levels; l++) {
        ktx_size_t off = 0;
        ktxTexture_GetImageOffset(ktxTexture(tex), l, 0, 0, &off);
        level_offsets[l] = off;
      }
    }
    *out_size = tex->dataSize;
  }
  ktxTexture_Destroy(ktxTexture(tex));
  return (int)rc;
}

/* A .ktx2 file image (UASTC) -> ktxTexture2_TranscodeBasis(fmt, flags) -> its pData. */
int drv_transcode(const unsigned char* file, unsigned long long file_size, unsigned fmt, unsigned flags, unsigned char* out,
                  unsigned long long* out_size) {
  ktxTexture2* tex = NULL;
  KTX_error_code rc = ktxTexture2_CreateFromMemory(file, (ktx_size_t)file_size, KTX_TEXTURE_CREATE_LOAD_IMAGE_DATA_BIT, &tex);
  if (rc != KTX_SUCCESS) return (int)rc;
  rc = ktxTexture2_TranscodeBasis(tex, (ktx_transcode_fmt_e)fmt, (ktx_transcode_flags)flags);
  if (rc == KTX_SUCCESS) {
    if (tex->dataSize > *out_size) rc = KTX_INVALID_VALUE; else memcpy(out, tex->pData, tex->dataSize);
    *out_size = tex->dataSize;
  }
  ktxTexture_Destroy(ktxTexture(tex));
  return (int)rc;
}

/* Encode then transcode in one texture object: the full round trip an application makes. */
int drv_compress_then_transcode(const unsigned char* pixels, unsigned w, unsigned h, unsigned uastc_flags, unsigned fmt,
                                unsigned char* out, unsigned long long* out_size) {
  ktxTextureCreateInfo ci;
  memset(&ci, 0, sizeof(ci));
  ci.vkFormat = 37; ci.baseWidth = w; ci.baseHeight = h; ci.baseDepth = 1; ci.numDimensions = 2;
  ci.numLevels = 1; ci.numLayers = 1; ci.numFaces = 1;
  ktxTexture2* tex = NULL;
  KTX_error_code rc = ktxTexture2_Create(&ci, KTX_TEXTURE_CREATE_ALLOC_STORAGE, &tex);
  if (rc != KTX_SUCCESS) return (int)rc;
  rc = ktxTexture_SetImageFromMemory(ktxTexture(tex), 0, 0, 0, pixels, (ktx_size_t)w * h * 4);
  ktxBasisParams bp;
  memset(&bp, 0, sizeof(bp));
  bp.structSize = sizeof(bp); bp.codec = KTX_BASIS_CODEC_UASTC_LDR_4x4; bp.threadCount = 1; bp.uastcFlags = uastc_flags;
  if (rc == KTX_SUCCESS) rc = ktxTexture2_CompressBasisEx(tex, &bp);
  if (rc == KTX_SUCCESS) rc = ktxTexture2_TranscodeBasis(tex, (ktx_transcode_fmt_e)fmt, 0);
  if (rc == KTX_SUCCESS) {
    if (tex->dataSize > *out_size) rc = KTX_INVALID_VALUE; else memcpy(out, tex->pData, tex->dataSize);
    *out_size = tex->dataSize;
  }
  ktxTexture_Destroy(ktxTexture(tex));
  return (int)rc;
}
