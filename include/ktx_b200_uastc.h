/* ktx_b200_uastc.h -- C ABI of the B200 (sm_100a) UASTC LDR 4x4 module.
 *
 * This is the drop-in boundary: the only symbols libktx's glue (lib/src/basis_encode.cpp,
 * lib/src/basis_transcode.cpp) needs in order to route UASTC LDR 4x4 work to the GPU.  Plain C types only.
 * Each entry point names the reference interface it replaces (paths relative to the KTX-Software checkout;
 * ENC/ = external/basis_universal/encoder/, XC/ = external/basis_universal/transcoder/).
 *
 * Threading: every call is re-entrant; each in-flight call owns its CUDA streams and staging buffers (a lane from a per-device pool).
 * Alignment: device pointers to blocks (UASTC in, any block target out, 64-byte texel blocks) must be 16-byte aligned
 * (B200_E_INVALID_VALUE otherwise: the kernels move them with 16-byte vector and bulk-copy accesses); image pointers may have any
 * 4-byte alignment and any pitch -- rows that are not 16-byte aligned take a narrower load path.
 * Errors: 0 = success; negative = B200_E_* below; b200_uastc_last_error() returns a thread-local message.
 * There is NO CPU fallback: if no sm_100 device is usable the calls fail with B200_E_NO_DEVICE.
 */
#ifndef KTX_B200_UASTC_H
#define KTX_B200_UASTC_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_E_NO_DEVICE (-1)     /* maps to KTX_UNSUPPORTED_FEATURE in the glue */
#define B200_E_INVALID_VALUE (-2) /* KTX_INVALID_VALUE */
#define B200_E_OUT_OF_MEMORY (-3) /* KTX_OUT_OF_MEMORY */
#define B200_E_CUDA (-4)          /* KTX_INVALID_OPERATION */
#define B200_E_BAD_BLOCK (-5)     /* KTX_TRANSCODE_FAILED: the input holds an invalid UASTC block */

/* Transcode targets; numeric values are ktx_transcode_fmt_e's (lib/include/ktx.h:1696-1703). */
#define B200_TTF_BC7_RGBA 6
#define B200_TTF_ASTC_4x4_RGBA 10
/* the hint-driven block targets (b200_uastc_transcode_ex): ktx_transcode_fmt_e values (lib/include/ktx.h:1654-1676);
 * ETC1_RGB, BC1_RGB and BC4_R write 8 bytes per block, the others 16 */
#define B200_TTF_ETC1_RGB 0
#define B200_TTF_ETC2_RGBA 1
#define B200_TTF_BC1_RGB 2
#define B200_TTF_BC3_RGBA 3
#define B200_TTF_BC4_R 4
#define B200_TTF_BC5_RG 5
#define B200_TTF_ETC2_EAC_R11 20  /* 8 bytes per block  (lib/include/ktx.h:1718) */
#define B200_TTF_ETC2_EAC_RG11 21 /* 16 bytes per block (lib/include/ktx.h:1722) */
/* decode flags the path honours: ktx_transcode_flag_bits_e KTX_TF_TRANSCODE_ALPHA_DATA_TO_OPAQUE_FORMATS and KTX_TF_HIGH_QUALITY
 * == basist cDecodeFlagsTranscodeAlphaDataToOpaqueFormats / cDecodeFlagsHighQuality (XC/basisu_transcoder.h:253,265) */
#define B200_DECODE_FLAG_ALPHA_TO_OPAQUE 4u
#define B200_DECODE_FLAG_HIGH_QUALITY 32u
/* uncompressed targets (b200_uastc_decode*): KTX_TTF_RGBA32 / RGB565 / BGR565 / RGBA4444 (lib/include/ktx.h:1728-1737) */
#define B200_TTF_RGBA32 13
#define B200_TTF_RGB565 14
#define B200_TTF_BGR565 15
#define B200_TTF_RGBA4444 16

/* Replaces basisu_encoder_init() (lib/src/basis_encode.cpp:625-633) and basisu_transcoder_init()
 * (lib/src/basis_transcode.cpp:427-436).  Idempotent and thread-safe.  Selects the current CUDA device
 * of the calling thread (or device 0), creates the stream pool. */
int b200_uastc_init(void);
/* Frees every idle stream set and staging buffer of every device; calls still in flight on other threads finish normally. */
void b200_uastc_shutdown(void);
/* Devices the HOST-buffer entry points (b200_uastc_encode*, b200_uastc_transcode*, b200_uastc_decode) spread a call over: bit d =
 * CUDA device ordinal d (d < 16).  0 (the default) = the calling thread's current device.  Replaces the reference's
 * ktxBasisParams::threadCount / job pool as the "how wide" knob (lib/src/basis_encode.cpp:1017-1029, ENC/basisu_comp.cpp:1503).
 * Every named device must be sm_100 (B200_E_NO_DEVICE otherwise).  Process-wide, thread-safe. */
int b200_uastc_set_devices(uint64_t device_mask);
/* Number of CUDA devices this library can run on (sm_100 class). */
int b200_uastc_device_count(void);
const char* b200_uastc_last_error(void);
/* Identifies the sources this binary was built from (SHA-1 prefix; bench.py quotes ncu counters only for a matching build). */
const char* b200_uastc_build_id(void);
/* Number of kernel launches this library has issued since init (bench.py's gpu_launches). */
uint64_t b200_uastc_launch_count(void);

/* One (level, layer, face/slice) image's worth of blocks inside the texture's pData. */
typedef struct b200_xcode_span {
  uint64_t src_offset; /* byte offset of the image's first UASTC block in src_base */
  uint64_t dst_offset; /* byte offset of its first output block in dst_base       */
  uint64_t n_blocks;   /* num_blocks_x * num_blocks_y                              */
} b200_xcode_span;

/* Replaces the per-image loop of transcodeUastcLDR4x4 (lib/src/basis_transcode.cpp:743-804), i.e. every call of
 * basisu_lowlevel_uastc_ldr_4x4_transcoder::transcode_image (XC/basisu_transcoder.cpp:10331) -> transcode_slice
 * (:10068) -> transcode_uastc_to_bc7 (:16172) / transcode_uastc_to_astc (:15649), for a whole texture in one call.
 * HOST pointers (pageable or pinned; pinned is DMA'd directly).  On B200_E_BAD_BLOCK *bad_span (if non-NULL)
 * receives the index of the first span holding an invalid block, as transcode_slice's `return false` would
 * have surfaced it (XC:10299-10303). */
int b200_uastc_transcode(uint32_t target, const b200_xcode_span* spans, uint32_t n_spans,
                         const uint8_t* src_base, uint8_t* dst_base, int64_t* bad_span);

/* Device-resident variant (roofline runs; no PCIe in the timed region).  d_src/d_dst are device pointers to
 * n_blocks contiguous 16-byte blocks; `stream` is a cudaStream_t (NULL = the library's stream for this thread).
 * d_first_bad (device uint64_t*, may be NULL) receives min(index of invalid blocks) or UINT64_MAX.
 * Asynchronous with respect to the host when a stream is given. */
int b200_uastc_transcode_dev(uint32_t target, const void* d_src, void* d_dst, uint64_t n_blocks, void* stream,
                             uint64_t* d_first_bad);

/* The same two entry points with what libktx forwards to transcode_image beside the target: transcodeFlags and
 * (alphaContent != eNone) (lib/src/basis_transcode.cpp:788-789), and with the hint-driven targets ETC1_RGB, ETC2_RGBA,
 * BC1_RGB, BC3_RGBA, BC4_R, BC5_RG, ETC2_EAC_R11, ETC2_EAC_RG11: transcode_slice's cETC1 / cETC2_RGBA / cBC1 / cBC3 / cBC4 / cBC5 /
 * cETC2_EAC_R11 / cETC2_EAC_RG11 cases (XC:10141-10207) -> transcode_uastc_to_etc1 (:16444), _etc2_rgba (:17341), _bc1 (:18365),
 * _bc3 (:18404), _bc4 (:18442), _bc5 (:18470), _etc2_eac_r11 (:18792), _etc2_eac_rg11 (:18819).
 * dst offsets and sizes count 8 bytes per block for ETC1_RGB, BC1_RGB, BC4_R and ETC2_EAC_R11.  BC4_R and ETC2_EAC_R11 take
 * alpha instead of red when has_alpha and B200_DECODE_FLAG_ALPHA_TO_OPAQUE are set (XC:10417, :10515); BC5_RG and
 * ETC2_EAC_RG11 take red and alpha (XC:10430, :10526); ETC1_RGB with has_alpha and B200_DECODE_FLAG_ALPHA_TO_OPAQUE encodes the
 * alpha channel alone (transcode_uastc_to_etc1(blk, dst, 3), XC:10142-10147, :16560).  Device pointers must be
 * 16-byte aligned (B200_E_INVALID_VALUE otherwise). */
int b200_uastc_transcode_ex(uint32_t target, uint32_t decode_flags, uint32_t has_alpha, const b200_xcode_span* spans,
                            uint32_t n_spans, const uint8_t* src_base, uint8_t* dst_base, int64_t* bad_span);
int b200_uastc_transcode_dev_ex(uint32_t target, uint32_t decode_flags, uint32_t has_alpha, const void* d_src, void* d_dst,
                                uint64_t n_blocks, void* stream, uint64_t* d_first_bad);

/* ---- decode to texels (the uncompressed transcode targets) ------------------------------------------ */

/* One (level, layer, face/slice) image: its UASTC blocks in raster order at src_base + src_offset, and where its
 * width x height texels go (tight rows of `width` texels, as libktx sizes the transcoded level: it passes
 * output_row_pitch_in_blocks_or_pixels = 0 and output_rows_in_pixels = 0, lib/src/basis_transcode.cpp:790-793). */
typedef struct b200_decode_image {
  uint64_t src_offset;
  uint64_t dst_offset;
  uint32_t width, height; /* texels; the image holds ceil(width/4) * ceil(height/4) blocks */
} b200_decode_image;

/* Replaces the same per-image loop of transcodeUastcLDR4x4 (lib/src/basis_transcode.cpp:743-804) for the targets
 * KTX_TTF_RGBA32, KTX_TTF_RGB565, KTX_TTF_BGR565 and KTX_TTF_RGBA4444, i.e. transcode_image (XC/basisu_transcoder.cpp:10331)
 * -> transcode_slice's cRGBA32 / cRGB565 / cBGR565 / cRGBA4444 cases (XC:10208-10292) -> unpack_uastc to pixels
 * (XC:15378-15529, srgb = false); texels past width / height are clipped as there.  HOST pointers.  On
 * B200_E_BAD_BLOCK *bad_image (if non-NULL) receives the index of the first image holding an invalid block. */
int b200_uastc_decode(uint32_t target, const b200_decode_image* images, uint32_t n_images, const uint8_t* src_base,
                      uint8_t* dst_base, int64_t* bad_image);

/* Device-resident variant, one image: d_src = ceil(w/4)*ceil(h/4) blocks, d_dst = height rows of
 * dst_row_pitch_texels (0 = width) texels of 4 (RGBA32) or 2 bytes.  d_first_bad as in b200_uastc_transcode_dev.
 * Asynchronous with respect to the host when a stream is given. */
int b200_uastc_decode_image_dev(uint32_t target, const void* d_src, void* d_dst, uint32_t width, uint32_t height,
                                uint32_t dst_row_pitch_texels, void* stream, uint64_t* d_first_bad);

/* ---- encode ---------------------------------------------------------------------------------------- */

typedef struct b200_image {
  const uint8_t* rgba;      /* RGBA8, post-swizzle (lib/src/basis_encode.cpp:806-833 did the copy/swizzle) */
  uint32_t width, height;   /* in pixels; edge texels are replicated to a multiple of 4 (ENC/basisu_comp.cpp:2449) */
  uint32_t row_pitch_bytes; /* 0 = width*4 */
} b200_image;

/* Mirrors the fields of ktxBasisParams that reach the UASTC LDR encoder (lib/include/ktx.h:1455-1625) after the
 * glue's mapping (lib/src/basis_encode.cpp:881-918). */
typedef struct b200_uastc_params {
  uint32_t struct_size;      /* sizeof(b200_uastc_params) */
  uint32_t pack_flags;       /* ktx_pack_uastc_flag_bits_e == basisu cPackUASTC* (ENC/basisu_uastc_enc.h:24-63) */
  uint32_t rdo_enable;       /* uastcRDO */
  float rdo_lambda;          /* uastcRDOQualityScalar, 0 -> 1.0 */
  uint32_t rdo_dict_size;    /* uastcRDODictSize, 0 -> 4096 */
  float rdo_max_smooth_block_error_scale; /* 0 -> 10 */
  float rdo_max_smooth_block_std_dev;     /* 0 -> 18 */
  uint32_t rdo_total_jobs;   /* min(4, threadCount) as ENC/basisu_comp.cpp:1574-1576; 0/1 = one serial chain */
  uint32_t device_mask;      /* bit d = use CUDA device d for this call; 0 = the set of b200_uastc_set_devices, else the calling
                              * thread's current device.  The call's images are cut into block-row slabs and spread over the
                              * devices the way the reference spreads 256-block jobs over its pool (ENC/basisu_comp.cpp:1503-1509);
                              * with rdo_enable whole images are spread (a job range never crosses an image, :1574-1576).
                              * Every device writes its slabs straight into out_blocks: the output bytes do not depend on the mask. */
  uint32_t reserved[3];
} b200_uastc_params;

/* Replaces basis_compressor::process -> encode_slices_to_uastc_4x4_ldr (ENC/basisu_comp.cpp:1474-1591) and the
 * .basis unpicking (lib/src/basis_encode.cpp:1084-1150) for codec KTX_BASIS_CODEC_UASTC_LDR_4x4.
 * out_blocks[i] receives ceil(w/4)*ceil(h/4)*16 bytes in raster block order.  *out_any_alpha = 1 if any input
 * texel had alpha < 255 (what cBASISHeaderFlagHasAlphaSlices told the glue, basis_encode.cpp:1284-1286). */
int b200_uastc_encode(const b200_image* images, uint32_t n_images, const b200_uastc_params* params,
                      uint8_t* const* out_blocks, uint32_t* out_any_alpha);

/* The same call taking the texture's own texel layout, so the glue can hand in pData + image offset without its
 * copy / swizzle pass (lib/src/basis_encode.cpp:806-833: copy_rgba_to_rgba, copy_rgb_to_rgba, swizzle_to_rgba :93-144): images[i].rgba
 * then points at num_components bytes per texel (row_pitch_bytes 0 = width * num_components) and swizzle[c] says what
 * RGBA channel c receives, in libktx's swizzle_e values (basis_encode.cpp:77-84): 1..4 = source component R, G, B, A,
 * 5 = 0x00, 6 = 0xFF.  The glue's defaults map to: 4 components {1,2,3,4}; 3 components {1,2,3,6}; 2 components (UASTC)
 * {1,2,5,6}; 1 component {1,1,1,6}; normal map {1,1,1,2}.  A swizzle naming a component the source does not have is
 * B200_E_INVALID_VALUE (the reference reads the neighbouring texel's bytes there).  format = NULL: RGBA8 as above.
 * The expansion runs on the device (k_expand_swizzle), so the upload is 1-4 bytes per texel. */
typedef struct b200_src_format {
  uint32_t num_components; /* 1..4 */
  uint8_t swizzle[4];      /* swizzle_e per output channel */
} b200_src_format;
int b200_uastc_encode_ex(const b200_image* images, uint32_t n_images, const b200_src_format* format,
                         const b200_uastc_params* params, uint8_t* const* out_blocks, uint32_t* out_any_alpha);

/* b200_uastc_encode_ex with a call per finished image: lets the host start the Zstd / ZLIB supercompression of a
 * level (ktxTexture2_DeflateZstd, lib/src/texture2.c; `ktx create --zstd`, tools/ktx/command_create.cpp:2077-2090) while later
 * levels are still being encoded.  on_done(i, user) is called once per image, after ALL bytes of out_blocks[i] are in host
 * memory and before the call returns.  Without RDO it runs between batch submissions on the calling thread, with up to two later
 * batches in flight on the GPU -- time spent inside it is hidden as long as the GPU has queued work; images of one device finish in
 * submission order.  When a device mask names several GPUs the call comes from whichever device's worker thread retires the
 * image's last slab: on_done must then be thread-safe.  With rdo_enable the calls come after the device's walk, in image order.
 * on_done must not call back into this library with the same lane-hungry entry points recursively (it may call any entry point
 * from another thread).  on_done = NULL: exactly b200_uastc_encode_ex.  No counterpart in the reference (its compress and
 * deflate steps are separate, serial API calls). */
typedef void (*b200_image_done_fn)(uint32_t image_index, void* user);
int b200_uastc_encode_stream(const b200_image* images, uint32_t n_images, const b200_src_format* format,
                             const b200_uastc_params* params, uint8_t* const* out_blocks, uint32_t* out_any_alpha,
                             b200_image_done_fn on_done, void* user);

/* Device-resident pipeline: only the BASE level crosses PCIe.  Levels 1 .. n_levels-1 (level l is
 * max(1, w >> l) x max(1, h >> l)) are filtered from the level above on the device with the 2x2 box filter, per channel
 * (a + b + c + d + 2) >> 2 with source coordinates clamped to the level, then every level is encoded.  This is the integer box
 * filter of the measurement harness; `ktx create --generate-mipmap` runs a float resampler on the host before libktx is
 * called (tools/ktx/command_create.cpp:1994-1996), which this does not replace.  base: tightly packed RGBA8.  out_blocks[l]
 * receives level l's blocks; out_level_rgba (may be NULL, entries may be NULL) receives the generated RGBA8 levels l >= 1.
 * No RDO (B200_E_INVALID_VALUE with rdo_enable).  Runs on the first device of the mask. */
int b200_uastc_encode_mipchain(const b200_image* base, uint32_t n_levels, const b200_uastc_params* params,
                               uint8_t* const* out_blocks, uint8_t* const* out_level_rgba, uint32_t* out_any_alpha);

/* Device-resident variant: d_block_pixels = n_blocks x 64 bytes (4x4 RGBA8 texels, raster within the block, as
 * basis_compressor::extract_source_blocks lays them out, ENC/basisu_comp.cpp:2705-2766); d_out = n_blocks x 16 B. */
int b200_uastc_encode_blocks_dev(const void* d_block_pixels, void* d_out, uint64_t n_blocks, uint32_t pack_flags,
                                 void* stream);

/* Device-resident image variant: d_rgba = one RGBA8 image in device memory (row_pitch_bytes 0 = width*4); the 4x4
 * gather with edge replication (image::extract_block_clamped, ENC/basisu_enc.h:3081) happens inside the kernel.
 * d_out = ceil(w/4)*ceil(h/4) x 16 B.  d_any_alpha (device uint32_t*, may be NULL) is OR-ed with 1 when a texel has
 * alpha < 255.  Asynchronous when a stream is given. */
int b200_uastc_encode_image_dev(const void* d_rgba, uint32_t width, uint32_t height, uint32_t row_pitch_bytes, void* d_out,
                                uint32_t pack_flags, void* stream, uint32_t* d_any_alpha);

/* Allocation-free variant of b200_uastc_encode_image_dev.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same arguments, same output bytes.  The encode kernel's scratch (cell results and candidate scores of the tiles in
 * flight, ~131 MiB) is a static array inside the library's device image: every GPU that loads the library holds it until the
 * process exits.  The call makes no cudaMalloc / cudaFree / cudaMallocAsync and creates no stream (stream = NULL runs on the
 * legacy default stream and synchronises it).  Limits: pack levels 0-3 only, and at most B200_NOALLOC_MAX_BLOCKS 4x4 blocks per
 * call (an 8192 x 8192 image); beyond either, B200_E_INVALID_VALUE.  EVERY call on a device shares the one array: calls whose
 * kernels overlap on the GPU (different streams, different host threads) race on it -- order them on one stream or with events.
 * (The scratch keeps the previous call's contents; the kernel clears what must start at zero.) */
#define B200_NOALLOC_MAX_BLOCKS 4194304u
int b200_uastc_encode_image_dev_noalloc(const void* d_rgba, uint32_t width, uint32_t height, uint32_t row_pitch_bytes, void* d_out,
                                        uint32_t pack_flags, void* stream, uint32_t* d_any_alpha);

/* Several device-resident images in ONE launch (a mip chain, cube faces, array layers): small images then share the
 * launch with big ones instead of each under-filling the GPU.  Same semantics as b200_uastc_encode_image_dev per image. */
typedef struct b200_dev_image {
  const void* d_rgba;       /* RGBA8 in device memory, 4-byte aligned */
  uint32_t width, height;
  uint32_t row_pitch_bytes; /* 0 = width*4; multiple of 4 */
  void* d_out;              /* ceil(w/4)*ceil(h/4) x 16 B */
} b200_dev_image;
int b200_uastc_encode_images_dev(const b200_dev_image* images, uint32_t n_images, uint32_t pack_flags, void* stream,
                                 uint32_t* d_any_alpha);

/* The RDO post-pass alone, device-resident: replaces uastc_rdo (ENC/basisu_uastc_enc.cpp:4109) as
 * encode_slices_to_uastc_4x4_ldr calls it per slice (ENC/basisu_comp.cpp:1564-1577).  d_blocks: n_blocks UASTC blocks of ONE
 * image, rewritten in place; d_block_pixels: their n_blocks x 64 B source texels.  Uses params->rdo_* and
 * params->pack_flags (for the hint recomputation); job ranges follow params->rdo_total_jobs.  Synchronous. */
int b200_uastc_rdo_blocks_dev(void* d_blocks, const void* d_block_pixels, uint64_t n_blocks, const b200_uastc_params* params,
                              void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KTX_B200_UASTC_H */
