// ref_shim.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// Thin extern "C" wrapper over the UNMODIFIED reference Basis Universal sources, compiled where
// they lie under /root/reference by oracle/Makefile into oracle/_ref/libref_uastc.so.
// Used by tests/ (parity checker), __graft_entry__.smoke() and bench.py's reference arm only.
//
// Reference entry points wrapped (file:line relative to /root/reference):
//   external/basis_universal/encoder/basisu_uastc_enc.cpp:3127  encode_uastc
//   external/basis_universal/encoder/basisu_uastc_enc.cpp:4109  uastc_rdo
//   external/basis_universal/encoder/basisu_comp.cpp:1503-1531  256-block job fan-out + extract_block_clamped
//   external/basis_universal/transcoder/basisu_transcoder.cpp:16172 transcode_uastc_to_bc7
//   external/basis_universal/transcoder/basisu_transcoder.cpp:15649 transcode_uastc_to_astc
//   external/basis_universal/transcoder/basisu_transcoder.cpp:15521 unpack_uastc (to pixels)
//   external/basis_universal/transcoder/basisu_transcoder.cpp:10331 basisu_lowlevel_uastc_ldr_4x4_transcoder::transcode_image
#include "basisu_enc.h"
#include "basisu_uastc_enc.h"
#include "basisu_bc7enc.h"
#include "basisu_gpu_texture.h"
#include "../transcoder/basisu_transcoder.h"
#include "../transcoder/basisu_transcoder_uastc.h"

#include <algorithm>
#include <cstring>
#include <mutex>

using namespace basisu;

#define REF_API extern "C" __attribute__((visibility("default")))

static std::once_flag g_init_once;

REF_API int ref_init(void)
{
	std::call_once(g_init_once, [] {
		basisu_encoder_init(false, false);  // also runs basist::basisu_transcoder_init() -> uastc_init()
	});
	return 0;
}

// Encode n pre-extracted 4x4 RGBA blocks (64 B each) -> n*16 B. Same 256-block jobs as basisu_comp.cpp:1503.
REF_API int ref_encode_blocks(const uint8_t* px, uint32_t n, uint32_t flags, uint8_t* out, uint32_t threads)
{
	ref_init();
	if (threads < 1) threads = 1;
	job_pool pool(threads);
	const uint32_t N = 256;
	for (uint32_t first = 0; first < n; first += N)
	{
		const uint32_t last = std::min<uint32_t>(n, first + N);
		pool.add_job([=] {
			for (uint32_t i = first; i < last; i++)
				encode_uastc(px + (size_t)i * 64, *(basist::uastc_block*)(out + (size_t)i * 16), flags);
		});
	}
	pool.wait_for_all();
	return 0;
}

// Edge-clamped 4x4 gather, as image::extract_block_clamped (basisu_enc.h:3081) does on a tight RGBA8 image.
REF_API void ref_extract_blocks(const uint8_t* rgba, uint32_t w, uint32_t h, uint8_t* out_px)
{
	const uint32_t bx = (w + 3) / 4, by = (h + 3) / 4;
	for (uint32_t y = 0; y < by; y++)
		for (uint32_t x = 0; x < bx; x++)
		{
			uint8_t* d = out_px + ((size_t)y * bx + x) * 64;
			for (uint32_t j = 0; j < 4; j++)
				for (uint32_t i = 0; i < 4; i++)
				{
					const uint32_t sx = std::min(x * 4 + i, w - 1), sy = std::min(y * 4 + j, h - 1);
					memcpy(d + (j * 4 + i) * 4, rgba + ((size_t)sy * w + sx) * 4, 4);
				}
		}
}

// Whole-image encode: gather + encode (+ optional RDO with the reference's job split).
REF_API int ref_encode_image(const uint8_t* rgba, uint32_t w, uint32_t h, uint32_t flags, uint8_t* out, uint32_t threads)
{
	const uint32_t n = ((w + 3) / 4) * ((h + 3) / 4);
	std::vector<uint8_t> px((size_t)n * 64);
	ref_extract_blocks(rgba, w, h, px.data());
	return ref_encode_blocks(px.data(), n, flags, out, threads);
}

// RDO post-pass in place. total_jobs as basisu_comp.cpp:1576 computes it: min(4, threads) or 0.
REF_API int ref_rdo(uint8_t* blocks, uint32_t n, const uint8_t* px, float lambda, uint32_t dict_size,
                    float max_rms_ratio, float skip_rms, float smooth_scale, float smooth_std_dev,
                    uint32_t flags, uint32_t threads, uint32_t total_jobs)
{
	ref_init();
	uastc_rdo_params p;
	p.m_lambda = lambda;
	p.m_lz_dict_size = dict_size;
	p.m_max_allowed_rms_increase_ratio = max_rms_ratio;
	p.m_skip_block_rms_thresh = skip_rms;
	p.m_smooth_block_max_error_scale = smooth_scale;
	p.m_max_smooth_block_std_dev = smooth_std_dev;
	if (threads < 1) threads = 1;
	job_pool pool(threads);
	bool ok = uastc_rdo(n, (basist::uastc_block*)blocks, (const color_rgba*)px, p, flags, total_jobs ? &pool : nullptr, total_jobs);
	return ok ? 0 : 1;
}

// fmt: 0 = BC7, 1 = ASTC 4x4. Returns -1 on success, else the index of the first invalid block.
REF_API int64_t ref_transcode(const uint8_t* blocks, uint64_t n, uint32_t fmt, uint8_t* out)
{
	ref_init();
	for (uint64_t i = 0; i < n; i++)
	{
		const basist::uastc_block& b = *(const basist::uastc_block*)(blocks + i * 16);
		bool ok = fmt ? basist::transcode_uastc_to_astc(b, out + i * 16) : basist::transcode_uastc_to_bc7(b, out + i * 16);
		if (!ok) return (int64_t)i;
	}
	return -1;
}

// Multi-threaded variant for the timed CPU baseline (one contiguous range per thread).
REF_API int64_t ref_transcode_mt(const uint8_t* blocks, uint64_t n, uint32_t fmt, uint8_t* out, uint32_t threads)
{
	ref_init();
	if (threads < 1) threads = 1;
	std::vector<std::thread> th;
	std::vector<int64_t> res(threads, -1);
	const uint64_t per = (n + threads - 1) / threads;
	for (uint32_t t = 0; t < threads; t++)
	{
		const uint64_t a = std::min<uint64_t>(n, t * per), b = std::min<uint64_t>(n, a + per);
		th.emplace_back([=, &res] {
			int64_t r = ref_transcode(blocks + a * 16, b - a, fmt, out + a * 16);
			res[t] = (r < 0) ? -1 : (int64_t)(a + r);
		});
	}
	for (auto& t : th) t.join();
	for (auto r : res) if (r >= 0) return r;
	return -1;
}

// UASTC -> 16 RGBA pixels per block (64 B). Returns -1 or the first invalid block.
REF_API int64_t ref_unpack_rgba(const uint8_t* blocks, uint64_t n, uint8_t* out_px, int srgb)
{
	ref_init();
	for (uint64_t i = 0; i < n; i++)
		if (!basist::unpack_uastc(*(const basist::uastc_block*)(blocks + i * 16), (basist::color32*)(out_px + i * 64), srgb != 0))
			return (int64_t)i;
	return -1;
}

// BC7 -> 16 RGBA pixels per block via the encoder-side decoder the reference uses for scoring.
REF_API void ref_unpack_bc7(const uint8_t* blocks, uint64_t n, uint8_t* out_px)
{
	ref_init();
	for (uint64_t i = 0; i < n; i++)
		unpack_block(texture_format::cBC7, blocks + i * 16, (color_rgba*)(out_px + i * 64), false);
}

// White-box tap: one color_cell_compression call in ASTC mode (basisu_bc7enc.cpp:1364), for stage-level diffing.
// out: [0..3] astc low, [4..7] astc high, [8..23] selectors; returns the error.
REF_API uint64_t ref_ccc(const uint8_t* px, uint32_t num_pixels, uint32_t num_weights, uint32_t endpoint_range,
                         int has_alpha, uint32_t uber_level, uint32_t ls_passes, uint8_t* out)
{
	ref_init();
	static const uint32_t* wtab[33] = { 0 };
	static const float* wtabx[33] = { 0 };
	wtab[2] = basist::g_bc7_weights1; wtab[4] = basist::g_bc7_weights2; wtab[8] = basist::g_bc7_weights3;
	wtab[16] = basist::g_astc_weights4; wtab[32] = basist::g_astc_weights5;
	wtabx[2] = g_bc7_weights1x; wtabx[4] = g_bc7_weights2x; wtabx[8] = g_bc7_weights3x;
	wtabx[16] = g_astc_weights4x; wtabx[32] = g_astc_weights5x;

	color_cell_compressor_params p;
	memset(&p, 0, sizeof(p));
	p.m_num_pixels = num_pixels;
	p.m_pPixels = (const basist::color_quad_u8*)px;
	p.m_num_selector_weights = num_weights;
	p.m_pSelector_weights = wtab[num_weights];
	p.m_pSelector_weightsx = (const bc7enc_vec4F*)wtabx[num_weights];
	p.m_astc_endpoint_range = endpoint_range;
	p.m_weights[0] = p.m_weights[1] = p.m_weights[2] = p.m_weights[3] = 1;
	p.m_has_alpha = has_alpha != 0;

	bc7enc_compress_block_params cp;
	memset(&cp, 0, sizeof(cp));
	cp.m_max_partitions_mode1 = 64;
	cp.m_least_squares_passes = ls_passes;
	cp.m_weights[0] = cp.m_weights[1] = cp.m_weights[2] = cp.m_weights[3] = 1;
	cp.m_uber_level = uber_level;

	color_cell_compressor_results r;
	uint8_t sel[16], sel_tmp[16];
	memset(&r, 0, sizeof(r));
	memset(sel, 0, sizeof(sel));
	r.m_pSelectors = sel;
	r.m_pSelectors_temp = sel_tmp;
	uint64_t err = color_cell_compression(255, &p, &r, &cp);
	memcpy(out, r.m_astc_low_endpoint.m_c, 4);
	memcpy(out + 4, r.m_astc_high_endpoint.m_c, 4);
	memcpy(out + 8, sel, 16);
	return err;
}

// Table taps for validating the product's own table builder.
REF_API void ref_get_astc_unquant(uint8_t* out /* 21*256*2: unquant,index */)
{
	ref_init();
	for (int r = 0; r < 21; r++)
		for (int i = 0; i < 256; i++)
		{
			out[(r * 256 + i) * 2 + 0] = basist::g_astc_unquant[r][i].m_unquant;
			out[(r * 256 + i) * 2 + 1] = basist::g_astc_unquant[r][i].m_index;
		}
}
REF_API void ref_get_astc_sorted_order_unquant(uint8_t* out /* 21*256*2 */)
{
	ref_init();
	for (int r = 0; r < 21; r++)
		for (int i = 0; i < 256; i++)
		{
			out[(r * 256 + i) * 2 + 0] = g_astc_sorted_order_unquant[r][i].m_unquant;
			out[(r * 256 + i) * 2 + 1] = g_astc_sorted_order_unquant[r][i].m_index;
		}
}
REF_API void ref_get_bc7_optimal(uint8_t* out /* mode6: 256*2*4 (err16,lo,hi) then mode5: 256*4 */)
{
	ref_init();
	uint8_t* p = out;
	for (int c = 0; c < 256; c++)
		for (int lp = 0; lp < 2; lp++)
		{
			const basist::endpoint_err& e = basist::g_bc7_mode_6_optimal_endpoints[c][lp];
			p[0] = e.m_error & 255; p[1] = e.m_error >> 8; p[2] = e.m_lo; p[3] = e.m_hi; p += 4;
		}
	for (int c = 0; c < 256; c++)
	{
		const basist::endpoint_err& e = basist::g_bc7_mode_5_optimal_endpoints[c];
		p[0] = e.m_error & 255; p[1] = e.m_error >> 8; p[2] = e.m_lo; p[3] = e.m_hi; p += 4;
	}
}

// One image through the low-level UASTC transcoder exactly as libktx calls it (lib/src/basis_transcode.cpp:776-798 ->
// basisu_lowlevel_uastc_ldr_4x4_transcoder::transcode_image, transcoder/basisu_transcoder.cpp:10331): row pitch and row
// count left at 0 (= the image's own width / height), so the uncompressed targets are clipped to w x h.
// tf: transcoder_texture_format (6 BC7, 10 ASTC, 13 RGBA32, 14 RGB565, 15 BGR565, 16 RGBA4444).  Returns 0 on success.
REF_API int ref_transcode_image(uint32_t tf, const uint8_t* blocks, uint32_t w, uint32_t h, uint8_t* out, uint32_t out_size_in_blocks_or_pixels)
{
	ref_init();
	basist::basisu_lowlevel_uastc_ldr_4x4_transcoder t;
	const uint32_t bx = (w + 3) / 4, by = (h + 3) / 4;
	return t.transcode_image((basist::transcoder_texture_format)tf, out, out_size_in_blocks_or_pixels, blocks, bx * by * 16, bx, by, w, h, 0,
		0, bx * by * 16, 0, true, false, 0, nullptr, 0, -1, -1) ? 0 : 1;
}

// The same with the caller's decode flags (cDecodeFlagsTranscodeAlphaDataToOpaqueFormats = 4, cDecodeFlagsHighQuality = 32)
// and has_alpha, as libktx forwards transcodeFlags and (alphaContent != eNone) (lib/src/basis_transcode.cpp:788-789).
// tf additionally: 0 ETC1_RGB, 1 ETC2_RGBA, 2 BC1_RGB, 3 BC3_RGBA, 4 BC4_R, 5 BC5_RG.
REF_API int ref_transcode_image_flags(uint32_t tf, const uint8_t* blocks, uint32_t w, uint32_t h, uint8_t* out, uint32_t out_size_in_blocks_or_pixels,
	uint32_t decode_flags, uint32_t has_alpha)
{
	ref_init();
	basist::basisu_lowlevel_uastc_ldr_4x4_transcoder t;
	const uint32_t bx = (w + 3) / 4, by = (h + 3) / 4;
	return t.transcode_image((basist::transcoder_texture_format)tf, out, out_size_in_blocks_or_pixels, blocks, bx * by * 16, bx, by, w, h, 0,
		0, bx * by * 16, decode_flags, has_alpha != 0, false, 0, nullptr, 0, -1, -1) ? 0 : 1;
}
