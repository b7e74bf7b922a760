// ref_tables_shim.cpp -- TEST/BUILD INFRASTRUCTURE ONLY.  Exposes the file-static init-time tables of the unmodified
// reference encoder (ENC/basisu_bc7enc.cpp:60-96, filled by bc7enc_compress_block_init :186-392) to
// tools/gen_tables.py by compiling that translation unit, where it lies, together with an accessor.
#include "basisu_bc7enc.cpp"
#include "../transcoder/basisu_transcoder.h"
#include <string.h>

extern "C" __attribute__((visibility("default"))) uint32_t ref_dump_encoder_tables(uint8_t* out)
{
	using namespace basisu;
	static bool inited = false;
	if (!inited) { basist::basisu_transcoder_init(); bc7enc_compress_block_init(); inited = true; }
	const endpoint_err* tabs[6] = { g_astc_4bit_3bit_optimal_endpoints, g_astc_4bit_2bit_optimal_endpoints, g_astc_range7_2bit_optimal_endpoints,
		g_astc_range13_2bit_optimal_endpoints, g_astc_range11_5bit_optimal_endpoints, g_astc_range13_4bit_optimal_endpoints };
	const uint32_t n = 21 * 256 + 6 * 256 * 4;
	if (!out) return n;
	memcpy(out, g_astc_nearest_sorted_index, 21 * 256);
	uint8_t* p = out + 21 * 256;
	for (int t = 0; t < 6; t++)
		for (int c = 0; c < 256; c++)
		{
			p[0] = tabs[t][c].m_error & 255; p[1] = tabs[t][c].m_error >> 8; p[2] = tabs[t][c].m_lo; p[3] = tabs[t][c].m_hi; p += 4;
		}
	return n;
}
