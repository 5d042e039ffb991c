/*
 * samples.c -- CPU restatement of the sample narrowing / widening Grok's applications do when
 * int32 image planes leave or enter the library.  TEST INFRASTRUCTURE ONLY.
 *
 * Restates (paths relative to /root/reference/src/bin):
 *   planar int32 -> interleaved int32   common/convert.cpp:397-405  (planarToInterleaved<N>)
 *   int32 -> 16-bit sample              common/convert.cpp:1158-1163 (convert_tif_32sto16u:
 *                                       dest[i] = (uint16_t)pSrc[i]); 8 bits: :564-end, cvtFrom32_LUT[8]
 *   16-bit sample -> int32              common/convert.cpp:547-556  (convert_16u32s_C1R; the
 *                                       reference reads big-endian file bytes, the value is the same)
 * Pinned by execution: oracle/Makefile compiles the reference's convert.cpp where it lies (with a
 * hand-written grk_apps_config.h stand-in, ref_config/) into oracle/_ref/libgrok_ref_apps.so, and
 * tests/test_oracle_vs_ref.py compares these loops with planarToInterleaved<N>,
 * convert_tif_32sto16u, cvtFrom32_LUT[8] and convert_16u32s_C1R called through
 * ref_apps_harness.cpp.
 */
#include "oracle.h"

/* planes[c][y * stride + x] -> out, samples of `bytes` (1 or 2) bytes in host order.
 * planar: plane c starts at c * H * row_bytes; interleaved: pixel x of a row at x * numcomps. */
void orc_pack_samples(const int32_t* const* planes, uint32_t numcomps, uint32_t stride, uint32_t W, uint32_t H,
					  uint32_t bytes, int interleaved, uint8_t* out, uint64_t row_bytes)
{
	for(uint32_t c = 0; c < numcomps; ++c)
		for(uint32_t y = 0; y < H; ++y)
			for(uint32_t x = 0; x < W; ++x)
			{
				const int32_t v = planes[c][(uint64_t)y * stride + x];
				uint8_t* row = out + (interleaved ? 0 : (uint64_t)c * H * row_bytes) + (uint64_t)y * row_bytes;
				const uint64_t i = interleaved ? (uint64_t)x * numcomps + c : x;
				if(bytes == 2)
					((uint16_t*)row)[i] = (uint16_t)v;
				else
					row[i] = (uint8_t)v;
			}
}

void orc_unpack_samples(const uint8_t* in, uint64_t row_bytes, uint32_t bytes, int interleaved, int sgnd,
						int32_t* const* planes, uint32_t numcomps, uint32_t stride, uint32_t W, uint32_t H)
{
	for(uint32_t c = 0; c < numcomps; ++c)
		for(uint32_t y = 0; y < H; ++y)
			for(uint32_t x = 0; x < W; ++x)
			{
				const uint8_t* row = in + (interleaved ? 0 : (uint64_t)c * H * row_bytes) + (uint64_t)y * row_bytes;
				const uint64_t i = interleaved ? (uint64_t)x * numcomps + c : x;
				int32_t v;
				if(bytes == 2)
					v = sgnd ? (int32_t)((const int16_t*)row)[i] : (int32_t)((const uint16_t*)row)[i];
				else
					v = sgnd ? (int32_t)((const int8_t*)row)[i] : (int32_t)row[i];
				planes[c][(uint64_t)y * stride + x] = v;
			}
}
