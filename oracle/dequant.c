/*
 * dequant.c -- CPU restatement of the dequantising copy between the block decoders and the tile
 * buffer.  TEST INFRASTRUCTURE ONLY.
 *
 * Restates src/lib/jp2/filters/PostDecompressFilters.h (paths relative to /root/reference), chosen
 * by TileComponent::postProcess / postProcessHT (tile/TileComponent.cpp:333-386):
 *   kind 0  ShiftFilter / RoiShiftFilter        :7-46    MQ, 5/3:  v / 2 (C truncation)
 *   kind 1  ScaleFilter / RoiScaleFilter        :48-89   MQ, 9/7:  (float)v * (stepsize / 2)
 *   kind 2  ShiftHTFilter / RoiShiftHTFilter    :91-129  HT, 5/3:  sign-magnitude >> (31 - (k_msbs + 1))
 *   kind 3  ScaleHTFilter / RoiScaleHTFilter    :131-177 HT, 9/7:  (float)magnitude * stepsize
 * The Roi* forms (roishift != 0) scale magnitudes at or above 2^roishift down first; the two HT
 * forms AND the shifted magnitude with the sign bit (:105, :148), which clears such samples --
 * kept as the reference computes it.
 * Pinned against the reference's own filter classes by tests/test_oracle_vs_ref.py (ref_dequant in
 * ref_harness.cpp instantiates them).
 */
#include "oracle.h"
#include <stdlib.h>

void orc_dequant(int kind, const int32_t* src, int32_t* dst, uint32_t n, uint32_t roishift,
				 uint32_t k_msbs, float stepsize)
{
	const int32_t thresh = (int32_t)(1u << roishift);
	const uint32_t shift = 31u - (k_msbs + 1u);
	const float half = stepsize / 2;
	for(uint32_t i = 0; i < n; ++i)
	{
		int32_t val = src[i];
		if(kind < 2)
		{
			if(roishift)
			{
				int32_t mag = abs(val);
				if(mag >= thresh)
				{
					mag >>= roishift;
					val = val < 0 ? -mag : mag;
				}
			}
			if(kind == 0)
				dst[i] = val / 2;
			else
				((float*)dst)[i] = (float)val * half;
		}
		else
		{
			if(roishift)
			{
				const int32_t mag = val & 0x7FFFFFFF;
				if(mag >= thresh)
					val = (int32_t)(((uint32_t)mag >> roishift) & ((uint32_t)val & 0x80000000u));
			}
			if(kind == 2)
			{
				const int32_t s = (val & 0x7FFFFFFF) >> shift;
				dst[i] = ((uint32_t)val & 0x80000000u) ? -s : s;
			}
			else
			{
				const float f = (float)(val & 0x7FFFFFFF) * stepsize;
				((float*)dst)[i] = ((uint32_t)val & 0x80000000u) ? -f : f;
			}
		}
	}
}
