// quant.cpp -- quantisation parameters on the host (no device work): the QCD step sizes an
// encoder writes and the band step size / bit-plane count both sides derive from them.
//
// Replaces, behind the C-ABI, the reference's
//   param_qcd::generate (Part-1 branch)   src/lib/jp2/codestream/HTParams.cpp:194-252
//   param_qcd::set_rev_quant (HT, 5/3)    src/lib/jp2/codestream/HTParams.cpp:253-272
//   Quantizer::setBandStepSizeAndBps      src/lib/jp2/codestream/Quantizer.cpp:26-66
// Pinned against the QCD markers the stock encoder writes and against the band parameters of
// the stock decoder (tests/test_quant.py, tests/test_e2e_oracle.py).
#include "../../include/grk_b200.h"
#include <math.h>
#include <stdint.h>

void grkb_set_error(const char* fmt, ...);

namespace
{
// L2 norms of the 9/7 synthesis basis functions (T1.cpp:244-248), [orientation][level]
const double NORMS_97[4][10] = {
	{1.000, 1.965, 4.177, 8.403, 16.90, 33.84, 67.69, 135.3, 270.6, 540.9},
	{2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
	{2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
	{2.080, 3.865, 8.307, 17.18, 34.71, 69.59, 139.3, 278.6, 557.2, 0}};

// BIBO gains of the 5/3 analysis bank per decomposition depth (HTParams.cpp:134-146)
const float BIBO_53_L[] = {1.0000f, 1.5000f, 1.6250f, 1.6875f, 1.6963f, 1.7067f, 1.7116f, 1.7129f,
						   1.7141f, 1.7145f, 1.7151f, 1.7152f, 1.7155f, 1.7155f, 1.7156f, 1.7156f,
						   1.7156f, 1.7156f, 1.7156f, 1.7156f, 1.7156f, 1.7156f, 1.7156f, 1.7156f,
						   1.7156f, 1.7156f, 1.7156f, 1.7156f, 1.7156f, 1.7156f, 1.7156f, 1.7156f,
						   1.7156f, 1.7156f};
const float BIBO_53_H[] = {2.0000f, 2.5000f, 2.7500f, 2.8047f, 2.8198f, 2.8410f, 2.8558f, 2.8601f,
						   2.8628f, 2.8656f, 2.8662f, 2.8667f, 2.8669f, 2.8670f, 2.8671f, 2.8671f,
						   2.8671f, 2.8671f, 2.8671f, 2.8671f, 2.8671f, 2.8671f, 2.8671f, 2.8671f,
						   2.8671f, 2.8671f, 2.8671f, 2.8671f, 2.8671f, 2.8671f, 2.8671f, 2.8671f,
						   2.8671f, 2.8671f};

int floorlog2(uint32_t v)
{
	int l = -1;
	while(v)
	{
		v >>= 1;
		++l;
	}
	return l;
}

int bibo_bits(float a, float b)
{
	return (int)ceil(log((double)(a * b * 1.1f)) / log(2.0));
}
} // namespace

extern "C" {

int grkb_qcd_generate(uint32_t numres, uint32_t prec, int reversible, int ht, uint32_t* expn,
					  uint32_t* mant)
{
	if(numres < 1 || numres > 33 || !expn || !mant || prec < 1 || prec > 38)
	{
		grkb_set_error("grkb_qcd_generate: bad arguments");
		return -1;
	}
	const uint32_t nbands = 3 * numres - 2;
	if(ht && reversible)
	{
		const uint32_t nd = numres - 1;
		expn[0] = prec + (uint32_t)bibo_bits(BIBO_53_L[nd], BIBO_53_L[nd]);
		mant[0] = 0;
		uint32_t k = 1;
		for(int d = (int)nd - 1; d >= 0; --d)
		{
			const float lo = BIBO_53_L[d + 1], hi = BIBO_53_H[d];
			expn[k] = expn[k + 1] = prec + (uint32_t)bibo_bits(hi, lo);
			expn[k + 2] = prec + (uint32_t)bibo_bits(hi, hi);
			mant[k] = mant[k + 1] = mant[k + 2] = 0;
			k += 3;
		}
		return 0;
	}
	if(ht)
	{
		grkb_set_error("grkb_qcd_generate: HT with the irreversible transform is not supported");
		return -1;
	}
	for(uint32_t bandno = 0; bandno < nbands; ++bandno)
	{
		const uint32_t resno = bandno == 0 ? 0 : (bandno - 1) / 3 + 1;
		const uint32_t orient = bandno == 0 ? 0 : (bandno - 1) % 3 + 1;
		uint32_t level = numres - 1 - resno;
		const uint32_t gain = !reversible ? 0 : (orient == 0 ? 0 : (orient == 3 ? 2 : 1));
		double stepsize = 1.0;
		if(!reversible)
		{
			if(orient == 0 && level > 9)
				level = 9;
			else if(orient > 0 && level > 8)
				level = 8;
			stepsize = (double)(1u << gain) / NORMS_97[orient][level];
		}
		const uint32_t step = (uint32_t)floor(stepsize * 8192.0);
		const int p = floorlog2(step) - 13;
		const int n = 11 - floorlog2(step);
		mant[bandno] = reversible ? 0 : ((n < 0 ? step >> -n : step << n) & 0x7FF);
		expn[bandno] = (uint32_t)((int)prec + (int)gain - p);
	}
	return 0;
}

int grkb_band_quant(uint32_t expn, uint32_t mant, uint32_t orient, uint32_t prec, int reversible,
					int compress, int ht, uint32_t numgbits, uint32_t roishift, float* stepsize,
					uint32_t* numbps)
{
	if(orient > 3 || !stepsize || !numbps)
	{
		grkb_set_error("grkb_band_quant: bad arguments");
		return -1;
	}
	const uint32_t log2_gain = (!compress && !reversible) ? 0 : orient == 0 ? 0 : orient == 3 ? 2 : 1;
	const uint32_t nb = prec + log2_gain;
	float s = (float)((1.0 + mant / 2048.0) * pow(2.0, (double)((int32_t)(nb - expn))));
	int8_t t = (int8_t)(expn + numgbits - 1u);
	if(t < 0)
		t = 0;
	const uint32_t band_numbps = roishift + (uint8_t)t;
	if(ht && !compress && !reversible)
	{
		if(band_numbps > 31)
		{
			grkb_set_error("grkb_band_quant: unsupported number of band bit-planes %u", band_numbps);
			return -1;
		}
		s /= (float)(1u << (31 - band_numbps));
	}
	*stepsize = s;
	*numbps = band_numbps;
	return 0;
}

} // extern "C"
