/*
 * mct.c -- CPU restatement of Grok's DC level shift and RCT / ICT multi-component
 * transforms.  TEST INFRASTRUCTURE ONLY.
 *
 * Restates (paths relative to /root/reference/src/lib/jp2):
 *   forward DC shift   tile/TileProcessor.cpp:506-535
 *   forward RCT / ICT  point_transform/mct.cpp:99-146, :147-219
 *   inverse RCT / ICT fused with DC shift + clamp  point_transform/mct.cpp:221-283, :284-364
 *   DC shift + clamp alone  point_transform/mct.cpp:35-97
 * Built with -ffp-contract=off: every multiply and add below is a separate IEEE single op,
 * as in the reference's ISO-C++ build.
 */
#include "oracle.h"
#include <math.h>

static inline int32_t clampi(int32_t v, int32_t lo, int32_t hi)
{
	return v < lo ? lo : (v > hi ? hi : v);
}

void orc_dc_shift_fwd(int32_t* c, uint64_t n, int32_t shift)
{
	for(uint64_t i = 0; i < n; ++i)
		c[i] -= shift;
}

void orc_rct_fwd(int32_t* c0, int32_t* c1, int32_t* c2, uint64_t n)
{
	for(uint64_t i = 0; i < n; ++i)
	{
		int32_t r = c0[i], g = c1[i], b = c2[i];
		c0[i] = (r + (g * 2) + b) >> 2;
		c1[i] = b - g;
		c2[i] = r - g;
	}
}

void orc_ict_fwd(int32_t* c0, int32_t* c1, int32_t* c2, uint64_t n)
{
	const float a_r = 0.299f, a_g = 0.587f, a_b = 0.114f;
	const float cb = 0.5f / (1.0f - a_b);
	const float cr = 0.5f / (1.0f - a_r);
	float* f0 = (float*)c0;
	float* f1 = (float*)c1;
	float* f2 = (float*)c2;
	for(uint64_t i = 0; i < n; ++i)
	{
		float r = (float)c0[i], g = (float)c1[i], b = (float)c2[i];
		float y = a_r * r + a_g * g + a_b * b;
		float u = cb * (b - y);
		float v = cr * (r - y);
		f0[i] = y;
		f1[i] = u;
		f2[i] = v;
	}
}

void orc_rct_inv(int32_t* c0, int32_t* c1, int32_t* c2, uint64_t n, const int32_t* shift,
				 const int32_t* lo, const int32_t* hi)
{
	for(uint64_t i = 0; i < n; ++i)
	{
		int32_t y = c0[i], u = c1[i], v = c2[i];
		int32_t g = y - ((u + v) >> 2);
		int32_t r = v + g;
		int32_t b = u + g;
		c0[i] = clampi(r + shift[0], lo[0], hi[0]);
		c1[i] = clampi(g + shift[1], lo[1], hi[1]);
		c2[i] = clampi(b + shift[2], lo[2], hi[2]);
	}
}

void orc_ict_inv(int32_t* c0, int32_t* c1, int32_t* c2, uint64_t n, const int32_t* shift,
				 const int32_t* lo, const int32_t* hi)
{
	const float* f0 = (const float*)c0;
	const float* f1 = (const float*)c1;
	const float* f2 = (const float*)c2;
	for(uint64_t i = 0; i < n; ++i)
	{
		float y = f0[i], u = f1[i], v = f2[i];
		float r = y + (v * 1.402f);
		float g = y - (u * 0.34413f) - (v * 0.71414f);
		float b = y + (u * 1.772f);
		/* lrintf: round to nearest even in the default rounding mode (util/simd.h:49-53) */
		c0[i] = clampi((int32_t)lrintf(r) + shift[0], lo[0], hi[0]);
		c1[i] = clampi((int32_t)lrintf(g) + shift[1], lo[1], hi[1]);
		c2[i] = clampi((int32_t)lrintf(b) + shift[2], lo[2], hi[2]);
	}
}

void orc_dc_shift_inv_rev(int32_t* c, uint64_t n, int32_t shift, int32_t lo, int32_t hi)
{
	for(uint64_t i = 0; i < n; ++i)
		c[i] = clampi(c[i] + shift, lo, hi);
}

void orc_dc_shift_inv_irrev(int32_t* c, uint64_t n, int32_t shift, int32_t lo, int32_t hi)
{
	const float* f = (const float*)c;
	for(uint64_t i = 0; i < n; ++i)
		c[i] = clampi((int32_t)lrintf(f[i]) + shift, lo, hi);
}


/* ---- array-based multi-component transform (Part 2 "custom MCT", row M7) ---------------------- */

/* mct::decompress_custom (point_transform/mct.cpp:660-686): planes read as float, out[j] =
 * sum_k M[j][k] * in[k], accumulated from 0 in k order (no contraction: -ffp-contract=off) */
void orc_mct_custom_inv(float* const* planes, uint32_t nc, uint64_t n, const float* matrix)
{
	float in[16], out[16];
	for(uint64_t i = 0; i < n; ++i)
	{
		for(uint32_t j = 0; j < nc; ++j)
			in[j] = planes[j][i];
		for(uint32_t j = 0; j < nc; ++j)
		{
			out[j] = 0;
			for(uint32_t k = 0; k < nc; ++k)
			{
				const float p = matrix[j * nc + k] * in[k];
				out[j] = out[j] + p;
			}
		}
		for(uint32_t j = 0; j < nc; ++j)
			planes[j][i] = out[j];
	}
}

/* mct::compress_custom (mct.cpp:622-658) with fix_mul (util/grk_intmath.h:66-79).  As the reference
 * computes it: the matrix pointer is reset to the FIRST row for every output component (:642), so
 * every output component is row 0 applied to the pixel (a reference defect, kept for parity). */
void orc_mct_custom_fwd(int32_t* const* planes, uint32_t nc, uint64_t n, const float* matrix)
{
	int32_t m[256], in[16];
	for(uint32_t i = 0; i < nc * nc; ++i)
		m[i] = (int32_t)(matrix[i] * (float)(1 << 13));
	for(uint64_t i = 0; i < n; ++i)
	{
		for(uint32_t j = 0; j < nc; ++j)
			in[j] = planes[j][i];
		for(uint32_t j = 0; j < nc; ++j)
		{
			int32_t acc = 0;
			for(uint32_t k = 0; k < nc; ++k)
				acc += (int32_t)(((int64_t)m[k] * (int64_t)in[k] + 4096) >> 13);
			planes[j][i] = acc;
		}
	}
}
