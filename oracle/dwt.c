/*
 * dwt.c -- CPU restatement of Grok's whole-tile wavelet transforms.  TEST INFRASTRUCTURE ONLY.
 *
 * Restates (paths relative to /root/reference/src/lib/jp2):
 *   inverse 5/3  transform/WaveletReverse.cpp:802-877 (decompress_tile_53), 1-D :373-663
 *   inverse 9/7  transform/WaveletReverse.cpp:1272-1347 (decompress_tile_97), 1-D :882-1022,
 *                boundary rule :1823-1849 (makeParams97)
 *   forward 5/3  transform/WaveletFwd.cpp:432-616 (encode_procedure), 1-D :635-960
 *   forward 9/7  transform/WaveletFwd.cpp:92-198, :964-1025
 *   layout       tile/TileComponentWindowBuffer.h:188-221 (Mallat in one buffer)
 *
 * The reference's hand-unrolled 1-D kernels are written here as plain lifting with
 * whole-sample symmetric extension, which is the same arithmetic: every boundary form in
 * the reference ((d+1)>>1, d+s, (c+c)*x, (2*x)*c) is the general step with the missing
 * neighbour mirrored, and scaling by 2 is exact in binary floating point.
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

static uint32_t ceildivpow2(uint32_t a, uint32_t b)
{
	return (uint32_t)(((uint64_t)a + ((uint64_t)1 << b) - 1) >> b);
}

uint32_t orc_stride(uint32_t width)
{
	return (uint32_t)((((uint64_t)width + 31) / 32) * 32);
}

orc_rect orc_band_rect(orc_rect tc, uint32_t nd, uint32_t orient)
{
	orc_rect r;
	if(nd == 0)
		return tc;
	uint32_t bx = (orient & 1) ? (1u << (nd - 1)) : 0;
	uint32_t by = (orient >> 1) ? (1u << (nd - 1)) : 0;
	r.x0 = (tc.x0 <= bx) ? 0 : ceildivpow2(tc.x0 - bx, nd);
	r.y0 = (tc.y0 <= by) ? 0 : ceildivpow2(tc.y0 - by, nd);
	r.x1 = (tc.x1 <= bx) ? 0 : ceildivpow2(tc.x1 - bx, nd);
	r.y1 = (tc.y1 <= by) ? 0 : ceildivpow2(tc.y1 - by, nd);
	return r;
}

/* mirrored neighbour index for a line of n >= 2 samples */
static inline int32_t mir(int32_t i, int32_t n)
{
	if(i < 0)
		return -i;
	if(i >= n)
		return 2 * (n - 1) - i;
	return i;
}

/* ---------------- 5/3 ---------------- */

/* x: interleaved line of n samples; parity = 1 when x[0] sits on an odd canvas coordinate.
 * low-pass samples are those at even canvas coordinates. */
static void inv53_line(int32_t* x, int32_t n, int32_t parity)
{
	/* n == 1 handled by the callers (WaveletReverse.cpp:573-584, :608-637) */
	for(int32_t i = parity; i < n; i += 2)
		x[i] -= (x[mir(i - 1, n)] + x[mir(i + 1, n)] + 2) >> 2;
	for(int32_t i = 1 - parity; i < n; i += 2)
		x[i] += (x[mir(i - 1, n)] + x[mir(i + 1, n)]) >> 1;
}

static void fwd53_line(int32_t* x, int32_t n, int32_t parity)
{
	for(int32_t i = 1 - parity; i < n; i += 2)
		x[i] -= (x[mir(i - 1, n)] + x[mir(i + 1, n)]) >> 1;
	for(int32_t i = parity; i < n; i += 2)
		x[i] += (x[mir(i - 1, n)] + x[mir(i + 1, n)] + 2) >> 2;
}

void orc_idwt53(int32_t* buf, uint32_t stride, orc_rect tc, uint32_t numres)
{
	if(numres <= 1)
		return;
	uint32_t maxdim = (tc.x1 - tc.x0) > (tc.y1 - tc.y0) ? (tc.x1 - tc.x0) : (tc.y1 - tc.y0);
	int32_t* tmp = (int32_t*)malloc(sizeof(int32_t) * (size_t)(maxdim + 2));
	orc_rect prev = orc_band_rect(tc, numres - 1, 0);
	for(uint32_t res = 1; res < numres; ++res)
	{
		orc_rect cur = orc_band_rect(tc, numres - 1 - res, 0);
		uint32_t rw = cur.x1 - cur.x0, rh = cur.y1 - cur.y0;
		uint32_t sw = prev.x1 - prev.x0, sh = prev.y1 - prev.y0;
		prev = cur;
		if(rw == 0 || rh == 0)
			continue; /* WaveletReverse.cpp:833-834 */
		uint32_t px = cur.x0 & 1, py = cur.y0 & 1;
		/* horizontal pass over all rows: [L | H] -> interleaved, in place */
		for(uint32_t y = 0; y < rh; ++y)
		{
			int32_t* row = buf + (size_t)y * stride;
			if(rw == 1)
			{
				if(px)
					row[0] = row[0] / 2; /* :583 C division, truncating */
				continue;
			}
			for(uint32_t i = 0; i < sw; ++i)
				tmp[px + 2 * i] = row[i];
			for(uint32_t i = 0; i < rw - sw; ++i)
				tmp[1 - px + 2 * i] = row[sw + i];
			inv53_line(tmp, (int32_t)rw, (int32_t)px);
			memcpy(row, tmp, sizeof(int32_t) * rw);
		}
		/* vertical pass over all columns */
		for(uint32_t x = 0; x < rw; ++x)
		{
			int32_t* col = buf + x;
			if(rh == 1)
			{
				if(py)
					col[0] = col[0] >> 1; /* :633-637 arithmetic shift, reads bandL */
				continue;
			}
			for(uint32_t i = 0; i < sh; ++i)
				tmp[py + 2 * i] = col[(size_t)i * stride];
			for(uint32_t i = 0; i < rh - sh; ++i)
				tmp[1 - py + 2 * i] = col[(size_t)(sh + i) * stride];
			inv53_line(tmp, (int32_t)rh, (int32_t)py);
			for(uint32_t i = 0; i < rh; ++i)
				col[(size_t)i * stride] = tmp[i];
		}
	}
	free(tmp);
}

void orc_fdwt53(int32_t* buf, uint32_t stride, orc_rect tc, uint32_t numres)
{
	if(numres <= 1)
		return;
	uint32_t maxdim = (tc.x1 - tc.x0) > (tc.y1 - tc.y0) ? (tc.x1 - tc.x0) : (tc.y1 - tc.y0);
	int32_t* tmp = (int32_t*)malloc(sizeof(int32_t) * (size_t)(maxdim + 2));
	for(uint32_t res = numres - 1; res >= 1; --res)
	{
		orc_rect cur = orc_band_rect(tc, numres - 1 - res, 0);
		orc_rect low = orc_band_rect(tc, numres - res, 0);
		uint32_t rw = cur.x1 - cur.x0, rh = cur.y1 - cur.y0;
		uint32_t sw = low.x1 - low.x0, sh = low.y1 - low.y0;
		uint32_t px = cur.x0 & 1, py = cur.y0 & 1;
		/* vertical first (WaveletFwd.cpp:496-548) */
		for(uint32_t x = 0; x < rw; ++x)
		{
			int32_t* col = buf + x;
			if(rh == 1)
			{
				if(py)
					col[0] *= 2; /* :650-659 */
				continue;
			}
			if(rh == 0)
				break;
			for(uint32_t i = 0; i < rh; ++i)
				tmp[i] = col[(size_t)i * stride];
			fwd53_line(tmp, (int32_t)rh, (int32_t)py);
			for(uint32_t i = 0; i < sh; ++i)
				col[(size_t)i * stride] = tmp[py + 2 * i];
			for(uint32_t i = 0; i < rh - sh; ++i)
				col[(size_t)(sh + i) * stride] = tmp[1 - py + 2 * i];
		}
		/* then horizontal (:550-606) */
		for(uint32_t y = 0; y < rh; ++y)
		{
			int32_t* row = buf + (size_t)y * stride;
			if(rw == 1)
			{
				if(px)
					row[0] *= 2; /* :934 */
				continue;
			}
			if(rw == 0)
				break;
			memcpy(tmp, row, sizeof(int32_t) * rw);
			fwd53_line(tmp, (int32_t)rw, (int32_t)px);
			for(uint32_t i = 0; i < sw; ++i)
				row[i] = tmp[px + 2 * i];
			for(uint32_t i = 0; i < rw - sw; ++i)
				row[sw + i] = tmp[1 - px + 2 * i];
		}
	}
	free(tmp);
}

/* ---------------- 9/7 ---------------- */

/* inverse constants WaveletReverse.cpp:365-371 */
static const float I_ALPHA = 1.586134342f, I_BETA = 0.052980118f, I_GAMMA = -0.882911075f,
				   I_DELTA = -0.443506852f, I_K = 1.230174105f, I_2INVK = 1.625732422f;
/* forward constants WaveletFwd.cpp:39-44 */
static const float F_ALPHA = -1.586134342f, F_BETA = -0.052980118f, F_GAMMA = 0.882911075f,
				   F_DELTA = 0.443506852f, F_K = 1.230174105f;

/* one lifting step on the samples of phase `ph` (ph = index of first updated sample):
 * x[i] += (x[i-1] + x[i+1]) * c, mirrored at the ends. Sum, then multiply, then add, as
 * separate IEEE single operations (WaveletReverse.cpp:914-917, WaveletFwd.cpp:141-160). */
static void lift97(float* x, int32_t n, int32_t ph, float c)
{
	for(int32_t i = ph; i < n; i += 2)
	{
		float s = x[mir(i - 1, n)] + x[mir(i + 1, n)];
		float p = s * c; /* built with -ffp-contract=off: never fused */
		x[i] = x[i] + p;
	}
}

static void inv97_line(float* x, int32_t n, int32_t parity)
{
	/* n == 1 is left untouched, unscaled (WaveletReverse.cpp:1012-1014) */
	if(n < 2)
		return;
	for(int32_t i = parity; i < n; i += 2)
		x[i] *= I_K;
	for(int32_t i = 1 - parity; i < n; i += 2)
		x[i] *= I_2INVK;
	lift97(x, n, parity, I_DELTA);
	lift97(x, n, 1 - parity, I_GAMMA);
	lift97(x, n, parity, I_BETA);
	lift97(x, n, 1 - parity, I_ALPHA);
}

static void fwd97_line(float* x, int32_t n, int32_t parity)
{
	if(n < 2)
		return; /* WaveletFwd.cpp:973,1018 */
	const float invK = (float)(1.0 / 1.230174105);
	lift97(x, n, 1 - parity, F_ALPHA);
	lift97(x, n, parity, F_BETA);
	lift97(x, n, 1 - parity, F_GAMMA);
	lift97(x, n, parity, F_DELTA);
	for(int32_t i = parity; i < n; i += 2)
		x[i] *= invK;
	for(int32_t i = 1 - parity; i < n; i += 2)
		x[i] *= F_K;
}

void orc_idwt97(float* buf, uint32_t stride, orc_rect tc, uint32_t numres)
{
	if(numres <= 1)
		return;
	uint32_t maxdim = (tc.x1 - tc.x0) > (tc.y1 - tc.y0) ? (tc.x1 - tc.x0) : (tc.y1 - tc.y0);
	float* tmp = (float*)malloc(sizeof(float) * (size_t)(maxdim + 2));
	orc_rect prev = orc_band_rect(tc, numres - 1, 0);
	for(uint32_t res = 1; res < numres; ++res)
	{
		orc_rect cur = orc_band_rect(tc, numres - 1 - res, 0);
		uint32_t rw = cur.x1 - cur.x0, rh = cur.y1 - cur.y0;
		uint32_t sw = prev.x1 - prev.x0, sh = prev.y1 - prev.y0;
		prev = cur;
		if(rw == 0 || rh == 0)
			continue;
		uint32_t px = cur.x0 & 1, py = cur.y0 & 1;
		for(uint32_t y = 0; y < rh; ++y)
		{
			float* row = buf + (size_t)y * stride;
			for(uint32_t i = 0; i < sw; ++i)
				tmp[px + 2 * i] = row[i];
			for(uint32_t i = 0; i < rw - sw; ++i)
				tmp[1 - px + 2 * i] = row[sw + i];
			inv97_line(tmp, (int32_t)rw, (int32_t)px);
			memcpy(row, tmp, sizeof(float) * rw);
		}
		for(uint32_t x = 0; x < rw; ++x)
		{
			float* col = buf + x;
			for(uint32_t i = 0; i < sh; ++i)
				tmp[py + 2 * i] = col[(size_t)i * stride];
			for(uint32_t i = 0; i < rh - sh; ++i)
				tmp[1 - py + 2 * i] = col[(size_t)(sh + i) * stride];
			inv97_line(tmp, (int32_t)rh, (int32_t)py);
			for(uint32_t i = 0; i < rh; ++i)
				col[(size_t)i * stride] = tmp[i];
		}
	}
	free(tmp);
}

void orc_fdwt97(float* buf, uint32_t stride, orc_rect tc, uint32_t numres)
{
	if(numres <= 1)
		return;
	uint32_t maxdim = (tc.x1 - tc.x0) > (tc.y1 - tc.y0) ? (tc.x1 - tc.x0) : (tc.y1 - tc.y0);
	float* tmp = (float*)malloc(sizeof(float) * (size_t)(maxdim + 2));
	for(uint32_t res = numres - 1; res >= 1; --res)
	{
		orc_rect cur = orc_band_rect(tc, numres - 1 - res, 0);
		orc_rect low = orc_band_rect(tc, numres - res, 0);
		uint32_t rw = cur.x1 - cur.x0, rh = cur.y1 - cur.y0;
		uint32_t sw = low.x1 - low.x0, sh = low.y1 - low.y0;
		uint32_t px = cur.x0 & 1, py = cur.y0 & 1;
		if(rh > 1)
			for(uint32_t x = 0; x < rw; ++x)
			{
				float* col = buf + x;
				for(uint32_t i = 0; i < rh; ++i)
					tmp[i] = col[(size_t)i * stride];
				fwd97_line(tmp, (int32_t)rh, (int32_t)py);
				for(uint32_t i = 0; i < sh; ++i)
					col[(size_t)i * stride] = tmp[py + 2 * i];
				for(uint32_t i = 0; i < rh - sh; ++i)
					col[(size_t)(sh + i) * stride] = tmp[1 - py + 2 * i];
			}
		if(rw > 1)
			for(uint32_t y = 0; y < rh; ++y)
			{
				float* row = buf + (size_t)y * stride;
				memcpy(tmp, row, sizeof(float) * rw);
				fwd97_line(tmp, (int32_t)rw, (int32_t)px);
				for(uint32_t i = 0; i < sw; ++i)
					row[i] = tmp[px + 2 * i];
				for(uint32_t i = 0; i < rw - sw; ++i)
					row[sw + i] = tmp[1 - px + 2 * i];
			}
	}
	free(tmp);
}
