// common.cuh -- shared device/host helpers for the grok_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <mutex>
#include <stdint.h>
#include <stdio.h>

#define GRKB_WARP 32

void grkb_set_error(const char* fmt, ...);

#define GRKB_CUDA_CHECK(expr)                                                             \
	do                                                                                    \
	{                                                                                     \
		cudaError_t _e = (expr);                                                          \
		if(_e != cudaSuccess)                                                             \
		{                                                                                 \
			grkb_set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__,  \
						   __LINE__, cudaGetErrorString(_e));                             \
			return -1;                                                                    \
		}                                                                                 \
	} while(0)

// Opt a kernel in to `bytes` of dynamic shared memory.  The attribute is per device, so what has
// been set is remembered per device (one process may drive several: plugin deviceId = -1).
#define GRKB_ENSURE_SMEM(kernel, bytes)                                                           \
	do                                                                                            \
	{                                                                                             \
		static std::mutex _mu;                                                                    \
		static size_t _set[64] = {};                                                              \
		int _d = 0;                                                                               \
		GRKB_CUDA_CHECK(cudaGetDevice(&_d));                                                      \
		std::lock_guard<std::mutex> _lk(_mu);                                                     \
		if(_d >= 0 && _d < 64 && (size_t)(bytes) > _set[_d])                                      \
		{                                                                                         \
			GRKB_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, \
												 (int)(bytes)));                                  \
			_set[_d] = (size_t)(bytes);                                                           \
		}                                                                                         \
	} while(0)

struct grkb_rect
{
	uint32_t x0, y0, x1, y1;
};

static inline __host__ __device__ uint32_t grkb_ceildivpow2(uint32_t a, uint32_t b)
{
	return (uint32_t)(((uint64_t)a + ((uint64_t)1 << b) - 1) >> b);
}

// Band rectangle of a tile-component, equation B-15
// (reference: src/lib/jp2/tile/TileComponentWindowBuffer.h:266-292).
static inline __host__ __device__ grkb_rect grkb_band_rect(grkb_rect tc, uint32_t nd,
															 uint32_t orient)
{
	if(nd == 0)
		return tc;
	uint32_t bx = (orient & 1) ? (1u << (nd - 1)) : 0;
	uint32_t by = (orient >> 1) ? (1u << (nd - 1)) : 0;
	grkb_rect r;
	r.x0 = (tc.x0 <= bx) ? 0 : grkb_ceildivpow2(tc.x0 - bx, nd);
	r.y0 = (tc.y0 <= by) ? 0 : grkb_ceildivpow2(tc.y0 - by, nd);
	r.x1 = (tc.x1 <= bx) ? 0 : grkb_ceildivpow2(tc.x1 - bx, nd);
	r.y1 = (tc.y1 <= by) ? 0 : grkb_ceildivpow2(tc.y1 - by, nd);
	return r;
}

// Row stride rule of the reference's tile buffers: width rounded up to 32 samples
// (src/lib/jp2/cache/MemManager.cpp:35-43).
static inline __host__ __device__ uint32_t grkb_stride(uint32_t w)
{
	return (uint32_t)((((uint64_t)w + 31) / 32) * 32);
}

// Whole-sample symmetric reflection of q into [p0, p0+n), n >= 1, any number of bounces.
static inline __host__ __device__ int grkb_mirror(int q, int p0, int n)
{
	if(n == 1)
		return p0;
	int period = 2 * (n - 1);
	int t = (q - p0) % period;
	if(t < 0)
		t += period;
	if(t >= n)
		t = period - t;
	return p0 + t;
}
