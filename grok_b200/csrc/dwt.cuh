// dwt.cuh -- lifting arithmetic and launch descriptors shared by the inverse and forward
// wavelet kernels (sm_100a).
//
// Arithmetic follows the reference exactly (paths under /root/reference/src/lib/jp2):
//   5/3 inverse  transform/WaveletReverse.cpp:373-663   s -= (d'+d+2)>>2 ; d += (s+s')>>1
//   5/3 forward  transform/WaveletFwd.cpp:635-960       d -= (s+s')>>1   ; s += (d'+d+2)>>2
//   9/7 inverse  transform/WaveletReverse.cpp:365-371, :882-1022  (K, 2/K, delta, gamma, beta, alpha)
//   9/7 forward  transform/WaveletFwd.cpp:39-44, :135-198         (alpha, beta, gamma, delta, 1/K, K)
// Every 9/7 step is x + ((a + b) * c) with three separately rounded IEEE single operations
// (the reference is built without FMA contraction), hence the __f*_rn intrinsics.
#pragma once
#include "common.cuh"

struct DwtPlane
{
	const void* ll; // inverse: low-resolution source (LL); forward: unused
	uint32_t ll_stride;
	const void* bands; // inverse: coefficient buffer, Mallat layout (HL,LH,HH taken from it)
	uint32_t b_stride;
	void* full; // inverse: destination of this level; forward: source of this level
	uint32_t f_stride;
	void* out_ll; // forward: where LL of this level goes (next level's source)
	uint32_t o_stride;
	void* out_bands; // forward: coefficient buffer receiving HL, LH, HH of this level
	uint32_t ob_stride;
};

struct DwtLevelArgs
{
	DwtPlane pl[3];
	uint32_t rw, rh; // size of the resolution produced (inverse) / consumed (forward)
	uint32_t sw, sh; // size of the low-pass part
	uint32_t px, py; // lifting parity = origin of this resolution & 1
	uint32_t nstrips, nchunks, chunk_rows;
	int32_t shift[3], lo[3], hi[3]; // DC shift + clamp of the finishing / ingest stage
};

struct MctArgs
{
	void* c[3];
	uint64_t n; // samples per component (4 at a time vectorised, tail scalar)
	int32_t shift[3], lo[3], hi[3];
};

struct K53
{
	typedef int32_t T;
	static constexpr int NS = 2;
	// inverse steps: k=0 on low-pass samples, k=1 on high-pass samples
	static __device__ __forceinline__ T iupd(int k, T x, T a, T b)
	{
		return k == 0 ? x - ((a + b + 2) >> 2) : x + ((a + b) >> 1);
	}
	// forward steps: k=0 on high-pass samples, k=1 on low-pass samples
	static __device__ __forceinline__ T fupd(int k, T x, T a, T b)
	{
		return k == 0 ? x - ((a + b) >> 1) : x + ((a + b + 2) >> 2);
	}
	static __device__ __forceinline__ T iscaleL(T x) { return x; }
	static __device__ __forceinline__ T iscaleH(T x) { return x; }
	static __device__ __forceinline__ T fscaleL(T x) { return x; }
	static __device__ __forceinline__ T fscaleH(T x) { return x; }
};

struct K97
{
	typedef float T;
	static constexpr int NS = 4;
	static __device__ __forceinline__ T lift(T x, T a, T b, float c)
	{
		return __fadd_rn(x, __fmul_rn(__fadd_rn(a, b), c));
	}
	static __device__ __forceinline__ T iupd(int k, T x, T a, T b)
	{
		const float c = k == 0 ? -0.443506852f : k == 1 ? -0.882911075f : k == 2 ? 0.052980118f
																				 : 1.586134342f;
		return lift(x, a, b, c);
	}
	static __device__ __forceinline__ T fupd(int k, T x, T a, T b)
	{
		const float c = k == 0 ? -1.586134342f : k == 1 ? -0.052980118f : k == 2 ? 0.882911075f
																				 : 0.443506852f;
		return lift(x, a, b, c);
	}
	static __device__ __forceinline__ T iscaleL(T x) { return __fmul_rn(x, 1.230174105f); }
	static __device__ __forceinline__ T iscaleH(T x) { return __fmul_rn(x, 1.625732422f); }
	static __device__ __forceinline__ T fscaleL(T x)
	{
		return __fmul_rn(x, (float)(1.0 / 1.230174105));
	}
	static __device__ __forceinline__ T fscaleH(T x) { return __fmul_rn(x, 1.230174105f); }
};

template<typename T, int N>
struct VecIO;
template<typename T>
struct VecIO<T, 4>
{
	static __device__ __forceinline__ void ld(const T* p, T (&v)[4])
	{
		int4 r = __ldg(reinterpret_cast<const int4*>(p));
		v[0] = *reinterpret_cast<T*>(&r.x);
		v[1] = *reinterpret_cast<T*>(&r.y);
		v[2] = *reinterpret_cast<T*>(&r.z);
		v[3] = *reinterpret_cast<T*>(&r.w);
	}
};
template<typename T>
struct VecIO<T, 2>
{
	static __device__ __forceinline__ void ld(const T* p, T (&v)[2])
	{
		int2 r = __ldg(reinterpret_cast<const int2*>(p));
		v[0] = *reinterpret_cast<T*>(&r.x);
		v[1] = *reinterpret_cast<T*>(&r.y);
	}
};

template<typename T>
static __device__ __forceinline__ void st4(T* p, T a, T b, T c, T d)
{
	int4 r;
	r.x = *reinterpret_cast<int*>(&a);
	r.y = *reinterpret_cast<int*>(&b);
	r.z = *reinterpret_cast<int*>(&c);
	r.w = *reinterpret_cast<int*>(&d);
	*reinterpret_cast<int4*>(p) = r;
}
template<typename T>
static __device__ __forceinline__ void st2(T* p, T a, T b)
{
	int2 r;
	r.x = *reinterpret_cast<int*>(&a);
	r.y = *reinterpret_cast<int*>(&b);
	*reinterpret_cast<int2*>(p) = r;
}

static __device__ __forceinline__ int32_t grkb_clamp(int32_t v, int32_t lo, int32_t hi)
{
	return min(max(v, lo), hi);
}
