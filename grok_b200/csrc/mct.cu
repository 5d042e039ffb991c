// mct.cu -- stand-alone DC level shift and RCT / ICT multi-component transforms (sm_100a).
//
// Used when no wavelet level exists to fuse them into (numresolutions == 1) and as the
// stage-level C-ABI entry points for
//   mct::compress_rev / compress_irrev            src/lib/jp2/point_transform/mct.cpp:583-596
//   mct::decompress_rev / decompress_irrev        src/lib/jp2/point_transform/mct.cpp:480-579
//   mct::decompress_dc_shift_rev / _irrev         src/lib/jp2/point_transform/mct.cpp:450-539
//   TileProcessor::dcLevelShiftCompress           src/lib/jp2/tile/TileProcessor.cpp:506-535
// Pure streaming kernels: 128-bit loads and stores, grid-stride, one read + one write per
// sample (HBM-bound; 8 B/sample/component algorithmic).
#include "dwt.cuh"
#include "../../include/grk_b200.h"
#include <algorithm>

namespace
{
static __device__ __forceinline__ void ict_inv(float y, float u, float v, int32_t& r, int32_t& g,
												int32_t& b)
{
	r = __float2int_rn(__fadd_rn(y, __fmul_rn(v, 1.402f)));
	g = __float2int_rn(
		__fsub_rn(__fsub_rn(y, __fmul_rn(u, 0.34413f)), __fmul_rn(v, 0.71414f)));
	b = __float2int_rn(__fadd_rn(y, __fmul_rn(u, 1.772f)));
}

static __device__ __forceinline__ void ict_fwd(int32_t r, int32_t g, int32_t b, float& y, float& u,
												float& v)
{
	const float fr = (float)r, fg = (float)g, fb = (float)b;
	y = __fadd_rn(__fadd_rn(__fmul_rn(0.299f, fr), __fmul_rn(0.587f, fg)), __fmul_rn(0.114f, fb));
	const float cb = 0.5f / (1.0f - 0.114f);
	const float cr = 0.5f / (1.0f - 0.299f);
	u = __fmul_rn(cb, __fsub_rn(fb, y));
	v = __fmul_rn(cr, __fsub_rn(fr, y));
}

// MODE: 0 inverse RCT, 1 inverse ICT, 2 forward RCT, 3 forward ICT (DC shift fused both ways)
template<int MODE>
static __device__ __forceinline__ void mct_one(int32_t& a, int32_t& b, int32_t& c, const MctArgs& p)
{
	if(MODE == 0)
	{
		const int32_t g = a - ((b + c) >> 2);
		const int32_t r = c + g;
		const int32_t bl = b + g;
		a = grkb_clamp(r + p.shift[0], p.lo[0], p.hi[0]);
		b = grkb_clamp(g + p.shift[1], p.lo[1], p.hi[1]);
		c = grkb_clamp(bl + p.shift[2], p.lo[2], p.hi[2]);
	}
	else if(MODE == 1)
	{
		int32_t r, g, bl;
		ict_inv(__int_as_float(a), __int_as_float(b), __int_as_float(c), r, g, bl);
		a = grkb_clamp(r + p.shift[0], p.lo[0], p.hi[0]);
		b = grkb_clamp(g + p.shift[1], p.lo[1], p.hi[1]);
		c = grkb_clamp(bl + p.shift[2], p.lo[2], p.hi[2]);
	}
	else if(MODE == 2)
	{
		const int32_t r = a - p.shift[0], g = b - p.shift[1], bl = c - p.shift[2];
		a = (r + (g * 2) + bl) >> 2;
		b = bl - g;
		c = r - g;
	}
	else
	{
		float y, u, v;
		ict_fwd(a - p.shift[0], b - p.shift[1], c - p.shift[2], y, u, v);
		a = __float_as_int(y);
		b = __float_as_int(u);
		c = __float_as_int(v);
	}
}

template<int MODE>
__global__ void __launch_bounds__(256) mct_kernel(const __grid_constant__ MctArgs p)
{
	const uint64_t nvec = p.n / 4;
	int4* c0 = (int4*)p.c[0];
	int4* c1 = (int4*)p.c[1];
	int4* c2 = (int4*)p.c[2];
	for(uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
		i += (uint64_t)gridDim.x * blockDim.x)
	{
		int4 a = c0[i], b = c1[i], c = c2[i];
		mct_one<MODE>(a.x, b.x, c.x, p);
		mct_one<MODE>(a.y, b.y, c.y, p);
		mct_one<MODE>(a.z, b.z, c.z, p);
		mct_one<MODE>(a.w, b.w, c.w, p);
		c0[i] = a;
		c1[i] = b;
		c2[i] = c;
	}
	if(blockIdx.x == 0)
	{
		for(uint64_t i = nvec * 4 + threadIdx.x; i < p.n; i += blockDim.x)
		{
			int32_t a = ((int32_t*)p.c[0])[i], b = ((int32_t*)p.c[1])[i], c = ((int32_t*)p.c[2])[i];
			mct_one<MODE>(a, b, c, p);
			((int32_t*)p.c[0])[i] = a;
			((int32_t*)p.c[1])[i] = b;
			((int32_t*)p.c[2])[i] = c;
		}
	}
}

// MODE: 0 inverse reversible (x+shift, clamp), 1 inverse irreversible (rne(x)+shift, clamp),
//       2 forward reversible (x-shift), 3 forward irreversible ((float)(x-shift))
template<int MODE>
static __device__ __forceinline__ int32_t dc_one(int32_t x, const MctArgs& p)
{
	if(MODE == 0)
		return grkb_clamp(x + p.shift[0], p.lo[0], p.hi[0]);
	if(MODE == 1)
		return grkb_clamp(__float2int_rn(__int_as_float(x)) + p.shift[0], p.lo[0], p.hi[0]);
	if(MODE == 2)
		return x - p.shift[0];
	return __float_as_int((float)(x - p.shift[0]));
}

template<int MODE>
__global__ void __launch_bounds__(256) dc_kernel(const __grid_constant__ MctArgs p)
{
	const uint64_t nvec = p.n / 4;
	int4* c0 = (int4*)p.c[0];
	for(uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
		i += (uint64_t)gridDim.x * blockDim.x)
	{
		int4 a = c0[i];
		a.x = dc_one<MODE>(a.x, p);
		a.y = dc_one<MODE>(a.y, p);
		a.z = dc_one<MODE>(a.z, p);
		a.w = dc_one<MODE>(a.w, p);
		c0[i] = a;
	}
	if(blockIdx.x == 0)
		for(uint64_t i = nvec * 4 + threadIdx.x; i < p.n; i += blockDim.x)
			((int32_t*)p.c[0])[i] = dc_one<MODE>(((int32_t*)p.c[0])[i], p);
}

static uint32_t grid_for(uint64_t n)
{
	uint64_t blocks = (n / 4 + 255) / 256;
	const uint64_t cap = 148ull * 8ull; // persistent-ish: 8 CTAs of 256 threads per SM
	if(blocks > cap)
		blocks = cap;
	if(blocks == 0)
		blocks = 1;
	return (uint32_t)blocks;
}
} // namespace

// mode: 0 inverse RCT, 1 inverse ICT, 2 forward RCT, 3 forward ICT; in place on device planes
int grkb_launch_mct(const MctArgs& p, int mode, cudaStream_t stream)
{
	const uint32_t g = grid_for(p.n);
	switch(mode)
	{
		case 0: mct_kernel<0><<<g, 256, 0, stream>>>(p); break;
		case 1: mct_kernel<1><<<g, 256, 0, stream>>>(p); break;
		case 2: mct_kernel<2><<<g, 256, 0, stream>>>(p); break;
		case 3: mct_kernel<3><<<g, 256, 0, stream>>>(p); break;
		default: return -1;
	}
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}

// mode: 0 inverse reversible, 1 inverse irreversible, 2 forward reversible, 3 forward irreversible
int grkb_launch_dc(const MctArgs& p, int mode, cudaStream_t stream)
{
	const uint32_t g = grid_for(p.n);
	switch(mode)
	{
		case 0: dc_kernel<0><<<g, 256, 0, stream>>>(p); break;
		case 1: dc_kernel<1><<<g, 256, 0, stream>>>(p); break;
		case 2: dc_kernel<2><<<g, 256, 0, stream>>>(p); break;
		case 3: dc_kernel<3><<<g, 256, 0, stream>>>(p); break;
		default: return -1;
	}
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}


// ---- array-based ("custom matrix") multi-component transform, Part 2 (row M7) ------------------
//   mct::decompress_custom  src/lib/jp2/point_transform/mct.cpp:660-686: planes are read as float,
//     out[j] = sum_k M[j][k] * in[k], products and sums rounded separately, accumulated in k order
//     from 0 (the DC shift follows per component, TileProcessor.cpp:490-504 with mct == 2)
//   mct::compress_custom    mct.cpp:622-658: integer planes, the matrix in 13-bit fixed point
//     ((int32_t)(m * 8192.f)), out[j] = sum_k fix_mul(M[j][k], in[k]) with fix_mul rounding
//     (util/grk_intmath.h:66-79) -- except that the reference resets its matrix pointer to the first
//     row for every output component (:642), so every component comes out as row 0 applied to the
//     pixel.  Kept as the reference computes it (parity); the defect is listed in DESIGN.md.
namespace
{
struct CustomMctArgs
{
	void* c[GRKB_MAX_COMPS];
	uint64_t n;
	uint32_t nc;
	float m[GRKB_MAX_COMPS * GRKB_MAX_COMPS];
	int32_t mi[GRKB_MAX_COMPS * GRKB_MAX_COMPS];
};

template<int FORWARD>
__global__ void __launch_bounds__(256) mct_custom_kernel(const __grid_constant__ CustomMctArgs p)
{
	for(uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.n; i += (uint64_t)gridDim.x * blockDim.x)
	{
		int32_t in[GRKB_MAX_COMPS], out[GRKB_MAX_COMPS];
#pragma unroll
		for(uint32_t j = 0; j < GRKB_MAX_COMPS; ++j)
			in[j] = j < p.nc ? ((const int32_t*)p.c[j])[i] : 0;
#pragma unroll
		for(uint32_t j = 0; j < GRKB_MAX_COMPS; ++j)
		{
			if(j >= p.nc)
				break;
			if(FORWARD)
			{
				int32_t acc = 0;
				for(uint32_t k = 0; k < p.nc; ++k)
					acc += (int32_t)(((int64_t)p.mi[k] * (int64_t)in[k] + 4096) >> 13); // row 0 for every j: see below
				out[j] = acc;
			}
			else
			{
				float acc = 0.0f;
				for(uint32_t k = 0; k < p.nc; ++k)
					acc = __fadd_rn(acc, __fmul_rn(p.m[j * p.nc + k], __int_as_float(in[k])));
				out[j] = __float_as_int(acc);
			}
		}
#pragma unroll
		for(uint32_t j = 0; j < GRKB_MAX_COMPS; ++j)
			if(j < p.nc)
				((int32_t*)p.c[j])[i] = out[j];
	}
}
} // namespace

int grkb_launch_mct_custom(void* const* planes, uint32_t nc, uint64_t n, const float* matrix, int forward,
						   cudaStream_t s)
{
	if(nc == 0 || nc > GRKB_MAX_COMPS || !matrix)
	{
		grkb_set_error("custom MCT: 1..%d components and a matrix are needed", GRKB_MAX_COMPS);
		return -1;
	}
	CustomMctArgs a{};
	a.n = n;
	a.nc = nc;
	for(uint32_t c = 0; c < nc; ++c)
		a.c[c] = planes[c];
	for(uint32_t i = 0; i < nc * nc; ++i)
	{
		a.m[i] = matrix[i];
		a.mi[i] = (int32_t)(matrix[i] * (float)(1 << 13));
	}
	const unsigned grid = (unsigned)std::min<uint64_t>((n + 255) / 256, 148u * 8u);
	if(forward)
		mct_custom_kernel<1><<<grid ? grid : 1, 256, 0, s>>>(a);
	else
		mct_custom_kernel<0><<<grid ? grid : 1, 256, 0, s>>>(a);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}

extern "C" int grkb_mct_custom_host(void* const* planes, uint32_t numcomps, uint64_t n, const float* matrix,
									int forward)
{
	if(n == 0)
		return 0;
	if(!planes || numcomps == 0 || numcomps > GRKB_MAX_COMPS || !matrix)
	{
		grkb_set_error("grkb_mct_custom_host: bad arguments");
		return -1;
	}
	void* d[GRKB_MAX_COMPS] = {};
	int rc = -1;
	do
	{
		bool ok = true;
		for(uint32_t c = 0; c < numcomps && ok; ++c)
			ok = cudaMalloc(&d[c], n * 4) == cudaSuccess &&
				 cudaMemcpy(d[c], planes[c], n * 4, cudaMemcpyHostToDevice) == cudaSuccess;
		if(!ok)
		{
			grkb_set_error("grkb_mct_custom_host: no CUDA device or allocation failed (there is no CPU fallback)");
			cudaGetLastError();
			break;
		}
		if(grkb_launch_mct_custom(d, numcomps, n, matrix, forward, 0))
			break;
		for(uint32_t c = 0; c < numcomps && ok; ++c)
			ok = cudaMemcpy(planes[c], d[c], n * 4, cudaMemcpyDeviceToHost) == cudaSuccess;
		if(!ok)
		{
			grkb_set_error("grkb_mct_custom_host: copy back failed");
			break;
		}
		rc = 0;
	} while(0);
	for(uint32_t c = 0; c < numcomps; ++c)
		cudaFree(d[c]);
	return rc;
}
