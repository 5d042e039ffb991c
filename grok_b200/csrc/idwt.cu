// idwt.cu -- inverse 5/3 and 9/7 wavelet transform, one resolution level per launch,
// horizontal + vertical synthesis fused, optional inverse MCT + DC shift + clamp fused into
// the store of the last level.  sm_100a.
//
// Replaces, behind the C-ABI, the reference's
//   decompress_tile_53  src/lib/jp2/transform/WaveletReverse.cpp:802-877
//   decompress_tile_97  src/lib/jp2/transform/WaveletReverse.cpp:1272-1347
//   mct::decompress_rev / decompress_irrev / decompress_dc_shift_*  point_transform/mct.cpp:450-579
//
// Shape of the kernel (HBM-bound, no tensor cores):
//   * one warp owns a strip of 30*V output columns and streams down a chunk of rows;
//   * per iteration it loads one vertical-low row (LL|HL) and one vertical-high row (LH|HH)
//     with 128-bit coalesced loads, V/2 low-pass + V/2 high-pass coefficients per lane;
//   * horizontal synthesis runs in registers, neighbours come from warp shuffles (lanes 0
//     and 31 are halo lanes: they recompute what the neighbouring warp owns);
//   * vertical synthesis is a streaming lifting pipeline: NS state registers per column,
//     two finished rows leave per iteration with 128-bit stores;
//   * whole-sample symmetric extension is done by reflecting the load index, so no boundary
//     special cases exist in the arithmetic (see oracle/dwt.c for why this is bit-identical).
// Each level therefore reads rw*rh and writes rw*rh samples once: the algorithmic bytes of
// SURVEY.md section 8d.
#include "dwt.cuh"

namespace
{
constexpr int WARPS_PER_CTA = 4;

template<class K, int V, bool AL>
struct RowLoader
{
	typedef typename K::T T;
	static constexpr int H = V / 2;
	// Load the horizontally de-interleaved samples of canvas row qy for columns
	// q0 .. q0+V-1 (canvas-aligned coordinates: even q = low-pass).
	static __device__ __forceinline__ void load(const DwtLevelArgs& a, const DwtPlane& p, int qy,
												 int q0, bool interior, T (&e)[H], T (&o)[H])
	{
		const int qm = grkb_mirror(qy, (int)a.py, (int)a.rh);
		const T* L;
		const T* Hp;
		if((qm & 1) == 0)
		{
			const int r = qm / 2 - (int)a.py;
			L = (const T*)p.ll + (size_t)r * p.ll_stride;
			Hp = (const T*)p.bands + (size_t)r * p.b_stride + a.sw;
		}
		else
		{
			const int r = (qm - 1) / 2;
			L = (const T*)p.bands + (size_t)(a.sh + r) * p.b_stride;
			Hp = L + a.sw;
		}
		if(AL && interior)
		{
			VecIO<T, H>::ld(L + q0 / 2, e);
			VecIO<T, H>::ld(Hp + q0 / 2, o);
		}
		else
		{
#pragma unroll
			for(int j = 0; j < H; ++j)
			{
				const int qe = grkb_mirror(q0 + 2 * j, (int)a.px, (int)a.rw);
				const int qo = grkb_mirror(q0 + 2 * j + 1, (int)a.px, (int)a.rw);
				e[j] = __ldg(L + max(qe / 2 - (int)a.px, 0)); // (sw == 0: value unused)
				o[j] = __ldg(Hp + ((qo - 1) / 2));
			}
		}
	}
};

// horizontal synthesis of one row held across the warp
template<class K, int H>
static __device__ __forceinline__ void hsynth(typename K::T (&e)[H], typename K::T (&o)[H])
{
	typedef typename K::T T;
#pragma unroll
	for(int j = 0; j < H; ++j)
	{
		e[j] = K::iscaleL(e[j]);
		o[j] = K::iscaleH(o[j]);
	}
#pragma unroll
	for(int k = 0; k < K::NS; ++k)
	{
		if((k & 1) == 0)
		{
			const T om1 = __shfl_up_sync(0xffffffffu, o[H - 1], 1);
#pragma unroll
			for(int j = 0; j < H; ++j)
				e[j] = K::iupd(k, e[j], j ? o[j - 1] : om1, o[j]);
		}
		else
		{
			const T ep1 = __shfl_down_sync(0xffffffffu, e[0], 1);
#pragma unroll
			for(int j = 0; j < H; ++j)
				o[j] = K::iupd(k, o[j], e[j], (j + 1 < H) ? e[j + 1] : ep1);
		}
	}
}

// FIN: 0 store raw samples, 1 DC shift + clamp of a single component, 2 inverse MCT + DC shift
// + clamp over three components.
template<class K, int V, int NC, int FIN, bool AL>
__global__ void __launch_bounds__(WARPS_PER_CTA* GRKB_WARP)
	idwt_level_kernel(const __grid_constant__ DwtLevelArgs a)
{
	typedef typename K::T T;
	constexpr int H = V / 2;
	constexpr int NS = K::NS;
	constexpr int STRIP = 30 * V;
	const int warp = blockIdx.x * WARPS_PER_CTA + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if(warp >= (int)(a.nstrips * a.nchunks))
		return;
	const int strip = warp % (int)a.nstrips;
	const int chunk = warp / (int)a.nstrips;
	const int q0 = strip * STRIP + (lane - 1) * V;
	const int px = (int)a.px, py = (int)a.py, rw = (int)a.rw, rh = (int)a.rh;
	const bool interior = (q0 >= px) && (q0 + V <= px + rw);
	const bool owner = lane >= 1 && lane <= 30;
	const bool hnorm = rw > 1, vnorm = rh > 1;

	// ---- emit one finished canvas row ------------------------------------------------------
	auto emit = [&](int qy, T(&xe)[NC][H], T(&xo)[NC][H]) {
		if(!owner || qy < py || qy >= py + rh)
			return;
		const size_t row = (size_t)(qy - py);
		if(FIN == 0)
		{
#pragma unroll
			for(int c = 0; c < NC; ++c)
			{
				T* dst = (T*)a.pl[c].full + row * a.pl[c].f_stride;
				if(AL && interior)
				{
					if(V == 8)
					{
						st4(dst + q0, xe[c][0], xo[c][0], xe[c][1], xo[c][1]);
						st4(dst + q0 + 4, xe[c][H - 2], xo[c][H - 2], xe[c][H - 1], xo[c][H - 1]);
					}
					else
						st4(dst + q0, xe[c][0], xo[c][0], xe[c][H - 1], xo[c][H - 1]);
				}
				else
				{
#pragma unroll
					for(int j = 0; j < H; ++j)
					{
						const int qe = q0 + 2 * j, qo = qe + 1;
						if(qe >= px && qe < px + rw)
							dst[qe - px] = xe[c][j];
						if(qo >= px && qo < px + rw)
							dst[qo - px] = xo[c][j];
					}
				}
			}
		}
		else
		{
			int32_t r[NC][V];
#pragma unroll
			for(int i = 0; i < V; ++i)
			{
				if(FIN == 2)
				{
					const T y = (i & 1) ? xo[0][i >> 1] : xe[0][i >> 1];
					const T u = (i & 1) ? xo[NC > 1 ? 1 : 0][i >> 1] : xe[NC > 1 ? 1 : 0][i >> 1];
					const T v = (i & 1) ? xo[NC > 2 ? 2 : 0][i >> 1] : xe[NC > 2 ? 2 : 0][i >> 1];
					int32_t c0, c1, c2;
					if(sizeof(T) == 4 && K::NS == 4)
					{
						// inverse ICT, point_transform/mct.cpp:284-364
						const float fy = *(const float*)&y, fu = *(const float*)&u,
									fv = *(const float*)&v;
						const float fr = __fadd_rn(fy, __fmul_rn(fv, 1.402f));
						const float fg = __fsub_rn(__fsub_rn(fy, __fmul_rn(fu, 0.34413f)),
												   __fmul_rn(fv, 0.71414f));
						const float fb = __fadd_rn(fy, __fmul_rn(fu, 1.772f));
						c0 = __float2int_rn(fr);
						c1 = __float2int_rn(fg);
						c2 = __float2int_rn(fb);
					}
					else
					{
						// inverse RCT, point_transform/mct.cpp:221-283
						const int32_t iy = *(const int32_t*)&y, iu = *(const int32_t*)&u,
									  iv = *(const int32_t*)&v;
						const int32_t g = iy - ((iu + iv) >> 2);
						c0 = iv + g;
						c1 = g;
						c2 = iu + g;
					}
					r[0][i] = grkb_clamp(c0 + a.shift[0], a.lo[0], a.hi[0]);
					if(NC > 1)
						r[NC > 1 ? 1 : 0][i] = grkb_clamp(c1 + a.shift[1], a.lo[1], a.hi[1]);
					if(NC > 2)
						r[NC > 2 ? 2 : 0][i] = grkb_clamp(c2 + a.shift[2], a.lo[2], a.hi[2]);
				}
				else
				{
#pragma unroll
					for(int c = 0; c < NC; ++c)
					{
						const T x = (i & 1) ? xo[c][i >> 1] : xe[c][i >> 1];
						int32_t iv;
						if(K::NS == 4)
							iv = __float2int_rn(*(const float*)&x);
						else
							iv = *(const int32_t*)&x;
						r[c][i] = grkb_clamp(iv + a.shift[c], a.lo[c], a.hi[c]);
					}
				}
			}
#pragma unroll
			for(int c = 0; c < NC; ++c)
			{
				int32_t* dst = (int32_t*)a.pl[c].full + row * a.pl[c].f_stride;
				if(AL && interior)
				{
#pragma unroll
					for(int i = 0; i < V; i += 4)
						st4(dst + q0 + i, r[c][i], r[c][i + 1], r[c][i + 2], r[c][i + 3]);
				}
				else
				{
#pragma unroll
					for(int i = 0; i < V; ++i)
					{
						const int q = q0 + i;
						if(q >= px && q < px + rw)
							dst[q - px] = r[c][i];
					}
				}
			}
		}
	};

	// ---- degenerate height: no vertical lifting (WaveletReverse.cpp:608-637, :1012-1014) ----
	if(!vnorm)
	{
		if(chunk != 0)
			return;
		T xe[NC][H], xo[NC][H];
#pragma unroll
		for(int c = 0; c < NC; ++c)
		{
			RowLoader<K, V, AL>::load(a, a.pl[c], py, q0, interior, xe[c], xo[c]);
			if(hnorm)
				hsynth<K, H>(xe[c], xo[c]);
			else if(K::NS == 2 && px == 1)
			{
#pragma unroll
				for(int j = 0; j < H; ++j)
					xo[c][j] = (T)((int32_t)xo[c][j] / 2);
			}
			if(K::NS == 2 && py == 1)
			{
#pragma unroll
				for(int j = 0; j < H; ++j)
				{
					xe[c][j] = (T)((int32_t)xe[c][j] >> 1);
					xo[c][j] = (T)((int32_t)xo[c][j] >> 1);
				}
			}
		}
		emit(py, xe, xo);
		return;
	}

	// ---- streaming vertical synthesis -------------------------------------------------------
	const int ya = chunk * (int)a.chunk_rows;
	int yb = ya + (int)a.chunk_rows;
	const int yend = (py + rh + 1) & ~1;
	if(yb > yend)
		yb = yend;
	const int m0 = ya / 2 - NS / 2;
	const int m1 = yb / 2 + NS / 2 - 1;

	T st[NC][NS][V];
#pragma unroll
	for(int c = 0; c < NC; ++c)
#pragma unroll
		for(int k = 0; k < NS; ++k)
#pragma unroll
			for(int i = 0; i < V; ++i)
				st[c][k][i] = (T)0;

	T ce[NC][2][H], co[NC][2][H]; // current pair: [.][0] even row 2m, [.][1] odd row 2m+1
#pragma unroll
	for(int c = 0; c < NC; ++c)
	{
		RowLoader<K, V, AL>::load(a, a.pl[c], 2 * m0, q0, interior, ce[c][0], co[c][0]);
		RowLoader<K, V, AL>::load(a, a.pl[c], 2 * m0 + 1, q0, interior, ce[c][1], co[c][1]);
	}
	for(int m = m0; m <= m1; ++m)
	{
		T ne[NC][2][H], no[NC][2][H];
		if(m < m1)
		{
#pragma unroll
			for(int c = 0; c < NC; ++c)
			{
				RowLoader<K, V, AL>::load(a, a.pl[c], 2 * m + 2, q0, interior, ne[c][0], no[c][0]);
				RowLoader<K, V, AL>::load(a, a.pl[c], 2 * m + 3, q0, interior, ne[c][1], no[c][1]);
			}
		}
		T fe_e[NC][H], fe_o[NC][H]; // finished even row
		T fo_e[NC][H], fo_o[NC][H]; // finished odd row
#pragma unroll
		for(int c = 0; c < NC; ++c)
		{
			if(hnorm)
			{
				hsynth<K, H>(ce[c][0], co[c][0]);
				hsynth<K, H>(ce[c][1], co[c][1]);
			}
			else if(K::NS == 2 && px == 1)
			{
#pragma unroll
				for(int j = 0; j < H; ++j)
				{
					co[c][0][j] = (T)((int32_t)co[c][0][j] / 2); // WaveletReverse.cpp:583
					co[c][1][j] = (T)((int32_t)co[c][1][j] / 2);
				}
			}
#pragma unroll
			for(int i = 0; i < V; ++i)
			{
				const int j = i >> 1;
				const T P = K::iscaleL((i & 1) ? co[c][0][j] : ce[c][0][j]);
				const T Q = K::iscaleH((i & 1) ? co[c][1][j] : ce[c][1][j]);
				T an[NS + 1];
				an[0] = Q;
				an[1] = K::iupd(0, P, st[c][0][i], Q);
#pragma unroll
				for(int k = 1; k < NS; ++k)
					an[k + 1] = K::iupd(k, st[c][k - 1][i], st[c][k][i], an[k]);
				// finished: even row value an[NS-1], odd row value an[NS]
				if(i & 1)
				{
					fe_o[c][j] = an[NS - 1];
					fo_o[c][j] = an[NS];
				}
				else
				{
					fe_e[c][j] = an[NS - 1];
					fo_e[c][j] = an[NS];
				}
#pragma unroll
				for(int k = 0; k < NS; ++k)
					st[c][k][i] = an[k];
			}
		}
		const int row_odd = 2 * m - (NS - 1);
		const int row_even = 2 * m - (NS - 2);
		if(row_odd >= ya && row_odd < yb)
			emit(row_odd, fo_e, fo_o);
		if(row_even >= ya && row_even < yb)
			emit(row_even, fe_e, fe_o);
#pragma unroll
		for(int c = 0; c < NC; ++c)
#pragma unroll
			for(int r = 0; r < 2; ++r)
#pragma unroll
				for(int j = 0; j < H; ++j)
				{
					ce[c][r][j] = ne[c][r][j];
					co[c][r][j] = no[c][r][j];
				}
	}
}

template<class K, int V, int NC, int FIN>
static int launch_idwt(const DwtLevelArgs& a, bool aligned, cudaStream_t stream)
{
	const uint32_t warps = a.nstrips * a.nchunks;
	const uint32_t ctas = (warps + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
	if(ctas == 0)
		return 0;
	if(aligned)
		idwt_level_kernel<K, V, NC, FIN, true><<<ctas, WARPS_PER_CTA * GRKB_WARP, 0, stream>>>(a);
	else
		idwt_level_kernel<K, V, NC, FIN, false><<<ctas, WARPS_PER_CTA * GRKB_WARP, 0, stream>>>(a);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
} // namespace

// Host-side dispatch.  fin: 0 raw, 1 dc-shift per component, 2 MCT over 3 components.
// ncomp planes of a.pl are processed (they must share geometry).
int grkb_launch_idwt_strip(const DwtLevelArgs& a, int reversible, int ncomp, int fin,
							cudaStream_t stream);

int grkb_launch_idwt_level(DwtLevelArgs a, int reversible, int ncomp, int fin, cudaStream_t stream)
{
	const int fast = grkb_launch_idwt_strip(a, reversible, ncomp, fin, stream);
	if(fast <= 0)
		return fast;
	const int V = (ncomp == 3) ? 4 : 8;
	const uint32_t H = V / 2;
	bool al = a.px == 0 && (a.sw % H) == 0;
	for(int c = 0; c < ncomp; ++c)
	{
		const DwtPlane& p = a.pl[c];
		al = al && (p.ll_stride % 4) == 0 && (p.b_stride % 4) == 0 && (p.f_stride % 4) == 0 &&
			 ((uintptr_t)p.ll % 16) == 0 && ((uintptr_t)p.bands % 16) == 0 &&
			 ((uintptr_t)p.full % 16) == 0;
	}
	const uint32_t strip = 30 * V;
	a.nstrips = (a.px + a.rw + strip - 1) / strip;
	// pick the chunk height: enough warps to fill 148 SMs x 16 warps, little halo overhead
	const uint32_t total_rows = a.py + a.rh;
	uint32_t ch = 256;
	while(ch > 16 && (uint64_t)a.nstrips * ((total_rows + ch - 1) / ch) < 148u * 16u)
		ch >>= 1;
	a.chunk_rows = ch;
	a.nchunks = (total_rows + ch - 1) / ch;
	if(a.rh <= 1)
		a.nchunks = 1;
	if(ncomp == 3)
	{
		if(reversible)
			return fin == 2 ? launch_idwt<K53, 4, 3, 2>(a, al, stream)
							: launch_idwt<K53, 4, 3, 0>(a, al, stream);
		return fin == 2 ? launch_idwt<K97, 4, 3, 2>(a, al, stream)
						: launch_idwt<K97, 4, 3, 0>(a, al, stream);
	}
	if(ncomp != 1)
		return -1;
	if(reversible)
		return fin == 1 ? launch_idwt<K53, 8, 1, 1>(a, al, stream)
						: launch_idwt<K53, 8, 1, 0>(a, al, stream);
	return fin == 1 ? launch_idwt<K97, 8, 1, 1>(a, al, stream)
					: launch_idwt<K97, 8, 1, 0>(a, al, stream);
}
