// idwt_strip.cu -- bandwidth path of the inverse 5/3 and 9/7 wavelet transform (sm_100a).
//
// Same arithmetic and same results as idwt.cu (which stays as the path for ragged geometries);
// this kernel is the one the large levels of a tile run on.  Replaces, behind the C-ABI,
//   decompress_tile_53  src/lib/jp2/transform/WaveletReverse.cpp:802-877
//   decompress_tile_97  src/lib/jp2/transform/WaveletReverse.cpp:1272-1347
//   mct::decompress_rev / decompress_irrev  src/lib/jp2/point_transform/mct.cpp:480-579 (fused)
//
// Shape:
//   * a CTA is CW consumer warps + 1 producer warp and owns CW adjacent strips of one chunk of
//     rows.  A consumer warp owns a strip of 30*V output columns (lanes 0 and 31 are halo
//     lanes) and streams down the chunk, two rows (one vertical-low, one vertical-high) per
//     iteration;
//   * the producer warp stages each row pair of the CTA into a shared-memory ring with
//     bulk asynchronous copies (cp.async.bulk, completion on an mbarrier), STAGES row pairs
//     ahead, so no consumer register or issue slot is spent on global loads or address math;
//     vertical mirroring (whole-sample symmetric extension) is a row-index reflection done by
//     the producer, horizontal mirroring an index reflection when the edge warps read the ring;
//   * horizontal synthesis in registers with warp shuffles, vertical synthesis as a register
//     lifting pipeline (NS state values per column), results leave with 128-bit stores; on the
//     last level inverse RCT/ICT + DC shift + clamp are applied to the registers being stored.
// HBM traffic per level: every coefficient read once, every sample written once (the
// algorithmic bytes of SURVEY.md 8d); halo rows/columns are re-read from L2.
#include "dwt.cuh"
#include "bulk.cuh"
#include <stdlib.h>

namespace
{
constexpr int STAGES = 4; // row pairs in flight per consumer warp

// CW: consumer warps per CTA.  A CTA takes CW adjacent strips, so CW is chosen per level to leave
// as few consumer warps without columns as possible (2048 columns are 8.5 strips of 240: three CTAs
// of three warps are 95 % busy, three of four 71 %).
template<int V, int CW>
struct StripGeom
{
	static constexpr int H = V / 2; // coefficient pairs per lane
	static constexpr int LEAD = 4; // staging starts LEAD pairs before the first owned pair
	static constexpr int LEN = CW * 30 * H + 2 * LEAD; // staged samples per part (low or high)
	static constexpr int LANE0 = LEAD - H; // ring index of warp 0 / lane 0's first pair
};

// horizontal synthesis of one row held across the warp (same step order as idwt.cu)
template<class K, int H>
static __device__ __forceinline__ void hsynth_row(typename K::T (&e)[H], typename K::T (&o)[H])
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

template<class T, int H>
static __device__ __forceinline__ void lds_vec(const T* p, T (&v)[H])
{
	if constexpr(H == 4)
	{
		const int4 r = *reinterpret_cast<const int4*>(p);
		v[0] = *reinterpret_cast<const T*>(&r.x);
		v[1] = *reinterpret_cast<const T*>(&r.y);
		v[2] = *reinterpret_cast<const T*>(&r.z);
		v[H - 1] = *reinterpret_cast<const T*>(&r.w);
	}
	else
	{
		const int2 r = *reinterpret_cast<const int2*>(p);
		v[0] = *reinterpret_cast<const T*>(&r.x);
		v[H - 1] = *reinterpret_cast<const T*>(&r.y);
	}
}

// Finishing stage on the V samples a lane holds of one finished row, then 128-bit stores.
// FIN: 0 store raw, 1 DC shift + clamp (one component), 2 inverse MCT + DC shift + clamp.
template<class K, int V, int NC, int FIN>
static __device__ __forceinline__ void finish_store(const DwtLevelArgs& a, int cbase, size_t row,
													 int q0, typename K::T (&x)[NC][V])
{
	typedef typename K::T T;
	if constexpr(FIN == 0)
	{
#pragma unroll
		for(int c = 0; c < NC; ++c)
		{
			T* dst = (T*)a.pl[cbase + c].full + row * a.pl[cbase + c].f_stride + q0;
#pragma unroll
			for(int i = 0; i < V; i += 4)
				st4(dst + i, x[c][i], x[c][i + 1], x[c][i + 2], x[c][i + 3]);
		}
	}
	else
	{
		int32_t r[NC][V];
#pragma unroll
		for(int i = 0; i < V; ++i)
		{
			if constexpr(FIN == 2)
			{
				int32_t c0, c1, c2;
				if constexpr(K::NS == 4)
				{
					// inverse ICT, point_transform/mct.cpp:284-364
					const float fy = x[0][i], fu = x[NC > 1 ? 1 : 0][i], fv = x[NC > 2 ? 2 : 0][i];
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
					const int32_t iy = x[0][i], iu = x[NC > 1 ? 1 : 0][i], iv = x[NC > 2 ? 2 : 0][i];
					const int32_t g = iy - ((iu + iv) >> 2);
					c0 = iv + g;
					c1 = g;
					c2 = iu + g;
				}
				r[0][i] = grkb_clamp(c0 + a.shift[0], a.lo[0], a.hi[0]);
				r[NC > 1 ? 1 : 0][i] = grkb_clamp(c1 + a.shift[1], a.lo[1], a.hi[1]);
				r[NC > 2 ? 2 : 0][i] = grkb_clamp(c2 + a.shift[2], a.lo[2], a.hi[2]);
			}
			else
			{
#pragma unroll
				for(int c = 0; c < NC; ++c)
				{
					int32_t iv;
					if constexpr(K::NS == 4)
						iv = __float2int_rn(x[c][i]);
					else
						iv = x[c][i];
					r[c][i] = grkb_clamp(iv + a.shift[cbase + c], a.lo[cbase + c], a.hi[cbase + c]);
				}
			}
		}
#pragma unroll
		for(int c = 0; c < NC; ++c)
		{
			int32_t* dst = (int32_t*)a.pl[cbase + c].full + row * a.pl[cbase + c].f_stride + q0;
#pragma unroll
			for(int i = 0; i < V; i += 4)
				st4(dst + i, r[c][i], r[c][i + 1], r[c][i + 2], r[c][i + 3]);
		}
	}
}

template<class K, int V, int NC, int FIN, int CW>
__global__ void __launch_bounds__((CW + 1) * GRKB_WARP)
	idwt_strip_kernel(const __grid_constant__ DwtLevelArgs a)
{
	typedef typename K::T T;
	typedef StripGeom<V, CW> G;
	constexpr int H = G::H, NS = K::NS, LEN = G::LEN;
	constexpr int ROW = 2 * LEN; // one staged row: low part then high part
	constexpr int STAGE = NC * 2 * ROW; // one stage: a row pair of NC components
	constexpr int NCOPY = NC * 4;
	extern __shared__ __align__(128) unsigned char smem_raw[];
	T* ring = reinterpret_cast<T*>(smem_raw); // [STAGES][STAGE], shared by the CW consumer warps
	uint64_t* bars = reinterpret_cast<uint64_t*>(ring + STAGES * STAGE); // full[s], empty[s]

	const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int cbase = (NC == 1) ? (int)blockIdx.y : 0;
	const int py = (int)a.py, rw = (int)a.rw, rh = (int)a.rh, sw = (int)a.sw, sh = (int)a.sh;
	// a CTA = CW adjacent strips of one chunk of rows
	const int ngroups = ((int)a.nstrips + CW - 1) / CW;
	const int group = blockIdx.x % ngroups;
	const int chunk = blockIdx.x / ngroups;
	const int nact = min(CW, (int)a.nstrips - group * CW); // consumer warps with work
	if(threadIdx.x == 0)
	{
		for(int i = 0; i < STAGES; ++i)
		{
			mbar_init(smem_u32(bars + 2 * i), NCOPY);
			mbar_init(smem_u32(bars + 2 * i + 1), nact);
		}
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();

	const int ya = chunk * (int)a.chunk_rows;
	const int yend = (py + rh + 1) & ~1;
	const int yb = min(ya + (int)a.chunk_rows, yend);
	const int m0 = ya / 2 - NS / 2;
	const int m1 = yb / 2 + NS / 2 - 1;
	const int A = group * CW * 30 * H - G::LEAD; // first staged pair index
	const int g0 = max(A, 0), g1 = min(A + LEN, sw); // staged pairs that exist

	if(wid == CW)
	{
		// ---- producer warp: lane k issues copy k of every stage ------------------------------
		const uint32_t bytes = (uint32_t)(g1 - g0) * 4u;
		const int c = lane >> 2, r = (lane >> 1) & 1, part = lane & 1;
		const DwtPlane& p = a.pl[cbase + (lane < NCOPY ? c : 0)];
		const T* ll = (const T*)p.ll;
		const T* bands = (const T*)p.bands;
		const uint32_t dst0 = smem_u32(ring + (c * 2 + r) * ROW + part * LEN + (g0 - A));
		int s = 0;
		uint32_t ph = 0;
		for(int m = m0; m <= m1; ++m)
		{
			const uint32_t full = smem_u32(bars + 2 * s), empty = smem_u32(bars + 2 * s + 1);
			mbar_wait(empty, ph ^ 1u);
			if(lane < NCOPY)
			{
				const int qm = grkb_mirror(2 * m + r, py, rh);
				const T* src;
				if((qm & 1) == 0)
				{
					const int rr = qm / 2 - py;
					src = part ? bands + (size_t)rr * p.b_stride + sw : ll + (size_t)rr * p.ll_stride;
				}
				else
				{
					const int rr = (qm - 1) / 2;
					src = bands + (size_t)(sh + rr) * p.b_stride + (part ? sw : 0);
				}
				mbar_arrive_tx(full, bytes);
				bulk_g2s(dst0 + (uint32_t)(s * STAGE * (int)sizeof(T)), src + g0, bytes, full);
			}
			if(++s == STAGES)
			{
				s = 0;
				ph ^= 1u;
			}
		}
		return;
	}

	// ---- consumer ----------------------------------------------------------------------------
	if(wid >= nact)
		return;
	const int li = G::LANE0 + (wid * 30 + lane) * H; // ring index of this lane's first pair
	const int q0 = 2 * (A + li); // first output column of this lane
	const bool owner = lane >= 1 && lane <= 30 && q0 + V <= rw;
	const bool edge = (A + li - lane * H < 0) || (A + li + (32 - lane) * H > sw);
	int ie[H], io[H];
#pragma unroll
	for(int j = 0; j < H; ++j)
	{
		const int n = A + li + j;
		int nl = n < 0 ? -n : (n >= sw ? 2 * sw - 1 - n : n);
		int nh = n < 0 ? -n - 1 : (n >= sw ? 2 * sw - 2 - n : n);
		nl = min(max(nl, g0), g1 - 1);
		nh = min(max(nh, g0), g1 - 1);
		ie[j] = nl - A;
		io[j] = nh - A + LEN;
	}

	T st[NC][NS][V];
#pragma unroll
	for(int c = 0; c < NC; ++c)
#pragma unroll
		for(int k = 0; k < NS; ++k)
#pragma unroll
			for(int i = 0; i < V; ++i)
				st[c][k][i] = (T)0;

	int s = 0;
	uint32_t ph = 0;
#pragma unroll 2
	for(int m = m0; m <= m1; ++m)
	{
		const uint32_t full = smem_u32(bars + 2 * s), empty = smem_u32(bars + 2 * s + 1);
		mbar_wait(full, ph);
		const T* stage = ring + s * STAGE;
		T fe[NC][V], fo[NC][V]; // finished even / odd row, natural column order
#pragma unroll
		for(int c = 0; c < NC; ++c)
		{
			T e0[H], o0[H], e1[H], o1[H];
			const T* r0 = stage + (c * 2) * ROW;
			const T* r1 = r0 + ROW;
			if(!edge)
			{
				lds_vec<T, H>(r0 + li, e0);
				lds_vec<T, H>(r0 + LEN + li, o0);
				lds_vec<T, H>(r1 + li, e1);
				lds_vec<T, H>(r1 + LEN + li, o1);
			}
			else
			{
#pragma unroll
				for(int j = 0; j < H; ++j)
				{
					e0[j] = r0[ie[j]];
					o0[j] = r0[io[j]];
					e1[j] = r1[ie[j]];
					o1[j] = r1[io[j]];
				}
			}
			if(c == NC - 1)
			{
				__syncwarp();
				if(lane == 0)
					mbar_arrive(empty);
			}
			hsynth_row<K, H>(e0, o0);
			hsynth_row<K, H>(e1, o1);
#pragma unroll
			for(int i = 0; i < V; ++i)
			{
				const int j = i >> 1;
				const T P = K::iscaleL((i & 1) ? o0[j] : e0[j]);
				const T Q = K::iscaleH((i & 1) ? o1[j] : e1[j]);
				T an[NS + 1];
				an[0] = Q;
				an[1] = K::iupd(0, P, st[c][0][i], Q);
#pragma unroll
				for(int k = 1; k < NS; ++k)
					an[k + 1] = K::iupd(k, st[c][k - 1][i], st[c][k][i], an[k]);
				fe[c][i] = an[NS - 1];
				fo[c][i] = an[NS];
#pragma unroll
				for(int k = 0; k < NS; ++k)
					st[c][k][i] = an[k];
			}
		}
		if(++s == STAGES)
		{
			s = 0;
			ph ^= 1u;
		}
		const int row_odd = 2 * m - (NS - 1);
		const int row_even = 2 * m - (NS - 2);
		if(owner)
		{
			if(row_odd >= ya && row_odd < yb && row_odd >= py && row_odd < py + rh)
				finish_store<K, V, NC, FIN>(a, cbase, (size_t)(row_odd - py), q0, fo);
			if(row_even >= ya && row_even < yb && row_even >= py && row_even < py + rh)
				finish_store<K, V, NC, FIN>(a, cbase, (size_t)(row_even - py), q0, fe);
		}
	}
}

template<class K, int V, int NC, int FIN, int CW>
static int launch_strip_cw(DwtLevelArgs a, int nplanes, cudaStream_t stream)
{
	typedef StripGeom<V, CW> G;
	constexpr size_t smem =
		(size_t)STAGES * NC * 2 * 2 * G::LEN * 4 + (size_t)STAGES * 2 * sizeof(uint64_t);
	GRKB_ENSURE_SMEM((idwt_strip_kernel<K, V, NC, FIN, CW>), smem);
	const uint32_t strip = 30 * V;
	a.nstrips = (a.rw + strip - 1) / strip;
	// Chunk height: all CTAs are resident at once (<= 4 per SM), so the grid is sized to fill
	// the 148 SMs evenly -- waves*148 CTAs or just under -- with chunks no shorter than
	// min_rows (every chunk re-runs NS rows of vertical warm-up).
	static int waves = -1, min_rows = 0;
	if(waves < 0)
	{
		const char* e = getenv("GRKB_IDWT_WAVES"); // tuning knobs, see DESIGN.md section 4
		waves = e ? atoi(e) : 0; // 0: per kernel (below)
		if(waves < 0 || waves > 4)
			waves = 0;
		e = getenv("GRKB_IDWT_MIN_ROWS");
		min_rows = e ? atoi(e) : 4; // small levels are latency-bound: short chunks, many CTAs
		if(min_rows < 4)
			min_rows = 4;
	}
	const uint32_t total_rows = a.py + a.rh;
	const uint32_t ngroups = (a.nstrips + CW - 1) / CW;
	const uint32_t per_chunk = ngroups * ((NC == 1) ? (uint32_t)nplanes : 1u);
	// CTAs per SM: the 9/7 kernels hide their longer dependency chains with a third CTA per SM
	// (3 x 160 threads x 125 registers and 3 x 47 KB still fit); 5/3 is fastest with two.
	// (Tried and dropped: the lifting steps on the packed FP32 pipe, FADD2 + FMUL2 + scalar FADD
	// -- ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2, so the last add must stay scalar;
	// bit-exact, but 3 % slower than this kernel: it is not bound by FP32 issue.)
	const uint32_t nwaves = waves ? (uint32_t)waves : (K::NS == 4 ? 3u : 2u);
	uint32_t nch = (148u * nwaves) / per_chunk;
	if(nch < 1)
		nch = 1;
	uint32_t ch = 2 * ((total_rows + 2 * nch - 1) / (2 * nch));
	if(ch < (uint32_t)min_rows)
		ch = (uint32_t)min_rows & ~1u;
	a.chunk_rows = ch;
	a.nchunks = (total_rows + ch - 1) / ch;
	dim3 grid(ngroups * a.nchunks, (NC == 1) ? nplanes : 1, 1);
	idwt_strip_kernel<K, V, NC, FIN, CW><<<grid, (CW + 1) * GRKB_WARP, smem, stream>>>(a);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}

// picks the CTA width that wastes the fewest consumer warps on this level's strips
template<class K, int V, int NC, int FIN>
static int launch_strip(DwtLevelArgs a, int nplanes, cudaStream_t stream)
{
	const uint32_t nstrips = (a.rw + 30 * V - 1) / (30 * V);
	static int forced = -1;
	if(forced < 0)
	{
		const char* e = getenv("GRKB_IDWT_CW");
		forced = e ? atoi(e) : 0;
	}
	// measured on B200 (4096x2160x3, 9/7): the 2048-wide level 18.5 -> 16.4 us with three warps;
	// the three-component kernel is fastest with four whatever the strip count (five: 45 -> 56 us)
	uint32_t best = 4;
	if(NC == 1 && ((nstrips + 2) / 3) * 3 < ((nstrips + 3) / 4) * 4)
		best = 3;
	if(forced == 3 || forced == 4)
		best = (uint32_t)forced;
	if(best == 3)
		return launch_strip_cw<K, V, NC, FIN, 3>(a, nplanes, stream);
	return launch_strip_cw<K, V, NC, FIN, 4>(a, nplanes, stream);
}
} // namespace

// Returns 1 when the level's geometry is not one this kernel takes (the caller then uses the
// general kernel of idwt.cu), 0 on success, -1 on a CUDA error.
int grkb_launch_idwt_strip(const DwtLevelArgs& a, int reversible, int ncomp, int fin,
							cudaStream_t stream)
{
	if(a.px != 0 || (a.rw % 8) != 0 || a.rw < 64 || a.rh < 2 || a.sw * 2 != a.rw)
		return 1;
	if(ncomp != 1 && ncomp != 3)
		return 1;
	for(int c = 0; c < ncomp; ++c)
	{
		const DwtPlane& p = a.pl[c];
		if((p.ll_stride % 4) || (p.b_stride % 4) || (p.f_stride % 4) || ((uintptr_t)p.ll % 16) ||
		   ((uintptr_t)p.bands % 16) || ((uintptr_t)p.full % 16))
			return 1;
	}
	if(ncomp == 3 && fin == 2)
		return reversible ? launch_strip<K53, 4, 3, 2>(a, 1, stream)
						  : launch_strip<K97, 4, 3, 2>(a, 1, stream);
	if(fin == 2)
		return 1;
	if(fin == 1)
	{
		if(ncomp != 1)
			return 1;
		return reversible ? launch_strip<K53, 8, 1, 1>(a, 1, stream)
						  : launch_strip<K97, 8, 1, 1>(a, 1, stream);
	}
	return reversible ? launch_strip<K53, 8, 1, 0>(a, ncomp, stream)
					  : launch_strip<K97, 8, 1, 0>(a, ncomp, stream);
}
