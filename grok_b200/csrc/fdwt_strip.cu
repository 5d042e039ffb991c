// fdwt_strip.cu -- bandwidth path of the forward 5/3 and 9/7 wavelet transform (sm_100a).
//
// Same arithmetic and same results as fdwt.cu (which stays as the path for ragged geometries);
// this kernel is the one the large levels of a tile run on.  Replaces, behind the C-ABI,
//   WaveletFwdImpl::encode_procedure   src/lib/jp2/transform/WaveletFwd.cpp:432-616
//   dcLevelShiftCompress / mct_encode  src/lib/jp2/tile/TileProcessor.cpp:506-573 (fused)
//   mct::compress_rev / compress_irrev src/lib/jp2/point_transform/mct.cpp:99-219 (fused)
//
// Shape (mirror of idwt_strip.cu):
//   * a CTA is CW consumer warps + 1 producer warp and owns CW adjacent strips of one chunk of
//     rows.  A consumer warp owns 30*V source columns (lanes 0 and 31 are halo lanes) and
//     streams down the chunk, two source rows (one odd, one even) per iteration;
//   * the producer warp stages every source row pair of the CTA into a shared-memory ring with
//     bulk asynchronous copies (cp.async.bulk, completion on an mbarrier), STAGES pairs ahead;
//     vertical mirroring is a row-index reflection done by the producer, horizontal mirroring
//     an index reflection when the edge warps read the ring;
//   * on the first level DC shift + forward RCT/ICT are applied to the samples as they leave
//     the ring; vertical analysis is a register lifting pipeline (NS state values per column),
//     horizontal analysis of each finished row runs in registers with warp shuffles, and the
//     de-interleaved halves go to the LL / HL / LH / HH windows with 128- or 64-bit stores.
// Forward order is vertical then horizontal, as in the reference (WaveletFwd.cpp:496-609).
// HBM traffic per level: every sample read once, every coefficient written once (the
// algorithmic bytes of SURVEY.md 8d); halo rows/columns are re-read from L2.
#include "dwt.cuh"
#include "bulk.cuh"
#include <stdlib.h>

namespace
{
constexpr int CW = 4; // consumer warps per CTA
constexpr int STAGES = 4; // row pairs in flight

template<class T, int V>
static __device__ __forceinline__ void lds_row(const int32_t* p, int32_t (&v)[V])
{
#pragma unroll
	for(int i = 0; i < V; i += 4)
	{
		const int4 r = *reinterpret_cast<const int4*>(p + i);
		v[i] = r.x;
		v[i + 1] = r.y;
		v[i + 2] = r.z;
		v[i + 3] = r.w;
	}
}

// Ingest stage on the V raw cells a lane holds of one source row.
// FIN: 0 cells are already of type T, 1 int32 image samples with DC shift (one component),
// 2 int32 image samples with DC shift + forward MCT over three components.
template<class K, int V, int NC, int FIN>
static __device__ __forceinline__ void ingest(const DwtLevelArgs& a, int cbase,
											   const int32_t (&raw)[NC][V], typename K::T (&x)[NC][V])
{
	typedef typename K::T T;
#pragma unroll
	for(int i = 0; i < V; ++i)
	{
		if constexpr(FIN == 0)
		{
#pragma unroll
			for(int c = 0; c < NC; ++c)
				x[c][i] = *reinterpret_cast<const T*>(&raw[c][i]);
		}
		else if constexpr(FIN == 1)
		{
#pragma unroll
			for(int c = 0; c < NC; ++c)
				x[c][i] = (T)(raw[c][i] - a.shift[cbase + c]);
		}
		else
		{
			const int32_t r = raw[0][i] - a.shift[0];
			const int32_t g = raw[NC > 1 ? 1 : 0][i] - a.shift[1];
			const int32_t b = raw[NC > 2 ? 2 : 0][i] - a.shift[2];
			if constexpr(K::NS == 4)
			{
				// forward ICT, point_transform/mct.cpp:147-219
				const float fr = (float)r, fg = (float)g, fb = (float)b;
				const float y = __fadd_rn(__fadd_rn(__fmul_rn(0.299f, fr), __fmul_rn(0.587f, fg)),
										  __fmul_rn(0.114f, fb));
				const float cb = 0.5f / (1.0f - 0.114f);
				const float cr = 0.5f / (1.0f - 0.299f);
				x[0][i] = y;
				x[NC > 1 ? 1 : 0][i] = __fmul_rn(cb, __fsub_rn(fb, y));
				x[NC > 2 ? 2 : 0][i] = __fmul_rn(cr, __fsub_rn(fr, y));
			}
			else
			{
				// forward RCT, point_transform/mct.cpp:99-146
				x[0][i] = (r + (g * 2) + b) >> 2;
				x[NC > 1 ? 1 : 0][i] = b - g;
				x[NC > 2 ? 2 : 0][i] = r - g;
			}
		}
	}
}

// horizontal analysis of one row held across the warp; x interleaved, V per lane
template<class K, int V>
static __device__ __forceinline__ void hanalysis_row(typename K::T (&x)[V],
													  typename K::T (&e)[V / 2],
													  typename K::T (&o)[V / 2])
{
	typedef typename K::T T;
	constexpr int H = V / 2;
#pragma unroll
	for(int j = 0; j < H; ++j)
	{
		e[j] = x[2 * j];
		o[j] = x[2 * j + 1];
	}
#pragma unroll
	for(int k = 0; k < K::NS; ++k)
	{
		if((k & 1) == 0)
		{
			const T ep1 = __shfl_down_sync(0xffffffffu, e[0], 1);
#pragma unroll
			for(int j = 0; j < H; ++j)
				o[j] = K::fupd(k, o[j], e[j], (j + 1 < H) ? e[j + 1] : ep1);
		}
		else
		{
			const T om1 = __shfl_up_sync(0xffffffffu, o[H - 1], 1);
#pragma unroll
			for(int j = 0; j < H; ++j)
				e[j] = K::fupd(k, e[j], j ? o[j - 1] : om1, o[j]);
		}
	}
#pragma unroll
	for(int j = 0; j < H; ++j)
	{
		e[j] = K::fscaleL(e[j]);
		o[j] = K::fscaleH(o[j]);
	}
}

template<class K, int V, int NC, int FIN>
__global__ void __launch_bounds__((CW + 1) * GRKB_WARP)
	fdwt_strip_kernel(const __grid_constant__ DwtLevelArgs a)
{
	typedef typename K::T T;
	constexpr int H = V / 2, NS = K::NS;
	constexpr int LEN = (CW * 30 + 2) * V; // staged samples of one row
	constexpr int STAGE = NC * 2 * LEN; // one stage: a row pair of NC components
	constexpr int NCOPY = NC * 2;
	extern __shared__ __align__(128) unsigned char smem_raw[];
	int32_t* ring = reinterpret_cast<int32_t*>(smem_raw); // [STAGES][STAGE]
	uint64_t* bars = reinterpret_cast<uint64_t*>(ring + STAGES * STAGE); // full[s], empty[s]

	const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int cbase = (NC == 1) ? (int)blockIdx.y : 0;
	const int py = (int)a.py, rw = (int)a.rw, rh = (int)a.rh, sw = (int)a.sw, sh = (int)a.sh;
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
	const int A = group * CW * 30 * V - V; // first staged column
	const int g0 = max(A, 0), g1 = min(A + LEN, rw); // staged columns that exist

	if(wid == CW)
	{
		// ---- producer warp: lane k issues copy k of every stage ------------------------------
		const uint32_t bytes = (uint32_t)(g1 - g0) * 4u;
		const int c = lane >> 1, r = lane & 1;
		const DwtPlane& p = a.pl[cbase + (lane < NCOPY ? c : 0)];
		const int32_t* full = (const int32_t*)p.full;
		const uint32_t dst0 = smem_u32(ring + (c * 2 + r) * LEN + (g0 - A));
		int s = 0;
		uint32_t ph = 0;
		for(int m = m0; m <= m1; ++m)
		{
			const uint32_t fullb = smem_u32(bars + 2 * s), empty = smem_u32(bars + 2 * s + 1);
			mbar_wait(empty, ph ^ 1u);
			if(lane < NCOPY)
			{
				// r = 0: odd row 2m-1, r = 1: even row 2m
				const int qm = grkb_mirror(2 * m - 1 + r, py, rh);
				const int32_t* src = full + (size_t)(qm - py) * p.f_stride + g0;
				mbar_arrive_tx(fullb, bytes);
				bulk_g2s(dst0 + (uint32_t)(s * STAGE * 4), src, bytes, fullb);
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
	const int li = (wid * 30 + lane) * V; // ring index of this lane's first sample
	const int q0 = A + li; // first source column of this lane
	const bool owner = lane >= 1 && lane <= 30 && q0 + V <= rw;
	const bool edge = (q0 - lane * V < 0) || (q0 + (32 - lane) * V > rw);
	int ix[V];
#pragma unroll
	for(int i = 0; i < V; ++i)
	{
		int n = q0 + i;
		n = n < 0 ? -n : (n >= rw ? 2 * (rw - 1) - n : n);
		n = min(max(n, g0), g1 - 1);
		ix[i] = n - A;
	}

	T st[NC][NS][V];
#pragma unroll
	for(int c = 0; c < NC; ++c)
#pragma unroll
		for(int k = 0; k < NS; ++k)
#pragma unroll
			for(int i = 0; i < V; ++i)
				st[c][k][i] = (T)0;

	// horizontal analysis + store of one vertically finished row (warp-uniform call sites:
	// every lane takes part in the shuffles)
	auto emit = [&](int qy, T(&x)[NC][V]) {
		const bool rowok = owner && qy >= py && qy < py + rh;
#pragma unroll
		for(int c = 0; c < NC; ++c)
		{
			T e[H], o[H];
			hanalysis_row<K, V>(x[c], e, o);
			if(!rowok)
				continue;
			const DwtPlane& p = a.pl[cbase + c];
			T* Lp;
			T* Hp;
			if((qy & 1) == 0)
			{
				const int r = qy / 2 - py;
				Lp = (T*)p.out_ll + (size_t)r * p.o_stride;
				Hp = (T*)p.out_bands + (size_t)r * p.ob_stride + sw;
			}
			else
			{
				const int r = (qy - 1) / 2;
				Lp = (T*)p.out_bands + (size_t)(sh + r) * p.ob_stride;
				Hp = Lp + sw;
			}
			if constexpr(H == 4)
			{
				st4(Lp + q0 / 2, e[0], e[1], e[H - 2], e[H - 1]);
				st4(Hp + q0 / 2, o[0], o[1], o[H - 2], o[H - 1]);
			}
			else
			{
				st2(Lp + q0 / 2, e[0], e[H - 1]);
				st2(Hp + q0 / 2, o[0], o[H - 1]);
			}
		}
	};

	int s = 0;
	uint32_t ph = 0;
#pragma unroll 1
	for(int m = m0; m <= m1; ++m)
	{
		const uint32_t fullb = smem_u32(bars + 2 * s), empty = smem_u32(bars + 2 * s + 1);
		mbar_wait(fullb, ph);
		const int32_t* stage = ring + s * STAGE;
		int32_t rp[NC][V], rq[NC][V]; // raw cells of the odd row 2m-1 and the even row 2m
#pragma unroll
		for(int c = 0; c < NC; ++c)
		{
			const int32_t* r0 = stage + (c * 2) * LEN;
			const int32_t* r1 = r0 + LEN;
			if(!edge)
			{
				lds_row<T, V>(r0 + li, rp[c]);
				lds_row<T, V>(r1 + li, rq[c]);
			}
			else
			{
#pragma unroll
				for(int i = 0; i < V; ++i)
				{
					rp[c][i] = r0[ix[i]];
					rq[c][i] = r1[ix[i]];
				}
			}
		}
		__syncwarp();
		if(lane == 0)
			mbar_arrive(empty);
		if(++s == STAGES)
		{
			s = 0;
			ph ^= 1u;
		}
		T cp[NC][V], cq[NC][V];
		ingest<K, V, NC, FIN>(a, cbase, rp, cp);
		ingest<K, V, NC, FIN>(a, cbase, rq, cq);
		T fo[NC][V], fe[NC][V]; // finished odd (vertical high) and even (vertical low) rows
#pragma unroll
		for(int c = 0; c < NC; ++c)
#pragma unroll
			for(int i = 0; i < V; ++i)
			{
				T an[NS + 1];
				an[0] = cq[c][i];
				an[1] = K::fupd(0, cp[c][i], st[c][0][i], cq[c][i]);
#pragma unroll
				for(int k = 1; k < NS; ++k)
					an[k + 1] = K::fupd(k, st[c][k - 1][i], st[c][k][i], an[k]);
				fo[c][i] = K::fscaleH(an[NS - 1]);
				fe[c][i] = K::fscaleL(an[NS]);
#pragma unroll
				for(int k = 0; k < NS; ++k)
					st[c][k][i] = an[k];
			}
		const int row_even = 2 * m - NS;
		const int row_odd = 2 * m - NS + 1;
		if(row_even >= ya && row_even < yb)
			emit(row_even, fe);
		if(row_odd >= ya && row_odd < yb)
			emit(row_odd, fo);
	}
}

template<class K, int V, int NC, int FIN>
static int launch_fstrip(DwtLevelArgs a, int nplanes, cudaStream_t stream)
{
	constexpr size_t smem =
		(size_t)STAGES * NC * 2 * (CW * 30 + 2) * V * 4 + (size_t)STAGES * 2 * sizeof(uint64_t);
	GRKB_ENSURE_SMEM((fdwt_strip_kernel<K, V, NC, FIN>), smem);
	const uint32_t strip = 30 * V;
	a.nstrips = (a.rw + strip - 1) / strip;
	// Chunk height: as in idwt_strip.cu -- all CTAs resident at once, grid sized to fill the
	// 148 SMs evenly, chunks no shorter than min_rows (every chunk re-runs NS row pairs of
	// vertical warm-up).
	static int waves = -1, min_rows = 0;
	if(waves < 0)
	{
		const char* e = getenv("GRKB_FDWT_WAVES"); // tuning knobs, see DESIGN.md section 4
		int w = e ? atoi(e) : 0; // 0: per kernel (below)
		if(w < 0 || w > 4)
			w = 0;
		e = getenv("GRKB_FDWT_MIN_ROWS");
		int mr = e ? atoi(e) : 4;
		min_rows = mr < 4 ? 4 : mr;
		waves = w;
	}
	const uint32_t total_rows = a.py + a.rh;
	const uint32_t ngroups = (a.nstrips + CW - 1) / CW;
	const uint32_t per_chunk = ngroups * ((NC == 1) ? (uint32_t)nplanes : 1u);
	const uint32_t nwaves = waves ? (uint32_t)waves : (K::NS == 4 ? 3u : 2u); // see idwt_strip.cu
	uint32_t nch = (148u * nwaves) / per_chunk;
	if(nch < 1)
		nch = 1;
	uint32_t ch = 2 * ((total_rows + 2 * nch - 1) / (2 * nch));
	if(ch < (uint32_t)min_rows)
		ch = (uint32_t)min_rows & ~1u;
	a.chunk_rows = ch;
	a.nchunks = (total_rows + ch - 1) / ch;
	dim3 grid(ngroups * a.nchunks, (NC == 1) ? nplanes : 1, 1);
	fdwt_strip_kernel<K, V, NC, FIN><<<grid, (CW + 1) * GRKB_WARP, smem, stream>>>(a);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
} // namespace

// Returns 1 when the level's geometry is not one this kernel takes (the caller then uses the
// general kernel of fdwt.cu), 0 on success, -1 on a CUDA error.
// fin: 0 raw, 1 DC shift of one component, 2 DC shift + forward MCT over three components
int grkb_launch_fdwt_strip(const DwtLevelArgs& a, int reversible, int ncomp, int fin,
							cudaStream_t stream)
{
	if(a.px != 0 || (a.rw % 8) != 0 || a.rw < 64 || a.rh < 2 || a.sw * 2 != a.rw)
		return 1;
	if(ncomp != 1 && ncomp != 3)
		return 1;
	for(int c = 0; c < ncomp; ++c)
	{
		const DwtPlane& p = a.pl[c];
		if((p.f_stride % 4) || (p.o_stride % 4) || (p.ob_stride % 4) || ((uintptr_t)p.full % 16) ||
		   ((uintptr_t)p.out_ll % 16) || ((uintptr_t)p.out_bands % 16))
			return 1;
	}
	if(ncomp == 3 && fin == 2)
		return reversible ? launch_fstrip<K53, 4, 3, 2>(a, 1, stream)
						  : launch_fstrip<K97, 4, 3, 2>(a, 1, stream);
	if(fin == 2)
		return 1;
	if(fin == 1)
	{
		if(ncomp != 1)
			return 1;
		return reversible ? launch_fstrip<K53, 8, 1, 1>(a, 1, stream)
						  : launch_fstrip<K97, 8, 1, 1>(a, 1, stream);
	}
	return reversible ? launch_fstrip<K53, 8, 1, 0>(a, ncomp, stream)
					  : launch_fstrip<K97, 8, 1, 0>(a, ncomp, stream);
}
