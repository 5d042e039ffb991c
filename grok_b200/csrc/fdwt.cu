// fdwt.cu -- forward 5/3 and 9/7 wavelet transform, one resolution level per launch,
// vertical + horizontal analysis fused, optional DC shift + forward RCT/ICT fused into the
// load of the first level.  sm_100a.
//
// Replaces, behind the C-ABI, the reference's
//   WaveletFwdImpl::encode_procedure  src/lib/jp2/transform/WaveletFwd.cpp:432-616
//   dcLevelShiftCompress / mct_encode src/lib/jp2/tile/TileProcessor.cpp:506-573
//   mct::compress_rev / compress_irrev src/lib/jp2/point_transform/mct.cpp:99-219
//
// Same shape as idwt.cu, mirrored: a warp owns 30*V columns, streams down the rows of the
// source resolution two at a time (128-bit loads), runs the vertical analysis as a streaming
// lifting pipeline in registers, then the horizontal analysis of each finished row with warp
// shuffles, and writes the de-interleaved halves to the LL / HL / LH / HH windows with
// 128-bit (V=8) or 64-bit (V=4) stores.  Forward order is vertical then horizontal, as in
// the reference (WaveletFwd.cpp:496-609).
#include "dwt.cuh"
#include <stdlib.h>

namespace
{
constexpr int WARPS_PER_CTA = 4;

// FIN: 0 raw samples of type T, 1 int32 image samples with DC shift (single component),
// 2 int32 image samples with DC shift + forward MCT over three components.
template<class K, int V, int NC, int FIN, bool AL>
struct SrcLoader
{
	typedef typename K::T T;
	static __device__ __forceinline__ void load(const DwtLevelArgs& a, int qy, int q0, bool interior,
												 T (&x)[NC][V])
	{
		const int qm = grkb_mirror(qy, (int)a.py, (int)a.rh);
		const size_t row = (size_t)(qm - (int)a.py);
		int32_t raw[NC][V];
#pragma unroll
		for(int c = 0; c < NC; ++c)
		{
			const int32_t* src = (const int32_t*)a.pl[c].full + row * a.pl[c].f_stride;
			if(AL && interior)
			{
#pragma unroll
				for(int i = 0; i < V; i += 4)
				{
					const int4 r = __ldg(reinterpret_cast<const int4*>(src + q0 + i));
					raw[c][i] = r.x;
					raw[c][i + 1] = r.y;
					raw[c][i + 2] = r.z;
					raw[c][i + 3] = r.w;
				}
			}
			else
			{
#pragma unroll
				for(int i = 0; i < V; ++i)
					raw[c][i] = __ldg(src + (grkb_mirror(q0 + i, (int)a.px, (int)a.rw) - (int)a.px));
			}
		}
#pragma unroll
		for(int i = 0; i < V; ++i)
		{
			if(FIN == 0)
			{
#pragma unroll
				for(int c = 0; c < NC; ++c)
					x[c][i] = *reinterpret_cast<const T*>(&raw[c][i]);
			}
			else if(FIN == 1)
			{
#pragma unroll
				for(int c = 0; c < NC; ++c)
					x[c][i] = (T)(raw[c][i] - a.shift[c]);
			}
			else
			{
				const int32_t r = raw[0][i] - a.shift[0];
				const int32_t g = raw[NC > 1 ? 1 : 0][i] - a.shift[1];
				const int32_t b = raw[NC > 2 ? 2 : 0][i] - a.shift[2];
				if(K::NS == 4)
				{
					// forward ICT, point_transform/mct.cpp:147-219
					const float fr = (float)r, fg = (float)g, fb = (float)b;
					const float y = __fadd_rn(
						__fadd_rn(__fmul_rn(0.299f, fr), __fmul_rn(0.587f, fg)),
						__fmul_rn(0.114f, fb));
					const float cb = 0.5f / (1.0f - 0.114f);
					const float cr = 0.5f / (1.0f - 0.299f);
					const float u = __fmul_rn(cb, __fsub_rn(fb, y));
					const float v = __fmul_rn(cr, __fsub_rn(fr, y));
					x[0][i] = *reinterpret_cast<const T*>(&y);
					x[NC > 1 ? 1 : 0][i] = *reinterpret_cast<const T*>(&u);
					x[NC > 2 ? 2 : 0][i] = *reinterpret_cast<const T*>(&v);
				}
				else
				{
					// forward RCT, point_transform/mct.cpp:99-146
					const int32_t y = (r + (g * 2) + b) >> 2;
					const int32_t u = b - g;
					const int32_t v = r - g;
					x[0][i] = *reinterpret_cast<const T*>(&y);
					x[NC > 1 ? 1 : 0][i] = *reinterpret_cast<const T*>(&u);
					x[NC > 2 ? 2 : 0][i] = *reinterpret_cast<const T*>(&v);
				}
			}
		}
	}
};

// horizontal analysis of one row held across the warp; x interleaved, V per lane
template<class K, int V>
static __device__ __forceinline__ void hanalysis(typename K::T (&x)[V], typename K::T (&e)[V / 2],
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

template<class K, int V, int NC, int FIN, bool AL>
__global__ void __launch_bounds__(WARPS_PER_CTA* GRKB_WARP)
	fdwt_level_kernel(const __grid_constant__ DwtLevelArgs a)
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
	const int sw = (int)a.sw, sh = (int)a.sh;
	const bool interior = (q0 >= px) && (q0 + V <= px + rw);
	const bool owner = lane >= 1 && lane <= 30;
	const bool hnorm = rw > 1, vnorm = rh > 1;

	// ---- horizontal analysis + store of one vertically finished canvas row -------------------
	auto emit = [&](int qy, T(&x)[NC][V]) {
		const bool rowok = qy >= py && qy < py + rh;
#pragma unroll
		for(int c = 0; c < NC; ++c)
		{
			T e[H], o[H];
			if(hnorm)
				hanalysis<K, V>(x[c], e, o); // all lanes take part in the shuffles
			else
			{
#pragma unroll
				for(int j = 0; j < H; ++j)
				{
					e[j] = x[c][2 * j];
					o[j] = x[c][2 * j + 1];
					if(K::NS == 2 && px == 1)
						o[j] = (T)((int32_t)o[j] * 2); // WaveletFwd.cpp:934
				}
			}
			if(!owner || !rowok)
				continue;
			T* Lp;
			T* Hp;
			if((qy & 1) == 0)
			{
				const int r = qy / 2 - py;
				Lp = (T*)a.pl[c].out_ll + (size_t)r * a.pl[c].o_stride;
				Hp = (T*)a.pl[c].out_bands + (size_t)r * a.pl[c].ob_stride + sw;
			}
			else
			{
				const int r = (qy - 1) / 2;
				Lp = (T*)a.pl[c].out_bands + (size_t)(sh + r) * a.pl[c].ob_stride;
				Hp = Lp + sw;
			}
			if(AL && interior)
			{
				if(H == 4)
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
			else
			{
#pragma unroll
				for(int j = 0; j < H; ++j)
				{
					const int qe = q0 + 2 * j, qo = qe + 1;
					if(qe >= px && qe < px + rw)
						Lp[qe / 2 - px] = e[j];
					if(qo >= px && qo < px + rw)
						Hp[(qo - 1) / 2] = o[j];
				}
			}
		}
	};

	if(!vnorm)
	{
		if(chunk != 0)
			return;
		T x[NC][V];
		SrcLoader<K, V, NC, FIN, AL>::load(a, py, q0, interior, x);
		if(K::NS == 2 && py == 1)
		{
#pragma unroll
			for(int c = 0; c < NC; ++c)
#pragma unroll
				for(int i = 0; i < V; ++i)
					x[c][i] = (T)((int32_t)x[c][i] * 2); // WaveletFwd.cpp:650-659
		}
		emit(py, x);
		return;
	}

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

	T cp[NC][V], cq[NC][V]; // current pair: P = odd row 2m-1, Q = even row 2m
	SrcLoader<K, V, NC, FIN, AL>::load(a, 2 * m0 - 1, q0, interior, cp);
	SrcLoader<K, V, NC, FIN, AL>::load(a, 2 * m0, q0, interior, cq);
	for(int m = m0; m <= m1; ++m)
	{
		T np[NC][V], nq[NC][V];
		if(m < m1)
		{
			SrcLoader<K, V, NC, FIN, AL>::load(a, 2 * m + 1, q0, interior, np);
			SrcLoader<K, V, NC, FIN, AL>::load(a, 2 * m + 2, q0, interior, nq);
		}
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
		// warp-uniform conditions: every lane runs the shuffles inside emit()
		if(row_even >= ya && row_even < yb)
			emit(row_even, fe);
		if(row_odd >= ya && row_odd < yb)
			emit(row_odd, fo);
#pragma unroll
		for(int c = 0; c < NC; ++c)
#pragma unroll
			for(int i = 0; i < V; ++i)
			{
				cp[c][i] = np[c][i];
				cq[c][i] = nq[c][i];
			}
	}
}

template<class K, int V, int NC, int FIN>
static int launch_fdwt(const DwtLevelArgs& a, bool aligned, cudaStream_t stream)
{
	const uint32_t warps = a.nstrips * a.nchunks;
	const uint32_t ctas = (warps + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
	if(ctas == 0)
		return 0;
	if(aligned)
		fdwt_level_kernel<K, V, NC, FIN, true><<<ctas, WARPS_PER_CTA * GRKB_WARP, 0, stream>>>(a);
	else
		fdwt_level_kernel<K, V, NC, FIN, false><<<ctas, WARPS_PER_CTA * GRKB_WARP, 0, stream>>>(a);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
} // namespace

// fin: 0 raw, 1 DC shift of one component, 2 DC shift + forward MCT over three components
int grkb_launch_fdwt_strip(const DwtLevelArgs& a, int reversible, int ncomp, int fin,
							cudaStream_t stream);

int grkb_launch_fdwt_level(DwtLevelArgs a, int reversible, int ncomp, int fin, cudaStream_t stream)
{
	static const bool use_strip = []() {
		const char* e = getenv("GRKB_FDWT_STRIP"); // 0: general kernel only (comparison runs)
		return !(e && e[0] == '0');
	}();
	if(use_strip)
	{
		const int fast = grkb_launch_fdwt_strip(a, reversible, ncomp, fin, stream);
		if(fast <= 0)
			return fast;
	}
	const int V = (ncomp == 3) ? 4 : 8;
	const uint32_t H = V / 2;
	bool al = a.px == 0 && (a.sw % H) == 0;
	for(int c = 0; c < ncomp; ++c)
	{
		const DwtPlane& p = a.pl[c];
		al = al && (p.f_stride % 4) == 0 && (p.o_stride % 4) == 0 && (p.ob_stride % 4) == 0 &&
			 ((uintptr_t)p.full % 16) == 0 && ((uintptr_t)p.out_ll % 16) == 0 &&
			 ((uintptr_t)p.out_bands % 16) == 0;
	}
	const uint32_t strip = 30 * V;
	a.nstrips = (a.px + a.rw + strip - 1) / strip;
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
			return fin == 2 ? launch_fdwt<K53, 4, 3, 2>(a, al, stream)
							: launch_fdwt<K53, 4, 3, 0>(a, al, stream);
		return fin == 2 ? launch_fdwt<K97, 4, 3, 2>(a, al, stream)
						: launch_fdwt<K97, 4, 3, 0>(a, al, stream);
	}
	if(ncomp != 1)
		return -1;
	if(reversible)
		return fin == 1 ? launch_fdwt<K53, 8, 1, 1>(a, al, stream)
						: launch_fdwt<K53, 8, 1, 0>(a, al, stream);
	return fin == 1 ? launch_fdwt<K97, 8, 1, 1>(a, al, stream)
					: launch_fdwt<K97, 8, 1, 0>(a, al, stream);
}
