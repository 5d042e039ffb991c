// t1_mq_dec.cu -- EBCOT (JPEG 2000 Part 1) code-block decoder with dequantisation fused into
// the store, sm_100a.  One code block per warp.
//
// Replaces, behind the C-ABI, the reference's
//   T1Part1::decompress            src/lib/jp2/t1/t1_part1/T1Part1.cpp:133-161
//   T1::decompress_cblk            src/lib/jp2/t1/t1_part1/T1.cpp:1365-1446
//   dec_clnpass / sigpass / refpass  T1.cpp:934-1364, MQ/RAW decoder mqc_dec_inl.h:25-180
//   ShiftFilter / ScaleFilter / Roi* src/lib/jp2/filters/PostDecompressFilters.h:7-89
//
// Shape: the MQ decoder is one serial dependency chain per block (every decision depends on
// a, c, ct and the adapted state of its context), so parallelism is across blocks.  A warp
// owns a block: all lanes clear and later dequantise + store the block (coalesced rows into
// the Mallat coefficient plane); lane 0 walks the coding passes over a shared-memory state
// array (one word per stripe column, layout in t1.cuh); samples are written straight into
// the block's rectangle of the (L2-resident) plane: a store when a sample becomes significant,
// a fire-and-forget red.add per refinement bit (the sign is kept in the state word).
// Warps pull blocks from a global work counter, longest blocks first.  Not HBM-bound: traffic
// is the compressed bytes in and 4 B/sample out; the bound is the serial decision rate.
#include "t1.cuh"
#include <stdlib.h>

namespace
{
struct Mq
{
	uint32_t a, c, ct;
	const uint8_t* p; // segment bytes
	uint32_t len, pos;
	uint32_t cur, nxt; // bytes at pos and pos+1 (0xFF past the end: mqc_dec.cpp:78-90)
};

static __device__ __forceinline__ uint32_t seg_byte(const Mq& q, uint32_t pos)
{
	return pos < q.len ? (uint32_t)__ldg(q.p + pos) : 0xFFu;
}

// bytein_dec_macro, mqc_dec_inl.h:82-108
static __device__ __forceinline__ void mq_bytein(Mq& q)
{
	if(q.cur == 0xFF)
	{
		if(q.nxt > 0x8F)
		{
			q.c += 0xFF00;
			q.ct = 8;
			return;
		}
		q.c += q.nxt << 9;
		q.ct = 7;
	}
	else
	{
		q.c += q.nxt << 8;
		q.ct = 8;
	}
	q.pos++;
	q.cur = q.nxt;
	q.nxt = seg_byte(q, q.pos + 1);
}

static __device__ __forceinline__ void mq_init(Mq& q, const uint8_t* p, uint32_t len)
{
	q.p = p;
	q.len = len;
	q.pos = 0;
	q.cur = seg_byte(q, 0);
	q.nxt = seg_byte(q, 1);
	q.c = q.cur << 16;
	mq_bytein(q);
	q.c <<= 7;
	q.ct -= 7;
	q.a = 0x8000;
}

static __device__ __forceinline__ void raw_init(Mq& q, const uint8_t* p, uint32_t len)
{
	q.p = p;
	q.len = len;
	q.pos = 0;
	q.c = 0;
	q.ct = 0;
}

// mqc_raw_decode, mqc_dec_inl.h:61-91
static __device__ __forceinline__ uint32_t raw_decode(Mq& q)
{
	if(q.ct == 0)
	{
		const uint32_t b = seg_byte(q, q.pos);
		if(q.c == 0xFF)
		{
			if(b > 0x8F)
			{
				q.c = 0xFF;
				q.ct = 8;
			}
			else
			{
				q.c = b;
				q.pos++;
				q.ct = 7;
			}
		}
		else
		{
			q.c = b;
			q.pos++;
			q.ct = 8;
		}
	}
	q.ct--;
	return (q.c >> q.ct) & 1u;
}

// renorm_dec_macro (mqc_dec_inl.h:111-121) with the shifts batched between byte fetches
static __device__ __forceinline__ void mq_renorm(Mq& q)
{
	uint32_t n = __clz(q.a) - 16; // shifts until bit 15 of a is set
	while(n)
	{
		if(q.ct == 0)
			mq_bytein(q);
		const uint32_t k = min(n, q.ct);
		q.a <<= k;
		q.c <<= k;
		q.ct -= k;
		n -= k;
	}
}

// decompress_macro, mqc_dec_inl.h:134-158
static __device__ __forceinline__ uint32_t mq_decode(Mq& q, uint8_t* cxs, const uint32_t* mqt,
													  uint32_t cx)
{
	uint32_t st = cxs[cx];
	const uint32_t e = mqt[st];
	const uint32_t qe = e & 0xFFFFu;
	uint32_t d = st & 1u;
	q.a -= qe;
	if(q.c < (qe << 16))
	{
		if(q.a < qe)
			st = (e >> 16) & 0xFFu;
		else
		{
			d ^= 1u;
			st = e >> 24;
		}
		q.a = qe;
		cxs[cx] = (uint8_t)st;
		mq_renorm(q);
	}
	else
	{
		q.c -= qe << 16;
		if(q.a < 0x8000u)
		{
			if(q.a < qe)
			{
				d ^= 1u;
				st = e >> 24;
			}
			else
				st = (e >> 16) & 0xFFu;
			cxs[cx] = (uint8_t)st;
			mq_renorm(q);
		}
	}
	return d;
}

static __device__ __forceinline__ void reset_states(uint8_t* cxs)
{
	// mqc_resetstates, mqc_dec.cpp:121-130
	for(int i = 0; i < T1_NCTX; ++i)
		cxs[i] = 0;
	cxs[T1_CTX_UNI] = 46 << 1;
	cxs[T1_CTX_AGG] = 3 << 1;
	cxs[T1_CTX_ZC] = 4 << 1;
}

struct Blk
{
	uint32_t* fl; // state words, (w+2) x (nstripes+2)
	int32_t* data; // the block's rectangle in the coefficient plane (x2 fixed point until dequantised)
	uint32_t ds; // row pitch of data
	uint32_t w, h, fs; // fs = w + 2
	const uint8_t* zc; // zero-coding table of this band orientation
	const uint8_t* sc;
	bool vsc;
};

// index into the sign-coding table from the centre word f (row r) and its west/east words
static __device__ __forceinline__ uint32_t sc_index(uint32_t f, uint32_t fw, uint32_t fe, uint32_t r)
{
	const uint32_t nb = f >> (3 * r); // 3x3 neighbourhood of row r at bits 0..8
	uint32_t i = ((nb >> 3) & 1u) | (((nb >> 5) & 1u) << 1) | (((nb >> 1) & 1u) << 2) |
				 (((nb >> 7) & 1u) << 3);
	i |= ((fw >> (18 + r + 1)) & 1u) << 4;
	i |= ((fe >> (18 + r + 1)) & 1u) << 5;
	i |= ((f >> (18 + r)) & 1u) << 6;
	i |= ((f >> (18 + r + 2)) & 1u) << 7;
	return i;
}

// sample (x, stripe row r) became significant with sign sgn: propagate to the neighbours
// (update_flags_macro, T1.cpp:209-233)
static __device__ __forceinline__ void mark_significant(const Blk& b, uint32_t* fp, uint32_t& f,
														 uint32_t r, uint32_t sgn)
{
	f |= T1_SIG_C(r) | (sgn << (18 + r + 1));
	fp[-1] |= 1u << (3 * (r + 1) + 2); // we are the east neighbour of the west column
	fp[1] |= 1u << (3 * (r + 1)); // and the west neighbour of the east column
	if(r == 0 && !b.vsc)
	{
		uint32_t* n = fp - b.fs; // row 4 of the stripe above
		n[0] |= (1u << 16) | (sgn << 23);
		n[-1] |= 1u << 17;
		n[1] |= 1u << 15;
	}
	if(r == 3)
	{
		uint32_t* s = fp + b.fs; // row -1 of the stripe below
		s[0] |= (1u << 1) | (sgn << 18);
		s[-1] |= 1u << 2;
		s[1] |= 1u << 0;
	}
}

// dec_clnpass, T1.cpp:934-1093
static __device__ void cleanup_pass(const Blk& b, Mq& q, uint8_t* cxs, const uint32_t* mqt,
									 int bpno_plus_one, bool segsym)
{
	const int32_t one = 1 << bpno_plus_one, oph = one | (one >> 1);
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
	{
		const uint32_t rows = min(4u, b.h - k);
		uint32_t* fp = b.fl + (s + 1) * b.fs + 1;
		int32_t* dp = b.data + (size_t)k * b.ds;
		for(uint32_t x = 0; x < b.w; ++x, ++fp, ++dp)
		{
			uint32_t f = *fp;
			uint32_t r = 0;
			bool sign_only = false;
			if(rows == 4 && f == 0)
			{
				if(!mq_decode(q, cxs, mqt, T1_CTX_AGG))
					continue;
				r = mq_decode(q, cxs, mqt, T1_CTX_UNI);
				r = (r << 1) | mq_decode(q, cxs, mqt, T1_CTX_UNI);
				sign_only = true;
			}
			for(; r < rows; ++r)
			{
				if(!sign_only)
				{
					if(f & (T1_SIG_C(r) | T1_PI(r)))
						continue;
					if(!mq_decode(q, cxs, mqt, b.zc[(f >> (3 * r)) & T1_NEIGH]))
						continue;
				}
				sign_only = false;
				const uint32_t e = b.sc[sc_index(f, fp[-1], fp[1], r)];
				const uint32_t v = mq_decode(q, cxs, mqt, e & 0x7Fu) ^ (e >> 7);
				dp[r * b.ds] = v ? -oph : oph;
				mark_significant(b, fp, f, r, v);
			}
			*fp = f & ~T1_PI_ALL;
		}
	}
	if(segsym)
	{
		// dec_clnpass_check_segsym, T1.cpp:1053-1073: the four decisions are consumed
		for(int i = 0; i < 4; ++i)
			(void)mq_decode(q, cxs, mqt, T1_CTX_UNI);
	}
}

// dec_sigpass_mqc / dec_sigpass_raw, T1.cpp:1094-1245
template<bool RAW>
static __device__ void sig_pass(const Blk& b, Mq& q, uint8_t* cxs, const uint32_t* mqt,
								 int bpno_plus_one)
{
	const int32_t one = 1 << bpno_plus_one, oph = one | (one >> 1);
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
	{
		const uint32_t rows = min(4u, b.h - k);
		uint32_t* fp = b.fl + (s + 1) * b.fs + 1;
		int32_t* dp = b.data + (size_t)k * b.ds;
		for(uint32_t x = 0; x < b.w; ++x, ++fp, ++dp)
		{
			uint32_t f = *fp;
			if(f == 0)
				continue;
			for(uint32_t r = 0; r < rows; ++r)
			{
				if((f & (T1_SIG_C(r) | T1_PI(r))) || !((f >> (3 * r)) & T1_NEIGH))
					continue;
				if(RAW)
				{
					if(raw_decode(q))
					{
						const uint32_t v = raw_decode(q);
						dp[r * b.ds] = v ? -oph : oph;
						mark_significant(b, fp, f, r, v);
					}
				}
				else if(mq_decode(q, cxs, mqt, b.zc[(f >> (3 * r)) & T1_NEIGH]))
				{
					const uint32_t e = b.sc[sc_index(f, fp[-1], fp[1], r)];
					const uint32_t v = mq_decode(q, cxs, mqt, e & 0x7Fu) ^ (e >> 7);
					dp[r * b.ds] = v ? -oph : oph;
					mark_significant(b, fp, f, r, v);
				}
				f |= T1_PI(r);
			}
			*fp = f;
		}
	}
}

// dec_refpass_mqc / dec_refpass_raw, T1.cpp:1246-1364
template<bool RAW>
static __device__ void ref_pass(const Blk& b, Mq& q, uint8_t* cxs, const uint32_t* mqt,
								 int bpno_plus_one)
{
	const int32_t poshalf = (1 << bpno_plus_one) >> 1;
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
	{
		const uint32_t rows = min(4u, b.h - k);
		uint32_t* fp = b.fl + (s + 1) * b.fs + 1;
		int32_t* dp = b.data + (size_t)k * b.ds;
		for(uint32_t x = 0; x < b.w; ++x, ++fp, ++dp)
		{
			uint32_t f = *fp;
			if(!(f & T1_SIGC_ALL))
				continue;
			for(uint32_t r = 0; r < rows; ++r)
			{
				if((f & (T1_SIG_C(r) | T1_PI(r))) != T1_SIG_C(r))
					continue;
				uint32_t v;
				if(RAW)
					v = raw_decode(q);
				else
				{
					const uint32_t cx = (f & T1_MU(r))					   ? T1_CTX_MAG + 2
										: ((f >> (3 * r)) & T1_NEIGH) ? T1_CTX_MAG + 1
																	   : T1_CTX_MAG;
					v = mq_decode(q, cxs, mqt, cx);
				}
				// the sign lives in the state word, so the update is a fire-and-forget reduction
				const uint32_t neg = (f >> (18 + r + 1)) & 1u;
				atomicAdd(&dp[r * b.ds], (v ^ neg) ? poshalf : -poshalf);
				f |= T1_MU(r);
			}
			*fp = f;
		}
	}
}

constexpr int MQ_MAX_THREADS = 256; // up to 8 warps = 8 code blocks in flight per CTA

__global__ void __launch_bounds__(MQ_MAX_THREADS)
	t1_mq_decode_kernel(const T1Launch l, const T1Tables* __restrict__ tab)
{
	extern __shared__ __align__(16) uint8_t smem[];
	// CTA-wide tables
	uint32_t* mqt = reinterpret_cast<uint32_t*>(smem);
	uint8_t* zc = smem + 94 * 4;
	uint8_t* sc = zc + 2048;
	uint8_t* warp_base = sc + 256;
	for(uint32_t i = threadIdx.x; i < 94; i += blockDim.x)
		mqt[i] = tab->mq[i];
	for(uint32_t i = threadIdx.x; i < 2048; i += blockDim.x)
		zc[i] = tab->zc[i];
	for(uint32_t i = threadIdx.x; i < 256; i += blockDim.x)
		sc[i] = tab->sc[i];
	__syncthreads();

	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t per_warp = 32 + (size_t)l.max_flag_words * 4;
	uint8_t* cxs = warp_base + warp * per_warp;
	uint32_t* fl = reinterpret_cast<uint32_t*>(cxs + 32);

	for(;;)
	{
		uint32_t item = 0;
		if(lane == 0)
			item = atomicAdd(l.counter, 1u);
		item = __shfl_sync(0xffffffffu, item, 0);
		if(item >= l.nblocks)
			break;
		const uint32_t bi = l.order ? l.order[item] : item;
		const grkb_cblk cb = l.blocks[bi];
		if(cb.cblk_sty & GRKB_CBLKSTY_HT)
			continue;
		const uint32_t w = cb.w, h = cb.h;
		int32_t* blk_data = reinterpret_cast<int32_t*>(l.planes[cb.comp]) + (size_t)cb.y * l.stride + cb.x;
		const uint32_t fs = w + 2, nstripes = (h + 3) >> 2, nfl = fs * (nstripes + 2);
		bool bad = (int)cb.numbps >= 31 - 6; // maxBitPlanesGRK, T1.cpp:1373-1377
		const bool coded = cb.numsegs != 0 && !bad;
		if(coded)
		{
			for(uint32_t i = lane; i < nfl; i += 32)
				fl[i] = 0;
			for(uint32_t y = 0; y < h; ++y)
				for(uint32_t x = lane; x < w; x += 32)
					blk_data[(size_t)y * l.stride + x] = 0;
		}
		__syncwarp();
		if(coded && lane == 0)
		{
			Blk b;
			b.fl = fl;
			b.data = blk_data;
			b.ds = l.stride;
			b.w = w;
			b.h = h;
			b.fs = fs;
			b.zc = zc + 512u * cb.orient;
			b.sc = sc;
			b.vsc = (cb.cblk_sty & GRKB_CBLKSTY_VSC) != 0;
			Mq q;
			reset_states(cxs);
			int bp1 = (int)cb.numbps;
			uint32_t passtype = 2;
			const uint8_t* p = l.bytes + cb.data_off;
			for(uint32_t s = 0; s < cb.numsegs; ++s)
			{
				const grkb_seg sg = l.segs[cb.seg_first + s];
				const bool raw = (bp1 <= (int)cb.numbps - 4) && passtype < 2 &&
								 (cb.cblk_sty & GRKB_CBLKSTY_LAZY);
				if(raw)
					raw_init(q, p, sg.len);
				else
					mq_init(q, p, sg.len);
				p += sg.len;
				for(uint32_t ps = 0; ps < sg.numpasses && bp1 >= 1; ++ps)
				{
					if(passtype == 0)
					{
						if(raw)
							sig_pass<true>(b, q, cxs, mqt, bp1);
						else
							sig_pass<false>(b, q, cxs, mqt, bp1);
					}
					else if(passtype == 1)
					{
						if(raw)
							ref_pass<true>(b, q, cxs, mqt, bp1);
						else
							ref_pass<false>(b, q, cxs, mqt, bp1);
					}
					else
						cleanup_pass(b, q, cxs, mqt, bp1, (cb.cblk_sty & GRKB_CBLKSTY_SEGSYM) != 0);
					if((cb.cblk_sty & GRKB_CBLKSTY_RESET) && !raw)
						reset_states(cxs);
					if(++passtype == 3)
					{
						passtype = 0;
						--bp1;
					}
				}
			}
		}
		__syncwarp();
		if(bad && lane == 0)
			atomicAdd(l.status, 1u);
		// dequantise + store: ShiftFilter (/2, C truncation) or ScaleFilter (x stepsize/2),
		// Roi* variants when roishift != 0 (PostDecompressFilters.h:7-89)
		const int32_t thresh = cb.roishift ? (1 << cb.roishift) : 0x7FFFFFFF;
		for(uint32_t y = 0; y < h; ++y)
		{
			int32_t* dst = blk_data + (size_t)y * l.stride;
			for(uint32_t x = lane; x < w; x += 32)
			{
				int32_t v = coded ? dst[x] : 0;
				int32_t mag = abs(v);
				if(mag >= thresh)
				{
					mag >>= cb.roishift;
					v = v < 0 ? -mag : mag;
				}
				if(l.reversible)
					dst[x] = v / 2;
				else
					dst[x] = __float_as_int(__fmul_rn((float)v, cb.scale));
			}
		}
		__syncwarp();
	}
}
} // namespace

int grkb_launch_t1_mq_decode_simt(const T1Launch& l, cudaStream_t s);

int grkb_launch_t1_mq_decode(const T1Launch& l, cudaStream_t s)
{
	if(l.nblocks == 0)
		return 0;
	// GRKB_T1_MQ_KERNEL=simt selects the one-block-per-lane kernel of t1_mq_dec_simt.cu (same
	// results; measured 3.5x slower on 64x64 blocks, where its 2.4 KB of state per lane leave
	// room for only two warps per SM -- see DESIGN.md section 4); the default is this file's
	// one-block-per-warp kernel
	static int simt = -1;
	if(simt < 0)
	{
		const char* e = getenv("GRKB_T1_MQ_KERNEL");
		simt = (e && e[0] == 's') ? 1 : 0;
	}
	if(simt)
	{
		const int rc = grkb_launch_t1_mq_decode_simt(l, s);
		if(rc <= 0)
			return rc;
	}
	const T1Tables* tab = grkb_t1_tables_device();
	if(!tab)
	{
		grkb_set_error("t1: cannot upload coding tables");
		return -1;
	}
	const size_t per_warp = 32 + (size_t)l.max_flag_words * 4;
	const size_t fixed = 94 * 4 + 2048 + 256;
	// warps per CTA: whatever packs most code blocks into the 227 KB of an SM
	int best_nw = 0, best_total = 0;
	for(int nw = 1; nw <= MQ_MAX_THREADS / 32; ++nw)
	{
		const size_t need = fixed + per_warp * nw;
		if(need > 227 * 1024)
			break;
		int per_sm = (int)((227 * 1024) / (need + 1024));
		if(per_sm > 32)
			per_sm = 32;
		if(nw * per_sm > 48)
			per_sm = 48 / nw;
		if(nw * per_sm >= best_total)
		{
			best_total = nw * per_sm;
			best_nw = nw;
		}
	}
	if(best_nw == 0)
	{
		grkb_set_error("t1: code block too large for shared memory (%zu bytes)", fixed + per_warp);
		return -1;
	}
	const size_t smem = fixed + per_warp * best_nw;
	static thread_local size_t configured = 0;
	if(smem > configured)
	{
		GRKB_CUDA_CHECK(cudaFuncSetAttribute(t1_mq_decode_kernel,
											 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		configured = smem;
	}
	int dev = 0, sms = 148;
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	uint32_t ctas = (uint32_t)(sms * (best_total / best_nw));
	const uint32_t need = (l.nblocks + best_nw - 1) / best_nw;
	if(ctas > need)
		ctas = need;
	GRKB_CUDA_CHECK(cudaMemsetAsync(l.counter, 0, sizeof(uint32_t), s));
	t1_mq_decode_kernel<<<ctas, best_nw * 32, smem, s>>>(l, tab);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
