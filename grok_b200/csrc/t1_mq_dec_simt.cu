// t1_mq_dec_simt.cu -- EBCOT (JPEG 2000 Part 1) code-block decoder, one code block per LANE.
//
// Same results as t1_mq_dec.cu (one block per warp, lane 0 decoding); replaces the same reference
// code: T1::decompress_cblk and its passes (src/lib/jp2/t1/t1_part1/T1.cpp:934-1446), the MQ/RAW
// decoder (t1/t1_part1/mqc_dec_inl.h:25-180) and the dequantisers
// (src/lib/jp2/filters/PostDecompressFilters.h:7-89).
//
// The MQ decoder is a serial chain per block, so a warp that decodes one block runs at 1/32 SIMT
// efficiency.  Here every lane owns a block and the warp runs one state machine in lockstep:
//   step A (lanes whose stripe column is finished): move to the next column -- slide a
//          three-column window over the block's state, rebuild the 3x6 neighbourhood word, find
//          the first sample of the column the current pass has to code;
//   step B (lanes with a pending decision): form the context of that decision (zero coding,
//          sign, refinement, run-length, uniform), decode ONE symbol, apply it, find the next.
// All lanes share the instruction stream of step B, which is where the time goes.
// Block state is non-redundant: 16 bits per stripe column (significance, sign, visited, refined
// of its four samples), interleaved across lanes in shared memory so that each lane's private
// array is bank-conflict free; the 3x6 neighbourhood word of t1.cuh is rebuilt per column from
// three such columns, so a sample that becomes significant touches only its own column.
// Whole-warp service sections (uniform control flow) hand out blocks from the global work
// counter, clear state and output rectangle, and dequantise + store a finished block.
#include "t1.cuh"

namespace
{
constexpr int SIMT_WARPS_MAX = 8;

enum
{
	K_NEXTCOL = 0,
	K_ZC,
	K_SC,
	K_MR,
	K_AGG,
	K_UNI1,
	K_UNI2,
	K_SEGSYM
};

struct Mq
{
	uint32_t a, c, ct;
	const uint8_t* p;
	uint32_t len, pos;
	uint32_t cur, nxt;
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

static __device__ __forceinline__ void mq_renorm(Mq& q)
{
	uint32_t n = __clz(q.a) - 16;
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

// decompress_macro, mqc_dec_inl.h:134-158; cx points at this lane's context state word
static __device__ __forceinline__ uint32_t mq_decode(Mq& q, uint32_t* cx, const uint32_t* mqt)
{
	uint32_t st = *cx;
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
		*cx = st;
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
			*cx = st;
			mq_renorm(q);
		}
	}
	return d;
}

// first row >= r of the column that the pass has to visit, or 4
static __device__ __forceinline__ uint32_t next_row(uint32_t passtype, uint32_t f, uint32_t r,
													 uint32_t rows)
{
	for(; r < rows; ++r)
	{
		const uint32_t sp = f & (T1_SIG_C(r) | T1_PI(r));
		if(passtype == 0)
		{
			if(!sp && ((f >> (3 * r)) & T1_NEIGH))
				break;
		}
		else if(passtype == 1)
		{
			if(sp == T1_SIG_C(r))
				break;
		}
		else if(!sp)
			break;
	}
	return r < rows ? r : 4u;
}

__global__ void __launch_bounds__(SIMT_WARPS_MAX* GRKB_WARP)
	t1_mq_decode_simt_kernel(const T1Launch l, const T1Tables* __restrict__ tab, uint32_t nw_state)
{
	extern __shared__ __align__(16) uint8_t smem[];
	uint32_t* mqt = reinterpret_cast<uint32_t*>(smem); // 94 (padded to 96)
	uint32_t* spread = mqt + 96; // 64: bit k of a 6-bit column vector -> bit 3k
	uint8_t* zc = reinterpret_cast<uint8_t*>(spread + 64);
	uint8_t* sc = zc + 2048;
	uint8_t* warp_base = sc + 256;
	for(uint32_t i = threadIdx.x; i < 94; i += blockDim.x)
		mqt[i] = tab->mq[i];
	for(uint32_t i = threadIdx.x; i < 64; i += blockDim.x)
	{
		uint32_t v = 0;
		for(uint32_t k = 0; k < 6; ++k)
			v |= ((i >> k) & 1u) << (3 * k);
		spread[i] = v;
	}
	for(uint32_t i = threadIdx.x; i < 2048; i += blockDim.x)
		zc[i] = tab->zc[i];
	for(uint32_t i = threadIdx.x; i < 256; i += blockDim.x)
		sc[i] = tab->sc[i];
	__syncthreads();

	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t per_warp = (size_t)T1_NCTX * 32 * 4 + (size_t)nw_state * 32 * 2;
	uint32_t* cxs = reinterpret_cast<uint32_t*>(warp_base + warp * per_warp) + lane; // [cx * 32]
	uint16_t* stw = reinterpret_cast<uint16_t*>(warp_base + warp * per_warp + T1_NCTX * 32 * 4);
	uint16_t* H = stw + lane; // this lane's state: H[idx * 32]

	// ---- per-lane block state --------------------------------------------------------------------
	bool have = false, finished = false, exhausted = false;
	grkb_cblk cb{};
	int32_t* blk_data = nullptr;
	uint32_t w = 0, h = 0, fs = 0, nstripes = 0;
	const uint8_t* zcb = zc;
	bool vsc = false;
	Mq q{};
	uint32_t seg_idx = 0, seg_left = 0, passtype = 2;
	const uint8_t* segp = nullptr;
	int bp1 = 0;
	bool raw = false;
	// position and column window
	uint32_t s = 0, x = 0, rows = 0, kind = K_NEXTCOL, r = 0, f = 0, hC = 0, runlen = 0, segsym = 0;
	uint32_t spW = 0, sgW = 0, spE = 0, sgE = 0, hE = 0; // spread significance / sign vectors
	bool fresh_stripe = true;

	auto reset_states = [&]() {
#pragma unroll
		for(int i = 0; i < T1_NCTX; ++i)
			cxs[i * 32] = 0;
		cxs[T1_CTX_UNI * 32] = 46 << 1;
		cxs[T1_CTX_AGG * 32] = 3 << 1;
		cxs[T1_CTX_ZC * 32] = 4 << 1;
	};
	// column vectors (significance / sign of rows -1..4) of column xx of stripe ss
	auto load_col = [&](uint32_t ss, uint32_t xx, uint32_t& hc, uint32_t& cv, uint32_t& sg) {
		const uint32_t idx = (ss + 1) * fs + xx + 1;
		const uint32_t ha = H[(idx - fs) * 32], hb = vsc ? 0u : (uint32_t)H[(idx + fs) * 32];
		hc = H[idx * 32];
		cv = ((ha >> 3) & 1u) | ((hc & 0xFu) << 1) | ((hb & 1u) << 5);
		sg = ((ha >> 7) & 1u) | (((hc >> 4) & 0xFu) << 1) | (((hb >> 4) & 1u) << 5);
	};
	// opens the next codeword segment with passes to run; false: the block is done
	auto next_segment = [&]() -> bool {
		while(seg_idx < cb.numsegs && bp1 >= 1)
		{
			const grkb_seg sg = l.segs[cb.seg_first + seg_idx];
			++seg_idx;
			raw = (bp1 <= (int)cb.numbps - 4) && passtype < 2 && (cb.cblk_sty & GRKB_CBLKSTY_LAZY);
			if(raw)
				raw_init(q, segp, sg.len);
			else
				mq_init(q, segp, sg.len);
			segp += sg.len;
			seg_left = sg.numpasses;
			if(seg_left)
				return true;
		}
		return false;
	};
	// end of a coding pass: false when the block is finished
	auto next_pass = [&]() -> bool {
		if((cb.cblk_sty & GRKB_CBLKSTY_RESET) && !raw)
			reset_states();
		if(++passtype == 3)
		{
			passtype = 0;
			--bp1;
		}
		--seg_left;
		if(bp1 < 1)
			return false;
		if(seg_left == 0 && !next_segment())
			return false;
		s = 0;
		x = 0xFFFFFFFFu; // so that the first advance lands on column 0
		fresh_stripe = true;
		return true;
	};

	for(;;)
	{
		// ---- service (uniform): finished blocks are dequantised, idle lanes get a new block ----
		uint32_t need = __ballot_sync(0xffffffffu, !have && !exhausted);
		while(need)
		{
			const int L = __ffs(need) - 1;
			need &= need - 1;
			if(__shfl_sync(0xffffffffu, finished ? 1 : 0, L))
			{
				// dequantise + store the block lane L has just decoded: ShiftFilter (/2, C
				// truncation) or ScaleFilter (x stepsize/2), Roi* variants
				// (PostDecompressFilters.h:7-89)
				const uint64_t ptr = __shfl_sync(0xffffffffu, (uint64_t)blk_data, L);
				const uint32_t bw = __shfl_sync(0xffffffffu, w, L), bh = __shfl_sync(0xffffffffu, h, L);
				const uint32_t roi = __shfl_sync(0xffffffffu, (uint32_t)cb.roishift, L);
				const float scale = __shfl_sync(0xffffffffu, cb.scale, L);
				int32_t* base = reinterpret_cast<int32_t*>(ptr);
				const int32_t thresh = roi ? (1 << roi) : 0x7FFFFFFF;
				__syncwarp();
				for(uint32_t y = 0; y < bh; ++y)
				{
					int32_t* dst = base + (size_t)y * l.stride;
					for(uint32_t xx = lane; xx < bw; xx += 32)
					{
						int32_t v = __ldcg(dst + xx); // written by another lane: read at L2
						int32_t mag = abs(v);
						if(mag >= thresh)
						{
							mag >>= roi;
							v = v < 0 ? -mag : mag;
						}
						if(l.reversible)
							dst[xx] = v / 2;
						else
							dst[xx] = __float_as_int(__fmul_rn((float)v, scale));
					}
				}
				if((int)lane == L)
					finished = false;
			}
			// hand lane L the next coded MQ block; blocks without data are zero-filled here
			for(;;)
			{
				uint32_t item = 0;
				if((int)lane == L)
					item = atomicAdd(l.counter, 1u);
				item = __shfl_sync(0xffffffffu, item, L);
				if(item >= l.nblocks)
				{
					if((int)lane == L)
						exhausted = true;
					break;
				}
				const uint32_t bi = l.order ? l.order[item] : item;
				const grkb_cblk nb = l.blocks[bi];
				if(nb.cblk_sty & GRKB_CBLKSTY_HT)
					continue;
				int32_t* data = reinterpret_cast<int32_t*>(l.planes[nb.comp]) + (size_t)nb.y * l.stride + nb.x;
				const bool bad = (int)nb.numbps >= 31 - 6; // maxBitPlanesGRK, T1.cpp:1373-1377
				for(uint32_t y = 0; y < nb.h; ++y)
					for(uint32_t xx = lane; xx < nb.w; xx += 32)
						data[(size_t)y * l.stride + xx] = 0;
				if(bad && (int)lane == L)
					atomicAdd(l.status, 1u);
				if(nb.numsegs == 0 || bad || nb.numbps == 0)
					continue;
				const uint32_t nfs = nb.w + 2u, nst = ((uint32_t)nb.h + 3) >> 2, nfl = nfs * (nst + 2);
				for(uint32_t i = lane; i < nfl; i += 32)
					stw[i * 32 + L] = 0;
				__syncwarp();
				if((int)lane == L)
				{
					cb = nb;
					blk_data = data;
					w = nb.w;
					h = nb.h;
					fs = nfs;
					nstripes = nst;
					zcb = zc + 512u * nb.orient;
					vsc = (nb.cblk_sty & GRKB_CBLKSTY_VSC) != 0;
					reset_states();
					bp1 = (int)nb.numbps;
					passtype = 2;
					seg_idx = 0;
					segp = l.bytes + nb.data_off;
					kind = K_NEXTCOL;
					s = 0;
					x = 0xFFFFFFFFu;
					fresh_stripe = true;
					if(next_segment())
						have = true;
					else
						finished = true; // nothing to run: goes through the dequantiser as zeros
				}
				if(!__shfl_sync(0xffffffffu, (have || finished) ? 1 : 0, L))
					continue;
				break;
			}
			// a lane that got a block without passes is served again in this loop
			if(__shfl_sync(0xffffffffu, (finished && !have) ? 1 : 0, L))
				need |= 1u << L;
		}
		if(__ballot_sync(0xffffffffu, have) == 0)
			break;

		if(have)
		{
			// ---- step A: next stripe column ---------------------------------------------------
			if(kind == K_NEXTCOL)
			{
				++x;
				if(x >= w)
				{
					x = 0;
					if(!fresh_stripe)
						++s;
					fresh_stripe = true;
				}
				if(s >= nstripes)
				{
					// end of pass
					if(passtype == 2 && (cb.cblk_sty & GRKB_CBLKSTY_SEGSYM) && segsym == 0)
					{
						segsym = 4; // dec_clnpass_check_segsym, T1.cpp:1053-1073
						kind = K_SEGSYM;
						s = nstripes;
						x = w; // stay put
					}
					else if(!next_pass())
					{
						have = false;
						finished = true;
					}
				}
				else
				{
					uint32_t cvC, sgC;
					if(fresh_stripe)
					{
						fresh_stripe = false;
						rows = min(4u, h - 4 * s);
						spW = 0;
						sgW = 0;
						load_col(s, 0, hC, cvC, sgC);
						cvC = spread[cvC];
					}
					else
					{
						cvC = spE;
						sgC = sgE;
						hC = hE;
					}
					uint32_t cvE;
					load_col(s, x + 1, hE, cvE, sgE);
					spE = spread[cvE];
					f = spW | (cvC << 1) | (spE << 2) | (sgC << 18) | ((hC >> 8) << 24);
					if(passtype == 2 && rows == 4 && f == 0)
						kind = K_AGG;
					else
					{
						r = next_row(passtype, f, 0, rows);
						if(r < 4)
							kind = passtype == 1 ? K_MR : K_ZC;
						else if(passtype == 2 && (f & T1_PI_ALL))
						{
							// cleanup clears the visited bits of every column
							f &= ~T1_PI_ALL;
							H[((s + 1) * fs + x + 1) * 32] = (uint16_t)(hC & ~0x0F00u);
							hC &= ~0x0F00u;
						}
						// window for the next column
						if(kind == K_NEXTCOL)
						{
							spW = (f >> 1) & 0x9249u;
							sgW = (f >> 18) & 0x3Fu;
						}
					}
				}
			}
			// ---- step B: one decision ---------------------------------------------------------
			if(kind != K_NEXTCOL && have)
			{
				uint32_t cx, xorbit = 0;
				if(kind == K_ZC)
					cx = zcb[(f >> (3 * r)) & T1_NEIGH];
				else if(kind == K_SC)
				{
					const uint32_t nb = f >> (3 * r);
					uint32_t i = ((nb >> 3) & 1u) | (((nb >> 5) & 1u) << 1) | (((nb >> 1) & 1u) << 2) |
								 (((nb >> 7) & 1u) << 3);
					i |= ((sgW >> (r + 1)) & 1u) << 4;
					i |= ((sgE >> (r + 1)) & 1u) << 5;
					i |= ((f >> (18 + r)) & 1u) << 6;
					i |= ((f >> (18 + r + 2)) & 1u) << 7;
					const uint32_t e = sc[i];
					cx = e & 0x7Fu;
					xorbit = e >> 7;
				}
				else if(kind == K_MR)
					cx = (f & T1_MU(r))					  ? T1_CTX_MAG + 2
						 : ((f >> (3 * r)) & T1_NEIGH) ? T1_CTX_MAG + 1
													   : T1_CTX_MAG;
				else if(kind == K_AGG)
					cx = T1_CTX_AGG;
				else
					cx = T1_CTX_UNI;
				uint32_t d;
				if(raw)
				{
					d = raw_decode(q);
					xorbit = 0;
				}
				else
					d = mq_decode(q, cxs + cx * 32, mqt);

				bool col_done = false;
				if(kind == K_ZC)
				{
					if(d)
						kind = K_SC;
					else
					{
						if(passtype == 0)
							f |= T1_PI(r);
						r = next_row(passtype, f, r + 1, rows);
						col_done = r >= 4;
					}
				}
				else if(kind == K_SC)
				{
					const uint32_t v = d ^ xorbit;
					const int32_t one = 1 << bp1, oph = one | (one >> 1);
					blk_data[(size_t)(4 * s + r) * l.stride + x] = v ? -oph : oph;
					f |= T1_SIG_C(r) | (v << (18 + r + 1));
					if(passtype == 0)
						f |= T1_PI(r);
					r = next_row(passtype, f, r + 1, rows);
					kind = K_ZC;
					col_done = r >= 4;
				}
				else if(kind == K_MR)
				{
					const int32_t poshalf = (1 << bp1) >> 1;
					const uint32_t neg = (f >> (18 + r + 1)) & 1u;
					atomicAdd(&blk_data[(size_t)(4 * s + r) * l.stride + x], (d ^ neg) ? poshalf : -poshalf);
					f |= T1_MU(r);
					r = next_row(1, f, r + 1, rows);
					col_done = r >= 4;
				}
				else if(kind == K_AGG)
				{
					if(d)
						kind = K_UNI1;
					else
						col_done = true;
				}
				else if(kind == K_UNI1)
				{
					runlen = d << 1;
					kind = K_UNI2;
				}
				else if(kind == K_UNI2)
				{
					r = runlen | d;
					kind = K_SC; // the sample at the end of the run is significant: sign only
				}
				else // K_SEGSYM
				{
					if(--segsym == 0)
					{
						kind = K_NEXTCOL;
						segsym = 0xFFFFFFFFu; // marks "consumed" for the end-of-pass test above
					}
				}
				if(col_done)
				{
					if(passtype == 2)
						f &= ~T1_PI_ALL;
					const uint32_t sig4 = ((f >> 4) & 1u) | ((f >> 6) & 2u) | ((f >> 8) & 4u) | ((f >> 10) & 8u);
					const uint32_t hn = sig4 | (((f >> 19) & 0xFu) << 4) | ((f >> 24) << 8);
					if(hn != hC)
						H[((s + 1) * fs + x + 1) * 32] = (uint16_t)hn;
					spW = (f >> 1) & 0x9249u;
					sgW = (f >> 18) & 0x3Fu;
					kind = K_NEXTCOL;
				}
			}
			if(segsym == 0xFFFFFFFFu && kind == K_NEXTCOL && have)
			{
				// the segmentation symbols are consumed: close the cleanup pass
				segsym = 0;
				if(!next_pass())
				{
					have = false;
					finished = true;
				}
			}
		}
	}
}
} // namespace

// Returns 1 when this kernel does not take the launch (state too large for shared memory), 0 on
// success, -1 on a CUDA error.
int grkb_launch_t1_mq_decode_simt(const T1Launch& l, cudaStream_t s)
{
	if(l.nblocks == 0)
		return 0;
	const T1Tables* tab = grkb_t1_tables_device();
	if(!tab)
	{
		grkb_set_error("t1: cannot upload coding tables");
		return -1;
	}
	const uint32_t nw_state = (l.max_flag_words + 1) & ~1u;
	const size_t per_warp = (size_t)T1_NCTX * 32 * 4 + (size_t)nw_state * 32 * 2;
	const size_t fixed = 96 * 4 + 64 * 4 + 2048 + 256;
	const size_t budget = 227 * 1024 - 1024;
	if(fixed + per_warp > budget)
		return 1;
	// one CTA per SM holding as many warps (32 blocks each) as fit
	int nw = (int)((budget - fixed) / per_warp);
	if(nw > SIMT_WARPS_MAX)
		nw = SIMT_WARPS_MAX;
	const size_t smem = fixed + per_warp * nw;
	GRKB_ENSURE_SMEM(t1_mq_decode_simt_kernel, smem);
	int dev = 0, sms = 148;
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	uint32_t ctas = (uint32_t)sms;
	const uint32_t need = (l.nblocks + nw * 32 - 1) / (nw * 32);
	if(ctas > need)
		ctas = need;
	GRKB_CUDA_CHECK(cudaMemsetAsync(l.counter, 0, sizeof(uint32_t), s));
	t1_mq_decode_simt_kernel<<<ctas, nw * 32, smem, s>>>(l, tab, nw_state);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
