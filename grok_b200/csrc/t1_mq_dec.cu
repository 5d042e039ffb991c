// t1_mq_dec.cu -- EBCOT (JPEG 2000 Part 1) code-block decoder with dequantisation fused into
// the store, sm_100a.  One code block per warp, the MQ decoder replicated in every lane.
//
// Replaces, behind the C-ABI, the reference's
//   T1Part1::decompress            src/lib/jp2/t1/t1_part1/T1Part1.cpp:133-161
//   T1::decompress_cblk            src/lib/jp2/t1/t1_part1/T1.cpp:1365-1446
//   dec_clnpass / sigpass / refpass  T1.cpp:934-1364, MQ/RAW decoder mqc_dec_inl.h:25-180
//   ShiftFilter / ScaleFilter / Roi* src/lib/jp2/filters/PostDecompressFilters.h:7-89
//
// Shape: the MQ decoder is one serial dependency chain per block (every decision depends on
// a, c, ct and the adapted state of its context), so parallelism is across blocks and the
// kernel is bound by warp instructions per decision, not by memory.  A warp owns a block and
// all 32 lanes run the SAME decision chain on identical register copies of the coder state
// (an instruction costs one issue slot whether one lane or 32 execute it, and identical
// values never diverge).  What that buys: a 4-row stripe is walked in segments of 32 columns
// with lane l holding the state word of column l in a register, so everything that does not
// depend on the decision just decoded runs 32 columns at a time --
//   * which columns / rows have anything to code in this pass (one ballot per segment; empty
//     columns and rows cost nothing in the serial chain),
//   * the refinement pass's contexts (they are fixed before the pass starts) and its sample
//     updates (each lane applies the decoded bits to its own column, coalesced, no atomics),
//   * neighbour updates when a sample becomes significant (the three lanes concerned update
//     their registers; rows above / below are touched by three lanes at once),
// and the serial chain is left with: pick the next (column, row), one table look-up for the
// context, the MQ decision itself.  Samples live in the block's rectangle of the (L2-resident)
// Mallat coefficient plane.  Warps pull blocks from a global work counter, longest first.
#include "t1.cuh"
#include <mutex>
#include <stdio.h>
#include <stdlib.h>

namespace
{
constexpr uint32_t FULL = 0xffffffffu;

struct Mq
{
	uint32_t a; // the A register in the upper half (A << 16): "A < 0x8000" is a sign test
	// C is kept 64 bits wide: c is the reference's 32-bit C register, cl holds the bytes that the
	// reference would add later, already added at the place they will have shifted to.  Every
	// update of C is an addition or a left shift, so adding a byte early (t shifts early, t bit
	// positions lower) gives the same register; ct = shifts left until the next byte is due.
	uint32_t c, cl;
	int32_t ct;
	const uint8_t* p; // segment bytes
	uint32_t len, pos;
	uint32_t cur; // byte at pos (0xFF past the end: mqc_dec.cpp:78-90)
	uint32_t w; // bytes pos+1 .. pos+4, the first in the top byte, 0xFF past the end
};

static __device__ __forceinline__ uint32_t seg_byte(const Mq& q, uint32_t pos)
{
	return pos < q.len ? (uint32_t)__ldg(q.p + pos) : 0xFFu;
}

// bytes pos .. pos+3 of the segment, the first in the top byte, 0xFF past the end (two aligned
// words and a funnel shift; the byte pool has 16 bytes of slack behind it)
static __device__ __forceinline__ uint32_t seg_word(const Mq& q, uint32_t pos)
{
	const uintptr_t a = reinterpret_cast<uintptr_t>(q.p) + pos;
	const uint32_t* wp = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
	const uint32_t le = __funnelshift_r(__ldg(wp), __ldg(wp + 1), 8u * ((uint32_t)a & 3u));
	const int32_t left = (int32_t)q.len - (int32_t)pos; // real bytes from pos on
	const uint32_t pad = left >= 4 ? 0u : (left <= 0 ? 0xFFFFFFFFu : 0xFFFFFFFFu >> (8 * left));
	return __byte_perm(le, 0, 0x0123) | pad;
}

// bytein_dec_macro (mqc_dec_inl.h:82-108), run ahead of time: called with ct < 16, bytes are
// added until at least 32 shifts are covered (a renormalisation shifts by at most 15).  Four
// bytes go in at once when none of the bytes that decide the stuffing rule (the current one and
// the next three) is 0xFF -- all but ~2 % of the calls.
static __device__ __forceinline__ void mq_fill(Mq& q)
{
	{
		const uint32_t y = ~((q.cur << 24) | (q.w >> 8));
		if(((y - 0x01010101u) & ~y & 0x80808080u) == 0) // no zero byte in y
		{
			const uint64_t cc = (((uint64_t)q.c << 32) | q.cl) + ((uint64_t)q.w << (16 - q.ct));
			q.c = (uint32_t)(cc >> 32);
			q.cl = (uint32_t)cc;
			q.ct += 32;
			q.pos += 4;
			q.cur = q.w & 0xFFu;
			q.w = seg_word(q, q.pos + 1);
			return;
		}
	}
	uint32_t nxt = q.w >> 24;
	do
	{
		uint32_t add = nxt, at = 40, step = 8;
		bool adv = true;
		if(q.cur == 0xFF)
		{
			if(nxt > 0x8F)
			{
				add = 0xFF; // c += 0xFF00, the position does not advance
				adv = false;
			}
			else
			{
				at = 41; // c += nxt << 9, seven new bits
				step = 7;
			}
		}
		const uint64_t w = (uint64_t)add << (at - q.ct);
		const uint32_t lo = q.cl + (uint32_t)w;
		q.c += (uint32_t)(w >> 32) + (lo < q.cl ? 1u : 0u);
		q.cl = lo;
		q.ct += step;
		if(adv)
		{
			q.pos++;
			q.cur = nxt;
			nxt = seg_byte(q, q.pos + 1);
		}
	} while(q.ct <= 32);
	q.w = seg_word(q, q.pos + 1);
}

static __device__ __forceinline__ void mq_init(Mq& q, const uint8_t* p, uint32_t len)
{
	q.p = p;
	q.len = len;
	q.pos = 0;
	q.cur = seg_byte(q, 0);
	q.w = seg_word(q, 1);
	q.c = q.cur << 16;
	q.cl = 0;
	q.ct = 0;
	mq_fill(q);
	q.c = __funnelshift_l(q.cl, q.c, 7);
	q.cl <<= 7;
	q.ct -= 7;
	q.a = 0x80000000u;
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

// Coding tables in constant memory (every lane reads the same entry, so a look-up is one LDC):
// MQ states with the MPS in bit 31, the context tables hold context numbers times four, and the
// sign-coding table is indexed the way sc_index() below gathers its bits.
struct DecTables
{
	uint32_t mq[96]; // state 2*i+mps: Qe | next-if-MPS << 16 | next-if-LPS << 24 | mps << 31
	uint8_t zc[4 * 512];
	uint8_t sc[256]; // context x 4 | xor bit << 7
};
__constant__ DecTables c_dec;

// decompress_macro + renorm_dec_macro, mqc_dec_inl.h:111-158.  cxs[] holds, per context, the
// table entry of its current state, so a decision costs one shared-memory read, plus a constant
// read and a shared write when the state moves; cx4 is the context number times four.  Written
// without branches: a decision that needs no renormalisation shifts by clz(A) = 0.
// The per-warp context words are addressed in the shared window (32-bit addresses, no generic
// address arithmetic in the decision chain).
static __device__ __forceinline__ uint32_t lds32(uint32_t sa)
{
	uint32_t v;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(sa) : "memory");
	return v;
}
static __device__ __forceinline__ void sts32(uint32_t sa, uint32_t v)
{
	asm volatile("st.shared.u32 [%0], %1;" ::"r"(sa), "r"(v) : "memory");
}
static __device__ __forceinline__ uint32_t lds16(uint32_t sa)
{
	uint32_t v;
	asm volatile("{ .reg .u16 t; ld.shared.u16 t, [%1]; cvt.u32.u16 %0, t; }" : "=r"(v) : "r"(sa) : "memory");
	return v;
}
static __device__ __forceinline__ void sts16(uint32_t sa, uint32_t v)
{
	asm volatile("{ .reg .u16 t; cvt.u16.u32 t, %1; st.shared.u16 [%0], t; }" ::"r"(sa), "r"(v) : "memory");
}
static __device__ __forceinline__ uint32_t shift_amount(uint32_t a) // clz for a != 0, one FLO
{
	uint32_t n;
	asm("bfind.shiftamt.u32 %0, %1;" : "=r"(n) : "r"(a));
	return n;
}

static __device__ __forceinline__ uint32_t mq_decode(Mq& q, uint32_t cxs, uint32_t cx4)
{
	const uint32_t slot = cxs + cx4;
	const uint32_t e = lds32(slot);
	const uint32_t qe = e << 16;
	uint32_t a = q.a - qe;
	const bool lt = q.c < qe;
	// c < qe: the MPS/LPS sense is exchanged when a < qe; otherwise LPS exactly when a < qe
	// (which cannot happen while a >= 0x8000, Qe being at most 0x5601)
	const uint32_t lps = (uint32_t)lt ^ (uint32_t)(a < qe);
	q.c -= lt ? 0u : qe;
	a = lt ? qe : a;
	const uint32_t n = shift_amount(a); // shifts until the top bit of A is set
	if(n)
	{
		const uint32_t next4 = (e >> (lps ? 22u : 14u)) & 0x1FCu; // next state x 4
		sts32(slot, *reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(c_dec.mq) + next4));
	}
	q.a = a << n;
	q.c = __funnelshift_l(q.cl, q.c, n);
	q.cl <<= n;
	q.ct -= (int32_t)n;
	if(q.ct < 16)
		mq_fill(q);
	return (e >> 31) ^ lps;
}

constexpr uint32_t CX4_MAG = T1_CTX_MAG * 4, CX4_AGG = T1_CTX_AGG * 4, CX4_UNI = T1_CTX_UNI * 4;

struct Blk
{
	uint32_t* fl; // state words, (w+2) x (nstripes+2)
	int32_t* data; // the block's rectangle in the coefficient plane (x2 fixed point until dequantised)
	uint32_t ds; // row pitch of data
	uint32_t w, h, fs; // fs = w + 2
	uint32_t zc; // offset of this band orientation's zero-coding table in c_dec.zc
	bool vsc;
	uint32_t lane; // lane within the group that owns the block (LW lanes: the whole warp or half of it)
	uint32_t gmask, gsh; // the group's lanes in the warp, and the first of them
};

// warp primitives restricted to the LW lanes that share a code block
template<uint32_t LW>
static __device__ __forceinline__ uint32_t g_ballot(const Blk& b, bool pred)
{
	if(LW == 32)
		return __ballot_sync(FULL, pred);
	return (__ballot_sync(b.gmask, pred) & b.gmask) >> b.gsh;
}
template<uint32_t LW>
static __device__ __forceinline__ uint32_t g_shfl(const Blk& b, uint32_t v, uint32_t col)
{
	return LW == 32 ? __shfl_sync(FULL, v, col) : __shfl_sync(b.gmask, v, col, LW);
}
template<uint32_t LW>
static __device__ __forceinline__ void g_sync(const Blk& b)
{
	if(LW == 32)
		__syncwarp();
	else
		__syncwarp(b.gmask);
}

template<uint32_t LW>
static __device__ __forceinline__ void reset_states(const Blk& b, uint32_t cxs)
{
	// mqc_resetstates, mqc_dec.cpp:121-130
	for(uint32_t i = b.lane; i < T1_NCTX; i += LW)
	{
		const uint32_t st = i == T1_CTX_UNI ? 46u << 1 : i == T1_CTX_AGG ? 3u << 1 : i == T1_CTX_ZC ? 4u << 1 : 0u;
		sts32(cxs + i * 4, c_dec.mq[st]);
	}
	g_sync<LW>(b);
}

// per-row 4-bit summaries of a state word
static __device__ __forceinline__ uint32_t sig4(uint32_t f)
{
	return ((f >> 4) & 1u) | ((f >> 6) & 2u) | ((f >> 8) & 4u) | ((f >> 10) & 8u);
}
static __device__ __forceinline__ uint32_t nz4(uint32_t f)
{
	uint32_t m = 0;
#pragma unroll
	for(uint32_t r = 0; r < 4; ++r)
		m |= ((f >> (3 * r)) & T1_NEIGH) ? 1u << r : 0u;
	return m;
}

// index into the sign-coding table from the centre word f (row r) and its west/east words
static __device__ __forceinline__ uint32_t sc_index(uint32_t f, uint32_t fw, uint32_t fe, uint32_t r)
{
	// 3x3 neighbourhood of row r at bits 0..8 of f >> 3r: N, W, E, S are its bits 1, 3, 5, 7 and
	// stay there; the signs of N and S go to bits 0 and 2, those of W and E to bits 4 and 6
	return ((f >> (3 * r)) & 0xAAu) | ((f >> (18 + r)) & 5u) | ((fw >> (15 + r)) & 0x10u) |
		   ((fe >> (13 + r)) & 0x40u);
}
// the same index for the table as T1Tables::sc orders it (W E N S significance, then signs)
static inline uint32_t sc_index_tables(uint32_t j)
{
	return ((j >> 3) & 1u) | (((j >> 5) & 1u) << 1) | (((j >> 1) & 1u) << 2) | (((j >> 7) & 1u) << 3) |
		   (((j >> 4) & 1u) << 4) | (((j >> 6) & 1u) << 5) | ((j & 1u) << 6) | (((j >> 2) & 1u) << 7);
}

// One LW-column segment of a stripe: lane l holds the state word of column x0 + l in `f`.
struct Seg
{
	uint32_t* fseg; // state word of the segment's first column (this stripe)
	int32_t* dseg; // sample of the segment's first column, stripe row 0
};

// state words of the west / east neighbours of segment column `col`, as of now
template<uint32_t LW>
static __device__ __forceinline__ void west_east(const Blk& b, const Seg& g, uint32_t f, uint32_t col,
												   uint32_t& fw, uint32_t& fe)
{
	fw = g_shfl<LW>(b, f, (col + LW - 1u) & (LW - 1u));
	fe = g_shfl<LW>(b, f, (col + 1u) & (LW - 1u));
	if(col == 0 || col == LW - 1)
	{
		g_sync<LW>(b); // the word may have been updated by another lane
		if(col == 0)
			fw = g.fseg[-1];
		else
			fe = g.fseg[LW];
	}
}

// sample (segment column col, stripe row r) became significant with sign sgn and value val:
// propagate to the neighbours (update_flags_macro, T1.cpp:209-233).  All lanes call this.
template<uint32_t LW>
static __device__ __forceinline__ void mark_significant(const Blk& b, const Seg& g, uint32_t& f,
														 uint32_t col, uint32_t r, uint32_t sgn,
														 int32_t val)
{
	const uint32_t lane = b.lane;
	const uint32_t sh = 3 * r + 3; // row r of the 3x6 neighbourhood: [west, centre, east] at sh..sh+2
	// the column itself, and the columns west / east of it (we are their east / west neighbour)
	uint32_t add = lane == col ? (2u << sh) | (sgn << (18 + r + 1)) : 0u;
	add = lane + 1 == col ? 4u << sh : add;
	add = lane == col + 1 ? 1u << sh : add;
	f |= add;
	if(lane == col)
		g.dseg[lane + r * b.ds] = val;
	// columns outside the segment are in shared memory
	if(col == 0 || col == LW - 1)
	{
		if(lane == col)
		{
			uint32_t* o = col ? g.fseg + LW : g.fseg - 1;
			*o |= (col ? 1u : 4u) << sh;
		}
	}
	// row 4 of the stripe above / row -1 of the stripe below: lanes 0..2 take columns col-1..col+1
	if((r == 0 && !b.vsc) || r == 3)
	{
		g_sync<LW>(b); // successive events update overlapping words from different lanes
		if(lane < 3)
		{
			uint32_t* n = (r ? g.fseg + b.fs : g.fseg - b.fs) + col + lane - 1;
			// [west, centre, east] of row -1 at bits 0..2 and its sign at bit 18; row 4: 15 higher
			const uint32_t bits = (4u >> lane) | (lane == 1 ? sgn << (r ? 18 : 8) : 0u);
			*n |= bits << (r ? 0 : 15);
		}
	}
}

// sign of a newly significant sample (T1.cpp dec_*pass "sign" step) + the bookkeeping
template<uint32_t LW, bool RAW>
static __device__ __forceinline__ void code_sign(const Blk& b, const Seg& g, Mq& q, uint32_t cxs,
												  uint32_t& f, uint32_t& fc,
												  uint32_t col, uint32_t r, int32_t oph)
{
	uint32_t v;
	if(RAW)
		v = raw_decode(q);
	else
	{
		uint32_t fw, fe;
		west_east<LW>(b, g, f, col, fw, fe);
		const uint32_t e = c_dec.sc[sc_index(fc, fw, fe, r)];
		v = mq_decode(q, cxs, e & 0x7Fu) ^ (e >> 7);
	}
	mark_significant<LW>(b, g, f, col, r, v, v ? -oph : oph);
	fc |= T1_SIG_C(r) | (v << (18 + r + 1));
}

// dec_clnpass, T1.cpp:934-1093
template<uint32_t LW>
static __device__ void cleanup_pass(const Blk& b, Mq& q, uint32_t cxs,
									 int bpno_plus_one, bool segsym)
{
	const int32_t one = 1 << bpno_plus_one, oph = one | (one >> 1);
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
	{
		const uint32_t rows = min(4u, b.h - k), rowmask = (1u << rows) - 1u;
		for(uint32_t x0 = 0; x0 < b.w; x0 += LW)
		{
			Seg g;
			g.fseg = b.fl + (s + 1) * b.fs + 1 + x0;
			g.dseg = b.data + (size_t)k * b.ds + x0;
			const bool act = x0 + b.lane < b.w;
			g_sync<LW>(b);
			uint32_t f = act ? g.fseg[b.lane] : 0u;
			// rows still to code: not significant, not visited in this bit-plane
			const uint32_t todo = act ? ~(sig4(f) | (f >> 24)) & rowmask : 0u;
			uint32_t colmask = g_ballot<LW>(b, todo != 0);
			while(colmask)
			{
				const uint32_t col = __ffs(colmask) - 1;
				colmask &= colmask - 1;
				uint32_t m = g_shfl<LW>(b, todo, col);
				uint32_t fc = g_shfl<LW>(b, f, col);
				bool sign_only = false;
				if(rows == 4 && fc == 0)
				{
					if(!mq_decode(q, cxs, CX4_AGG))
						continue;
					uint32_t r = mq_decode(q, cxs, CX4_UNI);
					r = (r << 1) | mq_decode(q, cxs, CX4_UNI);
					m = (0xFu << r) & 0xFu;
					sign_only = true;
				}
				while(m)
				{
					const uint32_t r = __ffs(m) - 1;
					m &= m - 1;
					if(!sign_only && !mq_decode(q, cxs, c_dec.zc[b.zc + ((fc >> (3 * r)) & T1_NEIGH)]))
						continue;
					sign_only = false;
					code_sign<LW, false>(b, g, q, cxs, f, fc, col, r, oph);
				}
			}
			if(act)
				g.fseg[b.lane] = f & ~T1_PI_ALL;
		}
	}
	if(segsym)
	{
		// dec_clnpass_check_segsym, T1.cpp:1053-1073: the four decisions are consumed
		for(int i = 0; i < 4; ++i)
			(void)mq_decode(q, cxs, CX4_UNI);
	}
}

// dec_sigpass_mqc / dec_sigpass_raw, T1.cpp:1094-1245
template<uint32_t LW, bool RAW>
static __device__ void sig_pass(const Blk& b, Mq& q, uint32_t cxs,
								 int bpno_plus_one)
{
	const int32_t one = 1 << bpno_plus_one, oph = one | (one >> 1);
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
	{
		const uint32_t rows = min(4u, b.h - k), rowmask = (1u << rows) - 1u;
		for(uint32_t x0 = 0; x0 < b.w; x0 += LW)
		{
			Seg g;
			g.fseg = b.fl + (s + 1) * b.fs + 1 + x0;
			g.dseg = b.data + (size_t)k * b.ds + x0;
			const bool act = x0 + b.lane < b.w;
			g_sync<LW>(b);
			uint32_t f = act ? g.fseg[b.lane] : 0u;
			// free = not significant and not visited; a free row is coded in this pass as soon
			// as it has a significant neighbour -- which can happen during the pass
			const uint32_t fr = act ? ~(sig4(f) | (f >> 24)) & rowmask : 0u;
			uint32_t cand = fr & nz4(f);
			uint32_t colmask = g_ballot<LW>(b, cand != 0);
			while(colmask)
			{
				const uint32_t col = __ffs(colmask) - 1;
				colmask &= colmask - 1;
				uint32_t m = g_shfl<LW>(b, cand, col);
				const uint32_t frc = g_shfl<LW>(b, fr, col);
				uint32_t fc = g_shfl<LW>(b, f, col);
				uint32_t pi = 0, grown = 0;
				while(m)
				{
					const uint32_t r = __ffs(m) - 1;
					m &= m - 1;
					pi |= 1u << r;
					const uint32_t bit = RAW ? raw_decode(q)
											 : mq_decode(q, cxs, c_dec.zc[b.zc + ((fc >> (3 * r)) & T1_NEIGH)]);
					if(!bit)
						continue;
					code_sign<LW, RAW>(b, g, q, cxs, f, fc, col, r, oph);
					m |= (2u << r) & frc; // the row below becomes a candidate
					grown |= (7u << r) >> 1; // and rows r-1..r+1 of the next column
				}
				if(b.lane == col)
					f |= pi << 24;
				if(grown)
				{
					if(b.lane == col + 1)
						cand |= grown & fr;
					colmask |= g_ballot<LW>(b, cand != 0) & (2u << col);
				}
			}
			if(act)
				g.fseg[b.lane] = f;
		}
	}
}

constexpr uint32_t CXS_LIST = 128; // byte offset of the refinement list behind the context words
constexpr uint32_t CXS_BYTES = CXS_LIST + 128 * 2; // 19 context words (padded) + 4 x LW list entries

// dec_refpass_mqc / dec_refpass_raw, T1.cpp:1246-1364.  Everything about this pass is known
// before it starts -- which samples are refined (significant before this bit-plane) and with
// which context -- so the lanes write the contexts of a segment's decisions, in scan order, to a
// list; the decision chain walks the list and leaves the decoded bits in it; the lanes then
// apply them to their own columns.
template<uint32_t LW, bool RAW>
static __device__ void ref_pass(const Blk& b, Mq& q, uint32_t cxs,
								 int bpno_plus_one)
{
	const int32_t poshalf = (1 << bpno_plus_one) >> 1;
	const uint32_t list = cxs + CXS_LIST, below = (1u << b.lane) - 1u;
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
	{
		const uint32_t rows = min(4u, b.h - k), rowmask = (1u << rows) - 1u;
		for(uint32_t x0 = 0; x0 < b.w; x0 += LW)
		{
			Seg g;
			g.fseg = b.fl + (s + 1) * b.fs + 1 + x0;
			g.dseg = b.data + (size_t)k * b.ds + x0;
			const bool act = x0 + b.lane < b.w;
			g_sync<LW>(b);
			const uint32_t f = act ? g.fseg[b.lane] : 0u;
			const uint32_t todo = sig4(f) & ~(f >> 24) & rowmask;
			// position of this column's first decision = decisions of the columns before it
			const uint32_t cnt = __popc(todo);
			const uint32_t b0 = g_ballot<LW>(b, cnt & 1u), b1 = g_ballot<LW>(b, cnt & 2u),
						   b2 = g_ballot<LW>(b, cnt & 4u);
			const uint32_t total = __popc(b0) + 2 * __popc(b1) + 4 * __popc(b2);
			if(total == 0)
				continue;
			const uint32_t mine = list + 2 * (__popc(b0 & below) + 2 * __popc(b1 & below) + 4 * __popc(b2 & below));
			if(!RAW)
			{
				const uint32_t nz = nz4(f);
				uint32_t at = mine;
#pragma unroll
				for(uint32_t r = 0; r < 4; ++r)
					if((todo >> r) & 1u)
					{
						sts16(at, (f & T1_MU(r)) ? CX4_MAG + 8 : ((nz >> r) & 1u) ? CX4_MAG + 4 : CX4_MAG);
						at += 2;
					}
				g_sync<LW>(b);
			}
			for(uint32_t at = list, end = list + 2 * total; at != end; at += 2)
				sts16(at, RAW ? raw_decode(q) : mq_decode(q, cxs, lds16(at)));
			g_sync<LW>(b);
			if(todo)
			{
				uint32_t at = mine;
#pragma unroll
				for(uint32_t r = 0; r < 4; ++r)
					if((todo >> r) & 1u)
					{
						const uint32_t neg = (f >> (18 + r + 1)) & 1u;
						int32_t* d = g.dseg + b.lane + r * b.ds;
						*d += (lds16(at) ^ neg) ? poshalf : -poshalf;
						at += 2;
					}
				g.fseg[b.lane] = f | (todo << 28);
			}
		}
	}
}

// up to 8 warps per CTA, four CTAs per SM: 63 registers.  (More registers per thread save 12 % of the
// instructions and lose more than that to the lower occupancy: (256,3) measured 537 vs 578 MP/s.)

// LW = 32: one code block per warp.  LW = 16: each half of a warp owns a block and the two chains
// share the warp's instruction issue wherever their control flow agrees (both halves start their
// blocks together, so they are always in the same kind of pass).
template<uint32_t LW, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB)
	t1_mq_decode_kernel(const T1Launch l)
{
	constexpr uint32_t G = 32 / LW;
	extern __shared__ __align__(16) uint8_t smem[];
	const uint32_t warp = threadIdx.x >> 5, wlane = threadIdx.x & 31;
	const uint32_t grp = wlane / LW, lane = wlane % LW;
	const size_t per_slot = CXS_BYTES + (size_t)l.max_flag_words * 4;
	uint8_t* mine = smem + (warp * G + grp) * per_slot;
	const uint32_t cxs = (uint32_t)__cvta_generic_to_shared(mine); // context words + refinement list
	uint32_t* fl = reinterpret_cast<uint32_t*>(mine + CXS_BYTES);
	Blk b;
	b.fl = fl;
	b.ds = l.stride;
	b.lane = lane;
	b.gsh = grp * LW;
	b.gmask = LW == 32 ? FULL : ((1u << LW) - 1u) << b.gsh;
	if(LW != 32) // kept in a register: rematerialised at every ballot / shuffle / syncwarp otherwise
		asm volatile("mov.u32 %0, %0;" : "+r"(b.gmask));

	for(;;)
	{
		__syncwarp(); // the groups of a warp start their blocks together
		uint32_t item = 0;
		if(wlane == 0)
			item = atomicAdd(l.counter, G);
		item = __shfl_sync(FULL, item, 0);
		if(item >= l.nblocks)
			break;
		item += grp;
		if(item >= l.nblocks)
			continue;
		const uint32_t bi = l.order ? l.order[item] : item;
		const grkb_cblk cb = l.blocks[bi];
		if(cb.cblk_sty & GRKB_CBLKSTY_HT)
			continue;
		const uint32_t w = cb.w, h = cb.h;
		int32_t* blk_data = reinterpret_cast<int32_t*>(l.planes[cb.comp]) + (size_t)cb.y * l.stride + cb.x;
		const uint32_t fs = w + 2, nstripes = (h + 3) >> 2, nfl = fs * (nstripes + 2);
		const bool bad = (int)cb.numbps >= 31 - 6; // maxBitPlanesGRK, T1.cpp:1373-1377
		const bool coded = cb.numsegs != 0 && !bad;
		if(coded)
		{
			for(uint32_t i = lane; i < nfl; i += LW)
				fl[i] = 0;
			for(uint32_t y = 0; y < h; ++y)
				for(uint32_t x = lane; x < w; x += LW)
					blk_data[(size_t)y * l.stride + x] = 0;
			g_sync<LW>(b);
			b.data = blk_data;
			b.w = w;
			b.h = h;
			b.fs = fs;
			b.zc = 512u * cb.orient;
			b.vsc = (cb.cblk_sty & GRKB_CBLKSTY_VSC) != 0;
			Mq q;
			reset_states<LW>(b, cxs);
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
							sig_pass<LW, true>(b, q, cxs, bp1);
						else
							sig_pass<LW, false>(b, q, cxs, bp1);
					}
					else if(passtype == 1)
					{
						if(raw)
							ref_pass<LW, true>(b, q, cxs, bp1);
						else
							ref_pass<LW, false>(b, q, cxs, bp1);
					}
					else
						cleanup_pass<LW>(b, q, cxs, bp1, (cb.cblk_sty & GRKB_CBLKSTY_SEGSYM) != 0);
					if((cb.cblk_sty & GRKB_CBLKSTY_RESET) && !raw)
					{
						g_sync<LW>(b);
						reset_states<LW>(b, cxs);
					}
					if(++passtype == 3)
					{
						passtype = 0;
						--bp1;
					}
				}
			}
		}
		g_sync<LW>(b);
		if(bad && lane == 0)
			atomicAdd(l.status, 1u);
		// dequantise + store: ShiftFilter (/2, C truncation) or ScaleFilter (x stepsize/2),
		// Roi* variants when roishift != 0 (PostDecompressFilters.h:7-89)
		const int32_t thresh = cb.roishift ? (1 << cb.roishift) : 0x7FFFFFFF;
		for(uint32_t y = 0; y < h; ++y)
		{
			int32_t* dst = blk_data + (size_t)y * l.stride;
			for(uint32_t x = lane; x < w; x += LW)
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
	}
}

// c_dec is per device: filled once per device from the host tables
int upload_tables()
{
	static std::mutex mu;
	static bool done[64] = {};
	int dev = 0;
	if(cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
		return -1;
	std::lock_guard<std::mutex> lk(mu);
	if(done[dev])
		return 0;
	const T1Tables& h = grkb_t1_tables_host();
	static DecTables t;
	for(uint32_t i = 0; i < 96; ++i)
		t.mq[i] = i < 94 ? h.mq[i] | ((i & 1u) << 31) : 0u;
	for(uint32_t i = 0; i < 4 * 512; ++i)
		t.zc[i] = (uint8_t)(h.zc[i] * 4u);
	for(uint32_t j = 0; j < 256; ++j)
	{
		const uint32_t e = h.sc[sc_index_tables(j)];
		t.sc[j] = (uint8_t)((e & 0x80u) | ((e & 0x7Fu) * 4u));
	}
	if(cudaMemcpyToSymbol(c_dec, &t, sizeof(t)) != cudaSuccess)
		return -1;
	done[dev] = true;
	return 0;
}
} // namespace

int grkb_launch_t1_mq_decode_simt(const T1Launch& l, cudaStream_t s);

int grkb_launch_t1_mq_decode(const T1Launch& l, cudaStream_t s)
{
	if(l.nblocks == 0)
		return 0;
	// GRKB_T1_MQ_KERNEL=simt selects the one-block-per-lane kernel of t1_mq_dec_simt.cu (same
	// results; measured 8x slower on 64x64 blocks, where its 2.4 KB of state per lane leave room
	// for only two warps per SM -- see DESIGN.md section 4); the default is this file's
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
	if(upload_tables() != 0)
	{
		grkb_set_error("t1: cannot upload coding tables");
		return -1;
	}
	const size_t per_slot = CXS_BYTES + (size_t)l.max_flag_words * 4;
	int dev = 0, sms = 148;
	if(cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
	{
		grkb_set_error("t1: no current device");
		return -1;
	}
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	// Lanes per code block.  16 (two blocks per warp, the halves sharing the instruction issue where
	// their control flow agrees) executes 0.70x the warp instructions of 32 but needs twice the blocks
	// in flight for the same number of warps: it wins when the launch has enough blocks to fill the
	// SMs' shared memory with pairs one and a half times over (a 4K frame of 32x32 blocks: 606 -> 745
	// MP/s), and loses on a single 4K frame of 64x64 blocks (6 480 chains for 9 472 slots: 15.3 ->
	// 18.7 ms).  GRKB_T1_MQ_GROUP = 32 | 16 forces one.
	// (read at every launch so that the tests can exercise both)
	int force_lw = 0, max_warps_sm = 32;
	{
		const char* e = getenv("GRKB_T1_MQ_GROUP");
		force_lw = e ? (atoi(e) == 32 ? 32 : atoi(e) == 16 ? 16 : 0) : 0;
		e = getenv("GRKB_T1_MQ_WARPS");
		max_warps_sm = e ? atoi(e) : 32;
		if(max_warps_sm < 1 || max_warps_sm > 64)
			max_warps_sm = 32;
	}
	// warps per CTA: whatever packs most code blocks into the 227 KB of an SM
	auto pack = [&](int G, int max_nw, int& nw_out) {
		int best_nw = 0, best_total = 0;
		for(int nw = 1; nw <= max_nw; ++nw)
		{
			const size_t need = per_slot * G * nw;
			if(need > 227 * 1024)
				break;
			int per_sm = (int)((227 * 1024) / (need + 1024));
			if(per_sm > 32)
				per_sm = 32;
			if(nw * per_sm > max_warps_sm)
				per_sm = max_warps_sm / nw;
			if(nw * per_sm >= best_total)
			{
				best_total = nw * per_sm;
				best_nw = nw;
			}
		}
		nw_out = best_nw;
		return best_total;
	};
	// two blocks per warp come in two builds: 63 registers and four CTAs of up to eight warps (small
	// blocks: 32 warps per SM), or 80 registers and two CTAs of up to twelve (blocks whose state lets
	// fewer than 28 warps of pairs onto an SM anyway: fewer instructions for the same occupancy)
	int nw16 = 0, nw16w = 0, nw32 = 0;
	int total16 = pack(2, 8, nw16);
	const int total16w = pack(2, 12, nw16w), total32 = pack(1, 8, nw32);
	const bool wide = total16 < 28 && total16w > 0;
	if(wide)
	{
		total16 = total16w;
		nw16 = nw16w;
	}
	int lw = force_lw;
	if(lw == 0)
	{
		// enough blocks to fill the SMs with pairs one and a half times over; or other frames in flight
		// on this device (throughput over this frame's latency) and room for at least 20 warps of pairs
		const bool many = (uint64_t)l.nblocks * 2 >= (uint64_t)3 * sms * total16 * 2;
		static const uint32_t busy_at = [] {
			const char* e = getenv("GRKB_T1_MQ_BUSY");
			return e ? (uint32_t)atoi(e) : 1u;
		}();
		const bool busy = l.concurrent >= busy_at && total16 >= 20;
		lw = (total16 > 0 && (many || busy)) ? 16 : 32;
	}
	if(lw == 16 && total16 == 0)
		lw = 32;
	if(getenv("GRKB_T1_MQ_DEBUG"))
		fprintf(stderr, "t1_mq: %u blocks, %u others in flight, %d lanes per block%s\n", l.nblocks, l.concurrent, lw,
				lw == 16 && wide ? " (wide)" : "");
	const int G = 32 / lw;
	const int best_nw = lw == 16 ? nw16 : nw32, best_total = lw == 16 ? total16 : total32;
	if(best_nw == 0)
	{
		grkb_set_error("t1: code block too large for shared memory (%zu bytes)", per_slot * G);
		return -1;
	}
	const size_t smem = per_slot * G * best_nw;
	void (*kern)(const T1Launch);
	if(lw == 32)
	{
		kern = t1_mq_decode_kernel<32, 256, 4>;
		GRKB_ENSURE_SMEM((t1_mq_decode_kernel<32, 256, 4>), smem);
	}
	else if(!wide)
	{
		kern = t1_mq_decode_kernel<16, 256, 4>;
		GRKB_ENSURE_SMEM((t1_mq_decode_kernel<16, 256, 4>), smem);
	}
	else
	{
		kern = t1_mq_decode_kernel<16, 384, 2>;
		GRKB_ENSURE_SMEM((t1_mq_decode_kernel<16, 384, 2>), smem);
	}
	int resident = 0; // CTAs per SM once registers are counted too
	GRKB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, kern, best_nw * 32, smem));
	if(resident < 1)
		resident = 1;
	if(resident > best_total / best_nw)
		resident = best_total / best_nw;
	uint32_t ctas = (uint32_t)(sms * resident);
	const uint32_t need = (l.nblocks + best_nw * G - 1) / (best_nw * G);
	if(ctas > need)
		ctas = need;
	GRKB_CUDA_CHECK(cudaMemsetAsync(l.counter, 0, sizeof(uint32_t), s));
	kern<<<ctas, best_nw * 32, smem, s>>>(l);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
