// t1_mq_enc.cu -- EBCOT (JPEG 2000 Part 1) code-block encoder with the quantiser fused into
// the load, sm_100a.  Two kernels: context modelling (one code block per warp) and the
// arithmetic coder (one code block per THREAD).
//
// Replaces, behind the C-ABI, the reference's
//   T1Part1::preCompress / compress   src/lib/jp2/t1/t1_part1/T1Part1.cpp:37-132
//   T1::compress_cblk                 src/lib/jp2/t1/t1_part1/T1.cpp:781-932
//   enc_sigpass / refpass / clnpass   T1.cpp:459-780, enc_is_term_pass :437-458
//   getwmsedec                        T1.cpp:418-436 (+ norms :239-277)
//   MQ / RAW encoder                  t1/t1_part1/mqc_enc.cpp:86-332, mqc_enc_inl.h:38-124
//
// Shape.  Unlike the decoder, the encoder knows every decision in advance, so only the
// arithmetic coder itself is a serial chain; the two halves have opposite shapes and are
// separate kernels with a symbol stream in global memory between them.
//
// t1_mq_model_kernel, one code block per warp:
//   * all lanes load the block's rows from the coefficient plane (coalesced), quantise to
//     sign-magnitude with 6 fractional bits into shared memory and reduce the block maximum;
//   * every coding pass walks the stripes in segments of 32 columns, one column per lane: the
//     lane finds which of its four samples the pass codes and their contexts as the
//     neighbourhood stands when the scan reaches them, and leaves its (context, decision)
//     symbols at its prefix-sum position of the block's symbol stream.  The membership of the
//     significance-propagation pass is recursive (the pass itself makes neighbours significant,
//     one column to the west or one row up) and is iterated to its fixed point with shuffles;
//     refinement and cleanup are known outright.  The state words (layout in t1.cuh) are
//     updated by all lanes at once afterwards;
//   * the stream is one byte per symbol in a chain of SYM_CHUNK-byte chunks taken from a
//     device pool with an atomic counter (the host sizes the pool per sample and retries with a
//     larger one if a frame runs it dry); the number of symbols of each pass is left in the
//     pass record, and the distortion estimate of a pass is a warp sum of table values
//     (integers, so the order of the additions does not matter) kept for the second kernel to
//     turn into the reference's double-precision figure.
//
// t1_mq_code_kernel, one code block per thread (32 coders per warp, in lockstep):
//   * reads its block's chain a word at a time with the next word prefetched, and runs the MQ
//     (or bypass) coder as one flat loop over symbols with the pass bookkeeping (termination,
//     rate, length, RESET / TERMALL / PTERM / SEGSYM) at the pass boundaries.  A branch taken
//     by any coder of the warp costs all of them, so the decision and byte-out are written
//     with selects; the context states of a thread are one column of a shared-memory array;
//   * bytes go into the block's slot in global memory with the byte under construction held
//     in a register (the coder's carry only ever touches that byte).
#include "t1.cuh"
#include <stdlib.h>

namespace
{
struct MqE
{
	uint32_t a, c, ct;
	uint8_t* out; // stream; out[-1] exists and is 0 (t1/Codeblock.h:156-172)
	int32_t bp; // index of the byte under construction, -1 at start
	uint32_t cur; // value of out[bp] (out[-1] = 0)
	int32_t cap;
	bool overflow;
};

static __device__ __forceinline__ uint32_t eget(const MqE& q, int32_t i)
{
	return i == q.bp ? q.cur : (i < 0 ? 0u : (uint32_t)q.out[i]);
}
// move the construction point to byte i (the current byte is written back first)
static __device__ __forceinline__ void eseek(MqE& q, int32_t i)
{
	if(q.bp >= 0)
		q.out[q.bp] = (uint8_t)q.cur;
	q.bp = i;
	q.cur = i < 0 ? 0u : (uint32_t)q.out[i];
}
// bp++ and start a new byte with value v
static __device__ __forceinline__ void epush(MqE& q, uint32_t v)
{
	if(q.bp >= 0)
		q.out[q.bp] = (uint8_t)q.cur;
	q.bp++;
	if(q.bp >= q.cap)
	{
		q.overflow = true;
		q.bp = q.cap - 1;
	}
	q.cur = v & 0xFFu;
}

// mqc_byteout, mqc_enc.cpp:86-125, its three cases in one branch-free form (the coders of a warp
// run in lockstep, one block per thread: a branch taken by any of them costs all of them): a
// carry out of C goes into the byte under construction unless that byte is 0xFF; the next byte
// takes 7 bits after a 0xFF, 8 otherwise
static __device__ __forceinline__ void mq_byteout(MqE& q)
{
	const uint32_t carry = q.cur != 0xFF ? (q.c >> 27) & 1u : 0u;
	q.cur += carry;
	q.c &= ~(carry << 27);
	const bool ff = q.cur == 0xFF;
	const uint32_t sh = ff ? 20u : 19u;
	const uint32_t nb = q.c >> sh;
	q.c &= (1u << sh) - 1u;
	q.ct = ff ? 7u : 8u;
	epush(q, nb);
}

// the shifts of a renormalisation that reach a byte boundary (about one decision in ten)
static __device__ __forceinline__ void mq_renorm_out(MqE& q, uint32_t n)
{
	while(n)
	{
		const uint32_t k = min(n, q.ct);
		q.c <<= k;
		q.ct -= k;
		n -= k;
		if(q.ct == 0)
			mq_byteout(q);
	}
}

// renorm of mqc_enc_inl.h:38-60 with the shifts batched between byte outputs: A is already
// shifted, C takes its n shifts, a byte leaving whenever ct reaches 0
static __device__ __forceinline__ void mq_renorm_c(MqE& q, uint32_t n)
{
	if(n < q.ct)
	{
		q.c <<= n;
		q.ct -= n;
		return;
	}
	mq_renorm_out(q, n);
}

// mqc_encode_macro, mqc_enc_inl.h:62-102.  cxs[cx] holds the table entry of the context's current
// state (qe | next-if-mps << 16 | next-if-lps << 24) with its MPS in bit 31, so a decision is one
// shared-memory read, plus a table read and a write when the state moves.  One formulation for
// the three cases of the reference: C takes Qe unless the (conditional) exchange gives A = Qe;
// the renormalisation shift is clz(A) - 16 = 0 for an MPS that leaves A >= 0x8000.
template<uint32_t STRIDE>
static __device__ __forceinline__ void mq_encode(MqE& q, uint32_t* cxs, const uint32_t* mqt,
												  uint32_t cx, uint32_t d)
{
	cxs += cx * STRIDE; // context words of one block are STRIDE words apart (one thread per block)
	cx = 0;
	const uint32_t e = cxs[cx];
	const uint32_t qe = e & 0xFFFFu;
	uint32_t a = q.a - qe;
	const bool ismps = d == (e >> 31);
	const bool lt = a < qe;
	if(ismps != lt)
		q.c += qe;
	else
		a = qe;
	const uint32_t n = __clz(a) - 16;
	if(n)
	{
		const uint32_t nx = (e >> (ismps ? 16u : 24u)) & 0x7Fu;
		cxs[cx] = mqt[nx];
		mq_renorm_c(q, n);
	}
	q.a = a << n;
}

// mqc_flush_enc, mqc_enc.cpp:213-227
static __device__ void mq_flush(MqE& q)
{
	const uint32_t tempc = q.c + q.a;
	q.c |= 0xFFFF;
	if(q.c >= tempc)
		q.c -= 0x8000;
	q.c <<= q.ct;
	mq_byteout(q);
	q.c <<= q.ct;
	mq_byteout(q);
	if(q.cur != 0xFF)
		eseek(q, min(q.bp + 1, q.cap - 1)); // bp++: the byte there is not part of the stream
}

#define BYPASS_CT_INIT 0xDEADBEEFu

// mqc_bypass_init_enc, mqc_enc.cpp:229-245
static __device__ __forceinline__ void raw_init(MqE& q)
{
	q.c = 0;
	q.ct = BYPASS_CT_INIT;
}

// mqc_bypass_enc_macro, mqc_enc_inl.h:104-124
static __device__ __forceinline__ void raw_encode(MqE& q, uint32_t d)
{
	if(q.ct == BYPASS_CT_INIT)
		q.ct = 8;
	q.ct--;
	q.c += d << q.ct;
	if(q.ct == 0)
	{
		q.cur = q.c & 0xFFu;
		q.ct = (q.cur == 0xFF) ? 7 : 8;
		eseek(q, min(q.bp + 1, q.cap - 1));
		if(q.bp >= q.cap - 1)
			q.overflow = true;
		q.c = 0;
	}
}

static __device__ __forceinline__ bool raw_needs_two(const MqE& q, bool erterm)
{
	return q.ct < 7 || (q.ct == 7 && (erterm || eget(q, q.bp - 1) != 0xFF));
}

// mqc_bypass_flush_enc, mqc_enc.cpp:252-292
static __device__ void raw_flush(MqE& q, bool erterm)
{
	if(raw_needs_two(q, erterm))
	{
		uint32_t bit = 0;
		while(q.ct > 0)
		{
			q.ct--;
			q.c += bit << q.ct;
			bit = 1 - bit;
		}
		q.cur = q.c & 0xFFu;
		eseek(q, min(q.bp + 1, q.cap - 1));
	}
	else if(q.ct == 7 && eget(q, q.bp - 1) == 0xFF)
		eseek(q, q.bp - 1);
	else if(q.ct == 8 && !erterm && eget(q, q.bp - 1) == 0x7F && eget(q, q.bp - 2) == 0xFF)
		eseek(q, q.bp - 2);
}

// mqc_restart_init_enc, mqc_enc.cpp:294-310
static __device__ __forceinline__ void mq_restart(MqE& q)
{
	q.a = 0x8000;
	q.c = 0;
	q.ct = 12;
	eseek(q, q.bp - 1);
	if(q.cur == 0xFF)
		q.ct = 13;
}

// mqc_erterm_enc, mqc_enc.cpp:312-325
static __device__ void mq_erterm(MqE& q)
{
	int32_t k = (int32_t)(11 - q.ct + 1);
	while(k > 0)
	{
		q.c <<= q.ct;
		q.ct = 0;
		mq_byteout(q);
		k -= (int32_t)q.ct;
	}
	if(q.cur != 0xFF)
		mq_byteout(q);
}

template<uint32_t STRIDE>
static __device__ __forceinline__ void reset_states(uint32_t* cxs, const uint32_t* mqt)
{
	// mqc_resetstates, mqc_dec.cpp:121-130 (shared by the encoder); all MPS are 0
	for(int i = 0; i < T1_NCTX; ++i)
		cxs[i * STRIDE] = mqt[0];
	cxs[T1_CTX_UNI * STRIDE] = mqt[46 << 1];
	cxs[T1_CTX_AGG * STRIDE] = mqt[3 << 1];
	cxs[T1_CTX_ZC * STRIDE] = mqt[4 << 1];
}

struct EBlk
{
	uint32_t* fl; // state words, (w+2) x (nstripes+2)
	const uint32_t* smp; // sign-magnitude samples, 6 fractional bits: [stripe][column][row of the stripe]
	uint32_t w, h, fs;
	const uint8_t* zc; // zero-coding table of this band orientation
	const uint8_t* sc;
	const int16_t* nms; // distortion tables [4][128]
	bool vsc;
};

// getnmsedec_sig / getnmsedec_ref, T1.cpp:194-208
static __device__ __forceinline__ int32_t nmsedec_sig(const EBlk& b, uint32_t mag, uint32_t bitpos)
{
	return bitpos > 0 ? b.nms[(mag >> bitpos) & 127] : b.nms[128 + (mag & 127)];
}
static __device__ __forceinline__ int32_t nmsedec_ref(const EBlk& b, uint32_t mag, uint32_t bitpos)
{
	return bitpos > 0 ? b.nms[256 + ((mag >> bitpos) & 127)] : b.nms[384 + (mag & 127)];
}

// ---- coding passes: context modelling by the whole warp, MQ coding by lane 0 ------------------
// The encoder knows every decision in advance, so only the arithmetic coder itself is serial.
// A stripe is walked in segments of 32 columns, lane l owning column x0 + l: the lane derives
// which of its four samples the pass codes, their contexts as the neighbourhood stands when the
// scan reaches them (state before the segment, plus the west column's and its own earlier rows'
// new significance of this pass), and leaves its (context, decision) symbols at its prefix-sum
// position of a list in shared memory; lane 0 then feeds the list to the MQ (or bypass) coder.
// The one recursive part is the membership of the significance-propagation pass (a sample is in
// the pass if a neighbour is significant, which the pass itself can cause one column to the
// west or one row up): lanes iterate it to its fixed point, usually in one or two rounds.
constexpr uint32_t CXS_BYTES = 80; // 19 context words, padded
constexpr uint32_t SYM_CAP = 320; // at most 10 symbols per column (run-length column of the cleanup pass)

// bit r of a row nibble -> bit 3(r+1)+c of the state word (c: 0 west, 1 centre, 2 east)
static __device__ __forceinline__ uint32_t spread4(uint32_t nib, uint32_t c)
{
	// b0 b1 . . . . b2 b3, then every bit to three times its index (two multiply-and-mask steps
	// instead of four shift-and-mask pairs: this runs several times per stripe column and pass)
	const uint32_t t = (nib * 17u) & 0xC3u;
	return ((t * 5u) & 0x249u) << (3u + c);
}
static __device__ __forceinline__ uint32_t sig4(uint32_t f)
{
	uint32_t g = (f >> 4) & 0x249u; // bits 0, 3, 6, 9
	g = (g | (g >> 2)) & 0xC3u;
	return (g | (g >> 4)) & 0xFu;
}
static __device__ __forceinline__ uint32_t nz4(uint32_t f)
{
	uint32_t m = 0;
#pragma unroll
	for(uint32_t r = 0; r < 4; ++r)
		m |= ((f >> (3 * r)) & T1_NEIGH) ? 1u << r : 0u;
	return m;
}
static __device__ __forceinline__ uint32_t warp_excl_scan(uint32_t v, uint32_t lane, uint32_t& total)
{
	uint32_t incl = v;
#pragma unroll
	for(int d = 1; d < 32; d <<= 1)
	{
		const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
		if((int)lane >= d)
			incl += t;
	}
	total = __shfl_sync(0xffffffffu, incl, 31);
	return incl - v;
}

// one stripe segment as a lane sees it
struct Col
{
	uint32_t* fp; // own state word
	uint32_t f; // its value before the segment is coded
	uint32_t mag[4], sgn4, bit4; // magnitudes, signs and the bits of this plane, rows 0..3
	uint32_t sgnL, sgnR; // sign nibbles of the west / east columns
	uint32_t rowmask;
	bool act;
};

static __device__ __forceinline__ Col load_col(const EBlk& b, uint32_t s, uint32_t k, uint32_t x0,
												uint32_t lane, uint32_t one)
{
	Col c;
	const uint32_t x = x0 + lane;
	const uint32_t rows = min(4u, b.h - k);
	c.act = x < b.w;
	c.rowmask = c.act ? (1u << rows) - 1u : 0u;
	c.fp = b.fl + (s + 1) * b.fs + 1 + x;
	c.f = c.act ? *c.fp : 0u;
	c.sgn4 = c.bit4 = 0;
	const uint4 q4 = c.act ? reinterpret_cast<const uint4*>(b.smp)[s * b.w + x] : make_uint4(0, 0, 0, 0);
	const uint32_t sm4[4] = {q4.x, q4.y, q4.z, q4.w};
#pragma unroll
	for(uint32_t r = 0; r < 4; ++r)
	{
		const uint32_t sm = sm4[r];
		c.mag[r] = sm & 0x7FFFFFFFu;
		c.sgn4 |= (sm >> 31) << r;
		c.bit4 |= ((sm & one) ? 1u : 0u) << r;
	}
	c.sgnL = __shfl_up_sync(0xffffffffu, c.sgn4, 1);
	c.sgnR = __shfl_down_sync(0xffffffffu, c.sgn4, 1);
	if(lane == 0 || lane == 31)
	{
		// the neighbour column belongs to another segment
		const bool west = lane == 0;
		const bool there = west ? x0 > 0 : x + 1 < b.w;
		uint32_t sg = 0;
		if(there && c.act)
		{
			const uint4 n4 = reinterpret_cast<const uint4*>(b.smp)[s * b.w + (west ? x - 1 : x + 1)];
			sg = (n4.x >> 31) | ((n4.y >> 31) << 1) | ((n4.z >> 31) << 2) | ((n4.w >> 31) << 3);
		}
		if(west)
			c.sgnL = sg;
		else
			c.sgnR = sg;
	}
	return c;
}

// sign-coding symbol of row r with the neighbourhood word fv as it stands at the visit
static __device__ __forceinline__ uint32_t sign_symbol(const EBlk& b, const Col& c, uint32_t fv, uint32_t r)
{
	const uint32_t nb = fv >> (3 * r);
	const uint32_t sw = (nb >> 3) & 1u, se = (nb >> 5) & 1u;
	uint32_t i = sw | (se << 1) | (((nb >> 1) & 1u) << 2) | (((nb >> 7) & 1u) << 3);
	i |= (sw & (c.sgnL >> r)) << 4;
	i |= (se & (c.sgnR >> r) & 1u) << 5;
	i |= ((fv >> (18 + r)) & 1u) << 6;
	i |= ((fv >> (18 + r + 2)) & 1u) << 7;
	i &= 0xFFu;
	const uint32_t e = b.sc[i];
	return (e & 0x7Fu) | ((((c.sgn4 >> r) & 1u) ^ (e >> 7)) << 5);
}

// the samples of this lane that turned significant (nibble ns): own word, the words of the
// stripes above and below, and -- for the lanes at the ends of the segment -- the neighbour
// column of the other segment (update_flags_macro, T1.cpp:209-233, for all of them at once)
static __device__ __forceinline__ void commit_significance(const EBlk& b, const Col& c, uint32_t f,
															uint32_t ns, uint32_t lane)
{
	uint32_t nsL = __shfl_up_sync(0xffffffffu, ns, 1), nsR = __shfl_down_sync(0xffffffffu, ns, 1);
	if(lane == 0)
		nsL = 0;
	if(lane == 31)
		nsR = 0;
	if(c.act)
	{
		*c.fp = f | spread4(nsL, 0) | spread4(nsR, 2);
		if(!b.vsc)
		{
			const uint32_t up = ((nsL & 1u) << 15) | ((ns & 1u) << 16) | ((nsR & 1u) << 17) |
								((ns & c.sgn4 & 1u) << 23);
			if(up)
				c.fp[-(int)b.fs] |= up;
		}
		const uint32_t dn = ((nsL >> 3) & 1u) | (((ns >> 3) & 1u) << 1) | (((nsR >> 3) & 1u) << 2) |
							(((ns & c.sgn4) >> 3) << 18);
		if(dn)
			c.fp[b.fs] |= dn;
	}
	__syncwarp();
	if((lane == 0 || lane == 31) && c.act && ns)
	{
		uint32_t* o = lane ? c.fp + 1 : c.fp - 1;
		const uint32_t col = lane ? 0u : 2u; // we are their west / east neighbour
		*o |= spread4(ns, col);
		if(!b.vsc && (ns & 1u))
			o[-(int)b.fs] |= 1u << (15 + col);
		if(ns & 8u)
			o[b.fs] |= 1u << col;
	}
	__syncwarp();
}

// ---- symbol stream of a block in the pool -------------------------------------------------------
// Chunks of SYM_CHUNK bytes: payload, then the number of payload bytes and the index of the next
// chunk.  A segment's symbols never straddle chunks.  All members are uniform across the warp.
constexpr uint32_t SYM_CHUNK = 1024, SYM_PAYLOAD = SYM_CHUNK - 8, SYM_NONE = 0xFFFFFFFFu;
struct SymWriter
{
	uint8_t* pool;
	uint64_t cap;
	uint32_t* ctrl;
	uint32_t chunk, fill;
	uint32_t count; // symbols of the current pass
	bool dead; // the pool ran out
};
static __device__ __forceinline__ uint32_t sym_alloc(SymWriter& w, uint32_t lane)
{
	uint32_t c = 0;
	if(lane == 0)
		c = atomicAdd(w.ctrl, 1u);
	c = __shfl_sync(0xffffffffu, c, 0);
	if(((uint64_t)c + 1) * SYM_CHUNK > w.cap)
	{
		if(lane == 0)
			w.ctrl[1] = 1;
		w.dead = true;
		return SYM_NONE;
	}
	return c;
}
static __device__ __forceinline__ void sym_close(const SymWriter& w, uint32_t next, uint32_t lane)
{
	if(lane == 0 && !w.dead && w.chunk != SYM_NONE)
	{
		uint32_t* t = reinterpret_cast<uint32_t*>(w.pool + (size_t)w.chunk * SYM_CHUNK + SYM_PAYLOAD);
		t[0] = w.fill;
		t[1] = next;
	}
}
// room for the `total` symbols of a stripe segment in the block's stream: the lanes write them
// straight to their prefix-sum positions there (`sink`: where they go once the pool has run out)
static __device__ __forceinline__ uint8_t* sym_reserve(SymWriter& w, uint32_t total, uint32_t lane,
														uint8_t* sink)
{
	w.count += total;
	if(w.dead || total == 0)
		return sink;
	if(w.fill + total > SYM_PAYLOAD)
	{
		const uint32_t nxt = sym_alloc(w, lane);
		if(nxt == SYM_NONE)
			return sink;
		sym_close(w, nxt, lane);
		w.chunk = nxt;
		w.fill = 0;
	}
	uint8_t* dst = w.pool + (size_t)w.chunk * SYM_CHUNK + w.fill;
	w.fill += total;
	return dst;
}

static __device__ __forceinline__ int32_t warp_sum(int32_t v)
{
#pragma unroll
	for(int o = 16; o; o >>= 1)
		v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

// enc_sigpass, T1.cpp:459-551
static __device__ int32_t enc_sig_pass(const EBlk& b, SymWriter& sw, uint8_t* list, int bpno, bool raw,
										uint32_t lane)
{
	const uint32_t one = 1u << (bpno + 6);
	int32_t nmsedec = 0;
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
		for(uint32_t x0 = 0; x0 < b.w; x0 += 32)
		{
			const Col c = load_col(b, s, k, x0, lane, one);
			const uint32_t fre = ~(sig4(c.f) | (c.f >> 24)) & c.rowmask; // not significant, not visited
			const uint32_t stat = nz4(c.f);
			// membership to its fixed point: a free sample is coded if it has a significant
			// neighbour when the scan reaches it
			uint32_t ns = 0, mem = 0;
			for(;;)
			{
				uint32_t nsL = __shfl_up_sync(0xffffffffu, ns, 1);
				if(lane == 0)
					nsL = 0;
				const uint32_t cand = (stat | nsL | (nsL << 1) | (nsL >> 1)) & fre;
				uint32_t n2 = 0, m2 = 0;
#pragma unroll
				for(uint32_t r = 0; r < 4; ++r)
				{
					const uint32_t m = ((cand >> r) | (r ? (n2 >> (r - 1)) & (fre >> r) : 0u)) & 1u;
					m2 |= m << r;
					n2 |= (m & (c.bit4 >> r)) << r;
				}
				const bool changed = n2 != ns;
				ns = n2;
				mem = m2;
				if(!__any_sync(0xffffffffu, changed))
					break;
			}
			uint32_t nsL = __shfl_up_sync(0xffffffffu, ns, 1);
			if(lane == 0)
				nsL = 0;
			uint32_t total;
			uint32_t at = warp_excl_scan((uint32_t)(__popc(mem) + __popc(ns)), lane, total);
			uint8_t* out = sym_reserve(sw, total, lane, list);
			const uint32_t fw = c.f | spread4(nsL, 0);
			const uint32_t sp1 = spread4(ns, 1); // own column; a visit sees the rows above it only
#pragma unroll
			for(uint32_t r = 0; r < 4; ++r)
				if((mem >> r) & 1u)
				{
					const uint32_t low = ns & ((1u << r) - 1u);
					const uint32_t fv = fw | (sp1 & ((1u << (3u * r + 4u)) - 1u)) | ((low & c.sgn4) << 19);
					const uint32_t v = (c.bit4 >> r) & 1u;
					out[at++] = (uint8_t)(b.zc[(fv >> (3 * r)) & T1_NEIGH] | (v << 5));
					if(v)
					{
						// the bypass coder takes the sign as it is (T1.cpp:505-508)
						out[at++] = raw ? (uint8_t)(((c.sgn4 >> r) & 1u) << 5) : (uint8_t)sign_symbol(b, c, fv, r);
						nmsedec += nmsedec_sig(b, c.mag[r], (uint32_t)bpno);
					}
				}
			const uint32_t f = c.f | spread4(ns, 1) | ((ns & c.sgn4) << 19) | (mem << 24);
			commit_significance(b, c, f, ns, lane);
		}
	return warp_sum(nmsedec);
}

// enc_refpass, T1.cpp:552-630
static __device__ int32_t enc_ref_pass(const EBlk& b, SymWriter& sw, uint8_t* list, int bpno, uint32_t lane)
{
	const uint32_t one = 1u << (bpno + 6);
	int32_t nmsedec = 0;
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
		for(uint32_t x0 = 0; x0 < b.w; x0 += 32)
		{
			const Col c = load_col(b, s, k, x0, lane, one);
			// significant before this bit-plane: significant and not coded by the pass before
			const uint32_t ref = sig4(c.f) & ~(c.f >> 24) & c.rowmask;
			uint32_t total;
			uint32_t at = warp_excl_scan((uint32_t)__popc(ref), lane, total);
			if(total == 0)
				continue;
			uint8_t* out = sym_reserve(sw, total, lane, list);
#pragma unroll
			for(uint32_t r = 0; r < 4; ++r)
				if((ref >> r) & 1u)
				{
					nmsedec += nmsedec_ref(b, c.mag[r], (uint32_t)bpno);
					const uint32_t cx = (c.f & T1_MU(r))					 ? T1_CTX_MAG + 2
										: ((c.f >> (3 * r)) & T1_NEIGH) ? T1_CTX_MAG + 1
																		 : T1_CTX_MAG;
					out[at++] = (uint8_t)(cx | (((c.bit4 >> r) & 1u) << 5));
				}
			if(ref)
				*c.fp = c.f | (ref << 28);
			__syncwarp();
		}
	return warp_sum(nmsedec);
}

// enc_clnpass, T1.cpp:631-780
static __device__ int32_t enc_cleanup_pass(const EBlk& b, SymWriter& sw, uint8_t* list, int bpno,
											uint32_t lane)
{
	const uint32_t one = 1u << (bpno + 6);
	int32_t nmsedec = 0;
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
		for(uint32_t x0 = 0; x0 < b.w; x0 += 32)
		{
			const Col c = load_col(b, s, k, x0, lane, one);
			// every sample that is neither significant nor coded by the significance pass
			const uint32_t todo = ~(sig4(c.f) | (c.f >> 24)) & c.rowmask;
			const uint32_t ns = todo & c.bit4;
			uint32_t nsL = __shfl_up_sync(0xffffffffu, ns, 1);
			if(lane == 0)
				nsL = 0;
			const uint32_t fw = c.f | spread4(nsL, 0);
			const uint32_t sp1 = spread4(ns, 1); // own column; a visit sees the rows above it only
			// run-length mode: four rows, nothing significant or visited in or around the column
			const bool rl = c.act && c.rowmask == 0xFu && fw == 0;
			const uint32_t r0 = ns ? (uint32_t)__ffs((int)ns) - 1u : 4u;
			uint32_t cnt = (uint32_t)(__popc(todo) + __popc(ns));
			if(rl)
				cnt = ns ? 3u + 1u + (uint32_t)(__popc(todo >> (r0 + 1)) + __popc(ns >> (r0 + 1))) : 1u;
			uint32_t total;
			uint32_t at = warp_excl_scan(cnt, lane, total);
			uint8_t* out = sym_reserve(sw, total, lane, list);
			uint32_t first = 0;
			if(rl)
			{
				out[at++] = (uint8_t)(T1_CTX_AGG | ((ns ? 1u : 0u) << 5));
				if(ns)
				{
					out[at++] = (uint8_t)(T1_CTX_UNI | ((r0 >> 1) << 5));
					out[at++] = (uint8_t)(T1_CTX_UNI | ((r0 & 1u) << 5));
				}
				first = r0; // rows above r0 are zero and need no symbol
			}
#pragma unroll
			for(uint32_t r = 0; r < 4; ++r)
				if(((todo >> r) & 1u) && r >= first)
				{
					const uint32_t low = ns & ((1u << r) - 1u);
					const uint32_t fv = fw | (sp1 & ((1u << (3u * r + 4u)) - 1u)) | ((low & c.sgn4) << 19);
					const uint32_t v = (c.bit4 >> r) & 1u;
					if(!(rl && r == r0))
						out[at++] = (uint8_t)(b.zc[(fv >> (3 * r)) & T1_NEIGH] | (v << 5));
					if(v)
					{
						out[at++] = (uint8_t)sign_symbol(b, c, fv, r);
						nmsedec += nmsedec_sig(b, c.mag[r], (uint32_t)bpno);
					}
				}
			const uint32_t f = (c.f | spread4(ns, 1) | ((ns & c.sgn4) << 19)) & ~T1_PI_ALL;
			commit_significance(b, c, f, ns, lane);
		}
	return warp_sum(nmsedec);
}

// T1.cpp:239-248: L2 norms of the synthesis basis functions, [orientation][level]
__constant__ double c_norms53[4][10] = {
	{1.000, 1.500, 2.750, 5.375, 10.68, 21.34, 42.67, 85.33, 170.7, 341.3},
	{1.038, 1.592, 2.919, 5.703, 11.33, 22.64, 45.25, 90.48, 180.9, 0},
	{1.038, 1.592, 2.919, 5.703, 11.33, 22.64, 45.25, 90.48, 180.9, 0},
	{.7186, .9218, 1.586, 3.043, 6.019, 12.01, 24.00, 47.97, 95.93, 0}};
__constant__ double c_norms97[4][10] = {
	{1.000, 1.965, 4.177, 8.403, 16.90, 33.84, 67.69, 135.3, 270.6, 540.9},
	{2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
	{2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
	{2.080, 3.865, 8.307, 17.18, 34.71, 69.59, 139.3, 278.6, 557.2, 0}};

// T1::enc_is_term_pass, T1.cpp:437-458
static __device__ __forceinline__ bool is_term_pass(uint32_t numbps, uint32_t sty, int bpno,
													 uint32_t passtype)
{
	if(passtype == 2 && bpno == 0)
		return true;
	if(sty & GRKB_CBLKSTY_TERMALL)
		return true;
	if(sty & GRKB_CBLKSTY_LAZY)
	{
		if(bpno == (int)numbps - 4 && passtype == 2)
			return true;
		if(bpno < (int)numbps - 4 && passtype > 0)
			return true;
	}
	return false;
}

constexpr int ENC_MAX_THREADS = 256;
constexpr uint32_t SLOT_LEAD = 16; // zero bytes in front of every slot (out[-1] must be 0)

// what the modelling kernel leaves for the coding kernel: in coded[bi] the number of bit-planes,
// the number of passes and (in data_off) the block's first chunk; in passes[bi][p] the number of
// symbols of pass p (rate) and its distortion sum (the int32 in the bits of `distortion`)
static __device__ __forceinline__ bool pass_is_raw(uint32_t numbps, uint32_t sty, int bpno, uint32_t passtype)
{
	return (bpno < (int)numbps - 4) && passtype < 2 && (sty & GRKB_CBLKSTY_LAZY);
}

// shared memory of one warp of the modelling kernel: symbol list, state words, samples (stripe
// by stripe, four rows per column: a block of height h takes 4 * ceil(h / 4) <= h + 3 rows)
static __host__ __device__ __forceinline__ size_t model_per_warp(const T1EncLaunch& l)
{
	const size_t flag_words = ((size_t)l.max_flag_words + 3) & ~(size_t)3;
	const size_t smp_words = ((size_t)l.max_area + 3 * (size_t)l.max_w + 3) & ~(size_t)3;
	return SYM_CAP + flag_words * 4 + smp_words * 4;
}

// ---- kernel 1: context modelling, one code block per warp ---------------------------------------
__global__ void __launch_bounds__(ENC_MAX_THREADS)
	t1_mq_model_kernel(const T1EncLaunch l, const T1Tables* __restrict__ tab)
{
	extern __shared__ __align__(16) uint8_t smem[];
	int16_t* nms = reinterpret_cast<int16_t*>(smem);
	uint8_t* zc = smem + 1024;
	uint8_t* sc = zc + 2048;
	uint8_t* warp_base = sc + 256;
	for(uint32_t i = threadIdx.x; i < 512; i += blockDim.x)
		nms[i] = tab->nmsedec[i >> 7][i & 127];
	for(uint32_t i = threadIdx.x; i < 2048; i += blockDim.x)
		zc[i] = tab->zc[i];
	for(uint32_t i = threadIdx.x; i < 256; i += blockDim.x)
		sc[i] = tab->sc[i];
	__syncthreads();

	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t per_warp = model_per_warp(l);
	uint8_t* list = warp_base + warp * per_warp; // (context, decision) symbols of one stripe segment
	uint32_t* fl = reinterpret_cast<uint32_t*>(list + SYM_CAP);
	uint32_t* smp = fl + ((l.max_flag_words + 3u) & ~3u); // 16-byte aligned: read as uint4

	for(;;)
	{
		uint32_t item = 0;
		if(lane == 0)
			item = atomicAdd(l.counter, 1u);
		item = __shfl_sync(0xffffffffu, item, 0);
		if(item >= l.nblocks)
			break;
		const uint32_t bi = item;
		const grkb_enc_cblk cb = l.blocks[bi];
		if(cb.cblk_sty & GRKB_CBLKSTY_HT)
		{
			if(lane == 0)
				l.sym_total[bi] = 0;
			continue;
		}
		const uint32_t w = cb.w, h = cb.h;
		const uint32_t fs = w + 2, nstripes = (h + 3) >> 2, nfl = fs * (nstripes + 2);
		// ---- quantise to sign-magnitude, 6 fractional bits (T1Part1.cpp:37-88) -----------------
		uint32_t mx = 0;
		// stored stripe by stripe with the four rows of a column side by side (one 128-bit load per
		// column and pass in load_col); rows past the block's height are zeros
		for(uint32_t s4 = 0; s4 < nstripes; ++s4)
		{
			const int32_t* src =
				reinterpret_cast<const int32_t*>(l.planes[cb.comp]) + (size_t)(cb.y + 4 * s4) * l.stride + cb.x;
			const uint32_t rows = min(4u, h - 4 * s4);
			for(uint32_t x = lane; x < w; x += 32)
			{
				uint32_t q[4] = {0, 0, 0, 0};
#pragma unroll
				for(uint32_t r = 0; r < 4; ++r)
					if(r < rows)
					{
						const int32_t v = src[(size_t)r * l.stride + x];
						int32_t t;
						if(l.reversible)
							t = (int32_t)((uint32_t)v << 6);
						else
							t = __float2int_rn(__fmul_rn(__fdiv_rn(__int_as_float(v), cb.stepsize), 64.0f));
						const uint32_t mag = (uint32_t)(t < 0 ? -t : t);
						q[r] = mag | (t < 0 ? 0x80000000u : 0u);
						mx = max(mx, mag);
					}
				reinterpret_cast<uint4*>(smp)[s4 * w + x] = make_uint4(q[0], q[1], q[2], q[3]);
			}
		}
		for(uint32_t i = lane; i < nfl; i += 32)
			fl[i] = 0;
		for(int o = 16; o; o >>= 1)
			mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
		__syncwarp();
		grkb_coded res;
		res.data_off = 0;
		res.data_len = 0;
		res.numbps = 0;
		res.numpasses = 0;
		res.failed = 0;
		res.distortion = 0.0;
		if(mx)
		{
			const uint32_t t = 32 - __clz(mx); // floorlog2(max) + 1
			res.numbps = (uint8_t)(t <= 6 ? 0 : t - 6);
		}
		grkb_pass* passes = l.passes + (size_t)bi * GRKB_ENC_MAX_PASSES;
		uint32_t nsym = 0;
		if(res.numbps)
		{
			EBlk b;
			b.fl = fl;
			b.smp = smp;
			b.w = w;
			b.h = h;
			b.fs = fs;
			b.zc = zc + 512u * cb.orient;
			b.sc = sc;
			b.nms = nms;
			b.vsc = (cb.cblk_sty & GRKB_CBLKSTY_VSC) != 0;
			SymWriter sw;
			sw.pool = l.sym_pool;
			sw.cap = l.sym_cap;
			sw.ctrl = l.sym_ctrl;
			sw.dead = false;
			sw.fill = 0;
			sw.count = 0;
			sw.chunk = SYM_NONE;
			sw.chunk = sym_alloc(sw, lane);
			res.data_off = sw.chunk;
			const uint32_t sty = cb.cblk_sty, numbps = res.numbps;
			int bpno = (int)numbps - 1;
			uint32_t passtype = 2, passno = 0;
			for(; bpno >= 0 && passno < GRKB_ENC_MAX_PASSES; ++passno)
			{
				const bool raw = pass_is_raw(numbps, sty, bpno, passtype);
				sw.count = 0;
				int32_t nmsedec;
				if(passtype == 0)
					nmsedec = enc_sig_pass(b, sw, list, bpno, raw, lane);
				else if(passtype == 1)
					nmsedec = enc_ref_pass(b, sw, list, bpno, lane);
				else
				{
					nmsedec = enc_cleanup_pass(b, sw, list, bpno, lane);
					if(sty & GRKB_CBLKSTY_SEGSYM)
					{
						// mqc_segmark_enc, mqc_enc.cpp:327-332: 1 0 1 0 in the uniform context
						uint8_t* out = sym_reserve(sw, 4, lane, list);
						if(lane < 4)
							out[lane] = (uint8_t)(T1_CTX_UNI | (((lane + 1) & 1u) << 5));
						__syncwarp();
					}
				}
				if(lane == 0)
				{
					grkb_pass ps;
					ps.distortion = __longlong_as_double((long long)nmsedec);
					ps.rate = sw.count;
					nsym += sw.count;
					ps.len = 0;
					ps.term = 0;
					ps.reserved = 0;
					passes[passno] = ps;
				}
				if(++passtype == 3)
				{
					passtype = 0;
					bpno--;
				}
			}
			sym_close(sw, SYM_NONE, lane);
			res.numpasses = (uint8_t)passno;
			res.failed = sw.dead ? 2 : 0;
		}
		if(lane == 0)
		{
			l.coded[bi] = res;
			l.sym_total[bi] = nsym;
		}
		__syncwarp();
	}
}

// ---- between the two: the order the coder takes the blocks in ------------------------------------
// The coder's running time is the length of its longest chain, and the 32 chains of a warp cost
// as much as the longest of them.  So the blocks are handed out longest first (bucket sort on the
// symbol count, 1024 buckets), which puts blocks of like length into the same warp, and the
// blocks at the very top -- longer than `frac`/256 of the longest one, at most `cap` of them --
// each get a warp to themselves, where the chain runs several times faster than in lockstep with
// 31 others on different branches.
constexpr int ORDER_THREADS = 1024;
__global__ void __launch_bounds__(ORDER_THREADS)
	t1_mq_order_kernel(const T1EncLaunch l, uint32_t frac, uint32_t cap, uint32_t min_long)
{
	__shared__ uint32_t cnt[ORDER_THREADS], start[ORDER_THREADS], red[32];
	const uint32_t t = threadIdx.x, lane = t & 31, warp = t >> 5;
	uint32_t mx = 0;
	for(uint32_t i = t; i < l.nblocks; i += ORDER_THREADS)
		mx = max(mx, l.sym_total[i]);
	for(int o = 16; o; o >>= 1)
		mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
	if(lane == 0)
		red[warp] = mx;
	cnt[t] = 0;
	__syncthreads();
	mx = red[lane];
	for(int o = 16; o; o >>= 1)
		mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
	const uint64_t scale = (uint64_t)mx + 1;
	// bucket 0 holds the longest blocks
	for(uint32_t i = t; i < l.nblocks; i += ORDER_THREADS)
		atomicAdd(&cnt[ORDER_THREADS - 1 - (uint32_t)(((uint64_t)l.sym_total[i] * ORDER_THREADS) / scale)], 1u);
	__syncthreads();
	// exclusive scan of the 1024 counts
	const uint32_t own = cnt[t];
	uint32_t inc = own;
	for(int o = 1; o < 32; o <<= 1)
	{
		const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
		if((int)lane >= o)
			inc += v;
	}
	__syncthreads();
	if(lane == 31)
		red[warp] = inc;
	__syncthreads();
	uint32_t wsum = red[lane];
	for(int o = 1; o < 32; o <<= 1)
	{
		const uint32_t v = __shfl_up_sync(0xffffffffu, wsum, o);
		if((int)lane >= o)
			wsum += v;
	}
	const uint32_t before = __shfl_sync(0xffffffffu, wsum, warp ? warp - 1 : 0);
	start[t] = inc - own + (warp ? before : 0);
	__syncthreads();
	if(t == 0)
	{
		// blocks in the buckets above the threshold, whole buckets only
		uint32_t nlong = 0;
		if(mx >= min_long)
		{
			const uint32_t nb = ORDER_THREADS - (ORDER_THREADS * frac >> 8); // buckets that qualify
			nlong = nb < ORDER_THREADS ? start[nb] : l.nblocks;
			if(nlong > cap)
				nlong = cap;
		}
		l.code_ctrl[0] = nlong;
	}
	__syncthreads();
	for(uint32_t i = t; i < l.nblocks; i += ORDER_THREADS)
	{
		const uint32_t b = ORDER_THREADS - 1 - (uint32_t)(((uint64_t)l.sym_total[i] * ORDER_THREADS) / scale);
		l.code_order[atomicAdd(&start[b], 1u)] = i;
	}
}

// ---- kernel 2: the arithmetic coder, one code block per THREAD ------------------------------------
// With the decisions on a list the MQ coder is a flat loop: 32 blocks per warp run it in lockstep,
// each lane on its own stream, coder registers and context words (shared memory, one column per
// thread).  Per pass the reference's bookkeeping follows: restart / bypass initialisation,
// termination (flush, predictable termination, bypass flush), the rate and the distortion record.
constexpr int CODE_THREADS = 64;

struct SymReader
{
	const uint8_t* pool;
	const uint8_t* base; // current chunk
	uint32_t pos, fill;
	uint32_t word, ahead; // the four symbols being handed out, and the next four (loaded early:
						  // the load would otherwise sit in the middle of the decision chain)
};
static __device__ __forceinline__ void sym_open(SymReader& r, uint32_t chunk)
{
	r.base = r.pool + (size_t)chunk * SYM_CHUNK;
	r.pos = 0;
	r.fill = *reinterpret_cast<const uint32_t*>(r.base + SYM_PAYLOAD);
	r.ahead = *reinterpret_cast<const uint32_t*>(r.base);
	r.word = 0;
	asm volatile("prefetch.global.L1 [%0];" ::"l"(r.base + 128));
	asm volatile("prefetch.global.L1 [%0];" ::"l"(r.base + 256));
	// and the start of the chunk after this one (its index is final: the chunks were closed in order)
	const uint32_t nxt = *reinterpret_cast<const uint32_t*>(r.base + SYM_PAYLOAD + 4);
	if(nxt != SYM_NONE)
		asm volatile("prefetch.global.L1 [%0];" ::"l"(r.pool + (size_t)nxt * SYM_CHUNK));
}
static __device__ __forceinline__ uint32_t sym_next(SymReader& r)
{
	if(r.pos == r.fill)
		sym_open(r, *reinterpret_cast<const uint32_t*>(r.base + SYM_PAYLOAD + 4));
	if((r.pos & 3u) == 0)
	{
		r.word = r.ahead;
		r.ahead = *reinterpret_cast<const uint32_t*>(r.base + r.pos + 4); // at worst the chunk's trailer
		// the pool is larger than L2 holds between the two kernels: pull the lines in well ahead
		if((r.pos & 127u) == 0 && r.pos + 384u < SYM_CHUNK)
			asm volatile("prefetch.global.L1 [%0];" ::"l"(r.base + r.pos + 384u));
	}
	const uint32_t v = r.word & 0xFFu;
	r.word >>= 8;
	r.pos++;
	return v;
}

__global__ void __launch_bounds__(CODE_THREADS)
	t1_mq_code_kernel(const T1EncLaunch l, const T1Tables* __restrict__ tab, const uint32_t group)
{
	__shared__ uint32_t mqt[96];
	__shared__ uint32_t cxs_all[T1_NCTX * CODE_THREADS];
	for(uint32_t i = threadIdx.x; i < 94; i += blockDim.x)
		mqt[i] = tab->mq[i] | (i << 31); // the MPS of state 2i+mps rides in bit 31
	__syncthreads();
	// the first warps take `group` of the nlong longest blocks each (one by default, on lane 0), the
	// others 32 blocks each
	const uint32_t nlong = l.code_ctrl[0];
	const uint32_t gw = blockIdx.x * (CODE_THREADS / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
	uint32_t idx;
	const uint32_t nlw = (nlong + group - 1) / group; // warps that hold `group` long blocks each
	if(gw < nlw)
	{
		idx = gw * group + lane;
		if(lane >= group || idx >= nlong)
			return;
	}
	else
		idx = nlong + (gw - nlw) * 32 + lane;
	if(idx >= l.nblocks)
		return;
	const uint32_t bi = l.code_order[idx];
	const grkb_enc_cblk cb = l.blocks[bi];
	if(cb.cblk_sty & GRKB_CBLKSTY_HT)
		return;
	grkb_coded res = l.coded[bi];
	if(res.numbps == 0)
		return; // nothing to code: the record is final
	const uint32_t npasses = res.numpasses, first_chunk = (uint32_t)res.data_off;
	res.data_off = 0;
	if(res.failed)
	{
		res.numpasses = 0;
		l.coded[bi] = res;
		return; // the symbol pool ran out: the host codes the tile again with a larger one
	}
	uint32_t* cxs = cxs_all + threadIdx.x;
	grkb_pass* passes = l.passes + (size_t)bi * GRKB_ENC_MAX_PASSES;
	SymReader rd;
	rd.pool = l.sym_pool;
	sym_open(rd, first_chunk);
	MqE q;
	q.out = l.slots + (size_t)bi * l.slot_bytes + SLOT_LEAD;
	q.cap = (int32_t)(l.slot_bytes - SLOT_LEAD);
	q.a = 0x8000;
	q.c = 0;
	q.bp = -1;
	q.cur = 0;
	q.ct = 12;
	q.overflow = false;
	reset_states<CODE_THREADS>(cxs, mqt);
	const uint32_t sty = cb.cblk_sty, numbps = res.numbps;
	const bool pterm = (sty & GRKB_CBLKSTY_PTERM) != 0;
	// getwmsedec weights: w1 * w2 * stepsize, T1.cpp:418-436
	uint32_t lvl = cb.level;
	if(cb.orient == 0 && lvl > 9)
		lvl = 9;
	else if(cb.orient > 0 && lvl > 8)
		lvl = 8;
	const double w1 = cb.comp < l.mct_numcomps ? l.mct_norms[cb.comp] : 1.0;
	const double w2 = l.reversible ? c_norms53[cb.orient][lvl] : c_norms97[cb.orient][lvl];
	const double wstep = __dmul_rn(__dmul_rn(w1, w2), (double)cb.stepsize);
	int bpno = (int)numbps - 1;
	uint32_t passtype = 2;
	double cum = 0.0;
	bool prev_term = false;
	for(uint32_t passno = 0; passno < npasses; ++passno)
	{
		const bool raw = pass_is_raw(numbps, sty, bpno, passtype);
		if(passno > 0 && prev_term)
		{
			if(raw)
				raw_init(q);
			else
				mq_restart(q);
		}
		const uint32_t count = passes[passno].rate;
		const int32_t nmsedec = (int32_t)__double_as_longlong(passes[passno].distortion);
		if(raw)
		{
#pragma unroll 1
			for(uint32_t i = 0; i < count; ++i)
				raw_encode(q, sym_next(rd) >> 5);
		}
		else
		{
#pragma unroll 1
			for(uint32_t i = 0; i < count; ++i)
			{
				const uint32_t v = sym_next(rd);
				mq_encode<CODE_THREADS>(q, cxs, mqt, v & 31u, v >> 5);
			}
		}
		grkb_pass ps;
		ps.distortion = 0.0;
		ps.reserved = 0;
		ps.len = 0;
		if(l.rate_control)
		{
			double wm = __dmul_rn(wstep, (double)(1 << bpno));
			wm = __dmul_rn(wm, __ddiv_rn(__dmul_rn(wm, (double)nmsedec), 8192.0));
			cum = __dadd_rn(cum, wm);
			ps.distortion = cum;
		}
		if(is_term_pass(numbps, sty, bpno, passtype))
		{
			if(raw)
				raw_flush(q, pterm);
			else if(pterm)
				mq_erterm(q);
			else
				mq_flush(q);
			ps.term = 1;
			ps.rate = (uint32_t)q.bp;
		}
		else
		{
			uint32_t extra;
			if(raw)
				extra = raw_needs_two(q, pterm) ? 2 : 1;
			else
				extra = q.ct < 5 ? 6 : 5;
			ps.term = 0;
			ps.rate = (uint32_t)q.bp + extra;
		}
		prev_term = ps.term != 0;
		passes[passno] = ps;
		if(++passtype == 3)
		{
			passtype = 0;
			bpno--;
		}
		if(sty & GRKB_CBLKSTY_RESET)
			reset_states<CODE_THREADS>(cxs, mqt);
	}
	eseek(q, q.bp); // write the byte under construction back
	// T1.cpp:905-931: rates made monotone, a trailing 0xFF is never included
	uint32_t last = (uint32_t)q.bp;
	for(uint32_t p = npasses; p > 0;)
	{
		--p;
		if(passes[p].rate > last)
			passes[p].rate = last;
		else
			last = passes[p].rate;
	}
	uint32_t prev = 0;
	for(uint32_t p = 0; p < npasses; ++p)
	{
		uint32_t r = passes[p].rate;
		if(r > 0 && q.out[r - 1] == 0xFF)
			r--;
		passes[p].rate = r;
		passes[p].len = (uint16_t)(r - prev);
		prev = r;
	}
	res.data_len = npasses ? prev : 0;
	res.distortion = cum;
	res.failed = q.overflow ? 1 : 0;
	l.coded[bi] = res;
}
} // namespace

static uint32_t env_u32(const char* name, uint32_t dflt)
{
	const char* v = getenv(name);
	return v && *v ? (uint32_t)strtoul(v, nullptr, 10) : dflt;
}

int grkb_launch_t1_mq_encode(const T1EncLaunch& l, cudaStream_t s)
{
	if(l.nblocks == 0)
		return 0;
	const T1Tables* tab = grkb_t1_tables_device();
	if(!tab)
	{
		grkb_set_error("t1: cannot upload coding tables");
		return -1;
	}
	if(!l.sym_pool || l.sym_cap < (uint64_t)SYM_CHUNK)
	{
		grkb_set_error("t1 encode: no symbol pool");
		return -1;
	}
	const size_t per_warp = model_per_warp(l);
	const size_t fixed = 1024 + 2048 + 256;
	int nw = ENC_MAX_THREADS / 32;
	while(nw > 1 && fixed + per_warp * nw > 100 * 1024)
		--nw;
	const size_t smem = fixed + per_warp * nw;
	if(smem > 227 * 1024)
	{
		grkb_set_error("t1 encode: code block too large for shared memory (%zu bytes)", smem);
		return -1;
	}
	GRKB_ENSURE_SMEM(t1_mq_model_kernel, smem);
	int dev = 0, sms = 148;
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	int per_sm = (int)((227 * 1024) / (smem + 1024));
	if(per_sm < 1)
		per_sm = 1;
	if(per_sm * nw > 64)
		per_sm = 64 / nw;
	uint32_t ctas = (uint32_t)(sms * per_sm);
	const uint32_t need = (l.nblocks + nw - 1) / nw;
	if(ctas > need)
		ctas = need;
	GRKB_CUDA_CHECK(cudaMemsetAsync(l.counter, 0, sizeof(uint32_t), s));
	t1_mq_model_kernel<<<ctas, nw * 32, smem, s>>>(l, tab);
	GRKB_CUDA_CHECK(cudaGetLastError());
	static const uint32_t frac = env_u32("GRKB_MQ_LONG_FRAC", 50); // of 256: 0.2 of the longest block
	static const uint32_t cap = env_u32("GRKB_MQ_LONG_CAP", 2048);
	static const uint32_t min_long = env_u32("GRKB_MQ_LONG_MIN", 1024);
	t1_mq_order_kernel<<<1, ORDER_THREADS, 0, s>>>(l, frac > 256 ? 256 : frac, cap, min_long);
	GRKB_CUDA_CHECK(cudaGetLastError());
	static const uint32_t group_env = env_u32("GRKB_MQ_LONG_GROUP", 1); // long blocks per warp
	const uint32_t group = group_env < 1 ? 1 : (group_env > 32 ? 32 : group_env);
	const uint32_t solo = cap < l.nblocks ? cap : l.nblocks;
	const uint32_t warps = (solo + group - 1) / group + (l.nblocks + 31) / 32;
	t1_mq_code_kernel<<<(warps + CODE_THREADS / 32 - 1) / (CODE_THREADS / 32), CODE_THREADS, 0, s>>>(l, tab, group);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
