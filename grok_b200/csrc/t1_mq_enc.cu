// t1_mq_enc.cu -- EBCOT (JPEG 2000 Part 1) code-block encoder with the quantiser fused into
// the load, sm_100a.  One code block per warp.
//
// Replaces, behind the C-ABI, the reference's
//   T1Part1::preCompress / compress   src/lib/jp2/t1/t1_part1/T1Part1.cpp:37-132
//   T1::compress_cblk                 src/lib/jp2/t1/t1_part1/T1.cpp:781-932
//   enc_sigpass / refpass / clnpass   T1.cpp:459-780, enc_is_term_pass :437-458
//   getwmsedec                        T1.cpp:418-436 (+ norms :239-277)
//   MQ / RAW encoder                  t1/t1_part1/mqc_enc.cpp:86-332, mqc_enc_inl.h:38-124
//
// Shape: like the decoder (t1_mq_dec.cu) the MQ coder is one serial chain per block.  All lanes
// of the warp load the block's rows from the coefficient plane (coalesced), quantise to
// sign-magnitude with 6 fractional bits into shared memory and reduce the block maximum; lane 0
// then walks the coding passes over the stripe-column state words (layout in t1.cuh), emitting
// bytes into the block's slot in global memory with the byte under construction held in a
// register (the coder's carry only ever touches that byte).  Per-pass rate / length /
// termination and the double-precision distortion estimate are written to the pass table.
#include "t1.cuh"

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

// mqc_byteout, mqc_enc.cpp:86-125
static __device__ __forceinline__ void mq_byteout(MqE& q)
{
	if(q.cur == 0xFF)
	{
		epush(q, q.c >> 20);
		q.c &= 0xFFFFF;
		q.ct = 7;
	}
	else if((q.c & 0x8000000) == 0)
	{
		epush(q, q.c >> 19);
		q.c &= 0x7FFFF;
		q.ct = 8;
	}
	else
	{
		q.cur++;
		if(q.cur == 0xFF)
		{
			q.c &= 0x7FFFFFF;
			epush(q, q.c >> 20);
			q.c &= 0xFFFFF;
			q.ct = 7;
		}
		else
		{
			epush(q, q.c >> 19);
			q.c &= 0x7FFFF;
			q.ct = 8;
		}
	}
}

// renorm of mqc_enc_inl.h:38-60 with the shifts batched between byte outputs
static __device__ __forceinline__ void mq_renorm(MqE& q)
{
	uint32_t n = __clz(q.a) - 16;
	while(n)
	{
		const uint32_t k = min(n, q.ct);
		q.a <<= k;
		q.c <<= k;
		q.ct -= k;
		n -= k;
		if(q.ct == 0)
			mq_byteout(q);
	}
}

// mqc_encode_macro, mqc_enc_inl.h:62-102; cxs[cx] = state index << 1 | mps as in the decoder
static __device__ __forceinline__ void mq_encode(MqE& q, uint8_t* cxs, const uint32_t* mqt,
												  uint32_t cx, uint32_t d)
{
	const uint32_t st = cxs[cx];
	const uint32_t e = mqt[st];
	const uint32_t qe = e & 0xFFFFu;
	q.a -= qe;
	if(d == (st & 1u))
	{
		if((q.a & 0x8000u) == 0)
		{
			if(q.a < qe)
				q.a = qe;
			else
				q.c += qe;
			cxs[cx] = (uint8_t)((e >> 16) & 0xFFu);
			mq_renorm(q);
		}
		else
			q.c += qe;
	}
	else
	{
		if(q.a < qe)
			q.c += qe;
		else
			q.a = qe;
		cxs[cx] = (uint8_t)(e >> 24);
		mq_renorm(q);
	}
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

static __device__ __forceinline__ void reset_states(uint8_t* cxs)
{
	// mqc_resetstates, mqc_dec.cpp:121-130 (shared by the encoder)
	for(int i = 0; i < T1_NCTX; ++i)
		cxs[i] = 0;
	cxs[T1_CTX_UNI] = 46 << 1;
	cxs[T1_CTX_AGG] = 3 << 1;
	cxs[T1_CTX_ZC] = 4 << 1;
}

struct EBlk
{
	uint32_t* fl; // state words, (w+2) x (nstripes+2)
	const uint32_t* smp; // sign-magnitude samples, 6 fractional bits, row pitch w
	uint32_t w, h, fs;
	const uint8_t* zc; // zero-coding table of this band orientation
	const uint8_t* sc;
	const int16_t* nms; // distortion tables [4][128]
	bool vsc;
};

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

// update_flags_macro, T1.cpp:209-233, on the stripe-column words
static __device__ __forceinline__ void mark_significant(const EBlk& b, uint32_t* fp, uint32_t& f,
														 uint32_t r, uint32_t sgn)
{
	f |= T1_SIG_C(r) | (sgn << (18 + r + 1));
	fp[-1] |= 1u << (3 * (r + 1) + 2);
	fp[1] |= 1u << (3 * (r + 1));
	if(r == 0 && !b.vsc)
	{
		uint32_t* n = fp - b.fs;
		n[0] |= (1u << 16) | (sgn << 23);
		n[-1] |= 1u << 17;
		n[1] |= 1u << 15;
	}
	if(r == 3)
	{
		uint32_t* s = fp + b.fs;
		s[0] |= (1u << 1) | (sgn << 18);
		s[-1] |= 1u << 2;
		s[1] |= 1u << 0;
	}
}

// getnmsedec_sig / getnmsedec_ref, T1.cpp:194-208
static __device__ __forceinline__ int32_t nmsedec_sig(const EBlk& b, uint32_t mag, uint32_t bitpos)
{
	return bitpos > 0 ? b.nms[(mag >> bitpos) & 127] : b.nms[128 + (mag & 127)];
}
static __device__ __forceinline__ int32_t nmsedec_ref(const EBlk& b, uint32_t mag, uint32_t bitpos)
{
	return bitpos > 0 ? b.nms[256 + ((mag >> bitpos) & 127)] : b.nms[384 + (mag & 127)];
}

// enc_sigpass, T1.cpp:459-551
template<bool RAW>
static __device__ int32_t enc_sig_pass(const EBlk& b, MqE& q, uint8_t* cxs, const uint32_t* mqt,
										int bpno)
{
	const uint32_t one = 1u << (bpno + 6);
	int32_t nmsedec = 0;
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
	{
		const uint32_t rows = min(4u, b.h - k);
		uint32_t* fp = b.fl + (s + 1) * b.fs + 1;
		const uint32_t* dp = b.smp + k * b.w;
		for(uint32_t x = 0; x < b.w; ++x, ++fp, ++dp)
		{
			uint32_t f = *fp;
			if(f == 0)
				continue;
			for(uint32_t r = 0; r < rows; ++r)
			{
				if((f & (T1_SIG_C(r) | T1_PI(r))) || !((f >> (3 * r)) & T1_NEIGH))
					continue;
				const uint32_t sm = dp[r * b.w];
				const uint32_t mag = sm & 0x7FFFFFFFu, sgn = sm >> 31;
				const uint32_t v = (mag & one) != 0;
				if(RAW)
					raw_encode(q, v);
				else
					mq_encode(q, cxs, mqt, b.zc[(f >> (3 * r)) & T1_NEIGH], v);
				if(v)
				{
					const uint32_t e = b.sc[sc_index(f, fp[-1], fp[1], r)];
					nmsedec += nmsedec_sig(b, mag, (uint32_t)bpno);
					if(RAW)
						raw_encode(q, sgn);
					else
						mq_encode(q, cxs, mqt, e & 0x7Fu, sgn ^ (e >> 7));
					mark_significant(b, fp, f, r, sgn);
				}
				f |= T1_PI(r);
			}
			*fp = f;
		}
	}
	return nmsedec;
}

// enc_refpass, T1.cpp:552-630
template<bool RAW>
static __device__ int32_t enc_ref_pass(const EBlk& b, MqE& q, uint8_t* cxs, const uint32_t* mqt,
										int bpno)
{
	const uint32_t one = 1u << (bpno + 6);
	int32_t nmsedec = 0;
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
	{
		const uint32_t rows = min(4u, b.h - k);
		uint32_t* fp = b.fl + (s + 1) * b.fs + 1;
		const uint32_t* dp = b.smp + k * b.w;
		for(uint32_t x = 0; x < b.w; ++x, ++fp, ++dp)
		{
			uint32_t f = *fp;
			if(!(f & T1_SIGC_ALL))
				continue;
			for(uint32_t r = 0; r < rows; ++r)
			{
				if((f & (T1_SIG_C(r) | T1_PI(r))) != T1_SIG_C(r))
					continue;
				const uint32_t mag = dp[r * b.w] & 0x7FFFFFFFu;
				nmsedec += nmsedec_ref(b, mag, (uint32_t)bpno);
				const uint32_t v = (mag & one) != 0;
				if(RAW)
					raw_encode(q, v);
				else
				{
					const uint32_t cx = (f & T1_MU(r))					   ? T1_CTX_MAG + 2
										: ((f >> (3 * r)) & T1_NEIGH) ? T1_CTX_MAG + 1
																	   : T1_CTX_MAG;
					mq_encode(q, cxs, mqt, cx, v);
				}
				f |= T1_MU(r);
			}
			*fp = f;
		}
	}
	return nmsedec;
}

// enc_clnpass, T1.cpp:631-780
static __device__ int32_t enc_cleanup_pass(const EBlk& b, MqE& q, uint8_t* cxs, const uint32_t* mqt,
											int bpno)
{
	const uint32_t one = 1u << (bpno + 6);
	int32_t nmsedec = 0;
	for(uint32_t k = 0, s = 0; k < b.h; k += 4, ++s)
	{
		const uint32_t rows = min(4u, b.h - k);
		uint32_t* fp = b.fl + (s + 1) * b.fs + 1;
		const uint32_t* dp = b.smp + k * b.w;
		for(uint32_t x = 0; x < b.w; ++x, ++fp, ++dp)
		{
			uint32_t f = *fp;
			uint32_t r = 0;
			bool sign_only = false;
			if(rows == 4 && f == 0)
			{
				while(r < 4 && !(dp[r * b.w] & one))
					++r;
				mq_encode(q, cxs, mqt, T1_CTX_AGG, r != 4);
				if(r == 4)
					continue;
				mq_encode(q, cxs, mqt, T1_CTX_UNI, r >> 1);
				mq_encode(q, cxs, mqt, T1_CTX_UNI, r & 1);
				sign_only = true;
			}
			for(; r < rows; ++r)
			{
				const uint32_t sm = dp[r * b.w];
				const uint32_t mag = sm & 0x7FFFFFFFu, sgn = sm >> 31;
				if(!sign_only)
				{
					if(f & (T1_SIG_C(r) | T1_PI(r)))
						continue;
					const uint32_t v = (mag & one) != 0;
					mq_encode(q, cxs, mqt, b.zc[(f >> (3 * r)) & T1_NEIGH], v);
					if(!v)
						continue;
				}
				sign_only = false;
				const uint32_t e = b.sc[sc_index(f, fp[-1], fp[1], r)];
				nmsedec += nmsedec_sig(b, mag, (uint32_t)bpno);
				mq_encode(q, cxs, mqt, e & 0x7Fu, sgn ^ (e >> 7));
				mark_significant(b, fp, f, r, sgn);
			}
			*fp = f & ~T1_PI_ALL;
		}
	}
	return nmsedec;
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

__global__ void __launch_bounds__(ENC_MAX_THREADS)
	t1_mq_encode_kernel(const T1EncLaunch l, const T1Tables* __restrict__ tab)
{
	extern __shared__ __align__(16) uint8_t smem[];
	uint32_t* mqt = reinterpret_cast<uint32_t*>(smem);
	int16_t* nms = reinterpret_cast<int16_t*>(smem + 94 * 4);
	uint8_t* zc = smem + 94 * 4 + 1024;
	uint8_t* sc = zc + 2048;
	uint8_t* warp_base = sc + 256;
	for(uint32_t i = threadIdx.x; i < 94; i += blockDim.x)
		mqt[i] = tab->mq[i];
	for(uint32_t i = threadIdx.x; i < 512; i += blockDim.x)
		nms[i] = tab->nmsedec[i >> 7][i & 127];
	for(uint32_t i = threadIdx.x; i < 2048; i += blockDim.x)
		zc[i] = tab->zc[i];
	for(uint32_t i = threadIdx.x; i < 256; i += blockDim.x)
		sc[i] = tab->sc[i];
	__syncthreads();

	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t per_warp = 32 + (size_t)l.max_flag_words * 4 + (size_t)l.max_area * 4;
	uint8_t* cxs = warp_base + warp * per_warp;
	uint32_t* fl = reinterpret_cast<uint32_t*>(cxs + 32);
	uint32_t* smp = fl + l.max_flag_words;

	for(;;)
	{
		uint32_t item = 0;
		if(lane == 0)
			item = atomicAdd(l.counter, 1u);
		item = __shfl_sync(0xffffffffu, item, 0);
		if(item >= l.nblocks)
			break;
		const uint32_t bi = l.order ? l.order[item] : item;
		const grkb_enc_cblk cb = l.blocks[bi];
		if(cb.cblk_sty & GRKB_CBLKSTY_HT)
			continue;
		const uint32_t w = cb.w, h = cb.h;
		const uint32_t fs = w + 2, nstripes = (h + 3) >> 2, nfl = fs * (nstripes + 2);
		// ---- quantise to sign-magnitude, 6 fractional bits (T1Part1.cpp:37-88) -----------------
		uint32_t mx = 0;
		for(uint32_t y = 0; y < h; ++y)
		{
			const int32_t* src =
				reinterpret_cast<const int32_t*>(l.planes[cb.comp]) + (size_t)(cb.y + y) * l.stride + cb.x;
			for(uint32_t x = lane; x < w; x += 32)
			{
				int32_t t;
				if(l.reversible)
					t = (int32_t)((uint32_t)src[x] << 6);
				else
					t = __float2int_rn(
						__fmul_rn(__fdiv_rn(__int_as_float(src[x]), cb.stepsize), 64.0f));
				const uint32_t mag = (uint32_t)(t < 0 ? -t : t);
				smp[y * w + x] = mag | (t < 0 ? 0x80000000u : 0u);
				mx = max(mx, mag);
			}
		}
		for(uint32_t i = lane; i < nfl; i += 32)
			fl[i] = 0;
		for(int o = 16; o; o >>= 1)
			mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
		__syncwarp();
		if(lane == 0)
		{
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
				MqE q;
				q.out = l.slots + (size_t)bi * l.slot_bytes + SLOT_LEAD;
				q.cap = (int32_t)(l.slot_bytes - SLOT_LEAD);
				q.a = 0x8000;
				q.c = 0;
				q.bp = -1;
				q.cur = 0;
				q.ct = 12;
				q.overflow = false;
				reset_states(cxs);
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
				uint32_t passtype = 2, passno = 0;
				double cum = 0.0;
				bool prev_term = false;
				for(; bpno >= 0 && passno < GRKB_ENC_MAX_PASSES; ++passno)
				{
					const bool raw = (bpno < (int)numbps - 4) && passtype < 2 && (sty & GRKB_CBLKSTY_LAZY);
					if(passno > 0 && prev_term)
					{
						if(raw)
							raw_init(q);
						else
							mq_restart(q);
					}
					int32_t nmsedec;
					if(passtype == 0)
						nmsedec = raw ? enc_sig_pass<true>(b, q, cxs, mqt, bpno)
									  : enc_sig_pass<false>(b, q, cxs, mqt, bpno);
					else if(passtype == 1)
						nmsedec = raw ? enc_ref_pass<true>(b, q, cxs, mqt, bpno)
									  : enc_ref_pass<false>(b, q, cxs, mqt, bpno);
					else
					{
						nmsedec = enc_cleanup_pass(b, q, cxs, mqt, bpno);
						if(sty & GRKB_CBLKSTY_SEGSYM)
							for(uint32_t i = 1; i < 5; ++i) // mqc_segmark_enc, mqc_enc.cpp:327-332
								mq_encode(q, cxs, mqt, T1_CTX_UNI, i & 1);
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
						reset_states(cxs);
				}
				eseek(q, q.bp); // write the byte under construction back
				// T1.cpp:905-931: rates made monotone, a trailing 0xFF is never included
				uint32_t last = (uint32_t)q.bp;
				for(uint32_t p = passno; p > 0;)
				{
					--p;
					if(passes[p].rate > last)
						passes[p].rate = last;
					else
						last = passes[p].rate;
				}
				uint32_t prev = 0;
				for(uint32_t p = 0; p < passno; ++p)
				{
					uint32_t r = passes[p].rate;
					if(r > 0 && q.out[r - 1] == 0xFF)
						r--;
					passes[p].rate = r;
					passes[p].len = (uint16_t)(r - prev);
					prev = r;
				}
				res.numpasses = (uint8_t)passno;
				res.data_len = passno ? prev : 0;
				res.distortion = cum;
				res.failed = q.overflow ? 1 : 0;
			}
			l.coded[bi] = res;
		}
		__syncwarp();
	}
}
} // namespace

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
	const size_t per_warp = 32 + (size_t)l.max_flag_words * 4 + (size_t)l.max_area * 4;
	const size_t fixed = 94 * 4 + 1024 + 2048 + 256;
	int nw = ENC_MAX_THREADS / 32;
	while(nw > 1 && fixed + per_warp * nw > 100 * 1024)
		--nw;
	const size_t smem = fixed + per_warp * nw;
	if(smem > 227 * 1024)
	{
		grkb_set_error("t1 encode: code block too large for shared memory (%zu bytes)", smem);
		return -1;
	}
	GRKB_ENSURE_SMEM(t1_mq_encode_kernel, smem);
	int dev = 0, sms = 148;
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	const int per_sm = (int)((227 * 1024) / (smem + 1024)) > 0 ? (int)((227 * 1024) / (smem + 1024)) : 1;
	uint32_t ctas = (uint32_t)(sms * per_sm);
	const uint32_t need = (l.nblocks + nw - 1) / nw;
	if(ctas > need)
		ctas = need;
	GRKB_CUDA_CHECK(cudaMemsetAsync(l.counter, 0, sizeof(uint32_t), s));
	t1_mq_encode_kernel<<<ctas, nw * 32, smem, s>>>(l, tab);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
