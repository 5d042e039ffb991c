/*
 * t1_mq.c -- CPU restatement of Grok's EBCOT (JPEG 2000 Part 1) block coder: context
 * formation, the three coding passes and the MQ / RAW coder, decode and encode.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * Restates (paths relative to /root/reference/src/lib/jp2/t1/t1_part1):
 *   T1::decompress_cblk          T1.cpp:1365-1446   (pass sequencing, LAZY/RESET/SEGSYM handling)
 *   dec_clnpass / sigpass / refpass  T1.cpp:934-1364
 *   T1::compress_cblk            T1.cpp:781-932     (rates, termination, distortion)
 *   enc_sigpass / refpass / clnpass  T1.cpp:459-780
 *   MQ decoder                   mqc_dec_inl.h:25-180, mqc_dec.cpp:78-130
 *   MQ encoder                   mqc_enc.cpp:86-332, mqc_enc_inl.h:38-124
 *   context rules                t1_generate_luts.cpp (ISO 15444-1 tables D.1-D.4)
 *
 * Deliberately NOT organised like the reference: the reference packs the state of a
 * 4-sample stripe column into one 32-bit word and indexes pre-computed tables; here every
 * sample has its own state bytes in a bordered array and the contexts are evaluated from the
 * rules of ISO 15444-1 Annex D each time.  Decisions, bytes, rates and distortions must come
 * out identical (tests/test_oracle_vs_ref.py pins that against the compiled reference).
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>
#include <math.h>

/* ---- MQ probability state machine, ISO 15444-1 table C.2 ---------------------------------- */

typedef struct
{
	uint16_t qe;
	uint8_t nmps, nlps, sw;
} mq_row;

static const mq_row MQ[47] = {
	{0x5601, 1, 1, 1},	 {0x3401, 2, 6, 0},	  {0x1801, 3, 9, 0},   {0x0AC1, 4, 12, 0},
	{0x0521, 5, 29, 0},	 {0x0221, 38, 33, 0}, {0x5601, 7, 6, 1},   {0x5401, 8, 14, 0},
	{0x4801, 9, 14, 0},	 {0x3801, 10, 14, 0}, {0x3001, 11, 17, 0}, {0x2401, 12, 18, 0},
	{0x1C01, 13, 20, 0}, {0x1601, 29, 21, 0}, {0x5601, 15, 14, 1}, {0x5401, 16, 14, 0},
	{0x5101, 17, 15, 0}, {0x4801, 18, 16, 0}, {0x3801, 19, 17, 0}, {0x3401, 20, 18, 0},
	{0x3001, 21, 19, 0}, {0x2801, 22, 19, 0}, {0x2401, 23, 20, 0}, {0x2201, 24, 21, 0},
	{0x1C01, 25, 22, 0}, {0x1801, 26, 23, 0}, {0x1601, 27, 24, 0}, {0x1401, 28, 25, 0},
	{0x1201, 29, 26, 0}, {0x1101, 30, 27, 0}, {0x0AC1, 31, 28, 0}, {0x09C1, 32, 29, 0},
	{0x08A1, 33, 30, 0}, {0x0521, 34, 31, 0}, {0x0441, 35, 32, 0}, {0x02A1, 36, 33, 0},
	{0x0221, 37, 34, 0}, {0x0141, 38, 35, 0}, {0x0111, 39, 36, 0}, {0x0085, 40, 37, 0},
	{0x0049, 41, 38, 0}, {0x0025, 42, 39, 0}, {0x0015, 43, 40, 0}, {0x0009, 44, 41, 0},
	{0x0005, 45, 42, 0}, {0x0001, 45, 43, 0}, {0x5601, 46, 46, 0}};

const uint16_t* orc_mq_table(void)
{
	/* flattened copy for the tests: qe, nmps, nlps, switch per row */
	static uint16_t flat[47 * 4];
	for(int i = 0; i < 47; ++i)
	{
		flat[4 * i] = MQ[i].qe;
		flat[4 * i + 1] = MQ[i].nmps;
		flat[4 * i + 2] = MQ[i].nlps;
		flat[4 * i + 3] = MQ[i].sw;
	}
	return flat;
}

enum
{
	CTX_ZC = 0,
	CTX_SC = 9,
	CTX_MAG = 14,
	CTX_AGG = 17,
	CTX_UNI = 18,
	NCTX = 19
};

typedef struct
{
	uint8_t idx[NCTX]; /* row of MQ[] */
	uint8_t mps[NCTX];
	/* decoder registers (mqc.h:30-67) */
	uint32_t a, c, ct;
	const uint8_t* in;
	uint32_t in_len, in_pos;
	/* encoder: out[-1] must exist and be 0 */
	uint8_t* out;
	int64_t bp; /* index into out of the byte being built; starts at -1 */
} mq_t;

/* mqc_resetstates, mqc_dec.cpp:121-130 */
static void mq_reset(mq_t* q)
{
	memset(q->idx, 0, sizeof(q->idx));
	memset(q->mps, 0, sizeof(q->mps));
	q->idx[CTX_UNI] = 46;
	q->idx[CTX_AGG] = 3;
	q->idx[CTX_ZC] = 4;
}

/* the reference plants 0xFF 0xFF after the segment (mqc_dec.cpp:78-90): reading at or past
 * the end yields 0xFF */
static inline uint32_t in_byte(const mq_t* q, uint32_t pos)
{
	return pos < q->in_len ? q->in[pos] : 0xFFu;
}

/* bytein_dec_macro, mqc_dec_inl.h:82-108 */
static void mq_bytein(mq_t* q)
{
	uint32_t next = in_byte(q, q->in_pos + 1);
	if(in_byte(q, q->in_pos) == 0xFF)
	{
		if(next > 0x8F)
		{
			q->c += 0xFF00;
			q->ct = 8;
		}
		else
		{
			q->in_pos++;
			q->c += next << 9;
			q->ct = 7;
		}
	}
	else
	{
		q->in_pos++;
		q->c += next << 8;
		q->ct = 8;
	}
}

/* mqc_init_dec, mqc_dec.cpp:91-106 */
static void mq_init_dec(mq_t* q, const uint8_t* data, uint32_t len)
{
	q->in = data;
	q->in_len = len;
	q->in_pos = 0;
	q->c = in_byte(q, 0) << 16;
	mq_bytein(q);
	q->c <<= 7;
	q->ct -= 7;
	q->a = 0x8000;
}

/* mqc_raw_init_dec, mqc_dec.cpp:108-113 */
static void raw_init_dec(mq_t* q, const uint8_t* data, uint32_t len)
{
	q->in = data;
	q->in_len = len;
	q->in_pos = 0;
	q->c = 0;
	q->ct = 0;
}

/* mqc_raw_decode, mqc_dec_inl.h:61-91 */
static uint32_t raw_decode(mq_t* q)
{
	if(q->ct == 0)
	{
		if(q->c == 0xFF)
		{
			if(in_byte(q, q->in_pos) > 0x8F)
			{
				q->c = 0xFF;
				q->ct = 8;
			}
			else
			{
				q->c = in_byte(q, q->in_pos);
				q->in_pos++;
				q->ct = 7;
			}
		}
		else
		{
			q->c = in_byte(q, q->in_pos);
			q->in_pos++;
			q->ct = 8;
		}
	}
	q->ct--;
	return (q->c >> q->ct) & 1u;
}

static void mq_renorm_dec(mq_t* q)
{
	do
	{
		if(q->ct == 0)
			mq_bytein(q);
		q->a <<= 1;
		q->c <<= 1;
		q->ct--;
	} while(q->a < 0x8000);
}

/* decompress_macro, mqc_dec_inl.h:134-158 (ISO 15444-1 C.3.2) */
static uint32_t mq_decode(mq_t* q, int cx)
{
	const mq_row* s = &MQ[q->idx[cx]];
	uint32_t d;
	q->a -= s->qe;
	if(q->c < ((uint32_t)s->qe << 16))
	{
		/* LPS exchange */
		if(q->a < s->qe)
		{
			d = q->mps[cx];
			q->idx[cx] = s->nmps;
		}
		else
		{
			d = q->mps[cx] ^ 1;
			if(s->sw)
				q->mps[cx] ^= 1;
			q->idx[cx] = s->nlps;
		}
		q->a = s->qe;
		mq_renorm_dec(q);
	}
	else
	{
		q->c -= (uint32_t)s->qe << 16;
		if(q->a < 0x8000)
		{
			/* MPS exchange */
			if(q->a < s->qe)
			{
				d = q->mps[cx] ^ 1;
				if(s->sw)
					q->mps[cx] ^= 1;
				q->idx[cx] = s->nlps;
			}
			else
			{
				d = q->mps[cx];
				q->idx[cx] = s->nmps;
			}
			mq_renorm_dec(q);
		}
		else
			d = q->mps[cx];
	}
	return d;
}

/* ---- MQ encoder ---------------------------------------------------------------------------- */

/* mqc_byteout, mqc_enc.cpp:86-125 */
static void mq_byteout(mq_t* q)
{
	if(q->out[q->bp] == 0xFF)
	{
		q->bp++;
		q->out[q->bp] = (uint8_t)(q->c >> 20);
		q->c &= 0xFFFFF;
		q->ct = 7;
	}
	else if((q->c & 0x8000000) == 0)
	{
		q->bp++;
		q->out[q->bp] = (uint8_t)(q->c >> 19);
		q->c &= 0x7FFFF;
		q->ct = 8;
	}
	else
	{
		q->out[q->bp]++;
		if(q->out[q->bp] == 0xFF)
		{
			q->c &= 0x7FFFFFF;
			q->bp++;
			q->out[q->bp] = (uint8_t)(q->c >> 20);
			q->c &= 0xFFFFF;
			q->ct = 7;
		}
		else
		{
			q->bp++;
			q->out[q->bp] = (uint8_t)(q->c >> 19);
			q->c &= 0x7FFFF;
			q->ct = 8;
		}
	}
}

static void mq_renorm_enc(mq_t* q)
{
	do
	{
		q->a <<= 1;
		q->c <<= 1;
		q->ct--;
		if(q->ct == 0)
			mq_byteout(q);
	} while((q->a & 0x8000) == 0);
}

/* mqc_encode_macro, mqc_enc_inl.h:38-102 */
static void mq_encode(mq_t* q, int cx, uint32_t d)
{
	const mq_row* s = &MQ[q->idx[cx]];
	q->a -= s->qe;
	if(q->mps[cx] == d)
	{
		if((q->a & 0x8000) == 0)
		{
			if(q->a < s->qe)
				q->a = s->qe;
			else
				q->c += s->qe;
			q->idx[cx] = s->nmps;
			mq_renorm_enc(q);
		}
		else
			q->c += s->qe;
	}
	else
	{
		if(q->a < s->qe)
			q->c += s->qe;
		else
			q->a = s->qe;
		if(s->sw)
			q->mps[cx] ^= 1;
		q->idx[cx] = s->nlps;
		mq_renorm_enc(q);
	}
}

/* mqc_init_enc, mqc_enc.cpp:182-203 */
static void mq_init_enc(mq_t* q, uint8_t* out)
{
	q->out = out;
	q->a = 0x8000;
	q->c = 0;
	q->bp = -1;
	q->ct = 12;
}

/* mqc_flush_enc, mqc_enc.cpp:213-227 */
static void mq_flush(mq_t* q)
{
	uint32_t tempc = q->c + q->a;
	q->c |= 0xFFFF;
	if(q->c >= tempc)
		q->c -= 0x8000;
	q->c <<= q->ct;
	mq_byteout(q);
	q->c <<= q->ct;
	mq_byteout(q);
	if(q->out[q->bp] != 0xFF)
		q->bp++;
}

#define BYPASS_CT_INIT 0xDEADBEEFu

/* mqc_bypass_init_enc, mqc_enc.cpp:229-245 */
static void raw_init_enc(mq_t* q)
{
	q->c = 0;
	q->ct = BYPASS_CT_INIT;
}

/* mqc_bypass_enc_macro, mqc_enc_inl.h:104-124 */
static void raw_encode(mq_t* q, uint32_t d)
{
	if(q->ct == BYPASS_CT_INIT)
		q->ct = 8;
	q->ct--;
	q->c += d << q->ct;
	if(q->ct == 0)
	{
		q->out[q->bp] = (uint8_t)q->c;
		q->ct = 8;
		if(q->out[q->bp] == 0xFF)
			q->ct = 7;
		q->bp++;
		q->c = 0;
	}
}

/* mqc_bypass_get_extra_bytes_enc, mqc_enc.cpp:247-250 */
static uint32_t raw_extra_bytes(const mq_t* q, int erterm)
{
	return (q->ct < 7 || (q->ct == 7 && (erterm || q->out[q->bp - 1] != 0xFF))) ? 2 : 1;
}

/* mqc_bypass_flush_enc, mqc_enc.cpp:252-292 */
static void raw_flush(mq_t* q, int erterm)
{
	if(q->ct < 7 || (q->ct == 7 && (erterm || q->out[q->bp - 1] != 0xFF)))
	{
		uint32_t bit = 0;
		while(q->ct > 0)
		{
			q->ct--;
			q->c += bit << q->ct;
			bit = 1 - bit;
		}
		q->out[q->bp] = (uint8_t)q->c;
		q->bp++;
	}
	else if(q->ct == 7 && q->out[q->bp - 1] == 0xFF)
		q->bp--;
	else if(q->ct == 8 && !erterm && q->out[q->bp - 1] == 0x7F && q->out[q->bp - 2] == 0xFF)
		q->bp -= 2;
}

/* mqc_restart_init_enc, mqc_enc.cpp:294-310 */
static void mq_restart_enc(mq_t* q)
{
	q->a = 0x8000;
	q->c = 0;
	q->ct = 12;
	q->bp--;
	if(q->out[q->bp] == 0xFF)
		q->ct = 13;
}

/* mqc_erterm_enc, mqc_enc.cpp:312-325 */
static void mq_erterm(mq_t* q)
{
	int32_t k = (int32_t)(11 - q->ct + 1);
	while(k > 0)
	{
		q->c <<= q->ct;
		q->ct = 0;
		mq_byteout(q);
		k -= (int32_t)q->ct;
	}
	if(q->out[q->bp] != 0xFF)
		mq_byteout(q);
}

/* ---- per-sample state and context rules ----------------------------------------------------- */

typedef struct
{
	uint32_t w, h, sw; /* sw = w + 2: row pitch of the bordered state arrays */
	uint8_t *sig, *neg, *pi, *mu; /* (w+2) x (h+2), sample (x,y) at [(y+1)*sw + x+1] */
	uint32_t orient;
	int vsc;
} blk_t;

static int blk_alloc(blk_t* b, uint32_t w, uint32_t h)
{
	size_t n = (size_t)(w + 2) * (h + 2);
	b->w = w;
	b->h = h;
	b->sw = w + 2;
	b->sig = (uint8_t*)calloc(4 * n, 1);
	if(!b->sig)
		return -1;
	b->neg = b->sig + n;
	b->pi = b->neg + n;
	b->mu = b->pi + n;
	return 0;
}

static void blk_free(blk_t* b)
{
	free(b->sig);
}

/* significance of the neighbour at (dx,dy) as the coding passes see it.  With vertically causal
 * contexts a sample in the last row of a stripe does not see the next stripe
 * (update_flags_macro, T1.cpp:209-233: the north propagation is skipped when vsc). */
static inline int nsig(const blk_t* b, uint32_t x, uint32_t y, int dx, int dy)
{
	if(dy > 0 && b->vsc && (y & 3) == 3)
		return 0;
	return b->sig[(size_t)(y + 1 + dy) * b->sw + (x + 1 + dx)];
}
static inline int nneg(const blk_t* b, uint32_t x, uint32_t y, int dx, int dy)
{
	return b->neg[(size_t)(y + 1 + dy) * b->sw + (x + 1 + dx)];
}

static int any_neighbour(const blk_t* b, uint32_t x, uint32_t y)
{
	return nsig(b, x, y, -1, -1) | nsig(b, x, y, 0, -1) | nsig(b, x, y, 1, -1) |
		   nsig(b, x, y, -1, 0) | nsig(b, x, y, 1, 0) | nsig(b, x, y, -1, 1) | nsig(b, x, y, 0, 1) |
		   nsig(b, x, y, 1, 1);
}

/* zero-coding context, ISO table D.1 (t1_generate_luts.cpp:27-138) */
static int ctx_zc(const blk_t* b, uint32_t x, uint32_t y)
{
	int hh = nsig(b, x, y, -1, 0) + nsig(b, x, y, 1, 0);
	int vv = nsig(b, x, y, 0, -1) + nsig(b, x, y, 0, 1);
	int d = nsig(b, x, y, -1, -1) + nsig(b, x, y, 1, -1) + nsig(b, x, y, -1, 1) + nsig(b, x, y, 1, 1);
	int n;
	if(b->orient == 3)
	{
		int hv = hh + vv;
		if(d == 0)
			n = hv == 0 ? 0 : hv == 1 ? 1 : 2;
		else if(d == 1)
			n = hv == 0 ? 3 : hv == 1 ? 4 : 5;
		else if(d == 2)
			n = hv == 0 ? 6 : 7;
		else
			n = 8;
	}
	else
	{
		if(b->orient == 1) /* HL: roles of horizontal and vertical swap */
		{
			int t = hh;
			hh = vv;
			vv = t;
		}
		if(hh == 0)
		{
			if(vv == 0)
				n = d == 0 ? 0 : d == 1 ? 1 : 2;
			else
				n = vv == 1 ? 3 : 4;
		}
		else if(hh == 1)
		{
			if(vv == 0)
				n = d == 0 ? 5 : 6;
			else
				n = 7;
		}
		else
			n = 8;
	}
	return CTX_ZC + n;
}

/* sign-coding context and xor bit, ISO table D.3 (t1_generate_luts.cpp:139-229) */
static int ctx_sc(const blk_t* b, uint32_t x, uint32_t y, uint32_t* xorbit)
{
	int hpos = 0, hneg = 0, vpos = 0, vneg = 0;
	static const int dxs[4] = {-1, 1, 0, 0}, dys[4] = {0, 0, -1, 1};
	for(int k = 0; k < 4; ++k)
	{
		if(!nsig(b, x, y, dxs[k], dys[k]))
			continue;
		int ng = nneg(b, x, y, dxs[k], dys[k]);
		if(k < 2)
		{
			hpos |= !ng;
			hneg |= ng;
		}
		else
		{
			vpos |= !ng;
			vneg |= ng;
		}
	}
	int hc = hpos - hneg, vc = vpos - vneg;
	*xorbit = (hc == 0 && vc == 0) ? 0 : !(hc > 0 || (hc == 0 && vc > 0));
	if(hc < 0)
	{
		hc = -hc;
		vc = -vc;
	}
	int n;
	if(hc == 0)
		n = vc == 0 ? 0 : 1;
	else
		n = vc == -1 ? 2 : vc == 0 ? 3 : 4;
	return CTX_SC + n;
}

/* magnitude-refinement context, ISO table D.4 (T1.cpp:187-193) */
static int ctx_mag(const blk_t* b, uint32_t x, uint32_t y)
{
	if(b->mu[(size_t)(y + 1) * b->sw + x + 1])
		return CTX_MAG + 2;
	return any_neighbour(b, x, y) ? CTX_MAG + 1 : CTX_MAG;
}

static inline void set_sig(blk_t* b, uint32_t x, uint32_t y, uint32_t negative)
{
	size_t i = (size_t)(y + 1) * b->sw + x + 1;
	b->sig[i] = 1;
	b->neg[i] = (uint8_t)negative;
}

/* ---- decoder passes -------------------------------------------------------------------------- */

/* dec_sigpass_mqc / dec_sigpass_raw, T1.cpp:1094-1245 */
static void dec_sigpass(blk_t* b, mq_t* q, int32_t* data, int32_t bpno_plus_one, int raw)
{
	const int32_t one = 1 << bpno_plus_one, oneplushalf = one | (one >> 1);
	for(uint32_t k = 0; k < b->h; k += 4)
		for(uint32_t x = 0; x < b->w; ++x)
			for(uint32_t y = k; y < k + 4 && y < b->h; ++y)
			{
				size_t i = (size_t)(y + 1) * b->sw + x + 1;
				if(b->sig[i] || b->pi[i] || !any_neighbour(b, x, y))
					continue;
				uint32_t v;
				if(raw)
				{
					if(raw_decode(q))
					{
						v = raw_decode(q);
						data[(size_t)y * b->w + x] = v ? -oneplushalf : oneplushalf;
						set_sig(b, x, y, v);
					}
				}
				else if(mq_decode(q, ctx_zc(b, x, y)))
				{
					uint32_t xr;
					int cx = ctx_sc(b, x, y, &xr);
					v = mq_decode(q, cx) ^ xr;
					data[(size_t)y * b->w + x] = v ? -oneplushalf : oneplushalf;
					set_sig(b, x, y, v);
				}
				b->pi[i] = 1;
			}
}

/* dec_refpass_mqc / dec_refpass_raw, T1.cpp:1246-1364 */
static void dec_refpass(blk_t* b, mq_t* q, int32_t* data, int32_t bpno_plus_one, int raw)
{
	const int32_t poshalf = (1 << bpno_plus_one) >> 1;
	for(uint32_t k = 0; k < b->h; k += 4)
		for(uint32_t x = 0; x < b->w; ++x)
			for(uint32_t y = k; y < k + 4 && y < b->h; ++y)
			{
				size_t i = (size_t)(y + 1) * b->sw + x + 1;
				if(!b->sig[i] || b->pi[i])
					continue;
				uint32_t v = raw ? raw_decode(q) : mq_decode(q, ctx_mag(b, x, y));
				int32_t* d = &data[(size_t)y * b->w + x];
				*d += (v ^ (uint32_t)(*d < 0)) ? poshalf : -poshalf;
				b->mu[i] = 1;
			}
}

/* dec_clnpass, T1.cpp:934-1093 */
static void dec_clnpass(blk_t* b, mq_t* q, int32_t* data, int32_t bpno_plus_one, uint32_t cblksty)
{
	const int32_t one = 1 << bpno_plus_one, oneplushalf = one | (one >> 1);
	for(uint32_t k = 0; k < b->h; k += 4)
		for(uint32_t x = 0; x < b->w; ++x)
		{
			uint32_t y = k;
			int first_is_sign_only = 0;
			if(k + 4 <= b->h)
			{
				/* run-length mode: four insignificant, not yet coded samples with empty
				 * neighbourhoods (the reference tests its packed word against 0) */
				int rl = 1;
				for(uint32_t yy = k; yy < k + 4 && rl; ++yy)
				{
					size_t i = (size_t)(yy + 1) * b->sw + x + 1;
					if(b->sig[i] || b->pi[i] || any_neighbour(b, x, yy))
						rl = 0;
				}
				if(rl)
				{
					if(!mq_decode(q, CTX_AGG))
						continue;
					uint32_t r = mq_decode(q, CTX_UNI);
					r = (r << 1) | mq_decode(q, CTX_UNI);
					y = k + r;
					first_is_sign_only = 1;
				}
			}
			for(; y < k + 4 && y < b->h; ++y)
			{
				size_t i = (size_t)(y + 1) * b->sw + x + 1;
				if(!first_is_sign_only)
				{
					if(b->sig[i] || b->pi[i])
						continue;
					if(!mq_decode(q, ctx_zc(b, x, y)))
						continue;
				}
				first_is_sign_only = 0;
				uint32_t xr;
				int cx = ctx_sc(b, x, y, &xr);
				uint32_t v = mq_decode(q, cx) ^ xr;
				data[(size_t)y * b->w + x] = v ? -oneplushalf : oneplushalf;
				set_sig(b, x, y, v);
			}
		}
	/* the cleanup pass ends a bit-plane: nothing stays marked "visited" */
	memset(b->pi, 0, (size_t)b->sw * (b->h + 2));
	if(cblksty & ORC_CBLKSTY_SEGSYM)
	{
		/* dec_clnpass_check_segsym, T1.cpp:1053-1073: four UNI decisions are consumed */
		for(int i = 0; i < 4; ++i)
			(void)mq_decode(q, CTX_UNI);
	}
}

int orc_t1_decode(const uint8_t* data, const uint32_t* seg_len, const uint32_t* seg_passes,
				  uint32_t numsegs, uint32_t numbps, uint32_t w, uint32_t h, uint32_t orient,
				  uint32_t cblksty, int32_t* out)
{
	if((int32_t)numbps >= 31 - 6) /* maxBitPlanesGRK, T1.cpp:1373-1377 */
		return -1;
	blk_t b;
	if(blk_alloc(&b, w, h))
		return -1;
	b.orient = orient;
	b.vsc = (cblksty & ORC_CBLKSTY_VSC) != 0;
	memset(out, 0, sizeof(int32_t) * (size_t)w * h);
	mq_t q;
	memset(&q, 0, sizeof(q));
	mq_reset(&q);
	int32_t bpno_plus_one = (int32_t)numbps;
	uint32_t passtype = 2, off = 0;
	for(uint32_t s = 0; s < numsegs; ++s)
	{
		int raw = (bpno_plus_one <= (int32_t)numbps - 4) && passtype < 2 &&
				  (cblksty & ORC_CBLKSTY_LAZY);
		if(raw)
			raw_init_dec(&q, data + off, seg_len[s]);
		else
			mq_init_dec(&q, data + off, seg_len[s]);
		off += seg_len[s];
		for(uint32_t p = 0; p < seg_passes[s] && bpno_plus_one >= 1; ++p)
		{
			if(passtype == 0)
				dec_sigpass(&b, &q, out, bpno_plus_one, raw);
			else if(passtype == 1)
				dec_refpass(&b, &q, out, bpno_plus_one, raw);
			else
				dec_clnpass(&b, &q, out, bpno_plus_one, cblksty);
			if((cblksty & ORC_CBLKSTY_RESET) && !raw)
				mq_reset(&q);
			if(++passtype == 3)
			{
				passtype = 0;
				bpno_plus_one--;
			}
		}
	}
	blk_free(&b);
	return 0;
}

/* ---- distortion tables (t1_generate_luts.cpp:300-340) ---------------------------------------- */

static int16_t g_nmsedec[4][128];
static int g_nmsedec_ready = 0;

__attribute__((constructor)) static void nmsedec_init(void)
{
	if(g_nmsedec_ready)
		return;
	for(int i = 0; i < 128; ++i)
	{
		double t = i / 64.0, u = t, v = t - 1.5;
		double r;
		r = floor((u * u - v * v) * 64.0 + 0.5) / 64.0 * 8192.0;
		g_nmsedec[0][i] = (int16_t)((int)r > 0 ? (int)r : 0);
		r = floor((u * u) * 64.0 + 0.5) / 64.0 * 8192.0;
		g_nmsedec[1][i] = (int16_t)((int)r > 0 ? (int)r : 0);
		u = t - 1.0;
		v = (i & 64) ? t - 1.5 : t - 0.5;
		r = floor((u * u - v * v) * 64.0 + 0.5) / 64.0 * 8192.0;
		g_nmsedec[2][i] = (int16_t)((int)r > 0 ? (int)r : 0);
		r = floor((u * u) * 64.0 + 0.5) / 64.0 * 8192.0;
		g_nmsedec[3][i] = (int16_t)((int)r > 0 ? (int)r : 0);
	}
	g_nmsedec_ready = 1;
}

const int16_t* orc_lut_nmsedec(int which)
{
	nmsedec_init();
	return g_nmsedec[which & 3];
}

/* getnmsedec_sig / getnmsedec_ref, T1.cpp:194-208 */
static int32_t nmsedec_sig(uint32_t mag, uint32_t bitpos)
{
	return bitpos > 0 ? g_nmsedec[0][(mag >> bitpos) & 127] : g_nmsedec[1][mag & 127];
}
static int32_t nmsedec_ref(uint32_t mag, uint32_t bitpos)
{
	return bitpos > 0 ? g_nmsedec[2][(mag >> bitpos) & 127] : g_nmsedec[3][mag & 127];
}

/* T1.cpp:239-248 */
static const double NORMS_53[4][10] = {
	{1.000, 1.500, 2.750, 5.375, 10.68, 21.34, 42.67, 85.33, 170.7, 341.3},
	{1.038, 1.592, 2.919, 5.703, 11.33, 22.64, 45.25, 90.48, 180.9, 0},
	{1.038, 1.592, 2.919, 5.703, 11.33, 22.64, 45.25, 90.48, 180.9, 0},
	{.7186, .9218, 1.586, 3.043, 6.019, 12.01, 24.00, 47.97, 95.93, 0}};
static const double NORMS_97[4][10] = {
	{1.000, 1.965, 4.177, 8.403, 16.90, 33.84, 67.69, 135.3, 270.6, 540.9},
	{2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
	{2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
	{2.080, 3.865, 8.307, 17.18, 34.71, 69.59, 139.3, 278.6, 557.2, 0}};

double orc_dwt_norm(uint32_t level, uint32_t orient, int reversible)
{
	/* T1::getnorm, T1.cpp:264-277 */
	if(orient == 0 && level > 9)
		level = 9;
	else if(orient > 0 && level > 8)
		level = 8;
	return reversible ? NORMS_53[orient][level] : NORMS_97[orient][level];
}

/* T1::getwmsedec, T1.cpp:418-436 */
static double wmsedec(int32_t nmsedec, uint32_t compno, uint32_t level, uint32_t orient, int32_t bpno,
					  uint32_t qmfbid, double stepsize, const double* mct_norms,
					  uint32_t mct_numcomps)
{
	double w1 = 1;
	if(mct_norms && compno < mct_numcomps)
		w1 = mct_norms[compno];
	double w2 = orc_dwt_norm(level, orient, qmfbid == 1);
	double wm = w1 * w2 * stepsize * (1 << bpno);
	wm *= wm * nmsedec / 8192.0;
	return wm;
}

/* ---- encoder passes ---------------------------------------------------------------------------- */

#define MAGOF(v) ((uint32_t)(v) & 0x7FFFFFFFu)
#define SGNOF(v) ((uint32_t)(v) >> 31)

/* enc_sigpass, T1.cpp:459-551 */
static void enc_sigpass(blk_t* b, mq_t* q, const int32_t* data, int32_t bpno, int32_t* nmsedec,
						int raw)
{
	const uint32_t one = 1u << (bpno + 6);
	*nmsedec = 0;
	for(uint32_t k = 0; k < b->h; k += 4)
		for(uint32_t x = 0; x < b->w; ++x)
			for(uint32_t y = k; y < k + 4 && y < b->h; ++y)
			{
				size_t i = (size_t)(y + 1) * b->sw + x + 1;
				if(b->sig[i] || b->pi[i] || !any_neighbour(b, x, y))
					continue;
				int32_t s = data[(size_t)y * b->w + x];
				uint32_t v = (MAGOF(s) & one) != 0;
				if(raw)
					raw_encode(q, v);
				else
					mq_encode(q, ctx_zc(b, x, y), v);
				if(v)
				{
					uint32_t xr;
					int cx = ctx_sc(b, x, y, &xr);
					*nmsedec += nmsedec_sig(MAGOF(s), (uint32_t)bpno);
					if(raw)
						raw_encode(q, SGNOF(s));
					else
						mq_encode(q, cx, SGNOF(s) ^ xr);
					set_sig(b, x, y, SGNOF(s));
				}
				b->pi[i] = 1;
			}
}

/* enc_refpass, T1.cpp:552-630 */
static void enc_refpass(blk_t* b, mq_t* q, const int32_t* data, int32_t bpno, int32_t* nmsedec,
						int raw)
{
	const uint32_t one = 1u << (bpno + 6);
	*nmsedec = 0;
	for(uint32_t k = 0; k < b->h; k += 4)
		for(uint32_t x = 0; x < b->w; ++x)
			for(uint32_t y = k; y < k + 4 && y < b->h; ++y)
			{
				size_t i = (size_t)(y + 1) * b->sw + x + 1;
				if(!b->sig[i] || b->pi[i])
					continue;
				int32_t s = data[(size_t)y * b->w + x];
				int cx = ctx_mag(b, x, y);
				*nmsedec += nmsedec_ref(MAGOF(s), (uint32_t)bpno);
				uint32_t v = (MAGOF(s) & one) != 0;
				if(raw)
					raw_encode(q, v);
				else
					mq_encode(q, cx, v);
				b->mu[i] = 1;
			}
}

/* enc_clnpass, T1.cpp:631-780 */
static void enc_clnpass(blk_t* b, mq_t* q, const int32_t* data, int32_t bpno, int32_t* nmsedec)
{
	const uint32_t one = 1u << (bpno + 6);
	*nmsedec = 0;
	for(uint32_t k = 0; k < b->h; k += 4)
		for(uint32_t x = 0; x < b->w; ++x)
		{
			uint32_t y = k;
			int first_is_sign_only = 0;
			if(k + 4 <= b->h)
			{
				int rl = 1;
				for(uint32_t yy = k; yy < k + 4 && rl; ++yy)
				{
					size_t i = (size_t)(yy + 1) * b->sw + x + 1;
					if(b->sig[i] || b->pi[i] || any_neighbour(b, x, yy))
						rl = 0;
				}
				if(rl)
				{
					uint32_t r = 0;
					while(r < 4 && !(MAGOF(data[(size_t)(k + r) * b->w + x]) & one))
						++r;
					mq_encode(q, CTX_AGG, r != 4);
					if(r == 4)
						continue;
					mq_encode(q, CTX_UNI, r >> 1);
					mq_encode(q, CTX_UNI, r & 1);
					y = k + r;
					first_is_sign_only = 1;
				}
			}
			for(; y < k + 4 && y < b->h; ++y)
			{
				size_t i = (size_t)(y + 1) * b->sw + x + 1;
				int32_t s = data[(size_t)y * b->w + x];
				if(!first_is_sign_only)
				{
					if(b->sig[i] || b->pi[i])
						continue;
					uint32_t v = (MAGOF(s) & one) != 0;
					mq_encode(q, ctx_zc(b, x, y), v);
					if(!v)
						continue;
				}
				first_is_sign_only = 0;
				uint32_t xr;
				int cx = ctx_sc(b, x, y, &xr);
				*nmsedec += nmsedec_sig(MAGOF(s), (uint32_t)bpno);
				mq_encode(q, cx, SGNOF(s) ^ xr);
				set_sig(b, x, y, SGNOF(s));
			}
		}
	memset(b->pi, 0, (size_t)b->sw * (b->h + 2));
}

/* T1::enc_is_term_pass, T1.cpp:437-458 */
static int is_term_pass(uint32_t numbps, uint32_t cblksty, int32_t bpno, uint32_t passtype)
{
	if(passtype == 2 && bpno == 0)
		return 1;
	if(cblksty & ORC_CBLKSTY_TERMALL)
		return 1;
	if(cblksty & ORC_CBLKSTY_LAZY)
	{
		if(bpno == (int32_t)numbps - 4 && passtype == 2)
			return 1;
		if(bpno < (int32_t)numbps - 4 && passtype > 0)
			return 1;
	}
	return 0;
}

static uint32_t floorlog2_u32(uint32_t v)
{
	uint32_t l = 0;
	while(v > 1)
	{
		v >>= 1;
		++l;
	}
	return l;
}

double orc_t1_encode(const int32_t* samples, uint32_t w, uint32_t h, uint32_t max, uint32_t orient,
					 uint32_t compno, uint32_t level, uint32_t qmfbid, double stepsize,
					 uint32_t cblksty, const double* mct_norms, uint32_t mct_numcomps,
					 int doRateControl, uint8_t* out, orc_enc_result* res)
{
	nmsedec_init();
	memset(res, 0, sizeof(*res));
	if(max)
	{
		uint32_t t = floorlog2_u32(max) + 1;
		res->numbps = t <= 6 ? 0 : t - 6;
	}
	if(res->numbps == 0)
		return 0;
	blk_t b;
	if(blk_alloc(&b, w, h))
		return 0;
	b.orient = orient;
	b.vsc = (cblksty & ORC_CBLKSTY_VSC) != 0;
	/* two zero bytes precede the stream (t1/Codeblock.h:156-172) */
	size_t cap = (size_t)w * h * 4 + 64 + 2;
	uint8_t* buf = (uint8_t*)calloc(cap, 1);
	mq_t q;
	memset(&q, 0, sizeof(q));
	mq_reset(&q);
	mq_init_enc(&q, buf + 2);

	int32_t bpno = (int32_t)res->numbps - 1;
	uint32_t passtype = 2, passno;
	double cum = 0;
	for(passno = 0; bpno >= 0; ++passno)
	{
		orc_pass* pass = &res->passes[passno];
		int raw = (bpno < (int32_t)res->numbps - 4) && passtype < 2 && (cblksty & ORC_CBLKSTY_LAZY);
		if(passno > 0 && res->passes[passno - 1].term)
		{
			if(raw)
				raw_init_enc(&q);
			else
				mq_restart_enc(&q);
		}
		int32_t nmsedec = 0;
		if(passtype == 0)
			enc_sigpass(&b, &q, samples, bpno, &nmsedec, raw);
		else if(passtype == 1)
			enc_refpass(&b, &q, samples, bpno, &nmsedec, raw);
		else
		{
			enc_clnpass(&b, &q, samples, bpno, &nmsedec);
			if(cblksty & ORC_CBLKSTY_SEGSYM)
			{
				/* mqc_segmark_enc, mqc_enc.cpp:327-332 */
				for(uint32_t i = 1; i < 5; ++i)
					mq_encode(&q, CTX_UNI, i % 2);
			}
		}
		if(doRateControl)
		{
			cum += wmsedec(nmsedec, compno, level, orient, bpno, qmfbid, stepsize, mct_norms,
						   mct_numcomps);
			pass->distortiondec = cum;
		}
		if(is_term_pass(res->numbps, cblksty, bpno, passtype))
		{
			if(raw)
				raw_flush(&q, (cblksty & ORC_CBLKSTY_PTERM) != 0);
			else if(cblksty & ORC_CBLKSTY_PTERM)
				mq_erterm(&q);
			else
				mq_flush(&q);
			pass->term = 1;
			pass->rate = (uint32_t)q.bp;
		}
		else
		{
			uint32_t extra;
			if(raw)
				extra = raw_extra_bytes(&q, (cblksty & ORC_CBLKSTY_PTERM) != 0);
			else
			{
				extra = 5;
				if(q.ct < 5)
					extra++;
			}
			pass->term = 0;
			pass->rate = (uint32_t)q.bp + extra;
		}
		if(++passtype == 3)
		{
			passtype = 0;
			bpno--;
		}
		if(cblksty & ORC_CBLKSTY_RESET)
			mq_reset(&q);
	}
	res->numPassesTotal = passno;
	uint32_t last = (uint32_t)q.bp;
	for(uint32_t p = passno; p > 0;)
	{
		orc_pass* pass = &res->passes[--p];
		if(pass->rate > last)
			pass->rate = last;
		else
			last = pass->rate;
	}
	const uint8_t* stream = buf + 2;
	for(uint32_t p = 0; p < passno; ++p)
	{
		orc_pass* pass = &res->passes[p];
		if(stream[pass->rate - 1] == 0xFF)
			pass->rate--;
		pass->len = pass->rate - (p == 0 ? 0 : res->passes[p - 1].rate);
	}
	res->num_bytes = passno ? res->passes[passno - 1].rate : 0;
	memcpy(out, stream, res->num_bytes);
	free(buf);
	blk_free(&b);
	return cum;
}

/* ---- context LUTs in the reference's packed indexing, for cross-checks only ---------------- */

/* index = 9 neighbour-significance bits (NW N NE W . E SW S SE at bits 0..8) + slot<<9 where
 * slot is the band orientation the decoder uses (t1_luts.h:3, slot 1 = HL swaps h and v) */
const uint8_t* orc_lut_ctxno_zc(void)
{
	static uint8_t lut[2048];
	for(uint32_t slot = 0; slot < 4; ++slot)
		for(uint32_t f = 0; f < 512; ++f)
		{
			blk_t b;
			uint8_t st[4][9];
			memset(st, 0, sizeof(st));
			b.w = 1;
			b.h = 1;
			b.sw = 3;
			b.sig = st[0];
			b.neg = st[1];
			b.pi = st[2];
			b.mu = st[3];
			b.orient = slot;
			b.vsc = 0;
			for(int bit = 0; bit < 9; ++bit)
				if(bit != 4)
					st[0][bit] = (f >> bit) & 1;
			lut[(slot << 9) | f] = (uint8_t)ctx_zc(&b, 0, 0);
		}
	return lut;
}

/* index bits: SGN_W 1, SIG_N 2, SGN_E 4, SIG_W 8, SGN_N 16, SIG_E 32, SGN_S 64, SIG_S 128
 * (T1.cpp:118-125); which = 0 context number, 1 xor bit */
static const uint8_t* sc_lut(int which)
{
	static uint8_t lut[2][256];
	for(uint32_t f = 0; f < 256; ++f)
	{
		blk_t b;
		uint8_t st[4][9];
		memset(st, 0, sizeof(st));
		b.w = 1;
		b.h = 1;
		b.sw = 3;
		b.sig = st[0];
		b.neg = st[1];
		b.pi = st[2];
		b.mu = st[3];
		b.orient = 0;
		b.vsc = 0;
		st[0][3] = (f >> 3) & 1;
		st[1][3] = f & 1;
		st[0][5] = (f >> 5) & 1;
		st[1][5] = (f >> 2) & 1;
		st[0][1] = (f >> 1) & 1;
		st[1][1] = (f >> 4) & 1;
		st[0][7] = (f >> 7) & 1;
		st[1][7] = (f >> 6) & 1;
		uint32_t xr;
		lut[0][f] = (uint8_t)ctx_sc(&b, 0, 0, &xr);
		lut[1][f] = (uint8_t)xr;
	}
	return lut[which];
}
const uint8_t* orc_lut_ctxno_sc(void)
{
	return sc_lut(0);
}
const uint8_t* orc_lut_spb(void)
{
	return sc_lut(1);
}
