/*
 * t1_ht.c -- CPU restatement of Grok's HTJ2K (ISO/IEC 15444-15) block coder: the cleanup pass
 * decoder (MEL, CxtVLC, U-VLC, MagSgn) and the cleanup pass encoder.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * Restates (paths relative to /root/reference/src/lib/jp2/t1/t1_ht):
 *   ojph_decode_codeblock     coding/ojph_block_decoder.cpp:989-1626 (cleanup pass)
 *     MEL decoder             :83-285      VLC backward reader   :291-453
 *     U-VLC                   :664-838     MagSgn forward reader :848-971
 *     decode tables           :585-651 (from ISO 15444-15 tables C.3/C.4, here ht_cxtvlc.inc)
 *   ojph_encode_codeblock     coding/ojph_block_encoder.cpp:470-945
 *   wrapper conventions       T1HT.cpp:56-187
 *
 * Deliberately NOT organised like the reference, which prefetches 32-bit words into 64-bit
 * accumulators, decodes runs ahead of time and unrolls the eight samples of a quad pair: here
 * the three bit streams are read lazily one byte at a time, MEL symbols are produced on demand,
 * and quads and samples are plain loops over per-column neighbour arrays.  Results on every
 * stream the reference accepts must be identical (tests/test_oracle_vs_ref.py).  SigProp and
 * MagRef passes (:1627-2209; readers :466-580, :875-975) are not reachable through Grok's host
 * (T1HT::decompress always passes lengths2 = 0, T1HT.cpp:171-173; the decoder then rejects
 * num_passes > 1, ojph_block_decoder.cpp:1014-1020); they are restated here as plain per-sample
 * loops and pinned by calling the reference's ojph_decode_codeblock directly with a second
 * segment (oracle/ref_harness.cpp ref_ht_decode).
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

#include "ht_cxtvlc.inc"

/* packed table word -> fields */
#define TW_CQ(w) ((w) & 7u)
#define TW_RHO(w) (((w) >> 3) & 15u)
#define TW_UOFF(w) (((w) >> 7) & 1u)
#define TW_EK(w) (((w) >> 8) & 15u)
#define TW_E1(w) (((w) >> 12) & 15u)
#define TW_CWD(w) (((w) >> 16) & 127u)
#define TW_LEN(w) (((w) >> 23) & 7u)

/* decode tables: [row0 / other rows][c_q << 7 | next 7 VLC bits] -> packed table word (0: none) */
static uint32_t g_dec[2][1024];
static int g_dec_ready = 0;

/* built once when the library is loaded (tests decode blocks from many threads) */
__attribute__((constructor)) static void build_dec_tables(void)
{
	if(g_dec_ready)
		return;
	for(int t = 0; t < 2; ++t)
	{
		const unsigned int* src = t ? HT_CXTVLC_ROWN : HT_CXTVLC_ROW0;
		const size_t n = t ? sizeof(HT_CXTVLC_ROWN) / 4 : sizeof(HT_CXTVLC_ROW0) / 4;
		for(uint32_t i = 0; i < 1024; ++i)
		{
			g_dec[t][i] = 0;
			for(size_t j = 0; j < n; ++j)
				if(TW_CQ(src[j]) == (i >> 7) &&
				   TW_CWD(src[j]) == ((i & 127u) & ((1u << TW_LEN(src[j])) - 1)))
					g_dec[t][i] = src[j]; /* the last match wins, as in the reference's loop */
		}
	}
	g_dec_ready = 1;
}

/* ---- the three bit streams of a cleanup segment ------------------------------------------ */

typedef struct
{
	const uint8_t* d; /* start of the cleanup segment */
	int lcup, scup;
	/* MEL: forward from lcup-scup, MSB first, 7 bits after a 0xFF byte */
	int mel_pos, mel_left, mel_unstuff;
	uint32_t mel_acc;
	int mel_bits;
	int mel_k, mel_run, mel_one; /* run-length state */
	/* VLC: backward from lcup-2, LSB first */
	int vlc_pos, vlc_real; /* next byte to read, real bytes still to deliver before zeros */
	int vlc_unstuff;
	uint64_t vlc_acc;
	int vlc_bits;
	/* MagSgn: forward from 0, LSB first */
	int ms_pos, ms_left, ms_unstuff;
	uint64_t ms_acc;
	int ms_bits;
} ht_streams;

static const int MEL_EXP[13] = {0, 0, 0, 1, 1, 1, 2, 2, 2, 3, 3, 4, 5};

static int mel_bit(ht_streams* s)
{
	if(s->mel_bits == 0)
	{
		/* mel_read, ojph_block_decoder.cpp:113-168: exhausted -> 0xFF; the last byte of the
		 * segment shares its low nibble with the segment length field -> forced to 1s */
		uint32_t b = 0xFF;
		if(s->mel_left > 0)
		{
			b = s->d[s->mel_pos];
			if(s->mel_left == 1)
				b |= 0xF;
			s->mel_pos++;
		}
		s->mel_left--;
		s->mel_bits = 8 - s->mel_unstuff;
		s->mel_acc = b;
		s->mel_unstuff = (b == 0xFF);
	}
	s->mel_bits--;
	return (int)((s->mel_acc >> s->mel_bits) & 1u);
}

/* one MEL symbol (ISO 15444-15 7.3.3; the reference pre-decodes runs, :170-220) */
static int mel_event(ht_streams* s)
{
	if(s->mel_run == 0 && !s->mel_one)
	{
		const int e = MEL_EXP[s->mel_k];
		if(mel_bit(s))
		{
			s->mel_run = 1 << e;
			s->mel_k = s->mel_k < 12 ? s->mel_k + 1 : 12;
		}
		else
		{
			int r = 0;
			for(int i = 0; i < e; ++i)
				r = (r << 1) | mel_bit(s);
			s->mel_run = r;
			s->mel_k = s->mel_k > 0 ? s->mel_k - 1 : 0;
			s->mel_one = 1;
		}
	}
	if(s->mel_run > 0)
	{
		s->mel_run--;
		return 0;
	}
	s->mel_one = 0;
	return 1;
}

/* rev_read / rev_init, ojph_block_decoder.cpp:327-421.  Bytes are delivered walking backwards;
 * the reference reads whole aligned 32-bit words while any byte of the segment is left, so a
 * few bytes in front of the VLC segment can be delivered before the zeros start. */
static void vlc_fill(ht_streams* s)
{
	while(s->vlc_bits <= 56)
	{
		uint32_t b = 0;
		if(s->vlc_real > 0)
		{
			b = s->vlc_pos >= 0 ? s->d[s->vlc_pos] : 0; /* 8 zero bytes precede the data */
			s->vlc_real--;
		}
		s->vlc_pos--;
		const int w = 8 - ((s->vlc_unstuff && (b & 0x7F) == 0x7F) ? 1 : 0);
		s->vlc_acc |= (uint64_t)b << s->vlc_bits;
		s->vlc_bits += w;
		s->vlc_unstuff = b > 0x8F;
	}
}

static uint32_t vlc_peek(ht_streams* s)
{
	if(s->vlc_bits < 32)
		vlc_fill(s);
	return (uint32_t)s->vlc_acc;
}

static void vlc_skip(ht_streams* s, uint32_t n)
{
	s->vlc_acc >>= n;
	s->vlc_bits -= (int)n;
}

/* frwd_read<0xFF>, ojph_block_decoder.cpp:871-900: a byte that follows 0xFF is 7 bits wide */
static void ms_fill(ht_streams* s)
{
	while(s->ms_bits <= 56)
	{
		uint32_t b = 0xFF;
		if(s->ms_left > 0)
			b = s->d[s->ms_pos];
		s->ms_pos++;
		s->ms_left--;
		s->ms_acc |= (uint64_t)b << s->ms_bits;
		s->ms_bits += 8 - s->ms_unstuff;
		s->ms_unstuff = (b == 0xFF);
	}
}

static uint32_t ms_peek(ht_streams* s)
{
	if(s->ms_bits < 32)
		ms_fill(s);
	return (uint32_t)s->ms_acc;
}

static void ms_skip(ht_streams* s, uint32_t n)
{
	s->ms_acc >>= n;
	s->ms_bits -= (int)n;
}

static void streams_init(ht_streams* s, const uint8_t* d, int lcup, int scup)
{
	memset(s, 0, sizeof(*s));
	s->d = d;
	s->lcup = lcup;
	s->scup = scup;
	s->mel_pos = lcup - scup;
	s->mel_left = scup - 1;
	/* VLC: the upper nibble of byte lcup-2 comes first (rev_init, :392-421) */
	const uint32_t first = d[lcup - 2];
	s->vlc_acc = first >> 4;
	s->vlc_bits = 4 - (((s->vlc_acc & 7) == 7) ? 1 : 0);
	s->vlc_unstuff = (first | 0xF) > 0x8F;
	s->vlc_pos = lcup - 3;
	{
		/* real bytes delivered: the reference takes 1 + (address & 3) single bytes (at most the
		 * segment), then 32-bit words while size > 0 (the segment start is 4-byte aligned in
		 * the reference's buffers, T1HT.cpp:142-153) */
		int size = scup - 2;
		int num = 1 + (s->vlc_pos & 3);
		int tnum = num < size ? num : size;
		if(tnum < 0)
			tnum = 0;
		size -= tnum;
		s->vlc_real = tnum + (size > 0 ? ((size + 3) / 4) * 4 : 0);
	}
	s->ms_left = lcup - scup;
}

/* ---- U-VLC (decode_init_uvlc / decode_noninit_uvlc, :664-838) ------------------------------ */

/* prefix of one u value from the low 3 bits: returns prefix length, *suf suffix length, *base */
static uint32_t uvlc_prefix(uint32_t bits, uint32_t* suf, uint32_t* base)
{
	if(bits & 1)
	{
		*suf = 0;
		*base = 1;
		return 1;
	}
	if(bits & 2)
	{
		*suf = 0;
		*base = 2;
		return 2;
	}
	if(bits & 4)
	{
		*suf = 1;
		*base = 3;
		return 3;
	}
	*suf = 5;
	*base = 5;
	return 3;
}

/* mode: u_off of quad 0 | u_off of quad 1 << 1, +1 when both are set on the initial row and the
 * MEL event was 1.  u[] receives u_q + kappa-offset exactly as the reference leaves them. */
static uint32_t uvlc_decode(uint32_t vlc, uint32_t mode, int initial, uint32_t* u)
{
	uint32_t used = 0, s0, b0, s1, b1, p;
	u[0] = u[1] = 1;
	if(mode == 0)
		return 0;
	if(mode <= 2)
	{
		p = uvlc_prefix(vlc, &s0, &b0);
		vlc >>= p;
		used = p + s0;
		b0 += vlc & ((1u << s0) - 1);
		u[mode == 1 ? 0 : 1] = b0 + 1;
		return used;
	}
	if(initial && mode == 3)
	{
		p = uvlc_prefix(vlc, &s0, &b0);
		vlc >>= p;
		used = p;
		if(p > 2)
		{
			/* the second quad's u is a single bit */
			u[1] = (vlc & 1) + 2;
			vlc >>= 1;
			used += 1 + s0;
			u[0] = b0 + (vlc & ((1u << s0) - 1)) + 1;
			return used;
		}
		p = uvlc_prefix(vlc, &s1, &b1);
		vlc >>= p;
		used += p + s0 + s1;
		u[0] = b0 + (vlc & ((1u << s0) - 1)) + 1;
		vlc >>= s0;
		u[1] = b1 + (vlc & ((1u << s1) - 1)) + 1;
		return used;
	}
	/* both present: two prefixes, then two suffixes; on the initial row with MEL event 1 both
	 * values are offset by 2 */
	p = uvlc_prefix(vlc, &s0, &b0);
	vlc >>= p;
	used = p;
	p = uvlc_prefix(vlc, &s1, &b1);
	vlc >>= p;
	used += p + s0 + s1;
	const uint32_t off = (initial && mode == 4) ? 3 : 1;
	u[0] = b0 + (vlc & ((1u << s0) - 1)) + off;
	vlc >>= s0;
	u[1] = b1 + (vlc & ((1u << s1) - 1)) + off;
	return used;
}

static uint32_t bitlen(uint32_t v)
{
	uint32_t n = 0;
	while(v)
	{
		++n;
		v >>= 1;
	}
	return n;
}

/* ---- SigProp / MagRef (ojph_block_decoder.cpp:1627-2209) -------------------------------------
 * The reference keeps nibble-packed significance / membership words for two stripes and runs the
 * refinement of a stripe interleaved with the cleanup pass of later rows.  The passes only read
 * what the cleanup pass left in their own stripe, the row above and the row below, so here they
 * run after it, stripe by stripe, straight on the sample array:
 *   MagRef  every sample the cleanup pass made significant takes one bit, stripe-column order;
 *   SigProp a zero sample with a significant 8-neighbour at the time it is visited (cleanup
 *           significance for samples not visited yet, plus what this pass has set so far) takes
 *           one bit; the signs of the samples a group of four columns made significant follow
 *           that group's significance bits. */
typedef struct
{
	const uint8_t* d;
	int pos, left, unstuff;
	uint64_t acc;
	int bits;
} ht_bits;

/* frwd_read<0> / frwd_init<0>, :875-945: forward, LSB first, 7 bits after 0xFF, zeros at the end */
static uint32_t spp_bit(ht_bits* s)
{
	while(s->bits <= 56)
	{
		uint32_t b = 0;
		if(s->left > 0)
			b = s->d[s->pos];
		s->pos++;
		s->left--;
		s->acc |= (uint64_t)b << s->bits; /* a set top bit of a 7-bit byte overlaps, as there */
		s->bits += 8 - s->unstuff;
		s->unstuff = (b == 0xFF);
	}
	const uint32_t r = (uint32_t)(s->acc & 1);
	s->acc >>= 1;
	s->bits--;
	return r;
}

/* rev_read_mrp / rev_init_mrp, :466-552: backward from the end of the second segment */
static uint32_t mrp_bit(ht_bits* s)
{
	while(s->bits <= 56)
	{
		uint32_t b = 0;
		if(s->left > 0)
			b = s->d[s->pos];
		s->pos--;
		s->left--;
		const int wd = 8 - ((s->unstuff && (b & 0x7F) == 0x7F) ? 1 : 0);
		s->acc |= (uint64_t)b << s->bits;
		s->bits += wd;
		s->unstuff = b > 0x8F;
	}
	const uint32_t r = (uint32_t)(s->acc & 1);
	s->acc >>= 1;
	s->bits--;
	return r;
}

static int ht_has_sig_neighbour(const uint32_t* out, uint32_t w, uint32_t h, uint32_t x, uint32_t y)
{
	for(int dy = -1; dy <= 1; ++dy)
		for(int dx = -1; dx <= 1; ++dx)
		{
			const int xx = (int)x + dx, yy = (int)y + dy;
			if((dx || dy) && xx >= 0 && yy >= 0 && xx < (int)w && yy < (int)h &&
			   out[(size_t)yy * w + xx] != 0)
				return 1;
		}
	return 0;
}

static void ht_refine(uint32_t* out, uint32_t w, uint32_t h, uint32_t p, uint32_t num_passes,
					  const uint8_t* seg2, int len2)
{
	ht_bits spp = {seg2, 0, len2, 0, 0, 0};
	ht_bits mrp = {seg2, len2 - 1, len2, 1, 0, 0};
	for(uint32_t y0 = 0; y0 < h; y0 += 4)
	{
		const uint32_t rows = h - y0 < 4 ? h - y0 : 4;
		if(num_passes > 2)
		{
			/* :1630-1701, :1924-1987 */
			for(uint32_t x = 0; x < w; ++x)
				for(uint32_t r = 0; r < rows; ++r)
				{
					uint32_t* v = out + (size_t)(y0 + r) * w + x;
					if(*v == 0)
						continue;
					const uint32_t sym = mrp_bit(&mrp);
					*v ^= (1 - sym) << (p - 1);
					*v |= 1u << (p - 2);
				}
		}
		/* :1731-1916, :1990-2203 */
		for(uint32_t xg = 0; xg < w; xg += 4)
		{
			const uint32_t xe = xg + 4 < w ? xg + 4 : w;
			uint32_t* fresh[16];
			uint32_t nfresh = 0;
			for(uint32_t x = xg; x < xe; ++x)
				for(uint32_t r = 0; r < rows; ++r)
				{
					uint32_t* v = out + (size_t)(y0 + r) * w + x;
					if(*v != 0 || !ht_has_sig_neighbour(out, w, h, x, y0 + r))
						continue;
					if(spp_bit(&spp))
					{
						*v = 3u << (p - 2); /* newly significant, bin centre included */
						fresh[nfresh++] = v;
					}
				}
			for(uint32_t i = 0; i < nfresh; ++i)
				*fresh[i] |= spp_bit(&spp) << 31;
		}
	}
}

/* ---- cleanup pass decoder ------------------------------------------------------------------- */

int orc_ht_decode(const uint8_t* data, uint32_t missing_msbs, uint32_t num_passes,
				  uint32_t lengths1, uint32_t lengths2, uint32_t w, uint32_t h, uint32_t* out)
{
	build_dec_tables();
	/* ojph_block_decoder.cpp:1014-1043 */
	if(num_passes > 1 && lengths2 == 0)
		return -1;
	if(num_passes > 3)
		return -1;
	if(missing_msbs > 29)
		return -1;
	if(missing_msbs == 29)
		num_passes = 1;
	const uint32_t p = 30 - missing_msbs;
	const int lcup = (int)lengths1;
	if(lcup < 2)
		return -1;
	const int scup = ((int)data[lcup - 1] << 4) + (data[lcup - 2] & 0xF);
	if(scup < 2 || scup > lcup || scup > 4079)
		return -1;
	ht_streams s;
	streams_init(&s, data, lcup, scup);

	/* bottom row of the previous quad row: significance and magnitude exponent per column,
	 * columns -1 .. w+2 (index + 1) */
	const uint32_t ncol = w + 8;
	uint8_t* sigN = (uint8_t*)calloc(2 * ncol, 1);
	uint8_t* expN = sigN + ncol;
	uint8_t* sigC = (uint8_t*)calloc(2 * ncol, 1);
	uint8_t* expC = sigC + ncol;
	memset(out, 0, sizeof(uint32_t) * (size_t)w * h);
	int rc = 0;
	for(uint32_t y = 0; y < h && rc == 0; y += 2)
	{
		const int initial = (y == 0);
		uint32_t cq_carry = 0; /* context contribution of the quad to the left */
		memset(sigC, 0, 2 * ncol);
		for(uint32_t x = 0; x < w && rc == 0; x += 4)
		{
			uint32_t tw[2] = {0, 0};
			const uint32_t nq = (x + 2 < w) ? 2 : 1;
			for(uint32_t q = 0; q < nq; ++q)
			{
				const uint32_t c0 = x + 2 * q; /* left column of this quad */
				uint32_t cq = cq_carry;
				if(!initial)
				{
					/* sigma^NW | sigma^N  and  sigma^NE | sigma^NF (:1376-1378, :1396-1397) */
					cq |= sigN[c0 - 1 + 1] | sigN[c0 + 1];
					cq |= (uint32_t)(sigN[c0 + 1 + 1] | sigN[c0 + 2 + 1]) << 2;
				}
				uint32_t t = g_dec[initial ? 0 : 1][(cq << 7) | (vlc_peek(&s) & 0x7F)];
				if(cq == 0 && !mel_event(&s))
					t = 0;
				vlc_skip(&s, TW_LEN(t));
				tw[q] = t;
				const uint32_t rho = TW_RHO(t);
				if(initial)
					cq_carry = ((rho & 1) | ((rho >> 1) & 1)) | (((rho >> 2) & 1) << 1) |
							   (((rho >> 3) & 1) << 2);
				else
					cq_carry = (((rho >> 2) & 1) | ((rho >> 3) & 1)) << 1;
			}
			/* U-VLC for the pair (:1180-1193, :1438-1456) */
			uint32_t mode = TW_UOFF(tw[0]) | (TW_UOFF(tw[1]) << 1);
			if(initial && mode == 3 && mel_event(&s))
				mode = 4;
			uint32_t U[2];
			vlc_skip(&s, uvlc_decode(vlc_peek(&s), mode, initial, U));
			if(!initial)
			{
				for(uint32_t q = 0; q < 2; ++q)
				{
					const uint32_t rho = TW_RHO(tw[q]);
					if(rho & (rho - 1)) /* gamma: more than one significant sample */
					{
						const uint32_t c0 = x + 2 * q;
						uint32_t E = 0;
						for(int dc = -1; dc <= 2; ++dc)
							if(expN[(int)c0 + dc + 1] > E)
								E = expN[(int)c0 + dc + 1];
						U[q] += E > 2 ? E - 2 : 0;
					}
				}
			}
			if(U[0] > missing_msbs + 1 || U[1] > missing_msbs + 1)
			{
				rc = -1;
				break;
			}
			/* MagSgn: quad by quad, samples in the order (c,r) = (0,0) (0,1) (1,0) (1,1) */
			for(uint32_t q = 0; q < 2; ++q)
			{
				const uint32_t t = tw[q], rho = TW_RHO(t);
				for(uint32_t n = 0; n < 4; ++n)
				{
					if(!((rho >> n) & 1))
						continue;
					const uint32_t ms = ms_peek(&s);
					const uint32_t m = U[q] - ((TW_EK(t) >> n) & 1);
					ms_skip(&s, m);
					uint32_t v = ms & ((1u << m) - 1);
					v |= ((TW_E1(t) >> n) & 1) << m;
					v |= 1;
					const uint32_t col = x + 2 * q + (n >> 1), row = y + (n & 1);
					if(col < w && row < h)
						out[(size_t)row * w + col] = (ms << 31) | ((v + 2) << (p - 1));
					if((n & 1) && col < w + 4)
					{
						sigC[col + 1] = 1;
						expC[col + 1] = (uint8_t)bitlen(v);
					}
				}
			}
		}
		/* this quad row becomes "north" */
		uint8_t* tmp = sigN;
		sigN = sigC;
		sigC = tmp;
		expN = sigN + ncol;
		expC = sigC + ncol;
	}
	free(sigN < sigC ? sigN : sigC);
	free(sigN < sigC ? sigC : sigN);
	if(rc == 0 && num_passes > 1)
		ht_refine(out, w, h, p, num_passes, data + lengths1, (int)lengths2);
	return rc;
}

/* ---- cleanup pass encoder (ojph_block_encoder.cpp:470-945) ----------------------------------- */

/* encode table: [row0 / other rows][c_q << 8 | rho << 4 | eps] -> packed table word (0: none),
 * chosen as the reference does (:74-215): without an EMB pattern the first u_off = 0 entry;
 * with one, among the u_off = 1 entries whose known bits agree ((eps & e_k) == e_1) the LAST one
 * with the most known bits */
static uint32_t g_enc[2][2048];
static int g_enc_ready = 0;

static int popcnt4(uint32_t v)
{
	return (int)((v & 1) + ((v >> 1) & 1) + ((v >> 2) & 1) + ((v >> 3) & 1));
}

__attribute__((constructor)) static void build_enc_tables(void)
{
	if(g_enc_ready)
		return;
	for(int t = 0; t < 2; ++t)
	{
		const unsigned int* src = t ? HT_CXTVLC_ROWN : HT_CXTVLC_ROW0;
		const size_t n = t ? sizeof(HT_CXTVLC_ROWN) / 4 : sizeof(HT_CXTVLC_ROW0) / 4;
		for(uint32_t i = 0; i < 2048; ++i)
		{
			const uint32_t cq = i >> 8, rho = (i >> 4) & 15, eps = i & 15;
			g_enc[t][i] = 0;
			if((eps & rho) != eps || (rho == 0 && cq == 0))
				continue;
			int best = -1;
			for(size_t j = 0; j < n; ++j)
			{
				const uint32_t w = src[j];
				if(TW_CQ(w) != cq || TW_RHO(w) != rho)
					continue;
				if(eps == 0)
				{
					if(TW_UOFF(w) == 0)
					{
						g_enc[t][i] = w | 0x80000000u;
						break;
					}
				}
				else if(TW_UOFF(w) == 1 && (eps & TW_EK(w)) == TW_E1(w) && popcnt4(TW_EK(w)) >= best)
				{
					best = popcnt4(TW_EK(w));
					g_enc[t][i] = w | 0x80000000u;
				}
			}
		}
	}
	g_enc_ready = 1;
}

typedef struct
{
	uint8_t ms[16384 + 64];
	uint32_t ms_pos, ms_tmp;
	int ms_used, ms_max;
	uint8_t mel[192];
	uint32_t mel_pos, mel_tmp;
	int mel_rem, mel_run, mel_k;
	uint8_t vlc[3072]; /* filled from the end */
	uint32_t vlc_pos, vlc_tmp; /* vlc_pos bytes written, the first being vlc[sizeof-1] */
	int vlc_used, vlc_gt8f;
	int overflow;
} ht_enc;

/* ms_encode, :397-425: LSB first, a byte after 0xFF holds 7 bits */
static void enc_ms(ht_enc* e, uint32_t cwd, int len)
{
	while(len > 0)
	{
		if(e->ms_pos >= 16384)
		{
			e->overflow = 1;
			return;
		}
		int t = e->ms_max - e->ms_used;
		if(t > len)
			t = len;
		e->ms_tmp |= (cwd & ((1u << t) - 1)) << e->ms_used;
		e->ms_used += t;
		cwd >>= t;
		len -= t;
		if(e->ms_used >= e->ms_max)
		{
			e->ms[e->ms_pos++] = (uint8_t)e->ms_tmp;
			e->ms_max = (e->ms_tmp == 0xFF) ? 7 : 8;
			e->ms_tmp = 0;
			e->ms_used = 0;
		}
	}
}

/* mel_emit_bit / mel_encode, :247-298 */
static void enc_mel_bit(ht_enc* e, int v)
{
	e->mel_tmp = (e->mel_tmp << 1) + (uint32_t)v;
	if(--e->mel_rem == 0)
	{
		if(e->mel_pos >= 128)
		{
			e->overflow = 1;
			return;
		}
		e->mel[e->mel_pos++] = (uint8_t)e->mel_tmp;
		e->mel_rem = (e->mel_tmp == 0xFF) ? 7 : 8;
		e->mel_tmp = 0;
	}
}

static void enc_mel(ht_enc* e, int bit)
{
	const int ex = MEL_EXP[e->mel_k];
	if(!bit)
	{
		if(++e->mel_run >= (1 << ex))
		{
			enc_mel_bit(e, 1);
			e->mel_run = 0;
			e->mel_k = e->mel_k < 12 ? e->mel_k + 1 : 12;
		}
	}
	else
	{
		enc_mel_bit(e, 0);
		for(int t = ex; t > 0;)
			enc_mel_bit(e, (e->mel_run >> --t) & 1);
		e->mel_run = 0;
		e->mel_k = e->mel_k > 0 ? e->mel_k - 1 : 0;
	}
}

/* vlc_encode, :325-358: written backwards, LSB first; after a byte > 0x8F the next byte may
 * use its 8th bit only if its low seven bits are not all ones */
static void enc_vlc(ht_enc* e, uint32_t cwd, int len)
{
	while(len > 0)
	{
		if(e->vlc_pos >= 3072 - 128)
		{
			e->overflow = 1;
			return;
		}
		int avail = 8 - e->vlc_gt8f - e->vlc_used;
		int t = avail < len ? avail : len;
		e->vlc_tmp |= (cwd & ((1u << t) - 1)) << e->vlc_used;
		e->vlc_used += t;
		avail -= t;
		len -= t;
		cwd >>= t;
		if(avail == 0)
		{
			if(e->vlc_gt8f && e->vlc_tmp != 0x7F)
			{
				e->vlc_gt8f = 0;
				continue;
			}
			e->vlc[sizeof(e->vlc) - 1 - e->vlc_pos] = (uint8_t)e->vlc_tmp;
			e->vlc_pos++;
			e->vlc_gt8f = e->vlc_tmp > 0x8F;
			e->vlc_tmp = 0;
			e->vlc_used = 0;
		}
	}
}

/* U-VLC codeword of u (uvlc_init_tables, :194-215): prefix value/length, suffix value/length */
static void uvlc_code(int u, uint32_t* pre, int* pre_len, uint32_t* suf, int* suf_len)
{
	if(u <= 0)
	{
		*pre = 0, *pre_len = 0, *suf = 0, *suf_len = 0;
	}
	else if(u <= 2)
	{
		*pre = (uint32_t)u, *pre_len = u, *suf = 0, *suf_len = 0;
	}
	else if(u <= 4)
	{
		*pre = 4, *pre_len = 3, *suf = (uint32_t)(u - 3), *suf_len = 1;
	}
	else
	{
		*pre = 0, *pre_len = 3, *suf = (uint32_t)(u - 5), *suf_len = 5;
	}
}

/* samples: w*h sign-magnitude words (sign bit 31, magnitude << (30 - missing_msbs)), row pitch w.
 * out: >= 16384 + 3072 bytes.  Returns the coded length (MagSgn | MEL | VLC, Scup in the last
 * twelve bits), or -1 if a stream buffer overflowed (the reference throws). */
int orc_ht_encode(const uint32_t* samples, uint32_t missing_msbs, uint32_t w, uint32_t h,
				  uint8_t* out)
{
	build_enc_tables();
	if(missing_msbs > 30)
		return -1;
	const uint32_t p = 30 - missing_msbs;
	ht_enc* e = (ht_enc*)calloc(1, sizeof(ht_enc));
	if(!e)
		return -1;
	e->ms_max = 8;
	e->mel_rem = 8;
	/* vlc_init, :313-323: the last byte is a placeholder, its neighbour starts with four 1s */
	e->vlc[sizeof(e->vlc) - 1] = 0xFF;
	e->vlc_pos = 1;
	e->vlc_used = 4;
	e->vlc_tmp = 0xF;
	e->vlc_gt8f = 1;

	const uint32_t ncol = w + 8;
	uint8_t* expN = (uint8_t*)calloc(4 * ncol, 1); /* exponent / significance, previous and current */
	uint8_t *sigN = expN + ncol, *expC = sigN + ncol, *sigC = expC + ncol;
	for(uint32_t y = 0; y < h; y += 2)
	{
		const int initial = (y == 0);
		uint32_t carry = 0;
		memset(expC, 0, ncol);
		memset(sigC, 0, ncol);
		for(uint32_t x = 0; x < w; x += 4)
		{
			int u[2] = {0, 0};
			const uint32_t nq = (x + 2 < w) ? 2 : 1;
			for(uint32_t q = 0; q < nq; ++q)
			{
				const uint32_t c0 = x + 2 * q;
				uint32_t rho = 0, v[4] = {0, 0, 0, 0};
				int eq[4] = {0, 0, 0, 0}, emax = 0;
				for(uint32_t n = 0; n < 4; ++n)
				{
					const uint32_t col = c0 + (n >> 1), row = y + (n & 1);
					if(col >= w || row >= h)
						continue;
					const uint32_t t = samples[(size_t)row * w + col];
					uint32_t val = ((t + t) >> p) & ~1u; /* 2 * mu_p */
					if(!val)
						continue;
					rho |= 1u << n;
					eq[n] = (int)bitlen(val - 1); /* exponent of 2 mu_p - 1 */
					if(eq[n] > emax)
						emax = eq[n];
					v[n] = (val - 2) + (t >> 31); /* 2 (mu_p - 1) + sign */
					if(n & 1)
					{
						expC[col + 1] = (uint8_t)eq[n];
						sigC[col + 1] = 1;
					}
				}
				uint32_t cq = carry;
				int kappa = 1;
				if(!initial)
				{
					cq |= (uint32_t)(sigN[c0] | sigN[c0 + 1]);
					cq |= (uint32_t)(sigN[c0 + 2] | sigN[c0 + 3]) << 2;
					if(rho & (rho - 1))
					{
						int me = 0;
						for(int i = 0; i < 4; ++i)
							if(expN[c0 + i] > me)
								me = expN[c0 + i];
						kappa = me - 1 > 1 ? me - 1 : 1;
					}
				}
				const int U = emax > kappa ? emax : kappa;
				u[q] = U - kappa;
				uint32_t eps = 0;
				if(u[q] > 0)
					for(uint32_t n = 0; n < 4; ++n)
						eps |= (uint32_t)(eq[n] == emax) << n;
				const uint32_t tw = g_enc[initial ? 0 : 1][(cq << 8) | (rho << 4) | eps];
				enc_vlc(e, TW_CWD(tw), (int)TW_LEN(tw));
				if(cq == 0)
					enc_mel(e, rho != 0);
				for(uint32_t n = 0; n < 4; ++n)
				{
					const int m = ((rho >> n) & 1) ? U - (int)((TW_EK(tw) >> n) & 1) : 0;
					enc_ms(e, v[n] & ((1u << m) - 1), m);
				}
				carry = initial ? ((rho >> 1) | (rho & 1)) : ((((rho >> 2) | (rho >> 3)) & 1u) << 1);
			}
			/* U-VLC of the pair (:693-716 initial row, :926-929 others) */
			uint32_t pre[2], suf[2];
			int prel[2], sufl[2];
			if(initial && u[0] > 0 && u[1] > 0)
				enc_mel(e, (u[0] < u[1] ? u[0] : u[1]) > 2);
			if(initial && u[0] > 2 && u[1] > 2)
			{
				uvlc_code(u[0] - 2, &pre[0], &prel[0], &suf[0], &sufl[0]);
				uvlc_code(u[1] - 2, &pre[1], &prel[1], &suf[1], &sufl[1]);
			}
			else if(initial && u[0] > 2 && u[1] > 0)
			{
				uvlc_code(u[0], &pre[0], &prel[0], &suf[0], &sufl[0]);
				pre[1] = (uint32_t)(u[1] - 1);
				prel[1] = 1;
				suf[1] = 0;
				sufl[1] = 0;
			}
			else
			{
				uvlc_code(u[0], &pre[0], &prel[0], &suf[0], &sufl[0]);
				uvlc_code(u[1], &pre[1], &prel[1], &suf[1], &sufl[1]);
			}
			enc_vlc(e, pre[0], prel[0]);
			enc_vlc(e, pre[1], prel[1]);
			enc_vlc(e, suf[0], sufl[0]);
			enc_vlc(e, suf[1], sufl[1]);
		}
		uint8_t* t1 = expN;
		expN = expC;
		expC = t1;
		t1 = sigN;
		sigN = sigC;
		sigC = t1;
	}
	/* terminate_mel_vlc, :363-392 */
	if(e->mel_run > 0)
		enc_mel_bit(e, 1);
	{
		e->mel_tmp <<= e->mel_rem;
		const uint32_t mel_mask = (0xFFu << e->mel_rem) & 0xFF;
		const uint32_t vlc_mask = 0xFFu >> (8 - e->vlc_used);
		if((mel_mask | vlc_mask) != 0)
		{
			const uint32_t fuse = e->mel_tmp | e->vlc_tmp;
			if((((fuse ^ e->mel_tmp) & mel_mask) | ((fuse ^ e->vlc_tmp) & vlc_mask)) == 0 &&
			   fuse != 0xFF && e->vlc_pos > 1)
				e->mel[e->mel_pos++] = (uint8_t)fuse;
			else
			{
				e->mel[e->mel_pos++] = (uint8_t)e->mel_tmp;
				e->vlc[sizeof(e->vlc) - 1 - e->vlc_pos] = (uint8_t)e->vlc_tmp;
				e->vlc_pos++;
			}
		}
	}
	/* ms_terminate, :427-461 */
	if(e->ms_used)
	{
		const int t = e->ms_max - e->ms_used;
		e->ms_tmp |= (0xFFu & ((1u << t) - 1)) << e->ms_used;
		if(e->ms_tmp != 0xFF)
			e->ms[e->ms_pos++] = (uint8_t)e->ms_tmp;
	}
	else if(e->ms_max == 7)
		e->ms_pos--;
	int total = -1;
	if(!e->overflow)
	{
		total = (int)(e->ms_pos + e->mel_pos + e->vlc_pos);
		memcpy(out, e->ms, e->ms_pos);
		memcpy(out + e->ms_pos, e->mel, e->mel_pos);
		memcpy(out + e->ms_pos + e->mel_pos, e->vlc + sizeof(e->vlc) - e->vlc_pos, e->vlc_pos);
		const uint32_t nb = e->mel_pos + e->vlc_pos;
		out[total - 1] = (uint8_t)(nb >> 4);
		out[total - 2] = (uint8_t)((out[total - 2] & 0xF0) | (nb & 0xF));
	}
	free(expN < expC ? expN : expC);
	free(e);
	return total;
}
