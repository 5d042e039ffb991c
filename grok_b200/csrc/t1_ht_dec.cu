// t1_ht_dec.cu -- HTJ2K (ISO/IEC 15444-15) cleanup-pass code-block decoder with the HT
// dequantisers fused into the store, sm_100a.  One code block per warp.
//
// Replaces, behind the C-ABI, the reference's
//   T1HT::decompress               src/lib/jp2/t1/t1_ht/T1HT.cpp:134-187
//   ojph_decode_codeblock          t1/t1_ht/coding/ojph_block_decoder.cpp:989-1626 (cleanup pass)
//   ShiftHTFilter / ScaleHTFilter / RoiShiftHT / RoiScaleHT   filters/PostDecompressFilters.h:91-177
//
// Shape: one code block per warp.  The cleanup pass has two halves with different structure and
// the kernel treats them differently, quad row by quad row (32 quads = 64 columns at a time):
//   1. MEL + CxtVLC + U-VLC: a serial chain (the context of a quad needs the previous quad's
//      significance pattern).  All lanes run it on identical register copies (one issue slot
//      either way) and lane q keeps the result of quad q: the table word and u.  It never needs
//      the MagSgn stream: the north context is significance only.
//   2. MagSgn: one quad per lane.  Each lane derives its exponent bound from the north row's
//      exponents, the bit counts of its four samples and, by a warp prefix sum, where its bits
//      start.  The MagSgn bytes are un-stuffed 32 at a time (one ballot gives every byte's bit
//      position) into a per-warp ring of bits in shared memory, from which the lanes pick their
//      fields; sample reconstruction, the HT dequantiser and the store (coalesced rows of the
//      L2-resident plane, written once) are fused here.
// Not HBM-bound: traffic is the coded bytes in and 4 B/sample out.
//
// SigProp / MagRef passes (ojph_block_decoder.cpp:1627-2209): a block whose descriptor carries a
// second codeword segment (cleanup | SigProp [+ MagRef], 2 or 3 passes) keeps the block decoder's
// sign-magnitude words in the plane after the cleanup pass, runs the refinement stripe by stripe
// (ht_refine below) and dequantises in place at the end.  Grok's own host never gets there
// (T1HT.cpp:171-173 passes lengths2 = 0, which ojph_block_decoder.cpp:1014-1020 rejects for more
// than one pass), so a block with more than one pass in a single segment is rejected here too
// and counted in the status word.
#include "t1.cuh"
#include <mutex>

#include "ht_cxtvlc.inc"

namespace
{
// Decode tables in constant memory (the VLC chain runs replicated, so every lane reads the same
// entry).  CxtVLC entry: len | u_off << 3 | rho << 4 | e_k << 8 | e_1 << 12 | context carried to
// the next quad << 16.  U-VLC entry (indexed by row kind, mode, six stream bits):
// prefix bits | width A << 3 | width B << 6 | base A << 9 | base B << 14 | swap << 19, where A and
// B are the two fields that follow the prefixes in stream order and swap says that A is u1.
struct HtTables
{
	uint32_t vlc[2][1024];
	uint32_t uvlc[2][8][64];
};
__constant__ HtTables c_ht;

// prefix of one u value: returns prefix length; suffix length and base value by reference
// (decode_init_uvlc / decode_noninit_uvlc, ojph_block_decoder.cpp:664-838)
static uint32_t uvlc_prefix(uint32_t bits, uint32_t& suf, uint32_t& base)
{
	if(bits & 1)
	{
		suf = 0;
		base = 1;
		return 1;
	}
	if(bits & 2)
	{
		suf = 0;
		base = 2;
		return 2;
	}
	suf = (bits & 4) ? 1 : 5;
	base = (bits & 4) ? 3 : 5;
	return 3;
}

static uint32_t uvlc_entry(uint32_t bits, uint32_t mode, bool initial)
{
	uint32_t plen = 0, wa = 0, wb = 0, ba = 1, bb = 1, swap = 0, s0, b0, s1, b1;
	if(mode == 1 || mode == 2)
	{
		plen = uvlc_prefix(bits, s0, b0);
		wa = s0;
		ba = b0 + 1;
		swap = mode == 2; // the only coded value is u1
	}
	else if(mode >= 3)
	{
		const uint32_t p0 = uvlc_prefix(bits, s0, b0);
		if(initial && mode == 3 && p0 > 2)
		{
			// one bit of u1 comes before u0's suffix
			plen = p0;
			wa = 1;
			ba = 2;
			wb = s0;
			bb = b0 + 1;
			swap = 1;
		}
		else
		{
			const uint32_t p1 = uvlc_prefix(bits >> p0, s1, b1);
			const uint32_t off = (initial && mode == 4) ? 3 : 1;
			plen = p0 + p1;
			wa = s0;
			ba = b0 + off;
			wb = s1;
			bb = b1 + off;
		}
	}
	return plen | (wa << 3) | (wb << 6) | (ba << 9) | (bb << 14) | (swap << 19);
}

std::mutex g_mu;
bool g_done[64] = {};

int ht_tables_upload()
{
	int dev = 0;
	if(cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
		return -1;
	std::lock_guard<std::mutex> lk(g_mu);
	if(g_done[dev])
		return 0;
	static HtTables h;
	for(int t = 0; t < 2; ++t)
	{
		const unsigned int* src = t ? HT_CXTVLC_ROWN : HT_CXTVLC_ROW0;
		const size_t n = t ? sizeof(HT_CXTVLC_ROWN) / 4 : sizeof(HT_CXTVLC_ROW0) / 4;
		for(uint32_t i = 0; i < 1024; ++i)
		{
			uint32_t e = 0;
			for(size_t j = 0; j < n; ++j)
			{
				const uint32_t w = src[j];
				const uint32_t len = (w >> 23) & 7u;
				if((w & 7u) == (i >> 7) && ((w >> 16) & 127u) == ((i & 127u) & ((1u << len) - 1)))
					e = len | (((w >> 7) & 1u) << 3) | (((w >> 3) & 15u) << 4) | (((w >> 8) & 15u) << 8) |
						(((w >> 12) & 15u) << 12);
			}
			const uint32_t rho = (e >> 4) & 15u;
			const uint32_t carry = t ? ((((rho >> 2) | (rho >> 3)) & 1u) << 1)
									 : (((rho | (rho >> 1)) & 1u) | ((rho >> 1) & 6u));
			h.vlc[t][i] = e | (carry << 16);
		}
		for(uint32_t mode = 0; mode < 8; ++mode)
			for(uint32_t bits = 0; bits < 64; ++bits)
				h.uvlc[t][mode][bits] = uvlc_entry(bits, mode > 4 ? 0 : mode, t == 0);
	}
	if(cudaMemcpyToSymbol(c_ht, &h, sizeof(h)) != cudaSuccess)
		return -1;
	g_done[dev] = true;
	return 0;
}

struct Streams
{
	const uint8_t* d;
	// MEL
	int mel_pos, mel_left;
	uint32_t mel_acc;
	int mel_bits, mel_unstuff, mel_k, mel_run, mel_one;
	// VLC (backward), as a ring of un-stuffed bits like the MagSgn stream below
	int vlc_a0, vlc_real; // offset of the first byte read; bytes delivered before zeros
	uint32_t vlc_k; // bytes read so far
	uint32_t vlc_unstuff; // the last byte read was > 0x8F
	uint32_t vlc_rp, vlc_wp; // read / fill position in bits
	uint32_t* vlc_ring;
};
constexpr uint32_t VLC_RING_WORDS = 128; // 4096 bits: 32 quads use at most 544
constexpr uint32_t VLC_AHEAD = 768; // bits made available before a group of 32 quads

static __device__ __forceinline__ int mel_bit(Streams& s)
{
	if(s.mel_bits == 0)
	{
		uint32_t b = 0xFF;
		if(s.mel_left > 0)
		{
			b = __ldg(s.d + s.mel_pos);
			if(s.mel_left == 1)
				b |= 0xF;
			s.mel_pos++;
		}
		s.mel_left--;
		s.mel_bits = 8 - s.mel_unstuff;
		s.mel_acc = b;
		s.mel_unstuff = (b == 0xFF);
	}
	s.mel_bits--;
	return (int)((s.mel_acc >> s.mel_bits) & 1u);
}

// one MEL symbol (ISO 15444-15 7.3.3; reference: mel_decode / mel_get_run, :170-285)
static __device__ __forceinline__ int mel_event(Streams& s)
{
	if(s.mel_run == 0 && !s.mel_one)
	{
		// exponents 0,0,0,1,1,1,2,2,2,3,3,4,5 packed 3 bits each
		const int e = (int)((05433222111000ull >> (3 * s.mel_k)) & 7ull);
		if(mel_bit(s))
		{
			s.mel_run = 1 << e;
			s.mel_k = min(s.mel_k + 1, 12);
		}
		else
		{
			int r = 0;
			for(int i = 0; i < e; ++i)
				r = (r << 1) | mel_bit(s);
			s.mel_run = r;
			s.mel_k = max(s.mel_k - 1, 0);
			s.mel_one = 1;
		}
	}
	if(s.mel_run > 0)
	{
		s.mel_run--;
		return 0;
	}
	s.mel_one = 0;
	return 1;
}

// rev_read, ojph_block_decoder.cpp:327-365: the VLC bytes are read backwards; a byte whose low
// seven bits are all ones is 7 bits wide when the byte read before it was > 0x8F; bytes before
// the block start, or past what the reference delivers, read as 0.  32 bytes per step, lane l
// takes the l-th; one ballot gives every byte's bit position.
static __device__ __forceinline__ void vlc_fill(Streams& s, uint32_t lane)
{
	const uint32_t k = s.vlc_k + lane;
	const int a = s.vlc_a0 - (int)k;
	const uint32_t b = ((int)k < s.vlc_real && a >= 0) ? (uint32_t)__ldg(s.d + a) : 0u;
	const uint32_t pb = __shfl_up_sync(0xffffffffu, b, 1);
	const bool un = lane ? pb > 0x8Fu : s.vlc_unstuff != 0;
	const uint32_t narrow = __ballot_sync(0xffffffffu, un && (b & 0x7Fu) == 0x7Fu);
	const uint32_t at = s.vlc_wp + 8 * lane - __popc(narrow & ((1u << lane) - 1u));
	if(lane < 9)
		s.vlc_ring[((s.vlc_wp >> 5) + 1 + lane) & (VLC_RING_WORDS - 1)] = 0;
	__syncwarp();
	const uint32_t sh = at & 31u;
	atomicOr(&s.vlc_ring[(at >> 5) & (VLC_RING_WORDS - 1)], b << sh);
	if(sh > 24)
		atomicOr(&s.vlc_ring[((at >> 5) + 1) & (VLC_RING_WORDS - 1)], b >> (32 - sh));
	__syncwarp();
	s.vlc_wp += 256 - __popc(narrow);
	s.vlc_unstuff = __shfl_sync(0xffffffffu, b, 31) > 0x8Fu;
	s.vlc_k += 32;
}
// The MagSgn stream (frwd_read<0xFF>, ojph_block_decoder.cpp:871-900) as a ring of un-stuffed
// bits, least significant first like the reference's accumulator: a byte that follows 0xFF is
// 7 bits wide, its eighth bit ORs with the next byte's first (kept that way for bit-exactness on
// malformed streams); past the end the stream reads as 0xFF.  All fields are warp-uniform.
constexpr uint32_t MS_RING_WORDS = 256; // 8192 bits: a group of 32 quads needs at most 3840
struct MagSgn
{
	const uint8_t* d;
	uint32_t len, pos; // stream length, next byte to fetch
	uint32_t prevff; // the byte before pos was 0xFF
	uint32_t rp, wp; // read / fill position in bits of the un-stuffed stream
	uint32_t* ring;
};

static __device__ __forceinline__ void ms_init(MagSgn& m, const uint8_t* d, uint32_t len, uint32_t* ring,
												uint32_t lane)
{
	m.d = d;
	m.len = len;
	m.pos = 0;
	m.prevff = 0;
	m.rp = m.wp = 0;
	m.ring = ring;
	if(lane == 0)
		ring[0] = 0;
	__syncwarp();
}

// un-stuff the next 32 bytes into the ring (lane l takes byte pos + l)
static __device__ __forceinline__ void ms_fill(MagSgn& m, uint32_t lane)
{
	const uint32_t i = m.pos + lane;
	const uint32_t b = i < m.len ? (uint32_t)__ldg(m.d + i) : 0xFFu;
	uint32_t pb = __shfl_up_sync(0xffffffffu, b, 1);
	if(lane == 0)
		pb = m.prevff ? 0xFFu : 0u;
	const uint32_t narrow = __ballot_sync(0xffffffffu, pb == 0xFFu); // bytes that are 7 bits wide
	const uint32_t at = m.wp + 8 * lane - __popc(narrow & ((1u << lane) - 1u));
	// the words this step enters (at most 8, plus the one the last byte's eighth bit may touch)
	if(lane < 9)
		m.ring[((m.wp >> 5) + 1 + lane) & (MS_RING_WORDS - 1)] = 0;
	__syncwarp();
	const uint32_t sh = at & 31u;
	atomicOr(&m.ring[(at >> 5) & (MS_RING_WORDS - 1)], b << sh);
	if(sh > 24)
		atomicOr(&m.ring[((at >> 5) + 1) & (MS_RING_WORDS - 1)], b >> (32 - sh));
	__syncwarp();
	m.wp += 256 - __popc(narrow);
	m.prevff = __shfl_sync(0xffffffffu, b, 31) == 0xFFu;
	m.pos += 32;
}

// 32 bits of the un-stuffed stream starting at bit `at`
static __device__ __forceinline__ uint32_t ms_bits_at(const MagSgn& m, uint32_t at)
{
	const uint32_t i = at >> 5;
	return __funnelshift_r(m.ring[i & (MS_RING_WORDS - 1)], m.ring[(i + 1) & (MS_RING_WORDS - 1)], at & 31u);
}

static __device__ __forceinline__ void streams_init(Streams& s, const uint8_t* d, int lcup, int scup,
													 uint32_t* vlc_ring, uint32_t lane)
{
	s.d = d;
	s.mel_pos = lcup - scup;
	s.mel_left = scup - 1;
	s.mel_acc = 0;
	s.mel_bits = s.mel_unstuff = s.mel_k = s.mel_run = s.mel_one = 0;
	// rev_init, :392-421: the first four bits come from the byte that also holds Scup's low half
	const uint32_t first = __ldg(d + lcup - 2);
	const uint32_t acc = first >> 4;
	s.vlc_ring = vlc_ring;
	if(lane == 0)
		vlc_ring[0] = acc;
	s.vlc_rp = 0;
	s.vlc_wp = 4 - (((acc & 7) == 7) ? 1 : 0);
	s.vlc_unstuff = (first | 0xF) > 0x8F;
	s.vlc_a0 = lcup - 3;
	s.vlc_k = 0;
	// real bytes the reference delivers before zeros: 1 + (address & 3) single bytes, then whole
	// 32-bit words while any byte of the segment is left
	int size = scup - 2;
	int tnum = min(1 + (s.vlc_a0 & 3), size);
	tnum = max(tnum, 0);
	size -= tnum;
	s.vlc_real = tnum + (size > 0 ? ((size + 3) >> 2) << 2 : 0);
	__syncwarp();
}

// the two u values of a quad pair from the table entry and the stream bits; returns bits used
static __device__ __forceinline__ uint32_t uvlc_decode(uint32_t vlc, uint32_t mode, uint32_t row,
														uint32_t& u0, uint32_t& u1)
{
	const uint32_t e = c_ht.uvlc[row][mode][vlc & 63u];
	const uint32_t plen = e & 7u, wa = (e >> 3) & 7u, wb = (e >> 6) & 7u;
	vlc >>= plen;
	const uint32_t fa = ((e >> 9) & 31u) + (vlc & ((1u << wa) - 1u));
	vlc >>= wa;
	const uint32_t fb = ((e >> 14) & 31u) + (vlc & ((1u << wb) - 1u));
	const bool swap = (e >> 19) & 1u;
	u0 = swap ? fb : fa;
	u1 = swap ? fa : fb;
	return plen + wa + wb;
}

// per-warp shared memory: the MagSgn ring, the VLC ring, the bottom-row significance masks of the
// previous / current quad row (one even-column and one odd-column word per 64 columns), the
// (table word, u) of the 32 quads of a group, then
// the north / current exponent rows (index = column + 1, a lane may look three columns past its
// quad)
constexpr uint32_t HT_MAX_GROUPS = 16; // code blocks are at most 1024 wide
static __host__ __device__ inline uint32_t ht_ncol(uint32_t max_w)
{
	return (((max_w + 63u) & ~63u) + 8u + 15u) & ~15u;
}
static __host__ __device__ inline size_t ht_warp_bytes(uint32_t max_w)
{
	return (MS_RING_WORDS + VLC_RING_WORDS + 4 * HT_MAX_GROUPS + 64) * 4 + 2 * (size_t)ht_ncol(max_w);
}

constexpr int HT_THREADS = 256; // 8 warps = 8 code blocks in flight per CTA

// what the store needs to turn a decoded sample into the plane's value: ShiftHTFilter /
// ScaleHTFilter and their ROI forms (PostDecompressFilters.h:91-177; the ROI forms AND the
// shifted magnitude with the sign bit exactly as the reference does)
struct Dequant
{
	uint32_t p; // 30 - missing_msbs: position of the magnitude's top bit
	uint32_t shift, roishift, thresh;
	int reversible;
	float scale;
};

// s: sign-magnitude, as the block decoder leaves it
static __device__ __forceinline__ uint32_t ht_dequant(const Dequant& dq, uint32_t s)
{
	const uint32_t mag0 = s & 0x7FFFFFFFu;
	if(dq.roishift && mag0 >= dq.thresh)
		s = (mag0 >> dq.roishift) & (s & 0x80000000u);
	const uint32_t mag = s & 0x7FFFFFFFu;
	if(dq.reversible)
	{
		const int32_t sv = (int32_t)(mag >> dq.shift);
		return (uint32_t)((s & 0x80000000u) ? -sv : sv);
	}
	const float f = __fmul_rn((float)(int32_t)mag, dq.scale);
	return __float_as_uint((s & 0x80000000u) ? -f : f);
}

template<bool RAW>
static __device__ __forceinline__ uint32_t ht_sample(const Dequant& dq, uint32_t bits, uint32_t m,
													  uint32_t e1, uint32_t& exponent)
{
	uint32_t v = bits & ((1u << m) - 1u);
	v |= e1 << m;
	v |= 1u;
	exponent = 32 - __clz(v);
	const uint32_t s = (bits << 31) | ((v + 2) << (dq.p - 1));
	return RAW ? s : ht_dequant(dq, s);
}

// ---- SigProp / MagRef ------------------------------------------------------------------------
// Raw-bit readers of the second segment, one byte at a time into a 64-bit window (lane 0 only).
// A 7-bit byte is OR-ed in whole, so a set top bit overlaps the next byte exactly as in the
// reference's accumulators (frwd_read<0> :875-945, rev_read_mrp :466-552); zeros once the segment
// is used up.
struct RawBits
{
	const uint8_t* d;
	int pos, left;
	uint32_t unstuff;
	uint64_t acc;
	int bits;
};
static __device__ __forceinline__ void spp_fill(RawBits& s)
{
	while(s.bits <= 56)
	{
		const uint32_t b = s.left > 0 ? (uint32_t)__ldg(s.d + s.pos) : 0u;
		s.pos++;
		s.left--;
		s.acc |= (uint64_t)b << s.bits;
		s.bits += 8 - (int)s.unstuff;
		s.unstuff = (b == 0xFFu);
	}
}
static __device__ __forceinline__ void mrp_fill(RawBits& s)
{
	while(s.bits <= 56)
	{
		const uint32_t b = s.left > 0 ? (uint32_t)__ldg(s.d + s.pos) : 0u;
		s.pos--;
		s.left--;
		const int wd = 8 - ((s.unstuff && (b & 0x7Fu) == 0x7Fu) ? 1 : 0);
		s.acc |= (uint64_t)b << s.bits;
		s.bits += wd;
		s.unstuff = b > 0x8Fu;
	}
}
static __device__ __forceinline__ uint32_t spp_bit(RawBits& s)
{
	if(s.bits < 1)
		spp_fill(s);
	const uint32_t r = (uint32_t)s.acc & 1u;
	s.acc >>= 1;
	s.bits--;
	return r;
}

constexpr uint32_t HT_ROW_WORDS = 32; // a block is at most 1024 columns wide

// Refinement of one block by one warp, after the cleanup pass left sign-magnitude words in
// `out`.  Stripe by stripe:
//   * the lanes turn six sample rows (the stripe, the row above, the row below) into bit rows of
//     "non-zero" in shared memory (one ballot per 32 columns);
//   * MagRef: every sample the cleanup pass made significant takes one bit, stripe-column order.
//     Lane 0 un-stuffs as many bits as the stripe has significant samples into shared memory,
//     then lane l applies them to columns l, l+32, ... (bit index = prefix sum of the per-column
//     counts): coalesced read-modify-write of the rows;
//   * SigProp: lane w derives the static membership words of 32 columns (zero samples with a
//     non-zero 8-neighbour); lane 0 walks the member columns in order, adding the members that
//     appear when a sample becomes significant (sample below, next column), reads the
//     significance bits and, after every group of four columns, the signs of that group.
// S: 6 x HT_ROW_WORDS words, M: 4 x HT_ROW_WORDS words, mbits: 128 words (per warp).
static __device__ __noinline__ void ht_refine(uint32_t* out, uint32_t ostride, uint32_t w, uint32_t h, uint32_t p,
								  uint32_t passes, const uint8_t* seg2, int len2, uint32_t* S,
								  uint32_t* M, uint32_t* mbits, uint32_t lane)
{
	RawBits spp{seg2, 0, len2, 0u, 0ull, 0};
	RawBits mrp{seg2, len2 - 1, len2, 1u, 0ull, 0};
	const uint32_t nw = (w + 31) >> 5;
	for(uint32_t y0 = 0; y0 < h; y0 += 4)
	{
		const uint32_t rows = min(4u, h - y0);
		// ---- bit rows: S[rr] = row y0 - 1 + rr
		for(uint32_t rr = 0; rr < 6; ++rr)
		{
			const int y = (int)y0 - 1 + (int)rr;
			const bool rowok = y >= 0 && y < (int)h;
			for(uint32_t c = 0; c < nw; ++c)
			{
				const uint32_t x = c * 32 + lane;
				const uint32_t v = (rowok && x < w) ? out[(size_t)y * ostride + x] : 0u;
				const uint32_t word = __ballot_sync(0xffffffffu, v != 0u);
				if(lane == 0)
					S[rr * HT_ROW_WORDS + c] = word;
			}
		}
		__syncwarp();
		// ---- MagRef (:1630-1701, :1924-1987)
		if(passes > 2)
		{
			// significant samples per 32-column chunk, summed over the stripe's rows
			uint32_t mine = 0; // lane c: count of chunk c
			if(lane < nw)
				for(uint32_t r = 0; r < rows; ++r)
					mine += (uint32_t)__popc(S[(r + 1) * HT_ROW_WORDS + lane]);
			uint32_t incl = mine;
#pragma unroll
			for(uint32_t d = 1; d < 32; d <<= 1)
			{
				const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
				if(lane >= d)
					incl += o;
			}
			const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
			if(lane == 0)
			{
				for(uint32_t i = 0; i * 32 < total; ++i)
				{
					if(mrp.bits < 32)
						mrp_fill(mrp);
					mbits[i] = (uint32_t)mrp.acc;
					const uint32_t take = min(32u, total - i * 32);
					mrp.acc >>= take;
					mrp.bits -= (int)take;
				}
			}
			__syncwarp();
			for(uint32_t c = 0; c < nw; ++c)
			{
				const uint32_t base = __shfl_sync(0xffffffffu, incl - mine, (int)c); // bits before chunk c
				const uint32_t x = c * 32 + lane;
				uint32_t nib = 0;
				for(uint32_t r = 0; r < rows; ++r)
					nib |= ((S[(r + 1) * HT_ROW_WORDS + c] >> lane) & 1u) << r;
				const uint32_t k = (uint32_t)__popc(nib);
				uint32_t at = k;
#pragma unroll
				for(uint32_t d = 1; d < 32; d <<= 1)
				{
					const uint32_t o = __shfl_up_sync(0xffffffffu, at, d);
					if(lane >= d)
						at += o;
				}
				at = base + at - k;
				for(uint32_t r = 0; r < rows; ++r)
					if((nib >> r) & 1u)
					{
						const uint32_t sym = (mbits[at >> 5] >> (at & 31u)) & 1u;
						++at;
						uint32_t* v = out + (size_t)(y0 + r) * ostride + x;
						*v = (*v ^ ((1u - sym) << (p - 1))) | (1u << (p - 2));
					}
			}
			__syncwarp();
		}
		// ---- SigProp (:1731-1916, :1990-2203): static membership, 32 columns per lane
		if(lane < nw)
		{
			const uint32_t last = (w & 31u) ? ((1u << (w & 31u)) - 1u) : 0xFFFFFFFFu;
			const uint32_t colmask = (lane == nw - 1) ? last : 0xFFFFFFFFu;
			uint32_t dil[6];
#pragma unroll
			for(uint32_t rr = 0; rr < 6; ++rr)
			{
				const uint32_t c = S[rr * HT_ROW_WORDS + lane];
				const uint32_t l = lane ? S[rr * HT_ROW_WORDS + lane - 1] : 0u;
				const uint32_t r = (lane + 1 < nw) ? S[rr * HT_ROW_WORDS + lane + 1] : 0u;
				dil[rr] = c | (c << 1) | (c >> 1) | (l >> 31) | (r << 31);
			}
#pragma unroll
			for(uint32_t r = 0; r < 4; ++r)
			{
				uint32_t m = 0;
				if(r < rows)
					m = (dil[r] | dil[r + 1] | dil[r + 2]) & ~S[(r + 1) * HT_ROW_WORDS + lane] & colmask;
				M[r * HT_ROW_WORDS + lane] = m;
			}
		}
		__syncwarp();
		if(lane == 0)
		{
			const uint32_t rowmask = (1u << rows) - 1u;
			const uint32_t val = 3u << (p - 2); // newly significant, bin centre included
			uint32_t dyn = 0; // members of column x added by new significant samples
			uint32_t nfresh = 0;
			uint64_t fresh = 0; // (x & 3) | r << 2, four bits per new sample of the group
			int group = -1;
			uint32_t x = 0;
			auto flush = [&]() {
				for(uint32_t i = 0; i < nfresh; ++i)
				{
					const uint32_t e = (uint32_t)(fresh >> (4 * i)) & 15u;
					const uint32_t fx = (uint32_t)group * 4 + (e & 3u), fr = e >> 2;
					out[(size_t)(y0 + fr) * ostride + fx] = val | (spp_bit(spp) << 31);
				}
				nfresh = 0;
				fresh = 0;
			};
			while(x < w)
			{
				if(dyn == 0)
				{
					// next column with a static member
					uint32_t c = x >> 5;
					uint32_t cm = (M[c] | M[HT_ROW_WORDS + c] | M[2 * HT_ROW_WORDS + c] | M[3 * HT_ROW_WORDS + c]) &
								  (0xFFFFFFFFu << (x & 31u));
					while(cm == 0 && ++c < nw)
						cm = M[c] | M[HT_ROW_WORDS + c] | M[2 * HT_ROW_WORDS + c] | M[3 * HT_ROW_WORDS + c];
					if(cm == 0)
						break;
					x = c * 32 + (uint32_t)__ffs((int)cm) - 1u;
				}
				if((int)(x >> 2) != group)
				{
					flush();
					group = (int)(x >> 2);
				}
				const uint32_t c = x >> 5, b = x & 31u;
				uint32_t m = dyn, sg = 0;
#pragma unroll
				for(uint32_t r = 0; r < 4; ++r)
				{
					m |= ((M[r * HT_ROW_WORDS + c] >> b) & 1u) << r;
					sg |= ((S[(r + 1) * HT_ROW_WORDS + c] >> b) & 1u) << r;
				}
				const uint32_t zero = ~sg & rowmask;
				m &= zero;
				uint32_t next = 0;
#pragma unroll
				for(uint32_t r = 0; r < 4; ++r)
					if((m >> r) & 1u)
					{
						if(spp_bit(spp))
						{
							fresh |= (uint64_t)((x & 3u) | (r << 2)) << (4 * nfresh);
							++nfresh;
							m |= (2u << r) & zero; // the sample below joins
							next |= ((7u << r) >> 1) & 15u; // rows r-1 .. r+1 of the next column
						}
					}
				dyn = next;
				++x;
			}
			flush();
		}
		__syncwarp();
	}
}

// Cleanup pass of one block by one warp (all lanes call this).  north/cur: per-column
// significance (bit 7) and magnitude exponent (low 7 bits) of the bottom sample row of the
// previous / current quad row, index = column + 1.  Returns false if the stream is rejected
// (the caller then zeroes the block).
template<bool RAW>
static __device__ bool ht_cleanup(const uint8_t* data, int lcup, uint32_t missing_msbs, uint32_t w,
								   uint32_t h, uint32_t* out, uint32_t ostride, const Dequant& dq,
								   uint8_t* north, uint8_t* cur, uint32_t ncol, uint32_t* ring,
								   uint32_t* vlc_ring, uint32_t* sig, uint2* quads, uint32_t lane)
{
	const int scup = ((int)__ldg(data + lcup - 1) << 4) + (__ldg(data + lcup - 2) & 0xF);
	if(scup < 2 || scup > lcup || scup > 4079)
		return false;
	Streams s;
	streams_init(s, data, lcup, scup, vlc_ring, lane);
	MagSgn ms;
	ms_init(ms, data, (uint32_t)(lcup - scup), ring, lane);
	for(uint32_t i = lane; i < 2 * ncol; i += 32)
		north[i] = 0; // north and cur are adjacent
	// significance of the bottom sample row, per 64 columns: [0] even columns, [1] odd columns
	uint32_t* psig = sig; // previous quad row
	uint32_t* csig = sig + 2 * HT_MAX_GROUPS;
	__syncwarp();
	for(uint32_t y = 0; y < h; y += 2)
	{
		const bool initial = (y == 0);
		const uint32_t row = initial ? 0u : 1u; // which CxtVLC / U-VLC table
		uint32_t carry = 0;
		for(uint32_t xg = 0; xg < w; xg += 64)
		{
			const uint32_t xend = min(xg + 64, w), g = xg >> 6;
			// ---- 1. MEL + CxtVLC + U-VLC of up to 32 quads, the same chain in every lane
			while(s.vlc_wp - s.vlc_rp < VLC_AHEAD)
				vlc_fill(s, lane);
			// the two ring words under the VLC read position stay in registers
			uint32_t vi = s.vlc_rp >> 5;
			uint32_t vw0 = s.vlc_ring[vi & (VLC_RING_WORDS - 1)], vw1 = s.vlc_ring[(vi + 1) & (VLC_RING_WORDS - 1)];
			auto vlc_skip = [&](uint32_t n) {
				s.vlc_rp += n;
				if((s.vlc_rp >> 5) != vi)
				{
					vi = s.vlc_rp >> 5;
					vw0 = vw1;
					vw1 = s.vlc_ring[(vi + 1) & (VLC_RING_WORDS - 1)];
				}
			};
			// north context of quad q: bit 0 = columns 2q-1 | 2q, bit 2 = columns 2q+1 | 2q+2
			uint32_t north0 = 0, north2 = 0;
			if(!initial)
			{
				const uint32_t ev = psig[2 * g], od = psig[2 * g + 1];
				north0 = (od << 1) | ev | (g ? psig[2 * g - 1] >> 31 : 0u);
				north2 = od | (ev >> 1) | (xg + 64 < w ? psig[2 * g + 2] << 31 : 0u);
			}
			for(uint32_t x = xg, qi = 0; x < xend; x += 4, qi += 2)
			{
				uint32_t tw[2] = {0, 0};
				const uint32_t nq = (x + 2 < w) ? 2 : 1;
				for(uint32_t q = 0; q < nq; ++q)
				{
					const uint32_t cq = carry | ((north0 >> (qi + q)) & 1u) | (((north2 >> (qi + q)) & 1u) << 2);
					uint32_t t = c_ht.vlc[row][(cq << 7) | (__funnelshift_r(vw0, vw1, s.vlc_rp) & 0x7F)];
					if(cq == 0 && !mel_event(s))
						t = 0;
					vlc_skip(t & 7u);
					tw[q] = t;
					carry = (t >> 16) & 7u;
				}
				uint32_t mode = ((tw[0] >> 3) & 1u) | (((tw[1] >> 3) & 1u) << 1);
				if(initial && mode == 3 && mel_event(s))
					mode = 4;
				uint32_t U[2];
				vlc_skip(uvlc_decode(__funnelshift_r(vw0, vw1, s.vlc_rp), mode, row, U[0], U[1]));
				quads[qi] = make_uint2(tw[0], U[0]);
				quads[qi + 1] = make_uint2(tw[1], U[1]);
			}
			__syncwarp();
			// ---- 2. MagSgn, one quad per lane
			const uint32_t c0 = xg + 2 * lane; // first column of this lane's quad
			const bool live = c0 < xend;
			const uint2 mine = quads[lane];
			__syncwarp();
			const uint32_t myt = mine.x, myu = mine.y;
			const uint32_t rho = live ? (myt >> 4) & 15u : 0u;
			uint32_t U = myu;
			if(!initial && (rho & (rho - 1)))
			{
				const uint32_t E = max(max(north[c0] & 0x7Fu, north[c0 + 1] & 0x7Fu),
									   max(north[c0 + 2] & 0x7Fu, north[c0 + 3] & 0x7Fu));
				U += E > 2 ? E - 2 : 0;
			}
			if(__any_sync(0xffffffffu, live && U > missing_msbs + 1))
				return false;
			uint32_t mn[4], total = 0;
#pragma unroll
			for(uint32_t n = 0; n < 4; ++n)
			{
				mn[n] = ((rho >> n) & 1u) ? U - ((myt >> (8 + n)) & 1u) : 0u;
				total += mn[n];
			}
			uint32_t at = total; // inclusive prefix sum over the lanes
#pragma unroll
			for(uint32_t d = 1; d < 32; d <<= 1)
			{
				const uint32_t o = __shfl_up_sync(0xffffffffu, at, d);
				if(lane >= d)
					at += o;
			}
			const uint32_t need = __shfl_sync(0xffffffffu, at, 31);
			while(ms.wp - ms.rp < need)
				ms_fill(ms, lane);
			at = ms.rp + at - total;
			ms.rp += need;
			uint32_t val[4], ex[4];
#pragma unroll
			for(uint32_t n = 0; n < 4; ++n)
			{
				val[n] = 0;
				ex[n] = 0;
				if((rho >> n) & 1u)
				{
					val[n] = ht_sample<RAW>(dq, ms_bits_at(ms, at), mn[n], (myt >> (12 + n)) & 1u, ex[n]);
					ex[n] |= 0x80u;
					at += mn[n];
				}
			}
			{
				const uint32_t ev = __ballot_sync(0xffffffffu, ex[1] != 0), od = __ballot_sync(0xffffffffu, ex[3] != 0);
				if(lane == 0)
				{
					csig[2 * g] = ev;
					csig[2 * g + 1] = od;
				}
			}
			if(live)
			{
				cur[c0 + 1] = (uint8_t)ex[1];
				cur[c0 + 2] = (uint8_t)ex[3];
#pragma unroll
				for(uint32_t r = 0; r < 2; ++r)
					if(y + r < h)
					{
						uint32_t* o = out + (size_t)(y + r) * ostride + c0;
						o[0] = val[r];
						if(c0 + 1 < w)
							o[1] = val[2 + r];
					}
			}
		}
		__syncwarp();
		uint8_t* tmp = north;
		north = cur;
		cur = tmp;
		uint32_t* tsig = psig;
		psig = csig;
		csig = tsig;
	}
	return true;
}

// REFINE: the table holds blocks with a second segment (known on the host); the instantiation
// without them keeps the one-pass decoder's registers and occupancy
template<bool REFINE>
__global__ void __launch_bounds__(HT_THREADS)
	t1_ht_decode_kernel(const T1Launch l, uint32_t max_w)
{
	extern __shared__ __align__(16) uint8_t smem[];
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t ncol = ht_ncol(max_w);
	uint8_t* mine = smem + (size_t)warp * ht_warp_bytes(max_w);
	uint32_t* ring = reinterpret_cast<uint32_t*>(mine);
	uint32_t* vlc_ring = ring + MS_RING_WORDS;
	uint32_t* sig = vlc_ring + VLC_RING_WORDS;
	uint2* quads = reinterpret_cast<uint2*>(sig + 4 * HT_MAX_GROUPS);
	uint8_t* north = reinterpret_cast<uint8_t*>(quads + 32);
	uint8_t* cur = north + ncol;
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
		if(!(cb.cblk_sty & GRKB_CBLKSTY_HT))
			continue;
		const uint32_t w = cb.w, h = cb.h;
		uint32_t* blk = reinterpret_cast<uint32_t*>(l.planes[cb.comp]) + (size_t)cb.y * l.stride + cb.x;
		// T1HT::decompress (T1HT.cpp:134-187): a block with segments but no passes or no bytes,
		// more than one pass (lengths2 is always 0 there), or k_msbs > 29 is an error
		bool decoded = false;
		if(cb.numsegs != 0)
		{
			uint32_t passes = 0, len = 0;
			for(uint32_t sgi = 0; sgi < cb.numsegs; ++sgi)
			{
				passes += l.segs[cb.seg_first + sgi].numpasses;
				len += l.segs[cb.seg_first + sgi].len;
			}
			const uint32_t mm = cb.k_msbs;
			// the first segment is the cleanup pass, what follows it is SigProp [+ MagRef]
			// (ojph_decode_codeblock's lengths1 / lengths2, :989-1043)
			const uint32_t len1 = l.segs[cb.seg_first].len, len2 = len - len1;
			const bool shape_ok = passes >= 1 && passes <= 3 && (passes == 1 || len2 != 0) &&
								  l.segs[cb.seg_first].numpasses == 1;
			if(shape_ok && len1 >= 2 && mm <= 29)
			{
				if(mm == 29)
					passes = 1; // :1031-1032
				Dequant dq;
				dq.p = 30 - mm;
				dq.shift = 31u - (mm + 1u);
				dq.roishift = cb.roishift;
				dq.thresh = cb.roishift ? (1u << cb.roishift) : 0xFFFFFFFFu;
				dq.reversible = l.reversible;
				dq.scale = cb.scale;
				if(passes == 1)
					decoded = ht_cleanup<false>(l.bytes + cb.data_off, (int)len1, mm, w, h, blk, l.stride, dq,
												north, cur, ncol, ring, vlc_ring, sig, quads, lane);
				else if(REFINE && ht_cleanup<true>(l.bytes + cb.data_off, (int)len1, mm, w, h, blk, l.stride,
													dq, north, cur, ncol, ring, vlc_ring, sig, quads, lane))
				{
					decoded = true;
					__syncwarp();
					// the cleanup pass's rings are free now: bit rows in the MagSgn ring, membership
					// words in (sig, quads), MagRef bits in the VLC ring
					ht_refine(blk, l.stride, w, h, dq.p, passes, l.bytes + cb.data_off + len1, (int)len2, ring,
							  sig, vlc_ring, lane);
					for(uint32_t y = 0; y < h; ++y)
						for(uint32_t x = lane; x < w; x += 32)
						{
							uint32_t* v = blk + (size_t)y * l.stride + x;
							*v = ht_dequant(dq, *v);
						}
				}
			}
			if(!decoded && lane == 0)
				atomicAdd(l.status, 1u);
		}
		__syncwarp();
		if(!decoded)
		{
			// empty or rejected block: zeros (0 and +0.0f are the same bits)
			for(uint32_t y = 0; y < h; ++y)
				for(uint32_t x = lane; x < w; x += 32)
					blk[(size_t)y * l.stride + x] = 0;
		}
		__syncwarp();
	}
}
} // namespace

int grkb_launch_t1_ht_decode(const T1Launch& l, cudaStream_t s)
{
	if(l.nblocks == 0)
		return 0;
	if(ht_tables_upload() != 0)
	{
		grkb_set_error("t1-ht: cannot upload decode tables");
		return -1;
	}
	const uint32_t max_w = l.max_ht_w ? l.max_ht_w : 64;
	const size_t smem = (size_t)(HT_THREADS / 32) * ht_warp_bytes(max_w);
	GRKB_ENSURE_SMEM(t1_ht_decode_kernel<false>, smem);
	GRKB_ENSURE_SMEM(t1_ht_decode_kernel<true>, smem);
	int dev = 0, sms = 148;
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	uint32_t ctas = (uint32_t)sms * 4;
	const uint32_t need = (l.nblocks + HT_THREADS / 32 - 1) / (HT_THREADS / 32);
	if(ctas > need)
		ctas = need;
	GRKB_CUDA_CHECK(cudaMemsetAsync(l.counter, 0, sizeof(uint32_t), s));
	if(l.ht_refine)
		t1_ht_decode_kernel<true><<<ctas, HT_THREADS, smem, s>>>(l, max_w);
	else
		t1_ht_decode_kernel<false><<<ctas, HT_THREADS, smem, s>>>(l, max_w);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
