// t1_ht_enc.cu -- HTJ2K (ISO/IEC 15444-15) cleanup-pass code-block encoder, quantiser fused into
// the load, sm_100a.  One code block per warp.
//
// Replaces, behind the C-ABI, the reference's
//   T1HT::preCompress / compress   src/lib/jp2/t1/t1_ht/T1HT.cpp:56-131
//   ojph_encode_codeblock          src/lib/jp2/t1/t1_ht/coding/ojph_block_encoder.cpp:470-945
//     (MEL :224-298, VLC :303-358, MagSgn :397-461, terminate_mel_vlc :363-392, tables :74-215)
//
// Shape: everything that does not depend on the position in the byte streams is done by the
// whole warp, one quad (2x2 samples) per lane and 32 quads per step: quantisation, significance
// pattern rho, exponents, the context c_q (significance of the quad to the left via a shuffle and
// of the row above from shared memory), kappa / U_q / u_q, the EMB pattern and the CxtVLC table
// look-up.  Lanes leave one record per quad in shared memory; lane 0 then appends the records
// to the three byte streams (MagSgn forwards into the block's slot in global memory, MEL and VLC
// in shared memory), which is serial because of bit stuffing.  The warp finally copies MEL and
// the reversed VLC stream behind MagSgn and writes Scup.
#include "t1.cuh"
#include "ht_cxtvlc.inc"

namespace
{
#define TW_CQ(w) ((w)&7u)
#define TW_RHO(w) (((w) >> 3) & 15u)
#define TW_UOFF(w) (((w) >> 7) & 1u)
#define TW_EK(w) (((w) >> 8) & 15u)
#define TW_E1(w) (((w) >> 12) & 15u)
#define TW_CWD(w) (((w) >> 16) & 127u)
#define TW_LEN(w) (((w) >> 23) & 7u)

// Encoder look-up: [initial row / other rows][c_q << 8 | rho << 4 | eps] -> table word.
// Selection rule of the reference (ojph_block_encoder.cpp:74-131): without an EMB pattern the
// first u_off = 0 entry; with one, among the u_off = 1 entries whose known bits agree
// ((eps & e_k) == e_1) the LAST one with the most known bits.
struct HtEncTables
{
	uint32_t enc[2][2048];
};

HtEncTables* g_dev[64] = {};

const HtEncTables* ht_enc_tables_device()
{
	int dev = 0;
	if(cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
		return nullptr;
	if(g_dev[dev])
		return g_dev[dev];
	static HtEncTables h;
	static bool built = false;
	if(!built)
	{
		for(int t = 0; t < 2; ++t)
		{
			const unsigned int* src = t ? HT_CXTVLC_ROWN : HT_CXTVLC_ROW0;
			const size_t n = t ? sizeof(HT_CXTVLC_ROWN) / 4 : sizeof(HT_CXTVLC_ROW0) / 4;
			for(uint32_t i = 0; i < 2048; ++i)
			{
				const uint32_t cq = i >> 8, rho = (i >> 4) & 15, eps = i & 15;
				h.enc[t][i] = 0;
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
							h.enc[t][i] = w;
							break;
						}
					}
					else if(TW_UOFF(w) == 1 && (eps & TW_EK(w)) == TW_E1(w) &&
							__builtin_popcount(TW_EK(w)) >= best)
					{
						best = __builtin_popcount(TW_EK(w));
						h.enc[t][i] = w;
					}
				}
			}
		}
		built = true;
	}
	HtEncTables* d = nullptr;
	if(cudaMalloc(&d, sizeof(HtEncTables)) != cudaSuccess)
		return nullptr;
	if(cudaMemcpy(d, &h, sizeof(HtEncTables), cudaMemcpyHostToDevice) != cudaSuccess)
	{
		cudaFree(d);
		return nullptr;
	}
	g_dev[dev] = d;
	return d;
}

constexpr uint32_t SLOT_LEAD = 16;
constexpr uint32_t MS_CAP = 16384, MEL_CAP = 192, VLC_CAP = 3072;
constexpr uint32_t MSW = 128; // words of the MagSgn step buffer (+4 behind it): 32 quads x 4 samples x <= 32 bits + 7
constexpr uint32_t VLW = 32; // words of the VLC step buffer: 32 quads x <= 15 bits + 7

// MEL coder state (uniform across the warp: every lane runs the same chain; lane 0 stores)
struct HtMel
{
	uint8_t* buf; // shared
	uint32_t pos, tmp;
	int rem, run, k;
	bool overflow;
};

// mel_emit_bit / mel_encode, ojph_block_encoder.cpp:247-298
static __device__ __forceinline__ void enc_mel_bit(HtMel& e, int v, uint32_t lane)
{
	e.tmp = (e.tmp << 1) + (uint32_t)v;
	if(--e.rem == 0)
	{
		if(e.pos >= 128)
		{
			e.overflow = true;
			e.rem = 8;
			e.tmp = 0;
			return;
		}
		if(lane == 0)
			e.buf[e.pos] = (uint8_t)e.tmp;
		e.pos++;
		e.rem = (e.tmp == 0xFF) ? 7 : 8;
		e.tmp = 0;
	}
}

static __device__ __forceinline__ void enc_mel(HtMel& e, int bit, uint32_t lane)
{
	// MEL exponents of table 2 of 15444-15: 0,0,0,1,1,1,2,2,2,3,3,4,5
	const int ex = (int)((0x5433222111000ull >> (4 * e.k)) & 15);
	if(!bit)
	{
		if(++e.run >= (1 << ex))
		{
			enc_mel_bit(e, 1, lane);
			e.run = 0;
			e.k = e.k < 12 ? e.k + 1 : 12;
		}
	}
	else
	{
		enc_mel_bit(e, 0, lane);
		for(int t = ex; t > 0;)
			enc_mel_bit(e, (e.run >> --t) & 1, lane);
		e.run = 0;
		e.k = e.k > 0 ? e.k - 1 : 0;
	}
}

// U-VLC codeword of u (uvlc tables, ojph_block_encoder.cpp:194-215): prefix | suffix packed as
// pre | pre_len << 8 | suf << 16 | suf_len << 24
static __device__ __forceinline__ uint32_t uvlc_code(int u)
{
	if(u <= 0)
		return 0;
	if(u <= 2)
		return (uint32_t)u | ((uint32_t)u << 8);
	if(u <= 4)
		return 4u | (3u << 8) | ((uint32_t)(u - 3) << 16) | (1u << 24);
	return 0u | (3u << 8) | ((uint32_t)(u - 5) << 16) | (5u << 24);
}

// OR `len` (<= 32) bits of `val` into a zeroed bit buffer at bit offset `off`
static __device__ __forceinline__ void put_bits(uint32_t* buf, uint32_t off, uint32_t val, uint32_t len)
{
	if(len == 0)
		return;
	const uint64_t x = (uint64_t)val << (off & 31u);
	const uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
	if(lo)
		atomicOr(buf + (off >> 5), lo);
	if(hi)
		atomicOr(buf + (off >> 5) + 1, hi);
}

// 8 bits of a bit buffer starting at bit offset `off`
static __device__ __forceinline__ uint32_t get_byte(const uint32_t* buf, uint32_t off)
{
	const uint32_t i = off >> 5;
	return __funnelshift_r(buf[i], buf[i + 1], off & 31u) & 0xFFu;
}

static __device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, uint32_t lane)
{
#pragma unroll
	for(int d = 1; d < 32; d <<= 1)
	{
		const uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
		if((int)lane >= d)
			v += t;
	}
	return v;
}

constexpr int HTE_THREADS = 256;

static __host__ __device__ __forceinline__ size_t hte_per_warp(uint32_t ncol)
{
	return (size_t)(MSW + 4) * 4 + (size_t)(VLW + 4) * 4 + MEL_CAP + VLC_CAP +
		   4 * (size_t)((ncol + 3) & ~3u);
}

__global__ void __launch_bounds__(HTE_THREADS)
	t1_ht_encode_kernel(const T1EncLaunch l, const HtEncTables* __restrict__ tab)
{
	extern __shared__ __align__(16) uint8_t smem[];
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t ncol = l.max_w + 8;
	uint8_t* base = smem + warp * hte_per_warp(ncol);
	uint32_t* msbuf = reinterpret_cast<uint32_t*>(base); // MagSgn bits of one step, un-stuffed
	uint32_t* vlbuf = msbuf + MSW + 4; // VLC bits of one step, un-stuffed
	uint8_t* mel = reinterpret_cast<uint8_t*>(vlbuf + VLW + 4);
	uint8_t* vlc = mel + MEL_CAP;
	uint8_t* rowbuf = vlc + VLC_CAP; // expN | sigN | expC | sigC, ncol bytes each

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
		if(!(cb.cblk_sty & GRKB_CBLKSTY_HT))
			continue;
		const uint32_t w = cb.w, h = cb.h;
		const uint32_t missing = cb.k_msbs;
		const bool bad = missing > 30;
		const uint32_t p = 30 - (bad ? 0 : missing);
		const int32_t* plane =
			reinterpret_cast<const int32_t*>(l.planes[cb.comp]) + (size_t)cb.y * l.stride + cb.x;
		uint8_t* const ms = l.slots + (size_t)bi * l.slot_bytes + SLOT_LEAD; // global: the block's slot
		const uint32_t ms_cap = min(MS_CAP, l.slot_bytes - SLOT_LEAD - MEL_CAP - VLC_CAP);
		// stream states, identical in all lanes.  MagSgn (ms_encode, :397-425): forward, LSB
		// first, the byte after 0xFF holds 7 bits.  ms_w0 = width of the next byte, ms_left /
		// ms_nleft = bits not yet forming a byte.
		uint32_t ms_pos = 0, ms_w0 = 8, ms_left = 0, ms_nleft = 0;
		// VLC (vlc_encode, :325-358): backwards, LSB first; after a byte > 0x8F the next byte
		// is 7 bits wide if those seven bits are all ones.  vlc_init, :313-323: the last byte
		// is a placeholder, its neighbour starts with four 1s.
		uint32_t vlc_pos = 1, vl_left = 0xF, vl_nleft = 4, vl_gt = 1;
		HtMel me;
		me.buf = mel;
		me.pos = 0;
		me.tmp = 0;
		me.rem = 8;
		me.run = 0;
		me.k = 0;
		me.overflow = false;
		bool overflow = bad;
		if(lane == 0)
			vlc[VLC_CAP - 1] = 0xFF;
		uint8_t *expN = rowbuf, *sigN = rowbuf + ncol, *expC = rowbuf + 2 * ncol, *sigC = rowbuf + 3 * ncol;
		const uint32_t nquads = (w + 1) >> 1;
		for(uint32_t y = 0; y < h && !bad; y += 2)
		{
			const bool initial = (y == 0);
			for(uint32_t i = lane; i < ncol; i += 32)
			{
				expC[i] = 0;
				sigC[i] = 0;
			}
			__syncwarp();
			uint32_t carry_in = 0;
			for(uint32_t g0 = 0; g0 < nquads; g0 += 32)
			{
				const uint32_t qi = g0 + lane, c0 = 2 * qi;
				const bool live = qi < nquads;
				uint32_t rho = 0, v[4] = {0, 0, 0, 0};
				int eq[4] = {0, 0, 0, 0}, emax = 0;
				// the step buffers are cleared while the samples load
				*reinterpret_cast<uint4*>(msbuf + 4 * lane) = make_uint4(0, 0, 0, 0);
				if(lane < (VLW + 4) / 4)
					*reinterpret_cast<uint4*>(vlbuf + 4 * lane) = make_uint4(0, 0, 0, 0);
				if(lane == 0)
					*reinterpret_cast<uint4*>(msbuf + MSW) = make_uint4(0, 0, 0, 0);
				if(live)
				{
#pragma unroll
					for(uint32_t n = 0; n < 4; ++n)
					{
						const uint32_t col = c0 + (n >> 1), row = y + (n & 1);
						if(col >= w || row >= h)
							continue;
						// T1HT::preCompress, T1HT.cpp:70-88: sign | abs << (31 - (k_msbs + 1))
						const int32_t x = plane[(size_t)row * l.stride + col];
						const uint32_t t = (x < 0 ? 0x80000000u : 0u) | ((uint32_t)(x < 0 ? -x : x) << p);
						const uint32_t val = ((t + t) >> p) & ~1u; // 2 * magnitude
						if(!val)
							continue;
						rho |= 1u << n;
						eq[n] = 32 - __clz(val - 1);
						emax = max(emax, eq[n]);
						v[n] = (val - 2) + (t >> 31);
						if(n & 1)
						{
							expC[col + 1] = (uint8_t)eq[n];
							sigC[col + 1] = 1;
						}
					}
				}
				// context: significance of the quad to the left (same row) and of the row above
				const uint32_t my_carry =
					initial ? ((rho >> 1) | (rho & 1)) : ((((rho >> 2) | (rho >> 3)) & 1u) << 1);
				uint32_t cq = __shfl_up_sync(0xffffffffu, my_carry, 1);
				if(lane == 0)
					cq = carry_in;
				carry_in = __shfl_sync(0xffffffffu, my_carry, 31);
				int kappa = 1;
				if(!initial && live)
				{
					cq |= (uint32_t)(sigN[c0] | sigN[c0 + 1]);
					cq |= (uint32_t)(sigN[c0 + 2] | sigN[c0 + 3]) << 2;
					if(rho & (rho - 1))
					{
						const int me2 = max(max((int)expN[c0], (int)expN[c0 + 1]),
											max((int)expN[c0 + 2], (int)expN[c0 + 3]));
						kappa = max(me2 - 1, 1);
					}
				}
				const int U = max(emax, kappa);
				const int u = live ? U - kappa : 0;
				uint32_t eps = 0;
				if(u > 0)
				{
#pragma unroll
					for(uint32_t n = 0; n < 4; ++n)
						eps |= (uint32_t)(eq[n] == emax) << n;
				}
				const uint32_t tw = live ? tab->enc[initial ? 0 : 1][(cq << 8) | (rho << 4) | eps] : 0u;

				// ---- MagSgn: every sample's field at its prefix-sum bit offset ----------------
				uint32_t mb[4], mtot = 0;
#pragma unroll
				for(uint32_t n = 0; n < 4; ++n)
				{
					const int m = ((rho >> n) & 1) ? U - (int)((TW_EK(tw) >> n) & 1) : 0;
					mb[n] = (uint32_t)m;
					mtot += (uint32_t)m;
				}
				const uint32_t mincl = warp_incl_scan(mtot, lane);
				const uint32_t ms_bits = ms_nleft + __shfl_sync(0xffffffffu, mincl, 31);

				// ---- VLC: CxtVLC codeword of the quad, U-VLC of the pair (prefixes, then suffixes;
				// :693-716 initial row, :926-929 others) -----------------------------------------
				const int up = __shfl_xor_sync(0xffffffffu, u, 1);
				const bool odd = lane & 1u;
				const int u0 = odd ? up : u, u1 = odd ? u : up;
				uint32_t uc;
				if(initial && u0 > 2 && u1 > 2)
					uc = uvlc_code(u - 2);
				else if(initial && u0 > 2 && u1 > 0 && odd)
					uc = (uint32_t)(u - 1) | (1u << 8);
				else
					uc = uvlc_code(u);
				const uint32_t Lc = TW_LEN(tw), Lp = (uc >> 8) & 0xFFu, Ls = uc >> 24;
				const uint32_t lens = Lc | (Lp << 8) | (Ls << 16);
				const uint32_t plens = __shfl_xor_sync(0xffffffffu, lens, 1);
				const uint32_t vtot = Lc + Lp + Ls;
				const uint32_t vincl = warp_incl_scan(vtot, lane);
				const uint32_t vl_bits = vl_nleft + __shfl_sync(0xffffffffu, vincl, 31);
				const uint32_t pLc = plens & 0xFFu, pLp = (plens >> 8) & 0xFFu, pLs = plens >> 16;
				// pair layout: cwd0 cwd1 pre0 pre1 suf0 suf1
				const uint32_t pbase = vl_nleft + vincl - vtot - (odd ? pLc + pLp + pLs : 0u);
				const uint32_t o_cwd = pbase + (odd ? pLc : 0u);
				const uint32_t o_pre = pbase + Lc + pLc + (odd ? pLp : 0u);
				const uint32_t o_suf = pbase + Lc + pLc + Lp + pLp + (odd ? pLs : 0u);

				// ---- MEL events, in lane order: the quad's (c_q == 0), then the pair's ------------
				const uint32_t evA = __ballot_sync(0xffffffffu, live && cq == 0);
				const uint32_t evAv = __ballot_sync(0xffffffffu, rho != 0);
				const uint32_t evP = __ballot_sync(0xffffffffu, initial && odd && u0 > 0 && u1 > 0);
				const uint32_t evPv = __ballot_sync(0xffffffffu, min(u0, u1) > 2);

				__syncwarp(); // buffers cleared
				if(lane == 0)
				{
					if(ms_nleft)
						atomicOr(msbuf, ms_left);
					if(vl_nleft)
						atomicOr(vlbuf, vl_left);
				}
				{
					uint32_t off = ms_nleft + mincl - mtot;
#pragma unroll
					for(uint32_t n = 0; n < 4; ++n)
					{
						put_bits(msbuf, off, v[n] & ((mb[n] >= 32u) ? 0xFFFFFFFFu : ((1u << mb[n]) - 1u)),
								 mb[n]);
						off += mb[n];
					}
				}
				put_bits(vlbuf, o_cwd, TW_CWD(tw), Lc);
				put_bits(vlbuf, o_pre, uc & 0xFFu, Lp);
				put_bits(vlbuf, o_suf, (uc >> 16) & 0xFFu, Ls);

				// MEL chain while the atomics land
				for(uint32_t M = evA | evP; M;)
				{
					const uint32_t b = (uint32_t)__ffs((int)M) - 1u;
					M &= M - 1u;
					if((evA >> b) & 1u)
						enc_mel(me, (int)((evAv >> b) & 1u), lane);
					if((evP >> b) & 1u)
						enc_mel(me, (int)((evPv >> b) & 1u), lane);
				}
				__syncwarp(); // bits in place

				// ---- MagSgn flush: whole bytes, up to and including the first 0xFF per round ----
				{
					uint32_t rd = 0, avail = ms_bits;
					while(avail >= ms_w0)
					{
						const uint32_t nb = min(32u, 1u + ((avail - ms_w0) >> 3));
						const uint32_t o = rd + (lane ? ms_w0 + 8u * (lane - 1u) : 0u);
						uint32_t byte = 0;
						if(lane < nb)
							byte = get_byte(msbuf, o) & (lane ? 0xFFu : ((1u << ms_w0) - 1u));
						const uint32_t ff = __ballot_sync(0xffffffffu, lane < nb && byte == 0xFFu);
						const uint32_t ne = ff ? (uint32_t)__ffs((int)ff) : nb;
						if(ms_pos + ne > ms_cap)
						{
							overflow = true;
							avail = 0;
							break;
						}
						if(lane < ne)
							ms[ms_pos + lane] = (uint8_t)byte;
						const uint32_t used = ms_w0 + 8u * (ne - 1u);
						rd += used;
						avail -= used;
						ms_pos += ne;
						ms_w0 = ff ? 7u : 8u;
					}
					ms_nleft = avail;
					ms_left = avail ? (get_byte(msbuf, rd) & ((1u << avail) - 1u)) : 0u;
				}
				// ---- VLC flush: whole bytes, up to and including the first 7-bit byte per round --
				{
					uint32_t rd = 0, avail = vl_bits;
					while(avail >= 8u)
					{
						const uint32_t nb = min(32u, avail >> 3);
						const uint32_t byte = (lane < nb) ? get_byte(vlbuf, rd + 8u * lane) : 0u;
						const uint32_t prev = __shfl_up_sync(0xffffffffu, byte, 1);
						const bool gt = lane ? (prev > 0x8Fu) : (vl_gt != 0u);
						const uint32_t sh =
							__ballot_sync(0xffffffffu, lane < nb && gt && (byte & 0x7Fu) == 0x7Fu);
						const uint32_t ne = sh ? (uint32_t)__ffs((int)sh) : nb;
						if(vlc_pos + ne > VLC_CAP - 128)
						{
							overflow = true;
							avail = 0;
							break;
						}
						const uint32_t outb = (sh && lane == ne - 1u) ? 0x7Fu : byte;
						if(lane < ne)
							vlc[VLC_CAP - 1 - (vlc_pos + lane)] = (uint8_t)outb;
						vl_gt = (__shfl_sync(0xffffffffu, outb, (int)(ne - 1u)) > 0x8Fu) ? 1u : 0u;
						const uint32_t used = 8u * ne - (sh ? 1u : 0u);
						rd += used;
						avail -= used;
						vlc_pos += ne;
					}
					vl_nleft = avail;
					vl_left = avail ? (get_byte(vlbuf, rd) & ((1u << avail) - 1u)) : 0u;
				}
				__syncwarp(); // buffers are cleared again at the top
			}
			uint8_t* t1 = expN;
			expN = expC;
			expC = t1;
			t1 = sigN;
			sigN = sigC;
			sigC = t1;
		}
		// ---- termination (every lane computes it, lane 0 stores), then the warp appends MEL
		// and VLC behind MagSgn ----------------------------------------------------------------
		uint32_t total = 0, mel_pos = 0;
		uint32_t failed = bad ? 1u : 0u;
		if(!bad)
		{
			// a 7-bit byte completed by the last bits (vlc_encode emits it as soon as it is full)
			if(vl_nleft == 7 && vl_gt && vl_left == 0x7Fu)
			{
				if(lane == 0)
					vlc[VLC_CAP - 1 - vlc_pos] = 0x7F;
				vlc_pos++;
				vl_nleft = 0;
				vl_left = 0;
			}
			// terminate_mel_vlc, :363-392
			if(me.run > 0)
				enc_mel_bit(me, 1, lane);
			me.tmp <<= me.rem;
			const uint32_t mel_mask = (0xFFu << me.rem) & 0xFF;
			const uint32_t vlc_mask = 0xFFu >> (8 - vl_nleft);
			if((mel_mask | vlc_mask) != 0)
			{
				const uint32_t fuse = me.tmp | vl_left;
				if((((fuse ^ me.tmp) & mel_mask) | ((fuse ^ vl_left) & vlc_mask)) == 0 && fuse != 0xFF &&
				   vlc_pos > 1)
				{
					if(lane == 0)
						mel[me.pos] = (uint8_t)fuse;
					me.pos++;
				}
				else
				{
					if(lane == 0)
					{
						mel[me.pos] = (uint8_t)me.tmp;
						vlc[VLC_CAP - 1 - vlc_pos] = (uint8_t)vl_left;
					}
					me.pos++;
					vlc_pos++;
				}
			}
			// ms_terminate, :427-461
			if(ms_nleft)
			{
				const uint32_t t = ms_w0 - ms_nleft;
				ms_left |= (0xFFu & ((1u << t) - 1u)) << ms_nleft;
				if(ms_left != 0xFF)
				{
					if(ms_pos < ms_cap)
					{
						if(lane == 0)
							ms[ms_pos] = (uint8_t)ms_left;
						ms_pos++;
					}
					else
						overflow = true;
				}
			}
			else if(ms_w0 == 7)
				ms_pos--;
			mel_pos = me.pos;
			failed = (overflow || me.overflow) ? 1u : 0u;
		}
		__syncwarp();
		if(!failed)
		{
			total = ms_pos + mel_pos + vlc_pos;
			uint8_t* out = ms;
			for(uint32_t i = lane; i < mel_pos; i += 32)
				out[ms_pos + i] = mel[i];
			for(uint32_t i = lane; i < vlc_pos; i += 32)
				out[ms_pos + mel_pos + i] = vlc[VLC_CAP - vlc_pos + i];
			__syncwarp();
			if(lane == 0)
			{
				// Scup in the last twelve bits (:931-942)
				const uint32_t nb = mel_pos + vlc_pos;
				out[total - 1] = (uint8_t)(nb >> 4);
				out[total - 2] = (uint8_t)((out[total - 2] & 0xF0) | (nb & 0xF));
			}
		}
		if(lane == 0)
		{
			grkb_coded res;
			res.data_off = 0;
			res.data_len = failed ? 0 : total;
			res.numbps = 1; // T1HT.cpp:128
			res.numpasses = 1;
			res.failed = (uint16_t)failed;
			res.distortion = 0.0;
			l.coded[bi] = res;
			grkb_pass ps;
			ps.distortion = 0.0;
			ps.rate = res.data_len;
			ps.len = (uint16_t)res.data_len;
			ps.term = 0;
			ps.reserved = 0;
			l.passes[(size_t)bi * GRKB_ENC_MAX_PASSES] = ps;
		}
		__syncwarp();
	}
}
} // namespace

int grkb_launch_t1_ht_encode(const T1EncLaunch& l, cudaStream_t s)
{
	if(l.nblocks == 0)
		return 0;
	const HtEncTables* tab = ht_enc_tables_device();
	if(!tab)
	{
		grkb_set_error("t1 ht encode: cannot upload coding tables");
		return -1;
	}
	if(l.slot_bytes < SLOT_LEAD + MEL_CAP + VLC_CAP + 64)
	{
		grkb_set_error("t1 ht encode: slot too small");
		return -1;
	}
	const uint32_t ncol = l.max_w + 8;
	const size_t per_warp = hte_per_warp(ncol);
	const int nw = HTE_THREADS / 32;
	const size_t smem = per_warp * nw;
	if(smem > 227 * 1024)
	{
		grkb_set_error("t1 ht encode: code block too wide for shared memory");
		return -1;
	}
	GRKB_ENSURE_SMEM(t1_ht_encode_kernel, smem);
	int dev = 0, sms = 148;
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	int per_sm = (int)((227 * 1024) / (smem + 1024));
	if(per_sm < 1)
		per_sm = 1;
	if(per_sm > 4)
		per_sm = 4;
	uint32_t ctas = (uint32_t)(sms * per_sm);
	const uint32_t need = (l.nblocks + nw - 1) / nw;
	if(ctas > need)
		ctas = need;
	GRKB_CUDA_CHECK(cudaMemsetAsync(l.counter, 0, sizeof(uint32_t), s));
	t1_ht_encode_kernel<<<ctas, nw * 32, smem, s>>>(l, tab);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
