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

struct HtE
{
	uint8_t* ms; // global: the block's slot
	uint32_t ms_pos, ms_tmp, ms_cap;
	int ms_used, ms_max;
	uint8_t* mel; // shared
	uint32_t mel_pos, mel_tmp;
	int mel_rem, mel_run, mel_k;
	uint8_t* vlc; // shared, filled from the end
	uint32_t vlc_pos, vlc_tmp;
	int vlc_used, vlc_gt8f;
	bool overflow;
};

// ms_encode, ojph_block_encoder.cpp:397-425: LSB first, a byte after 0xFF holds 7 bits
static __device__ __forceinline__ void enc_ms(HtE& e, uint32_t cwd, int len)
{
	while(len > 0)
	{
		if(e.ms_pos >= e.ms_cap)
		{
			e.overflow = true;
			return;
		}
		int t = e.ms_max - e.ms_used;
		if(t > len)
			t = len;
		e.ms_tmp |= (cwd & ((1u << t) - 1)) << e.ms_used;
		e.ms_used += t;
		cwd >>= t;
		len -= t;
		if(e.ms_used >= e.ms_max)
		{
			e.ms[e.ms_pos++] = (uint8_t)e.ms_tmp;
			e.ms_max = (e.ms_tmp == 0xFF) ? 7 : 8;
			e.ms_tmp = 0;
			e.ms_used = 0;
		}
	}
}

// mel_emit_bit / mel_encode, ojph_block_encoder.cpp:247-298
static __device__ __forceinline__ void enc_mel_bit(HtE& e, int v)
{
	e.mel_tmp = (e.mel_tmp << 1) + (uint32_t)v;
	if(--e.mel_rem == 0)
	{
		if(e.mel_pos >= 128)
		{
			e.overflow = true;
			return;
		}
		e.mel[e.mel_pos++] = (uint8_t)e.mel_tmp;
		e.mel_rem = (e.mel_tmp == 0xFF) ? 7 : 8;
		e.mel_tmp = 0;
	}
}

static __device__ __forceinline__ void enc_mel(HtE& e, int bit)
{
	// MEL exponents of table 2 of 15444-15: 0,0,0,1,1,1,2,2,2,3,3,4,5
	const int ex = (int)((0x5433222111000ull >> (4 * e.mel_k)) & 15);
	if(!bit)
	{
		if(++e.mel_run >= (1 << ex))
		{
			enc_mel_bit(e, 1);
			e.mel_run = 0;
			e.mel_k = e.mel_k < 12 ? e.mel_k + 1 : 12;
		}
	}
	else
	{
		enc_mel_bit(e, 0);
		for(int t = ex; t > 0;)
			enc_mel_bit(e, (e.mel_run >> --t) & 1);
		e.mel_run = 0;
		e.mel_k = e.mel_k > 0 ? e.mel_k - 1 : 0;
	}
}

// vlc_encode, ojph_block_encoder.cpp:325-358: written backwards, LSB first; after a byte > 0x8F
// the next byte may use its 8th bit only if its low seven bits are not all ones
static __device__ __forceinline__ void enc_vlc(HtE& e, uint32_t cwd, int len)
{
	while(len > 0)
	{
		if(e.vlc_pos >= VLC_CAP - 128)
		{
			e.overflow = true;
			return;
		}
		int avail = 8 - e.vlc_gt8f - e.vlc_used;
		const int t = avail < len ? avail : len;
		e.vlc_tmp |= (cwd & ((1u << t) - 1)) << e.vlc_used;
		e.vlc_used += t;
		avail -= t;
		len -= t;
		cwd >>= t;
		if(avail == 0)
		{
			if(e.vlc_gt8f && e.vlc_tmp != 0x7F)
			{
				e.vlc_gt8f = 0;
				continue;
			}
			e.vlc[VLC_CAP - 1 - e.vlc_pos] = (uint8_t)e.vlc_tmp;
			e.vlc_pos++;
			e.vlc_gt8f = e.vlc_tmp > 0x8F;
			e.vlc_tmp = 0;
			e.vlc_used = 0;
		}
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

constexpr int HTE_THREADS = 256;
constexpr uint32_t REC_WORDS = 8;

__global__ void __launch_bounds__(HTE_THREADS)
	t1_ht_encode_kernel(const T1EncLaunch l, const HtEncTables* __restrict__ tab)
{
	extern __shared__ __align__(16) uint8_t smem[];
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t ncol = l.max_w + 8;
	const size_t per_warp = 32 * REC_WORDS * 4 + MEL_CAP + VLC_CAP + 4 * (size_t)((ncol + 3) & ~3u);
	uint8_t* base = smem + warp * per_warp;
	uint32_t* rec = reinterpret_cast<uint32_t*>(base);
	uint8_t* mel = base + 32 * REC_WORDS * 4;
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
		const uint32_t bi = l.order ? l.order[item] : item;
		const grkb_enc_cblk cb = l.blocks[bi];
		if(!(cb.cblk_sty & GRKB_CBLKSTY_HT))
			continue;
		const uint32_t w = cb.w, h = cb.h;
		const uint32_t missing = cb.k_msbs;
		const bool bad = missing > 30;
		const uint32_t p = 30 - (bad ? 0 : missing);
		const int32_t* plane =
			reinterpret_cast<const int32_t*>(l.planes[cb.comp]) + (size_t)cb.y * l.stride + cb.x;
		HtE e;
		e.ms = l.slots + (size_t)bi * l.slot_bytes + SLOT_LEAD;
		e.ms_cap = min(MS_CAP, l.slot_bytes - SLOT_LEAD - MEL_CAP - VLC_CAP);
		e.ms_pos = 0;
		e.ms_tmp = 0;
		e.ms_used = 0;
		e.ms_max = 8;
		e.mel = mel;
		e.mel_pos = 0;
		e.mel_tmp = 0;
		e.mel_rem = 8;
		e.mel_run = 0;
		e.mel_k = 0;
		e.vlc = vlc;
		// vlc_init, :313-323: the last byte is a placeholder, its neighbour starts with four 1s
		e.vlc_pos = 1;
		e.vlc_tmp = 0xF;
		e.vlc_used = 4;
		e.vlc_gt8f = 1;
		e.overflow = bad;
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
				uint32_t rho = 0, v[4] = {0, 0, 0, 0};
				int eq[4] = {0, 0, 0, 0}, emax = 0;
				if(qi < nquads)
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
				if(!initial && qi < nquads)
				{
					cq |= (uint32_t)(sigN[c0] | sigN[c0 + 1]);
					cq |= (uint32_t)(sigN[c0 + 2] | sigN[c0 + 3]) << 2;
					if(rho & (rho - 1))
					{
						const int me = max(max((int)expN[c0], (int)expN[c0 + 1]),
										   max((int)expN[c0 + 2], (int)expN[c0 + 3]));
						kappa = max(me - 1, 1);
					}
				}
				const int U = max(emax, kappa);
				const int u = U - kappa;
				uint32_t eps = 0;
				if(u > 0)
				{
#pragma unroll
					for(uint32_t n = 0; n < 4; ++n)
						eps |= (uint32_t)(eq[n] == emax) << n;
				}
				if(qi < nquads)
				{
					const uint32_t tw = tab->enc[initial ? 0 : 1][(cq << 8) | (rho << 4) | eps];
					uint32_t* r = rec + lane * REC_WORDS;
					r[0] = tw;
					r[1] = (cq == 0 ? 1u : 0u) | (rho ? 2u : 0u) | ((uint32_t)U << 8) | ((uint32_t)u << 16) |
						   (rho << 24);
					r[2] = v[0];
					r[3] = v[1];
					r[4] = v[2];
					r[5] = v[3];
				}
				__syncwarp();
				if(lane == 0)
				{
					const uint32_t ng = min(32u, nquads - g0);
					for(uint32_t j = 0; j < ng; j += 2)
					{
						int uq[2] = {0, 0};
						const uint32_t nq = (j + 1 < ng) ? 2 : 1;
						for(uint32_t q = 0; q < nq; ++q)
						{
							const uint32_t* r = rec + (j + q) * REC_WORDS;
							const uint32_t tw = r[0], fl = r[1];
							const uint32_t qrho = fl >> 24;
							const int qU = (int)((fl >> 8) & 0xFF);
							uq[q] = (int)((fl >> 16) & 0xFF);
							enc_vlc(e, TW_CWD(tw), (int)TW_LEN(tw));
							if(fl & 1u)
								enc_mel(e, (fl >> 1) & 1);
#pragma unroll
							for(uint32_t n = 0; n < 4; ++n)
							{
								const int m = ((qrho >> n) & 1) ? qU - (int)((TW_EK(tw) >> n) & 1) : 0;
								enc_ms(e, r[2 + n] & ((1u << m) - 1), m);
							}
						}
						// U-VLC of the pair (:693-716 initial row, :926-929 others)
						uint32_t c0w, c1w;
						if(initial && uq[0] > 0 && uq[1] > 0)
							enc_mel(e, min(uq[0], uq[1]) > 2);
						if(initial && uq[0] > 2 && uq[1] > 2)
						{
							c0w = uvlc_code(uq[0] - 2);
							c1w = uvlc_code(uq[1] - 2);
						}
						else if(initial && uq[0] > 2 && uq[1] > 0)
						{
							c0w = uvlc_code(uq[0]);
							c1w = (uint32_t)(uq[1] - 1) | (1u << 8);
						}
						else
						{
							c0w = uvlc_code(uq[0]);
							c1w = uvlc_code(uq[1]);
						}
						enc_vlc(e, c0w & 0xFF, (int)((c0w >> 8) & 0xFF));
						enc_vlc(e, c1w & 0xFF, (int)((c1w >> 8) & 0xFF));
						enc_vlc(e, (c0w >> 16) & 0xFF, (int)(c0w >> 24));
						enc_vlc(e, (c1w >> 16) & 0xFF, (int)(c1w >> 24));
					}
				}
				__syncwarp();
			}
			uint8_t* t1 = expN;
			expN = expC;
			expC = t1;
			t1 = sigN;
			sigN = sigC;
			sigC = t1;
		}
		// ---- termination (lane 0), then the warp appends MEL and VLC behind MagSgn -------------
		uint32_t total = 0, ms_pos = 0, mel_pos = 0, vlc_pos = 0;
		uint32_t failed = bad ? 1u : 0u;
		if(lane == 0 && !bad)
		{
			// terminate_mel_vlc, :363-392
			if(e.mel_run > 0)
				enc_mel_bit(e, 1);
			e.mel_tmp <<= e.mel_rem;
			const uint32_t mel_mask = (0xFFu << e.mel_rem) & 0xFF;
			const uint32_t vlc_mask = 0xFFu >> (8 - e.vlc_used);
			if((mel_mask | vlc_mask) != 0)
			{
				const uint32_t fuse = e.mel_tmp | e.vlc_tmp;
				if((((fuse ^ e.mel_tmp) & mel_mask) | ((fuse ^ e.vlc_tmp) & vlc_mask)) == 0 &&
				   fuse != 0xFF && e.vlc_pos > 1)
					e.mel[e.mel_pos++] = (uint8_t)fuse;
				else
				{
					e.mel[e.mel_pos++] = (uint8_t)e.mel_tmp;
					e.vlc[VLC_CAP - 1 - e.vlc_pos] = (uint8_t)e.vlc_tmp;
					e.vlc_pos++;
				}
			}
			// ms_terminate, :427-461
			if(e.ms_used)
			{
				const int t = e.ms_max - e.ms_used;
				e.ms_tmp |= (0xFFu & ((1u << t) - 1)) << e.ms_used;
				if(e.ms_tmp != 0xFF)
				{
					if(e.ms_pos < e.ms_cap)
						e.ms[e.ms_pos++] = (uint8_t)e.ms_tmp;
					else
						e.overflow = true;
				}
			}
			else if(e.ms_max == 7)
				e.ms_pos--;
			ms_pos = e.ms_pos;
			mel_pos = e.mel_pos;
			vlc_pos = e.vlc_pos;
			failed = e.overflow ? 1u : 0u;
		}
		ms_pos = __shfl_sync(0xffffffffu, ms_pos, 0);
		mel_pos = __shfl_sync(0xffffffffu, mel_pos, 0);
		vlc_pos = __shfl_sync(0xffffffffu, vlc_pos, 0);
		failed = __shfl_sync(0xffffffffu, failed, 0);
		__syncwarp();
		if(!failed)
		{
			total = ms_pos + mel_pos + vlc_pos;
			uint8_t* out = e.ms;
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
	const size_t per_warp = 32 * REC_WORDS * 4 + MEL_CAP + VLC_CAP + 4 * (size_t)((ncol + 3) & ~3u);
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
