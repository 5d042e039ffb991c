// pcrd.cu -- rate allocation of one tile's single quality layer on the device.
//
// Replaces, for the codestream-level encoder (compress.cu), the reference's
//   TileProcessor::pcrdBisectSimple        src/lib/jp2/tile/TileProcessor.cpp:1195-1357
//   TileProcessor::makeLayerSimple         tile/TileProcessor.cpp:1366-1440
//   T2Compress::compressPacketsSimulate    t2/T2Compress.cpp:58-110, :331-420 (packet accounting)
//   T2Compress::compressHeader (simulated) t2/T2Compress.cpp:130-260, TagTree::compress, BitIO
// with the same arithmetic (IEEE doubles, the same comparisons, the same packet-by-packet byte
// accounting including the 0xFF bit stuffing of the packet headers), so the pass counts it chooses --
// and with them the codestream -- are the reference's.
//
// Shape: the bisection is a chain of up to 128 dependent steps, each of which needs the exact size of
// every packet of the layer for one slope threshold.  A step is one kernel, a CTA per packet, and the
// state of the bisection stays in device memory, so the host enqueues the whole chain at once and
// never looks at it (a step that finds the chain finished returns immediately).  Inside a step, the
// threads of a packet's CTA take its code blocks:
//   A  per code block   : passes that clear the threshold (makeLayerSimple), bytes they bring
//   B1 per (band, precinct): the two tag trees as reductions, level by level (any block included /
//                         fewest missing bit planes / first included block in raster order, per node)
//   B2 per code block   : length of its part of the packet header.  In a single layer a tag-tree node
//                         speaks exactly once, when the first leaf below it is coded, so a block's
//                         bits follow from the node values alone
//   B3 per packet       : offsets of the blocks' bits (warp scan), body bytes
//   Z, C per block      : the header bits themselves, written into the packet's bit buffer
//   D  per packet       : bytes of the header once the bit after every 0xFF byte is stuffed (only
//                         walked when eight ones in a row occur at all)
//   E  last CTA         : the reference's running budget over the packets in progression order (a
//                         prefix sum: in its unsigned arithmetic the checks of a packet depend on the
//                         bytes before it only), then the bisection's next threshold.
// On the host the same allocation costs 68 ms per 4K frame and thread.
#include "common.cuh"
#include "pcrd.h"
#include <algorithm>
#include <float.h>
#include <string.h>

namespace
{
constexpr uint32_t INF = 0xFFFFFFFFu;
constexpr uint32_t MAX_LEVELS = 18;

// a block's share of the layer at the current threshold and of its packet's header
struct Sel
{
	uint32_t llen; // body bytes
	uint32_t seglen; // bytes of its first codeword segment (the only one outside the TERMALL / LAZY modes)
	uint16_t tailbits; // header bits after the tag trees: passes, Lblock increment, segment lengths
	uint8_t np; // passes in the layer
	uint8_t inc; // Lblock increment
	uint8_t nsegs; // codeword segments in the layer
	uint8_t lognump; // floorlog2(passes of the first segment)
	uint16_t pad;
};

struct Args
{
	const grkb_coded* coded;
	const grkb_pass* passes;
	uint32_t nblocks;
	const PcrdPacket* packets;
	uint32_t npackets;
	const PcrdBand* bands;
	uint32_t nbands;
	const int32_t* leaf_block;
	const uint32_t* leaf_band;
	uint32_t nleaves;
	const uint64_t* pk_bitbase; // first bit of every packet's region of the bit buffer
	const uint32_t* pk_bitcap;
	uint32_t* npass; // per block: passes of the layer (what the host gets back)
	struct Sel* sel; // per block: the same and what its packet header needs
	uint32_t* node_min; // per node
	uint32_t* node_first;
	uint32_t* leaf_bits; // per leaf: bits, then offset in the packet
	uint32_t* pk_bits; // per packet: raw header bits
	uint32_t* pk_hdr; // per packet: header bytes
	uint64_t* pk_body; // per packet: body bytes
	uint32_t* bitbuf;
	uint64_t max_bytes;
	uint32_t csty;
	uint32_t dyn_words; // 32-bit words of dynamic shared memory a step CTA has
	PcrdResult* result;
};

static __device__ __forceinline__ uint32_t floorlog2u(uint32_t v)
{
	return v <= 1 ? 0u : 31u - (uint32_t)__clz((int)v);
}

static __device__ __forceinline__ uint32_t numpasses_bits(uint32_t np)
{
	return np == 1 ? 1u : np == 2 ? 2u : np <= 5 ? 4u : np <= 36 ? 9u : 16u;
}

// makeLayerSimple for one block -- the passes of the layer at this slope threshold -- and, in the same
// walk over its pass records, what T2Compress.cpp:205-256 will write for them: the codeword segments
// end at terminated passes and at the last pass of the layer, the Lblock increment is the largest
// excess of a segment's length over the bits it has by default.
static __device__ Sel select_block(const grkb_pass* ps, uint32_t total, uint32_t data_len, double thresh)
{
	uint32_t rate_inc = 0;
	double dist_inc = 0;
	uint32_t open_len = 0, open_nump = 0, closed = 0, sumlog = 0, first_len = 0, first_log = 0;
	int need_max = 0;
	Sel r{};
	uint32_t sel_sumlog = 0;
	const uint4* raw = reinterpret_cast<const uint4*>(ps);
#pragma unroll 4
	for(uint32_t q = 0; q < total; ++q)
	{
		const uint4 v = __ldg(raw + q); // distortion (lo, hi), rate, len | term << 16
		const double dist = __hiloint2double((int)v.y, (int)v.x);
		const uint32_t rate = v.z, len = v.w & 0xFFFFu, term = (v.w >> 16) & 0xFFu;
		++open_nump;
		open_len += len;
		bool take = true;
		if(thresh != 0)
		{
			const uint32_t dr = r.np ? rate - rate_inc : rate;
			const double dd = r.np ? __dsub_rn(dist, dist_inc) : dist;
			if(!dr)
				take = dd != 0;
			else
				take = __dsub_rn(thresh, __ddiv_rn(dd, (double)dr)) < DBL_EPSILON;
		}
		const int need = (int)floorlog2u(open_len) + 1 - (int)(3 + floorlog2u(open_nump));
		if(take)
		{
			r.np = (uint8_t)(q + 1);
			rate_inc = rate;
			dist_inc = dist;
			r.inc = (uint8_t)max(need_max, need);
			r.nsegs = (uint8_t)(closed + 1);
			sel_sumlog = sumlog + floorlog2u(open_nump);
			r.seglen = closed ? first_len : open_len;
			r.lognump = (uint8_t)(closed ? first_log : floorlog2u(open_nump));
		}
		if(term)
		{
			need_max = max(need_max, need);
			if(!closed)
			{
				first_len = open_len;
				first_log = floorlog2u(open_nump);
			}
			++closed;
			sumlog += floorlog2u(open_nump);
			open_len = 0;
			open_nump = 0;
		}
	}
	if(r.np)
	{
		r.llen = min(rate_inc, data_len);
		r.tailbits = (uint16_t)(numpasses_bits(r.np) + r.inc + 1u + (uint32_t)r.nsegs * (3u + r.inc) + sel_sumlog);
	}
	return r;
}

// Sinks for the bits of a packet header.  tag(true) .. tag(false) brackets the fields whose writes the
// reference does not check against the remaining budget (number of passes, Lblock comma code).
struct CountSink
{
	static constexpr bool WRITES = false;
	uint32_t n = 0;
	__device__ void tag(bool) {}
	__device__ void put(uint32_t, uint32_t k)
	{
		n += k;
	}
	__device__ void zeros(uint32_t k)
	{
		n += k;
	}
	__device__ void ones(uint32_t k)
	{
		n += k;
	}
};
// bit i of a packet's header is bit 31 - (i % 32) of word i / 32 (the buffer is zeroed beforehand)
struct WriteSink
{
	static constexpr bool WRITES = true;
	uint32_t* buf;
	uint64_t pos;
	__device__ void tag(bool) {}
	__device__ void put(uint32_t v, uint32_t k) // k in 1..32, v < 2^k
	{
		const uint32_t sh = (uint32_t)(pos & 31);
		const uint64_t w = (((uint64_t)v) << (64 - k)) >> sh;
		const uint32_t hi = (uint32_t)(w >> 32), lo = (uint32_t)w;
		uint32_t* p = buf + (pos >> 5);
		if(hi)
			atomicOr(p, hi);
		if(lo)
			atomicOr(p + 1, lo);
		pos += k;
	}
	__device__ void zeros(uint32_t k)
	{
		pos += k;
	}
	__device__ void ones(uint32_t k)
	{
		while(k)
		{
			const uint32_t t = k > 32 ? 32 : k;
			put(t == 32 ? 0xFFFFFFFFu : (1u << t) - 1u, t);
			k -= t;
		}
	}
};

// which kind of field holds bit `target` of a leaf's header bits: 1 unchecked, 0 checked
struct ClassifySink
{
	static constexpr bool WRITES = true;
	uint32_t target, pos = 0;
	int result = 0;
	bool un = false;
	__device__ void tag(bool u)
	{
		un = u;
	}
	__device__ void put(uint32_t, uint32_t k)
	{
		if(target >= pos && target < pos + k)
			result = un ? 1 : 0;
		pos += k;
	}
	__device__ void zeros(uint32_t k)
	{
		put(0, k);
	}
	__device__ void ones(uint32_t k)
	{
		put(0, k);
	}
};

// what follows the tag-tree bits of an included block: number of passes, Lblock increment, one length
// per codeword segment (T2Compress.cpp:205-256)
template<class Sink>
static __device__ void put_numpasses(Sink& s, uint32_t np)
{
	if(np == 1)
		s.put(0, 1);
	else if(np == 2)
		s.put(2, 2);
	else if(np <= 5)
		s.put(0xC | (np - 3), 4);
	else if(np <= 36)
		s.put(0x1E0 | (np - 6), 9);
	else
		s.put(0xFF80 | (np - 37), 16);
}
template<class Sink>
static __device__ void block_tail(Sink& s, const grkb_pass* ps, uint32_t np)
{
	s.tag(true);
	put_numpasses(s, np);
	s.tag(false);
	uint32_t numlenbits = 3, inc = 0, nump = 0, len = 0;
	for(uint32_t q = 0; q < np; ++q)
	{
		++nump;
		len += ps[q].len;
		if(ps[q].term || q == np - 1)
		{
			const int need = (int)floorlog2u(len) + 1 - (int)(numlenbits + floorlog2u(nump));
			if(need > (int)inc)
				inc = (uint32_t)need;
			len = 0;
			nump = 0;
		}
	}
	s.tag(true);
	s.ones(inc);
	s.put(0, 1);
	s.tag(false);
	numlenbits += inc;
	for(uint32_t q = 0; q < np; ++q)
	{
		++nump;
		len += ps[q].len;
		if(ps[q].term || q == np - 1)
		{
			uint32_t n = numlenbits + floorlog2u(nump);
			if(n > 32)
			{
				s.zeros(n - 32);
				n = 32;
			}
			s.put(n == 32 ? len : len & ((1u << n) - 1u), n);
			len = 0;
			nump = 0;
		}
	}
}

// the part of a packet header that belongs to leaf li of band b (block bi, -1: degenerate)
template<class Sink>
static __device__ void leaf_header(Sink& s, const Args& a, const uint32_t* node_min, const uint32_t* node_first,
								   const PcrdBand& b, uint32_t li, int32_t bi)
{
	const uint32_t x = li % b.cgw, y = li / b.cgw;
	// the nodes above the leaf, leaf first (all loaded before any is looked at: they come from L2)
	uint32_t nfirst[MAX_LEVELS], nmin[MAX_LEVELS];
	int depth = 0;
	{
		uint32_t off = 0, lw = b.cgw, lh = b.cgh;
		while(true)
		{
			const uint32_t node = b.node_first + off + (y >> depth) * lw + (x >> depth);
			nfirst[depth] = node_first[node];
			nmin[depth] = node_min[node];
			++depth;
			if((lw <= 1 && lh <= 1) || depth == (int)MAX_LEVELS)
				break;
			off += lw * lh;
			lw = (lw + 1) >> 1;
			lh = (lh + 1) >> 1;
		}
	}
	Sel sel{};
	if(bi >= 0)
		sel = a.sel[bi];
	const uint32_t np = sel.np;
	// inclusion tree, threshold 1: a node is coded by the first leaf below it (its top-left one), and
	// only while every node above it holds an included block
	bool alive = true;
	for(int l = depth - 1; l >= 0 && alive; --l)
	{
		const bool any = nfirst[l] != INF;
		const uint32_t m = (1u << l) - 1u;
		if((x & m) == 0 && (y & m) == 0)
			s.put(any ? 1u : 0u, 1);
		alive = any;
	}
	if(!np)
		return;
	// zero-bit-plane tree, coded outright: a node is coded by the first INCLUDED leaf below it
	uint32_t above = 0;
	for(int l = depth - 1; l >= 0; --l)
	{
		if(nfirst[l] == li)
		{
			s.zeros(nmin[l] - above);
			s.put(1, 1);
		}
		above = nmin[l];
	}
	if(!Sink::WRITES)
		s.zeros(sel.tailbits);
	else if(sel.nsegs == 1 && 3u + sel.inc + sel.lognump <= 32)
	{
		s.tag(true);
		put_numpasses(s, np);
		s.ones(sel.inc);
		s.put(0, 1);
		s.tag(false);
		const uint32_t n = 3u + sel.inc + sel.lognump;
		s.put(n == 32 ? sel.seglen : sel.seglen & ((1u << n) - 1u), n);
	}
	else
		block_tail(s, a.passes + (size_t)bi * GRKB_ENC_MAX_PASSES, np);
}

// the bisection's state between the steps (device memory)
struct State
{
	double lower, upper, prev, thresh;
	uint64_t layer_bytes;
	uint32_t iter;
	uint32_t go; // 1: evaluate the layer at thresh; 0: make the final layer at thresh; 2: finished;
				 // 3: the step waits for the classification of one header bit (pend_*)
	uint32_t pend_packet;
	uint64_t pend_limit;
	uint32_t err;
	uint32_t counter; // CTAs that finished the current step
	uint32_t counter0; // same for the slope-range kernel
};

constexpr int INIT_CTAS = 64, INIT_T = 256;

// ---- slope range of the tile (pcrdBisectSimple :1215-1260) and the first threshold
__global__ void __launch_bounds__(INIT_T) pcrd_init_kernel(const Args a, State* st, double* part)
{
	__shared__ double s_lo[INIT_T / 32], s_hi[INIT_T / 32];
	__shared__ bool s_last;
	const uint32_t tid = threadIdx.x;
	double lo = DBL_MAX, hi = -1;
	for(uint32_t bi = blockIdx.x * INIT_T + tid; bi < a.nblocks; bi += INIT_CTAS * INIT_T)
	{
		const grkb_pass* ps = a.passes + (size_t)bi * GRKB_ENC_MAX_PASSES;
		const uint32_t total = a.coded[bi].numpasses;
		uint32_t prate = 0;
		double pdist = 0;
		for(uint32_t q = 0; q < total; ++q)
		{
			const uint32_t rate = ps[q].rate;
			const double dist = ps[q].distortion;
			const int32_t dr = q ? (int32_t)(rate - prate) : (int32_t)rate;
			const double dd = q ? __dsub_rn(dist, pdist) : dist;
			prate = rate;
			pdist = dist;
			if(dr == 0)
				continue;
			const double sl = __ddiv_rn(dd, (double)dr);
			if(sl < lo)
				lo = sl;
			if(sl > hi)
				hi = sl;
		}
	}
	for(int o = 16; o; o >>= 1)
	{
		const double l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
		lo = l2 < lo ? l2 : lo;
		hi = h2 > hi ? h2 : hi;
	}
	if((tid & 31) == 0)
	{
		s_lo[tid >> 5] = lo;
		s_hi[tid >> 5] = hi;
	}
	__syncthreads();
	if(tid == 0)
	{
		for(int i = 1; i < INIT_T / 32; ++i)
		{
			lo = s_lo[i] < lo ? s_lo[i] : lo;
			hi = s_hi[i] > hi ? s_hi[i] : hi;
		}
		part[2 * blockIdx.x] = lo;
		part[2 * blockIdx.x + 1] = hi;
		__threadfence();
		s_last = atomicAdd(&st->counter0, 1u) == INIT_CTAS - 1;
	}
	__syncthreads();
	if(s_last && tid == 0)
	{
		__threadfence();
		double lower = DBL_MAX, upper = -1;
		for(int i = 0; i < INIT_CTAS; ++i)
		{
			const double l2 = __ldcg(part + 2 * i), h2 = __ldcg(part + 2 * i + 1);
			lower = l2 < lower ? l2 : lower;
			upper = h2 > upper ? h2 : upper;
		}
		st->lower = lower;
		st->upper = upper;
		st->prev = -1;
		st->thresh = (upper == -1) ? lower : __ddiv_rn(__dadd_rn(lower, upper), 2.0);
		st->layer_bytes = 0;
		st->iter = 0;
		st->go = 1;
		st->err = 0;
		st->counter = 0;
		st->counter0 = 0;
	}
}

// ---- A: every code block's share of the layer at the current threshold (all SMs)
constexpr int SEL_T = 128;
__global__ void __launch_bounds__(SEL_T) pcrd_select_kernel(const Args a, const State* st)
{
	if(st->go >= 2)
		return;
	const uint32_t bi = blockIdx.x * SEL_T + threadIdx.x;
	if(bi >= a.nblocks)
		return;
	const grkb_coded cb = a.coded[bi];
	const Sel r = select_block(a.passes + (size_t)bi * GRKB_ENC_MAX_PASSES, cb.numpasses, cb.data_len, st->thresh);
	*reinterpret_cast<uint4*>(a.sel + bi) = *reinterpret_cast<const uint4*>(&r);
	a.npass[bi] = r.np;
}

// Bytes of a packet header once the bit after every 0xFF byte is stuffed (BitIO: seven bits in the byte
// after 0xFF; flush writes the open byte and one more if that was 0xFF -- the same count as stuffing a
// zero after EVERY 0xFF byte, the last included): (total + stuffed bits + 7) / 8.  A byte can only be
// 0xFF inside a run of eight or more ones, and whether such a run yields one depends on where it
// falls among the byte boundaries, i.e. on the number of bits stuffed before it modulo 8.  So the
// words of the header are cut into one piece per warp; a warp works out, for each of the eight
// possible counts modulo 8 on entry, how many bits get stuffed in the runs that START in its piece
// (32 words at a time: every lane lists the runs starting in its word; per entry value, the first run
// that yields a 0xFF byte is walked uniformly and the rest re-checked) and the pieces are chained.
static __device__ void stuffed_in_piece(const uint32_t* w, uint32_t total, uint32_t wbeg, uint32_t wend,
										 uint32_t lane, uint32_t out[8])
{
	uint32_t stuffed[8], cur[8];
#pragma unroll
	for(int e = 0; e < 8; ++e)
	{
		stuffed[e] = (uint32_t)e;
		cur[e] = wbeg * 32;
	}
	for(uint32_t i0 = wbeg; i0 < wend; i0 += 32)
	{
		const uint32_t i = i0 + lane;
		uint32_t starts = 0; // bit 31 - k: a run of eight or more ones starts at bit k of word i
		if(i < wend)
		{
			const uint32_t w0 = w[i];
			const uint64_t v = ((uint64_t)w0 << 32) | w[i + 1];
			uint64_t r = v & (v << 1);
			r &= r << 2;
			r &= r << 4;
			const uint32_t before = (w0 >> 1) | (i ? w[i - 1] << 31 : 0u); // the bit before each bit
			starts = (uint32_t)(r >> 32) & ~before;
		}
		if(!__any_sync(0xffffffffu, starts != 0))
			continue;
		// the (at most four) runs that start in my word
		uint32_t hpos[4], hend[4];
		int n = 0;
#pragma unroll
		for(int j = 0; j < 4; ++j)
		{
			hpos[j] = 0xFFFFFFFFu;
			hend[j] = 0;
			if(starts)
			{
				const uint32_t k = __clz(starts);
				starts &= ~(0x80000000u >> k);
				const uint32_t h = i * 32 + k;
				uint32_t end = h + 8;
				while(end < total && ((w[end >> 5] >> (31 - (end & 31))) & 1u))
					++end;
				if(h + 8 <= total) // (a prefix of a longer header may end inside the eight ones)
				{
					hpos[j] = h;
					hend[j] = end;
					n = j + 1;
				}
			}
		}
#pragma unroll
		for(int e = 0; e < 8; ++e)
		{
			for(;;)
			{
				// my first run at or after cur that makes a 0xFF byte with stuffed[e] bits before it
				uint32_t hit = 0xFFFFFFFFu, hit_end = 0;
#pragma unroll
				for(int j = 3; j >= 0; --j)
					if(j < n && hpos[j] != 0xFFFFFFFFu && hpos[j] >= cur[e] &&
					   hend[j] - hpos[j] >= ((0u - (hpos[j] + stuffed[e])) & 7u) + 8u)
					{
						hit = hpos[j];
						hit_end = hend[j];
					}
				const uint32_t m = __ballot_sync(0xffffffffu, hit != 0xFFFFFFFFu);
				if(!m)
					break;
				const uint32_t src = __ffs(m) - 1;
				const uint32_t h = __shfl_sync(0xffffffffu, hit, src), end = __shfl_sync(0xffffffffu, hit_end, src);
				uint32_t len = end - h, o = h + stuffed[e]; // o: position in the stuffed stream
				for(;;)
				{
					const uint32_t b = (o + 7) & ~7u; // next byte boundary
					if(len < b - o + 8)
						break;
					len -= b - o + 8; // the byte [b, b + 8) is 0xFF: a zero follows it
					++stuffed[e];
					o = b + 9;
				}
				cur[e] = end;
			}
		}
	}
#pragma unroll
	for(int e = 0; e < 8; ++e)
		out[e] = stuffed[e] - (uint32_t)e;
}

// shared memory of a packet's CTA
template<int T>
struct PacketShared
{
	uint64_t sum[T], total_bytes;
	uint32_t scan[T / 32 + 1];
	uint32_t piece[T / 32][8];
	uint32_t off, fail, key, word;
	uint64_t limit;
	bool last;
};

// B1 .. C for packet p: tag trees, bits of every leaf and their offsets (a.leaf_bits), the header bits
// themselves at w (shared memory when it fits).  All threads of the CTA; false: no room for the bits.
template<int T>
static __device__ bool build_header(const Args& a, const PcrdPacket& pk, uint32_t p, uint32_t* dyn, PacketShared<T>& sh,
									uint32_t*& w, uint32_t& total, uint64_t& body, uint32_t& err)
{
	const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	const uint32_t gend = pk.leaf_first + pk.nleaves;
	// the packet's tag-tree nodes and header bits live in shared memory when they fit (the phases below
	// hand them from thread to thread: from L2 every hand-over costs its latency)
	uint32_t* node_min = a.node_min;
	uint32_t* node_first = a.node_first;
	uint32_t dyn_used = 0;
	if(2 * pk.nnodes <= a.dyn_words)
	{
		node_min = dyn - pk.node_first;
		node_first = dyn + pk.nnodes - pk.node_first;
		dyn_used = 2 * pk.nnodes;
	}
	// B1: the two tag trees of every band as reductions, level by level (all bands of the packet in the
	// same round)
	for(uint32_t bj = pk.band_first; bj < pk.band_first + pk.nbands; ++bj)
	{
		const PcrdBand b = a.bands[bj];
		const uint32_t n0 = b.cgw * b.cgh;
		for(uint32_t li = tid; li < n0; li += T)
		{
			const int32_t bi = a.leaf_block[b.leaf_first + li];
			const uint32_t bps = bi >= 0 ? a.coded[bi].numbps : 0u;
			if(bps > b.band_bps)
				err = 1;
			node_min[b.node_first + li] = b.band_bps - bps;
			node_first[b.node_first + li] = (bi >= 0 && a.sel[bi].np) ? li : INF;
		}
	}
	for(uint32_t lvl = 0;; ++lvl)
	{
		bool more = false;
		__syncthreads();
		for(uint32_t bj = pk.band_first; bj < pk.band_first + pk.nbands; ++bj)
		{
			const PcrdBand b = a.bands[bj];
			uint32_t off = 0, lw = b.cgw, lh = b.cgh;
			for(uint32_t l = 0; l < lvl && !(lw <= 1 && lh <= 1); ++l)
			{
				off += lw * lh;
				lw = (lw + 1) >> 1;
				lh = (lh + 1) >> 1;
			}
			if(lw <= 1 && lh <= 1)
				continue;
			more = true;
			const uint32_t pw = (lw + 1) >> 1, ph = (lh + 1) >> 1, poff = off + lw * lh;
			for(uint32_t n = tid; n < pw * ph; n += T)
			{
				const uint32_t nx = n % pw, ny = n / pw;
				uint32_t mn = INF, fi = INF;
#pragma unroll
				for(uint32_t d = 0; d < 4; ++d)
				{
					const uint32_t cx = 2 * nx + (d & 1), cy = 2 * ny + (d >> 1);
					if(cx < lw && cy < lh)
					{
						const uint32_t c = b.node_first + off + cy * lw + cx;
						mn = min(mn, node_min[c]);
						fi = min(fi, node_first[c]);
					}
				}
				node_min[b.node_first + poff + n] = mn;
				node_first[b.node_first + poff + n] = fi;
			}
		}
		if(!more)
			break;
	}
	if(tid == 0)
		sh.off = 1; // bit 0 is the "packet not empty" bit
	__syncthreads();
	// B2 + B3: bits of every leaf, their offsets in the packet, body bytes
	body = 0;
	for(uint32_t g0 = pk.leaf_first; g0 < gend; g0 += T)
	{
		const uint32_t g = g0 + tid;
		uint32_t bits = 0;
		if(g < gend)
		{
			const PcrdBand& b = a.bands[a.leaf_band[g]];
			const int32_t bi = a.leaf_block[g];
			CountSink cs;
			leaf_header(cs, a, node_min, node_first, b, g - b.leaf_first, bi);
			bits = cs.n;
			if(bi >= 0)
				body += a.sel[bi].llen;
		}
		uint32_t inc = bits;
		for(int o = 1; o < 32; o <<= 1)
		{
			const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
			if((int)lane >= o)
				inc += t;
		}
		if(lane == 31)
			sh.scan[wid] = inc;
		__syncthreads();
		if(tid == 0)
		{
			uint32_t run = sh.off;
			for(int i = 0; i < T / 32; ++i)
			{
				const uint32_t t = sh.scan[i];
				sh.scan[i] = run;
				run += t;
			}
			sh.scan[T / 32] = run;
		}
		__syncthreads();
		if(g < gend)
			a.leaf_bits[g] = sh.scan[wid] + inc - bits;
		__syncthreads();
		if(tid == 0)
			sh.off = sh.scan[T / 32];
	}
	__syncthreads();
	for(int o = 16; o; o >>= 1)
		body += __shfl_xor_sync(0xffffffffu, body, o);
	if(lane == 0)
		sh.sum[wid] = body;
	__syncthreads();
	body = 0;
	for(int i = 0; i < T / 32; ++i)
		body += sh.sum[i];
	total = sh.off;
	if(total + 64 > a.pk_bitcap[p])
		return false;
	// Z: clear the bits in use; C: write them
	const uint32_t words = (total + 63) / 32 + 1;
	w = dyn_used + words <= a.dyn_words ? dyn + dyn_used : a.bitbuf + (a.pk_bitbase[p] >> 5);
	for(uint32_t i = tid; i < words; i += T)
		w[i] = 0;
	__syncthreads();
	for(uint32_t g = pk.leaf_first + tid; g < gend; g += T)
	{
		const PcrdBand& b = a.bands[a.leaf_band[g]];
		WriteSink ws{w, a.leaf_bits[g]};
		leaf_header(ws, a, node_min, node_first, b, g - b.leaf_first, a.leaf_block[g]);
	}
	if(tid == 0)
		atomicOr(w, 0x80000000u);
	__threadfence_block();
	__syncthreads();
	return true;
}

// D: bits stuffed into the first `total` bits at w, a piece per warp (all threads; same value in all)
template<int T>
static __device__ uint32_t cta_stuffed(const uint32_t* w, uint32_t total, PacketShared<T>& sh)
{
	const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	const uint32_t hw = (total + 31) / 32, per = ((hw + T / 32 - 1) / (T / 32) + 31) & ~31u;
	uint32_t cnt[8];
	stuffed_in_piece(w, total, min(wid * per, hw), min((wid + 1) * per, hw), lane, cnt);
	__syncthreads();
	if(lane == 0)
		for(int e = 0; e < 8; ++e)
			sh.piece[wid][e] = cnt[e];
	__syncthreads();
	uint32_t stuffed = 0;
	for(int i = 0; i < T / 32; ++i)
		stuffed += sh.piece[i][stuffed & 7];
	return stuffed;
}

// the bisection's next move once the layer at `th` is known to fit or not (one thread)
static __device__ void finish_step(State* st, double th, bool fits, uint64_t layer_bytes)
{
	double lower = st->lower, upper = st->upper, thresh = th;
	const uint32_t iter = st->iter + 1;
	if(fits)
	{
		st->layer_bytes = layer_bytes;
		upper = thresh;
	}
	else
		lower = thresh;
	uint32_t next_go = 1;
	if(iter == 128 || __ldcg(&st->err))
		next_go = 0;
	else
	{
		// the reference makes the layer at the next threshold and then stops if it moved by less than
		// 0.001; the layer it leaves is the one at `upper`
		const double prev = thresh;
		thresh = (upper == -1) ? lower : __ddiv_rn(__dadd_rn(lower, upper), 2.0);
		st->prev = prev;
		if(fabs(__dsub_rn(prev, thresh)) < 0.001)
			next_go = 0;
	}
	st->lower = lower;
	st->upper = upper;
	st->thresh = next_go ? thresh : ((upper == -1) ? thresh : upper);
	st->iter = iter;
	st->go = next_go;
	st->counter = 0;
}

// ---- one step of the bisection: a CTA per packet, then the last CTA to finish closes the step
template<int T>
__global__ void __launch_bounds__(T) pcrd_step_kernel(const Args a, State* st)
{
	__shared__ PacketShared<T> sh;
	extern __shared__ uint32_t dyn[];
	const uint32_t go = st->go;
	if(go >= 2)
		return;
	const double th = st->thresh;
	const uint32_t tid = threadIdx.x;
	const uint32_t p = blockIdx.x;
	uint32_t err = 0;
	if(go == 1)
	{
		const PcrdPacket pk = a.packets[p];
		uint32_t* w = nullptr;
		uint32_t total = 0, nbytes = 0;
		uint64_t body = 0;
		if(!build_header<T>(a, pk, p, dyn, sh, w, total, body, err))
			err = 2;
		else
			nbytes = (total + cta_stuffed<T>(w, total, sh) + 7) / 8;
		if(tid == 0)
		{
			a.pk_hdr[p] = nbytes;
			a.pk_body[p] = body;
		}
	}
	if(err)
		atomicOr(&st->err, err);
	// ---- the last CTA to get here closes the step
	__syncthreads();
	if(tid == 0)
	{
		__threadfence();
		sh.last = atomicAdd(&st->counter, 1u) == gridDim.x - 1;
		sh.fail = 0;
		sh.key = 0xFFFFFFFFu;
	}
	__syncthreads();
	if(!sh.last)
		return;
	__threadfence();
	if(go == 0)
	{
		if(tid == 0)
		{
			a.result->iterations = st->iter;
			a.result->error = __ldcg(&st->err);
			a.result->thresh = th;
			a.result->layer_bytes = st->layer_bytes;
			st->counter = 0;
			st->go = 2;
		}
		return;
	}
	// E: the running budget over the packets in progression order (compressPacketsSimulate).  In the
	// reference's unsigned arithmetic what is left before packet p is max_bytes minus the bytes of the
	// packets before it, so the packets can be checked independently once those sums are known; the
	// first packet that runs into the budget decides.  A header that reaches the limit gives the layer
	// up -- unless the byte that reaches it is completed by a bit of an unchecked field (see
	// pcrd_classify_kernel), in which case the budget wraps and the whole layer "fits".
	const uint32_t chunk = (a.npackets + T - 1) / T;
	const uint32_t p0 = min(tid * chunk, a.npackets), p1 = min(p0 + chunk, a.npackets);
	const uint64_t fixed = ((a.csty & 2) ? 6u : 0u) + ((a.csty & 4) ? 2u : 0u);
	uint64_t sum = 0;
	for(uint32_t q = p0; q < p1; ++q)
		sum += fixed + __ldcg(a.pk_hdr + q) + __ldcg(a.pk_body + q);
	sh.sum[tid] = sum;
	__syncthreads();
	if(tid == 0)
	{
		uint64_t run = 0;
		const uint32_t used = min((uint32_t)T, (a.npackets + chunk - 1) / chunk); // threads that have packets
		for(uint32_t i = 0; i < used; ++i)
		{
			const uint64_t t = sh.sum[i];
			sh.sum[i] = run;
			run += t;
		}
		sh.total_bytes = run;
	}
	__syncthreads();
	uint64_t before = sh.sum[tid], my_limit = 0;
	uint32_t my_key = 0xFFFFFFFFu; // packet * 4 + kind of its event: 1 does not fit, 2 header limit inside the header
	for(uint32_t q = p0; q < p1; ++q)
	{
		const uint64_t hdr = __ldcg(a.pk_hdr + q), body = __ldcg(a.pk_body + q);
		uint64_t avail = a.max_bytes - before;
		if(a.csty & 2)
			avail -= 6;
		uint32_t ev = 0;
		if(avail != 0 && hdr >= avail)
			ev = hdr > avail ? 2 : 1; // reached by the last byte: at the flush, which is checked
		else
		{
			uint64_t left = avail - hdr;
			if(a.csty & 4)
				left -= 2;
			if(body > left)
				ev = 1;
		}
		if(ev && my_key == 0xFFFFFFFFu)
		{
			my_key = q * 4 + ev;
			my_limit = avail;
		}
		before += fixed + hdr + body;
	}
	if(my_key != 0xFFFFFFFFu)
		atomicMin(&sh.key, my_key);
	__syncthreads();
	if(my_key != 0xFFFFFFFFu && my_key == sh.key)
		sh.limit = my_limit;
	__syncthreads();
	if(tid == 0)
	{
		const uint32_t key = sh.key;
		if(key == 0xFFFFFFFFu)
			finish_step(st, th, true, sh.total_bytes);
		else if((key & 3) == 1)
			finish_step(st, th, false, 0);
		else
		{
			st->pend_packet = key >> 2;
			st->pend_limit = sh.limit;
			st->counter = 0;
			st->go = 3;
		}
	}
}

// ---- a header reached the end of the budget before its last byte.  BitIO::writeByte fails on the
// byte that makes the count reach the limit, and it is called when the first bit of the NEXT byte
// arrives; T2Compress::compressHeader looks at the result for every field except the number of passes
// and the Lblock comma code (T2Compress.cpp:208, :228).  There the failed byte is simply written again
// with the following bit, the count runs past the limit, the budget left wraps around and every later
// check passes: the layer "fits".  So: rebuild that packet's header, find the raw bit that opens byte
// limit + 1 of the stuffed stream, and see which field it belongs to.
template<int T>
__global__ void __launch_bounds__(T) pcrd_classify_kernel(const Args a, State* st)
{
	__shared__ PacketShared<T> sh;
	__shared__ int s_kind;
	extern __shared__ uint32_t dyn[];
	if(st->go != 3)
		return;
	const double th = st->thresh;
	const uint32_t tid = threadIdx.x, p = st->pend_packet;
	const uint64_t limit = st->pend_limit;
	const PcrdPacket pk = a.packets[p];
	uint32_t* w = nullptr;
	uint32_t total = 0, err = 0;
	uint64_t body = 0;
	if(tid == 0)
		s_kind = 0;
	if(!build_header<T>(a, pk, p, dyn, sh, w, total, body, err))
	{
		if(tid == 0)
		{
			atomicOr(&st->err, 2u);
			finish_step(st, th, false, 0);
		}
		return;
	}
	// smallest raw bit r whose place in the stuffed stream, r + (bits stuffed before it), is in byte
	// limit + 1 or later
	uint32_t lo = 1, hi = total; // r in [lo, hi]; hi == total: there is none
	while(lo < hi)
	{
		const uint32_t mid = lo + (hi - lo) / 2;
		const uint32_t s = cta_stuffed<T>(w, mid, sh);
		if((uint64_t)mid + s >= 8 * limit)
			hi = mid;
		else
			lo = mid + 1;
	}
	const uint32_t r = lo;
	if(r < total)
	{
		// the leaf that holds bit r says what kind of field it is
		const uint32_t gend = pk.leaf_first + pk.nleaves;
		uint32_t* node_min = a.node_min;
		uint32_t* node_first = a.node_first;
		if(2 * pk.nnodes <= a.dyn_words)
		{
			node_min = dyn - pk.node_first;
			node_first = dyn + pk.nnodes - pk.node_first;
		}
		for(uint32_t g = pk.leaf_first + tid; g < gend; g += T)
		{
			const uint32_t off = a.leaf_bits[g], next = g + 1 < gend ? a.leaf_bits[g + 1] : total;
			if(r >= off && r < next)
			{
				const PcrdBand& b = a.bands[a.leaf_band[g]];
				ClassifySink cs;
				cs.target = r - off;
				leaf_header(cs, a, node_min, node_first, b, g - b.leaf_first, a.leaf_block[g]);
				s_kind = cs.result;
			}
		}
	}
	__syncthreads();
	if(tid == 0)
		finish_step(st, th, r < total && s_kind == 1, 0);
}
} // namespace

struct grkb_pcrd
{
	int device = 0;
	uint32_t nblocks = 0, npackets = 0, nbands = 0, nleaves = 0, nnodes = 0;
	void* mem = nullptr; // one allocation, carved below
	Args a{};
	State* state = nullptr;
	double* part = nullptr;
	uint32_t max_leaves = 0;
	size_t dyn_bytes = 0; // dynamic shared memory of a step CTA
};

int grkb_pcrd_create(const PcrdLayout& lay, uint32_t nblocks, grkb_pcrd** out)
{
	*out = nullptr;
	auto* p = new grkb_pcrd();
	if(cudaGetDevice(&p->device) != cudaSuccess)
	{
		grkb_set_error("pcrd: no current device");
		delete p;
		return -1;
	}
	p->nblocks = nblocks;
	p->npackets = (uint32_t)lay.packets.size();
	p->nbands = (uint32_t)lay.bands.size();
	p->nleaves = (uint32_t)lay.leaf_block.size();
	p->nnodes = lay.nnodes;
	// bit-buffer capacity per packet: what a block can need outside the all-segments-terminated modes
	// (1 + 17 + 64 tag-tree bits, 16 + 33 + 33 for passes, Lblock and one length) with room to spare, and
	// 40 bits for each further codeword segment a block could have
	std::vector<uint64_t> base(p->npackets);
	std::vector<uint32_t> cap(p->npackets);
	uint64_t bits = 0;
	for(uint32_t i = 0; i < p->npackets; ++i)
	{
		p->max_leaves = std::max(p->max_leaves, lay.packets[i].nleaves);
		// shared memory for the nodes and about 128 header bits per block of the largest packet
		p->dyn_bytes = std::max<size_t>(p->dyn_bytes, (size_t)lay.packets[i].nnodes * 8 +
															(size_t)lay.packets[i].nleaves * 16 + 1024);
		const uint64_t c = ((64 + (uint64_t)lay.packets[i].nleaves * (192 + 40 * GRKB_ENC_MAX_PASSES) + 128 + 31) / 32) * 32;
		base[i] = bits;
		cap[i] = (uint32_t)std::min<uint64_t>(c, 0xFFFFFF00u);
		bits += cap[i];
	}
	p->dyn_bytes = std::min<size_t>((p->dyn_bytes + 15) & ~(size_t)15, 200 * 1024);
	auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
	size_t off = 0;
	auto carve = [&](size_t bytes) {
		const size_t o = off;
		off += up(bytes ? bytes : 1);
		return o;
	};
	const size_t o_packets = carve(p->npackets * sizeof(PcrdPacket)), o_bands = carve(p->nbands * sizeof(PcrdBand)),
				 o_lblock = carve(p->nleaves * 4), o_lband = carve(p->nleaves * 4), o_base = carve(p->npackets * 8),
				 o_cap = carve(p->npackets * 4), o_npass = carve((size_t)nblocks * 4), o_llen = carve((size_t)nblocks * sizeof(Sel)),
				 o_nmin = carve((size_t)p->nnodes * 4), o_nfirst = carve((size_t)p->nnodes * 4),
				 o_lbits = carve(p->nleaves * 4), o_pbits = carve(p->npackets * 4), o_phdr = carve(p->npackets * 4),
				 o_pbody = carve(p->npackets * 8), o_res = carve(sizeof(PcrdResult)), o_state = carve(sizeof(State)),
				 o_part = carve(2 * INIT_CTAS * sizeof(double)), o_bit = carve((size_t)(bits / 8) + 64);
	if(cudaMalloc(&p->mem, off) != cudaSuccess)
	{
		grkb_set_error("pcrd: cannot allocate %zu bytes of device memory", off);
		delete p;
		return -1;
	}
	uint8_t* m = (uint8_t*)p->mem;
	auto put = [&](size_t o, const void* src, size_t bytes) {
		return bytes == 0 || cudaMemcpy(m + o, src, bytes, cudaMemcpyHostToDevice) == cudaSuccess;
	};
	if(!put(o_packets, lay.packets.data(), p->npackets * sizeof(PcrdPacket)) ||
	   !put(o_bands, lay.bands.data(), p->nbands * sizeof(PcrdBand)) || !put(o_lblock, lay.leaf_block.data(), p->nleaves * 4) ||
	   !put(o_lband, lay.leaf_band.data(), p->nleaves * 4) || !put(o_base, base.data(), p->npackets * 8) ||
	   !put(o_cap, cap.data(), p->npackets * 4))
	{
		grkb_set_error("pcrd: cannot upload the packet tables");
		cudaFree(p->mem);
		delete p;
		return -1;
	}
	Args& a = p->a;
	a.nblocks = nblocks;
	a.packets = (const PcrdPacket*)(m + o_packets);
	a.npackets = p->npackets;
	a.bands = (const PcrdBand*)(m + o_bands);
	a.nbands = p->nbands;
	a.leaf_block = (const int32_t*)(m + o_lblock);
	a.leaf_band = (const uint32_t*)(m + o_lband);
	a.nleaves = p->nleaves;
	a.pk_bitbase = (const uint64_t*)(m + o_base);
	a.pk_bitcap = (const uint32_t*)(m + o_cap);
	a.npass = (uint32_t*)(m + o_npass);
	a.sel = (Sel*)(m + o_llen);
	a.node_min = (uint32_t*)(m + o_nmin);
	a.node_first = (uint32_t*)(m + o_nfirst);
	a.leaf_bits = (uint32_t*)(m + o_lbits);
	a.pk_bits = (uint32_t*)(m + o_pbits);
	a.pk_hdr = (uint32_t*)(m + o_phdr);
	a.pk_body = (uint64_t*)(m + o_pbody);
	a.result = (PcrdResult*)(m + o_res);
	a.bitbuf = (uint32_t*)(m + o_bit);
	p->state = (State*)(m + o_state);
	p->part = (double*)(m + o_part);
	if(cudaMemset(p->state, 0, sizeof(State)) != cudaSuccess)
	{
		grkb_set_error("pcrd: cannot clear the state");
		cudaFree(p->mem);
		delete p;
		return -1;
	}
	*out = p;
	return 0;
}

void grkb_pcrd_destroy(grkb_pcrd* p)
{
	if(!p)
		return;
	cudaFree(p->mem);
	delete p;
}

int grkb_pcrd_enqueue(grkb_pcrd* p, const grkb_coded* coded, const grkb_pass* passes, uint64_t max_bytes,
					  uint32_t csty, uint32_t* npass_host, PcrdResult* result_host, cudaStream_t s)
{
	Args a = p->a;
	a.coded = coded;
	a.passes = passes;
	a.max_bytes = max_bytes;
	a.csty = csty;
	pcrd_init_kernel<<<INIT_CTAS, INIT_T, 0, s>>>(a, p->state, p->part);
	// at most 128 evaluations and the final layer; a step that finds the bisection finished returns at
	// once (the host does not wait to find out how many it takes: about 40 for a 4K frame)
	// a CTA per packet, as wide as the largest packet has code blocks
	const uint32_t ctas = p->npackets;
	const size_t dyn = p->dyn_bytes;
	a.dyn_words = (uint32_t)(dyn / 4);
	if(p->max_leaves <= 64)
	{
		GRKB_ENSURE_SMEM(pcrd_step_kernel<64>, dyn);
		GRKB_ENSURE_SMEM(pcrd_classify_kernel<64>, dyn);
	}
	else if(p->max_leaves <= 256)
	{
		GRKB_ENSURE_SMEM(pcrd_step_kernel<256>, dyn);
		GRKB_ENSURE_SMEM(pcrd_classify_kernel<256>, dyn);
	}
	else
	{
		GRKB_ENSURE_SMEM(pcrd_step_kernel<1024>, dyn);
		GRKB_ENSURE_SMEM(pcrd_classify_kernel<1024>, dyn);
	}
	if(ctas)
		for(int i = 0; i < 129; ++i)
		{
			pcrd_select_kernel<<<(p->nblocks + SEL_T - 1) / SEL_T, SEL_T, 0, s>>>(a, p->state);
			if(p->max_leaves <= 64)
			{
				pcrd_step_kernel<64><<<ctas, 64, dyn, s>>>(a, p->state);
				pcrd_classify_kernel<64><<<1, 64, dyn, s>>>(a, p->state);
			}
			else if(p->max_leaves <= 256)
			{
				pcrd_step_kernel<256><<<ctas, 256, dyn, s>>>(a, p->state);
				pcrd_classify_kernel<256><<<1, 256, dyn, s>>>(a, p->state);
			}
			else
			{
				pcrd_step_kernel<1024><<<ctas, 1024, dyn, s>>>(a, p->state);
				pcrd_classify_kernel<1024><<<1, 1024, dyn, s>>>(a, p->state);
			}
		}
	else
	{
		GRKB_CUDA_CHECK(cudaMemsetAsync(a.result, 0, sizeof(PcrdResult), s));
		GRKB_CUDA_CHECK(cudaMemsetAsync(a.npass, 0, (size_t)(p->nblocks ? p->nblocks : 1) * 4, s));
	}
	GRKB_CUDA_CHECK(cudaGetLastError());
	GRKB_CUDA_CHECK(cudaMemcpyAsync(npass_host, a.npass, (size_t)p->nblocks * 4, cudaMemcpyDeviceToHost, s));
	GRKB_CUDA_CHECK(cudaMemcpyAsync(result_host, a.result, sizeof(PcrdResult), cudaMemcpyDeviceToHost, s));
	return 0;
}
