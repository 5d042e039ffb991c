// t2_parse.cpp -- T2 of one tile on our side of the plugin (host only, no CUDA):
// walks the tile's packets in progression order, decodes every packet header (inclusion and
// zero-bit-plane tag trees, number of passes, Lblock, segment lengths) and leaves the block
// decoders' descriptor table, segment table and byte pool (include/grk_b200.h: grkb_cblk, grkb_seg).
//
// Follows ISO/IEC 15444-1 Annex B as the reference implements it (paths under
// /root/reference/src/lib/jp2):
//   packet order            t2/PacketIter.cpp:55-292 (the five progressions), :306-355 (which
//                           positions start a precinct), t2/PacketManager.cpp:150-176 (POC bounds)
//   packet header           t2/T2Decompress.cpp:225-480 (readPacketHeader), :29-56 (initSegment:
//                           passes per codeword segment under TERMALL / LAZY)
//   packet body             t2/T2Decompress.cpp:482-566 (readPacketData)
//   tag trees               t2/TagTree.h:188-245
//   bit reader              t2/BitIO.cpp:52-77 (bytein: 7 bits after 0xFF), :150-155 (inalign),
//                           :170-209 (comma code, number of passes)
//   band bit planes / step  codestream/Quantizer.cpp:26-66 (grkb_band_quant)
// HT blocks take the same segment rule as the reference gives them (no Part-15 cleanup | refinement
// split: t2/T2Decompress.cpp:52-55), so an HT block with refinement passes arrives as one segment and
// is rejected by the block decoder exactly as T1HT::decompress rejects it.
#include "codestream.h"
#include "geometry.h"
#include "packet_order.h"
#include <algorithm>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int grkb_band_quant(uint32_t expn, uint32_t mant, uint32_t orient, uint32_t prec, int reversible,
					int compress, int ht, uint32_t numgbits, uint32_t roishift, float* stepsize,
					uint32_t* numbps);

namespace grkb_cs
{
namespace
{
constexpr uint32_t MAX_PASSES_PER_SEG = 112; // maxPassesPerSegmentJ2K (codestream/CodeStreamLimits.h:26)
constexpr uint32_t MAX_BITPLANES = 25; // maxBitPlanesGRK (:42)

struct Truncated
{
};
struct Corrupt
{
	const char* what;
};

// packet-header bit reader: MSB first, the byte after 0xFF carries 7 bits
struct BitReader
{
	const uint8_t* p;
	const uint8_t* end;
	uint32_t buf = 0, ct = 0;
	BitReader(const uint8_t* b, const uint8_t* e) : p(b), end(e) {}
	inline void bytein()
	{
		if(p >= end)
			throw Truncated();
		const bool ff = buf == 0xFF;
		buf = *p++;
		ct = ff ? 7 : 8;
		if(ff && buf >= 0x90)
			throw Corrupt{"marker inside a packet header"};
	}
	inline uint32_t bit()
	{
		if(ct == 0)
			bytein();
		--ct;
		return (buf >> ct) & 1;
	}
	inline uint32_t bits(uint32_t n)
	{
		uint32_t v = 0;
		while(n)
		{
			if(ct == 0)
				bytein();
			const uint32_t take = n < ct ? n : ct;
			ct -= take;
			v = (v << take) | ((buf >> ct) & ((1u << take) - 1u));
			n -= take;
		}
		return v;
	}
	// end of the header: the stuffing byte after a final 0xFF belongs to it
	inline const uint8_t* align()
	{
		if(buf == 0xFF)
			bytein();
		ct = 0;
		return p;
	}
	inline uint32_t comma()
	{
		uint32_t n = 0;
		while(bit())
			++n;
		return n;
	}
	inline uint32_t numpasses()
	{
		if(!bit())
			return 1;
		if(!bit())
			return 2;
		uint32_t n = bits(2);
		if(n != 3)
			return n + 3;
		n = bits(5);
		if(n != 31)
			return n + 6;
		return bits(7) + 37;
	}
};

// All tag trees of a tile live in one arena.  A tree over a w x h grid of leaves: level 0 the leaves,
// each next level halves (rounding up) until 1 x 1.
struct TagNode
{
	uint16_t value, low;
};
struct TagTree
{
	uint32_t first = 0; // node index of leaf (0, 0) in the arena
	uint32_t w = 0, h = 0;
};
struct TagArena
{
	std::vector<TagNode> nodes;
	TagTree make(uint32_t w, uint32_t h)
	{
		TagTree t;
		t.first = (uint32_t)nodes.size();
		t.w = w;
		t.h = h;
		size_t n = 0;
		uint32_t lw = w, lh = h;
		while(true)
		{
			n += (size_t)lw * lh;
			if(lw <= 1 && lh <= 1)
				break;
			lw = (lw + 1) >> 1;
			lh = (lh + 1) >> 1;
		}
		nodes.resize(nodes.size() + n, TagNode{0xFFFF, 0});
		return t;
	}
	// true when the leaf's value is known to be < threshold after reading what that takes
	bool decode(BitReader& br, const TagTree& t, uint32_t leaf, uint32_t threshold)
	{
		uint32_t path[34];
		int depth = 0;
		uint32_t x = leaf % t.w, y = leaf / t.w, lw = t.w, lh = t.h, base = t.first;
		while(true)
		{
			path[depth++] = base + y * lw + x;
			if(lw <= 1 && lh <= 1)
				break;
			base += lw * lh;
			x >>= 1;
			y >>= 1;
			lw = (lw + 1) >> 1;
			lh = (lh + 1) >> 1;
		}
		uint32_t low = 0;
		for(int i = depth - 1; i >= 0; --i)
		{
			TagNode& nd = nodes[path[i]];
			if(low > nd.low)
				nd.low = (uint16_t)low;
			else
				low = nd.low;
			while(low < threshold && low < nd.value)
			{
				if(br.bit())
					nd.value = (uint16_t)low;
				else
					++low;
			}
			nd.low = (uint16_t)low;
		}
		return nodes[path[0]].value < threshold;
	}
	// The leaf's value outright: what `i = 1; while(!decode(leaf, i)) ++i;` reads (the reference's
	// zero-bit-plane loop, T2Decompress.cpp:333-349), in one walk from the root.  limit: give up
	// (corrupt stream) beyond this value.
	uint32_t decode_value(BitReader& br, const TagTree& t, uint32_t leaf, uint32_t limit)
	{
		uint32_t path[34];
		int depth = 0;
		uint32_t x = leaf % t.w, y = leaf / t.w, lw = t.w, lh = t.h, base = t.first;
		while(true)
		{
			path[depth++] = base + y * lw + x;
			if(lw <= 1 && lh <= 1)
				break;
			base += lw * lh;
			x >>= 1;
			y >>= 1;
			lw = (lw + 1) >> 1;
			lh = (lh + 1) >> 1;
		}
		uint32_t low = 0;
		for(int i = depth - 1; i >= 0; --i)
		{
			TagNode& nd = nodes[path[i]];
			if(low < nd.low)
				low = nd.low;
			while(nd.value == 0xFFFF)
			{
				if(br.bit())
					nd.value = (uint16_t)low;
				else if(++low > limit)
					throw Corrupt{"zero bit-plane count out of range"};
			}
			low = nd.value;
			nd.low = (uint16_t)low;
		}
		return low;
	}
};

// a few elements inline, more on the heap: almost every block has one segment and one chunk
template<typename T, unsigned N>
struct SmallVec
{
	T inl[N];
	uint32_t n = 0;
	std::vector<T> more;
	bool empty() const { return n == 0; }
	size_t size() const { return n; }
	T& operator[](size_t i) { return i < N ? inl[i] : more[i - N]; }
	const T& operator[](size_t i) const { return i < N ? inl[i] : more[i - N]; }
	T& back() { return (*this)[n - 1]; }
	T& emplace_back()
	{
		if(n < N)
			inl[n] = T();
		else
			more.emplace_back();
		return (*this)[n++];
	}
	void push_back(const T& v) { emplace_back() = v; }
};

// what T2 keeps per code block between packets
struct BlockState
{
	uint32_t numlenbits = 0; // 0: not yet included
	uint32_t passes_in_packet = 0;
	uint32_t first_seg_in_packet = 0; // first segment the current packet contributes to
	bool dropped = false;
};
struct SegState
{
	uint32_t numpasses = 0, maxpasses = 0; // structure: every packet seen, kept or not
	uint32_t klen = 0, kpasses = 0; // what the kept packets (resolution / layer limits) contributed
	uint32_t bytes_in_packet = 0, passes_in_packet = 0;
};
struct Chunk
{
	size_t off; // in the caller's buffer
	uint32_t len;
};

uint32_t seg_maxpasses(uint8_t sty, bool first, uint32_t last_max)
{
	if(sty & GRKB_CBLKSTY_TERMALL)
		return 1;
	if(sty & GRKB_CBLKSTY_LAZY)
	{
		if(first)
			return 10;
		return (last_max == 1 || last_max == 10) ? 2 : 1;
	}
	return MAX_PASSES_PER_SEG;
}
uint32_t floorlog2(uint32_t v)
{
	uint32_t l = 0;
	while(v > 1)
	{
		v >>= 1;
		++l;
	}
	return l;
}

// reader over the tile-part bodies of a tile: a packet never straddles two tile parts
struct BodyReader
{
	const uint8_t* data;
	const std::vector<Range>& parts;
	size_t part = 0, pos = 0; // pos: offset inside parts[part]
	BodyReader(const uint8_t* d, const std::vector<Range>& r) : data(d), parts(r) {}
	void settle()
	{
		while(part < parts.size() && pos >= parts[part].len)
		{
			++part;
			pos = 0;
		}
	}
	bool empty()
	{
		settle();
		return part >= parts.size();
	}
	const uint8_t* cur() const { return data + parts[part].off + pos; }
	const uint8_t* end() const { return data + parts[part].off + parts[part].len; }
	size_t left() const { return parts[part].len - pos; }
	size_t abs() const { return parts[part].off + pos; }
};
} // namespace

bool parse_tile(const Codestream& cs, uint32_t tileno, uint32_t reduce, uint32_t max_layers, TileTables& out,
				std::string& err, const uint32_t* window)
{
	out = TileTables();
	if(tileno >= cs.tiles.size())
	{
		err = "tile index out of range";
		return false;
	}
	const Tile& tile = cs.tiles[tileno];
	const TileHeader& h = tile.seen ? tile.hdr : cs.main;
	const uint32_t nc = cs.numcomps;
	// every component must share the coding geometry: the device decoder runs one launch over
	// components of identical geometry
	const Coding& cod = h.cod[0];
	for(uint32_t c = 1; c < nc; ++c)
	{
		const Coding& o = h.cod[c];
		if(o.numres != cod.numres || o.cblkw_exp != cod.cblkw_exp || o.cblkh_exp != cod.cblkh_exp ||
		   o.cblk_sty != cod.cblk_sty || o.reversible != cod.reversible ||
		   memcmp(o.pw, cod.pw, cod.numres) != 0 || memcmp(o.ph, cod.ph, cod.numres) != 0)
		{
			err = "component-specific coding styles (COC differing from COD) are not supported";
			return false;
		}
	}
	if(reduce >= cod.numres)
	{
		err = "cannot drop more resolutions than the codestream holds";
		return false;
	}
	const uint32_t numres = cod.numres, keep_res = numres - reduce;
	const bool ht = (cod.cblk_sty & GRKB_CBLKSTY_HT) != 0;
	// ---- tile rectangle
	const uint32_t tp = tileno % cs.ntx, tq = tileno / cs.ntx;
	const uint32_t tx0 = std::max<uint64_t>((uint64_t)cs.tx0 + (uint64_t)tp * cs.tw, cs.x0);
	const uint32_t ty0 = std::max<uint64_t>((uint64_t)cs.ty0 + (uint64_t)tq * cs.th, cs.y0);
	const uint32_t tx1 = (uint32_t)std::min<uint64_t>((uint64_t)cs.tx0 + (uint64_t)(tp + 1) * cs.tw, cs.x1);
	const uint32_t ty1 = (uint32_t)std::min<uint64_t>((uint64_t)cs.ty0 + (uint64_t)(tq + 1) * cs.th, cs.y1);
	// ---- geometry (shared by all components: no sub-sampling)
	grkb_layout_params lp{};
	lp.x0 = tx0;
	lp.y0 = ty0;
	lp.x1 = tx1;
	lp.y1 = ty1;
	lp.numres = numres;
	lp.cblkw_exp = cod.cblkw_exp;
	lp.cblkh_exp = cod.cblkh_exp;
	for(uint32_t r = 0; r < 33; ++r)
	{
		lp.prec_w_exp[r] = r < numres ? cod.pw[r] : 15;
		lp.prec_h_exp[r] = r < numres ? cod.ph[r] : 15;
	}
	uint32_t ng = 0;
	grkb_layout_index idx;
	if(grkb_layout_walk(&lp, nullptr, 0, &ng, nullptr))
	{
		err = "bad tile geometry";
		return false;
	}
	std::vector<grkb_cblk_geom> geom(ng);
	if(ng && grkb_layout_walk(&lp, geom.data(), ng, &ng, &idx))
	{
		err = "bad tile geometry";
		return false;
	}
	else if(!ng)
		grkb_layout_walk(&lp, nullptr, 0, &ng, &idx);
	// the reduced tile the decoder works on: rectangle / 2^reduce, the kept resolutions keep their
	// band-relative block positions (band rectangles of level nd equal those of level nd - reduce of
	// the reduced rectangle)
	auto cdp2 = [](uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a + ((uint64_t)1 << b) - 1) >> b); };
	grkb_tile_params& P = out.params;
	memset(&P, 0, sizeof(P));
	P.x0 = cdp2(tx0, reduce);
	P.y0 = cdp2(ty0, reduce);
	P.x1 = cdp2(tx1, reduce);
	P.y1 = cdp2(ty1, reduce);
	P.numcomps = nc;
	P.numres = keep_res;
	P.reversible = cod.reversible;
	P.mct = (h.mct == 1 && nc >= 3) ? 1 : 0;
	for(uint32_t c = 0; c < nc; ++c)
	{
		const Comp& k = cs.comps[c];
		P.shift[c] = k.sgnd ? 0 : 1 << (k.prec - 1);
		P.lo[c] = k.sgnd ? -(1 << (k.prec - 1)) : 0;
		P.hi[c] = k.sgnd ? (1 << (k.prec - 1)) - 1 : (int32_t)((1u << k.prec) - 1);
	}
	// blocks of the kept resolutions, component-major
	uint32_t ng_keep = 0;
	for(uint32_t i = 0; i < ng; ++i)
		if(geom[i].res < keep_res)
			ng_keep = i + 1; // geom is ordered by resolution
	// per band quantisation of each component
	struct BandQ
	{
		float step;
		uint32_t numbps;
	};
	std::vector<BandQ> bandq((size_t)nc * (3 * numres - 2));
	for(uint32_t c = 0; c < nc; ++c)
		for(uint32_t b = 0; b < 3 * numres - 2; ++b)
		{
			const uint32_t orient = b == 0 ? 0 : (b - 1) % 3 + 1;
			BandQ q{};
			if(grkb_band_quant(h.q[c].expn[b], h.q[c].mant[b], orient, cs.comps[c].prec, cod.reversible, 0, ht,
							   h.q[c].numgbits, h.roishift[c], &q.step, &q.numbps))
			{
				err = "unsupported number of band bit planes";
				return false;
			}
			bandq[(size_t)c * (3 * numres - 2) + b] = q;
		}
	// ---- T2 state
	std::vector<BlockState> st((size_t)ng * nc);
	typedef SmallVec<SegState, 2> Segs;
	typedef SmallVec<Chunk, 1> Chunks;
	std::vector<Segs> bsegs; // per block that was ever included (slot[])
	std::vector<Chunks> bchunks;
	bsegs.reserve((size_t)ng * nc);
	bchunks.reserve((size_t)ng * nc);
	std::vector<uint32_t> slot((size_t)ng * nc, UINT32_MAX);
	std::vector<uint8_t> numbps((size_t)ng * nc, 0);
	TagArena arena;
	arena.nodes.reserve((size_t)ng * nc * 3);
	std::vector<TagTree> incl((size_t)idx.pb.size() * nc), imsb((size_t)idx.pb.size() * nc);
	std::vector<uint8_t> tree_made((size_t)idx.pb.size() * nc, 0);
	// ---- packet sequence
	const uint32_t numlayers = h.numlayers;
	const uint32_t lay_keep = max_layers ? std::min<uint32_t>(max_layers, numlayers) : numlayers;
	std::vector<Poc> progs = h.pocs;
	if(progs.empty())
		progs.push_back(Poc{0, (uint8_t)numres, h.prog, 0, (uint16_t)nc, (uint16_t)numlayers});
	// which packets were already sent (only needed with progression order changes)
	uint64_t max_prec = 0;
	for(const auto& r : idx.res)
		max_prec = std::max<uint64_t>(max_prec, (uint64_t)r.gw * r.gh);
	std::vector<uint8_t> sent;
	const bool track = progs.size() > 1;
	if(track)
	{
		const uint64_t n = (uint64_t)numlayers * numres * nc * max_prec;
		if(n > ((uint64_t)1 << 31))
		{
			err = "progression order changes over more than 2^31 packets";
			return false;
		}
		sent.assign((size_t)n, 0);
	}
	BodyReader body(cs.data, tile.parts);
	bool stop = false;
	uint32_t packet_counter = 0;

	auto do_packet = [&](uint32_t layno, uint32_t resno, uint32_t compno, uint64_t precinct) {
		if(stop)
			return;
		if(track)
		{
			uint8_t& s = sent[(size_t)((((uint64_t)layno * numres + resno) * nc + compno) * max_prec + precinct)];
			if(s)
				return;
			s = 1;
		}
		if(body.empty())
		{
			out.truncated = true;
			stop = true;
			return;
		}
		const grkb_layout_res& R = idx.res[resno];
		const bool keep = resno < keep_res && layno < lay_keep;
		try
		{
			const uint8_t* p0 = body.cur();
			const uint8_t* pe = body.end();
			const uint8_t* p = p0;
			if(h.csty & 2) // SOP
			{
				if(pe - p < 6)
					throw Truncated();
				if(p[0] != 0xFF || p[1] != 0x91)
					throw Corrupt{"expected an SOP marker"};
				// Nsop is not checked: the reference's own encoder writes wrong counters after the first
				// tile part of a tile (they jump by the packets already written), which its decoder then
				// refuses (T2Decompress.cpp:247-256); the packet sequence does not depend on the counter
				p += 6;
			}
			if(p >= pe)
				throw Truncated();
			BitReader br(p, pe);
			uint32_t data_bytes = 0;
			const uint32_t present = br.bit();
			static thread_local std::vector<uint32_t> touched;
			touched.clear();
			if(present)
			{
				for(uint32_t band = 0; band < R.nbands; ++band)
				{
					const size_t pbi = (size_t)R.first_pb + (size_t)band * R.gw * R.gh + (size_t)precinct;
					const grkb_layout_pb& pb = idx.pb[pbi];
					// an empty band is skipped; a non-empty band's precinct takes its code-block grid from
					// floor / ceil of its bounds even when those bounds are empty, as the reference does
					// (t1/Precinct.h:46-52), so its packets carry bits for such degenerate blocks
					const uint32_t ncb = pb.cgw * pb.cgh;
					if(ncb == 0 || pb.band_empty)
						continue;
					const size_t ti = pbi * nc + compno;
					if(!tree_made[ti])
					{
						incl[ti] = arena.make(pb.cgw, pb.cgh);
						imsb[ti] = arena.make(pb.cgw, pb.cgh);
						tree_made[ti] = 1;
					}
					const uint32_t qb = (resno == 0 ? 0 : 3 * resno - 2) + band;
					const uint32_t band_numbps = bandq[(size_t)compno * (3 * numres - 2) + qb].numbps;
					for(uint32_t cb = 0; cb < ncb; ++cb)
					{
						const uint32_t bi = compno * ng + pb.first + cb;
						BlockState& B = st[bi];
						bool included;
						if(!B.numlenbits)
							included = arena.decode(br, incl[ti], cb, layno + 1);
						else
							included = br.bit() != 0;
						if(!included)
						{
							B.passes_in_packet = 0;
							continue;
						}
						if(!B.numlenbits)
						{
							const uint32_t kmsbs = arena.decode_value(br, imsb[ti], cb, MAX_BITPLANES + 50);
							uint32_t nb = kmsbs > band_numbps ? MAX_BITPLANES : band_numbps - kmsbs;
							if(nb > MAX_BITPLANES)
								nb = MAX_BITPLANES;
							numbps[bi] = (uint8_t)nb;
							B.numlenbits = 3;
							slot[bi] = (uint32_t)bsegs.size();
							bsegs.emplace_back();
							bchunks.emplace_back();
						}
						B.passes_in_packet = br.numpasses();
						B.numlenbits += br.comma();
						Segs& S = bsegs[slot[bi]];
						if(S.empty())
						{
							S.emplace_back();
							S.back().maxpasses = seg_maxpasses(cod.cblk_sty, true, 0);
						}
						else if(S.back().numpasses == S.back().maxpasses)
						{
							const uint32_t last = S.back().maxpasses;
							S.emplace_back();
							S.back().maxpasses = seg_maxpasses(cod.cblk_sty, false, last);
						}
						int32_t left = (int32_t)B.passes_in_packet;
						size_t si = S.size() - 1;
						B.first_seg_in_packet = (uint32_t)si;
						do
						{
							SegState& sg = S[si];
							if(sg.maxpasses == MAX_PASSES_PER_SEG)
								sg.passes_in_packet = std::min<uint32_t>((uint32_t)left, MAX_PASSES_PER_SEG);
							else
								sg.passes_in_packet = (uint32_t)std::min<int32_t>((int32_t)(sg.maxpasses - sg.numpasses), left);
							const uint32_t nbits = B.numlenbits + floorlog2(sg.passes_in_packet);
							if(nbits > 32)
								throw Corrupt{"segment length field wider than 32 bits"};
							sg.bytes_in_packet = br.bits(nbits);
							data_bytes += sg.bytes_in_packet;
							left -= (int32_t)sg.passes_in_packet;
							if(left > 0)
							{
								const uint32_t last = sg.maxpasses;
								S.emplace_back();
								S.back().maxpasses = seg_maxpasses(cod.cblk_sty, false, last);
								++si;
							}
						} while(left > 0);
						touched.push_back(bi);
					}
				}
			}
			p = br.align();
			if(h.csty & 4) // EPH
			{
				if(pe - p < 2)
					throw Truncated();
				if(p[0] != 0xFF || p[1] != 0x92)
					throw Corrupt{"expected an EPH marker"};
				p += 2;
			}
			// ---- packet body
			size_t avail = (size_t)(pe - p);
			size_t at = body.abs() + (size_t)(p - p0);
			for(uint32_t bi : touched)
			{
				BlockState& B = st[bi];
				Segs& S = bsegs[slot[bi]];
				for(size_t k = B.first_seg_in_packet; k < S.size(); ++k)
				{
					SegState& sg = S[k];
					if(sg.bytes_in_packet > avail)
					{
						// truncated codestream: HT decoding runs from both ends of a segment, so such a
						// block is dropped; a Part-1 block keeps what it had (T2Decompress.cpp:508-518)
						out.truncated = true;
						stop = true;
						if(ht)
							B.dropped = true;
						sg.bytes_in_packet = 0;
						sg.passes_in_packet = 0;
						break;
					}
					if(sg.bytes_in_packet && keep && !B.dropped)
					{
						Chunks& C = bchunks[slot[bi]];
						if(!C.empty() && C.back().off + C.back().len == at)
							C.back().len += sg.bytes_in_packet;
						else
							C.push_back(Chunk{at, sg.bytes_in_packet});
					}
					at += sg.bytes_in_packet;
					avail -= sg.bytes_in_packet;
					// a skipped packet still advances the segment structure the next headers build on
					sg.numpasses += sg.passes_in_packet;
					if(keep)
					{
						sg.klen += sg.bytes_in_packet;
						sg.kpasses += sg.passes_in_packet;
					}
					sg.bytes_in_packet = sg.passes_in_packet = 0;
				}
				if(stop)
					break;
			}
			body.pos += (at - body.abs());
			++packet_counter;
			++out.packets;
		}
		catch(const Truncated&)
		{
			out.truncated = true;
			stop = true;
		}
		catch(const Corrupt& c)
		{
			if(getenv("GRKB_T2_DEBUG"))
				fprintf(stderr, "[grkb t2] tile %u packet %u (l %u r %u c %u p %llu): %s\n", tileno, packet_counter,
						layno, resno, compno, (unsigned long long)precinct, c.what);
			err = c.what;
			out.truncated = true;
			stop = true;
		}
	};

	for_each_packet(progs, numres, nc, numlayers, idx, tx0, ty0, tx1, ty1,
					[&](uint32_t l, uint32_t r, uint32_t c, uint64_t p) {
						do_packet(l, r, c, p);
						return !stop;
					});
	if(!err.empty() && out.packets == 0)
		return false;
	err.clear();
	// Windowed decode (tile/TileProcessor.cpp:288-317, transform/WaveletReverse.cpp:1374-2230 in the
	// reference): a block is needed when its band samples reach the window through the synthesis
	// filters.  At resolution r (nd levels below the decoded one) the window covers
	// [w0 >> nd, ceil(w1 / 2^nd)); every level above adds four samples of filter reach at its own
	// scale (less than eight in total at this one); band coordinates are half of that.
	auto in_window = [&](const grkb_cblk_geom& g) -> bool {
		if(!window)
			return true;
		const uint32_t nd = keep_res - 1 - g.res;
		const int64_t M = 10;
		int64_t rx0 = (int64_t)(window[0] >> nd) - M, rx1 = (int64_t)cdp2(window[2], nd) + M;
		int64_t ry0 = (int64_t)(window[1] >> nd) - M, ry1 = (int64_t)cdp2(window[3], nd) + M;
		if(g.res > 0)
		{
			rx0 = (rx0 - 1) >> 1;
			ry0 = (ry0 - 1) >> 1;
			rx1 = (rx1 + 2) >> 1;
			ry1 = (ry1 + 2) >> 1;
		}
		return !((int64_t)g.cx0 + g.w <= rx0 || (int64_t)g.cx0 >= rx1 || (int64_t)g.cy0 + g.h <= ry0 ||
				 (int64_t)g.cy0 >= ry1);
	};
	// ---- tables
	size_t lo = SIZE_MAX, hi = 0, total = 0;
	bool single = true;
	for(uint32_t c = 0; c < nc; ++c)
		for(uint32_t i = 0; i < ng_keep; ++i)
		{
			const uint32_t bi = c * ng + i;
			if(slot[bi] == UINT32_MAX || st[bi].dropped || !in_window(geom[i]))
				continue;
			const Chunks& C = bchunks[slot[bi]];
			if(C.size() > 1)
				single = false;
			for(size_t ki = 0; ki < C.size(); ++ki)
			{
				const Chunk& k = C[ki];
				lo = std::min(lo, k.off);
				hi = std::max(hi, k.off + k.len);
				total += k.len;
			}
		}
	if(hi - (lo == SIZE_MAX ? 0 : lo) > 0xFFFFFFF0ull || total > 0xFFFFFFF0ull)
	{
		err = "more than 4 GB of code-block bytes in one tile";
		return false;
	}
	// one contiguous run per block lets the codestream itself be the byte pool -- unless the tile's
	// parts are interleaved with other tiles' and the span is mostly foreign bytes
	const size_t span = lo == SIZE_MAX ? 0 : hi - lo;
	if(single && span > total + total / 2 + 65536)
		single = false;
	out.zero_copy = single;
	if(single)
	{
		out.src_off = lo == SIZE_MAX ? 0 : lo;
		out.src_len = lo == SIZE_MAX ? 0 : hi - lo;
	}
	else
		out.pool.reserve(total + 16);
	out.blocks.reserve((size_t)ng_keep * nc);
	for(uint32_t c = 0; c < nc; ++c)
		for(uint32_t i = 0; i < ng_keep; ++i)
		{
			const uint32_t bi = c * ng + i;
			const grkb_cblk_geom& g = geom[i];
			if(g.w == 0 || g.h == 0)
				continue; // degenerate blocks exist for T2 only: they hold no samples
			if(!in_window(g))
				continue;
			out.blocks.emplace_back();
			grkb_cblk& b = out.blocks.back();
			memset(&b, 0, sizeof(b));
			b.x = g.x;
			b.y = g.y;
			b.w = g.w;
			b.h = g.h;
			b.comp = (uint8_t)c;
			b.orient = g.orient;
			b.cblk_sty = cod.cblk_sty;
			b.roishift = h.roishift[c];
			const uint32_t qb = (g.res == 0 ? 0 : 3 * g.res - 2) + g.band;
			const BandQ& q = bandq[(size_t)c * (3 * numres - 2) + qb];
			b.scale = ht ? q.step : q.step / 2; // PostDecompressFilters.h:51, :139
			b.seg_first = (uint32_t)out.segs.size();
			if(slot[bi] == UINT32_MAX || st[bi].dropped)
				continue;
			b.numbps = numbps[bi];
			const uint32_t band_numbps = q.numbps;
			b.k_msbs = (uint8_t)(band_numbps > b.numbps ? band_numbps - b.numbps : 0);
			const Chunks& C = bchunks[slot[bi]];
			uint32_t nbytes = 0;
			if(single)
			{
				if(!C.empty())
				{
					b.data_off = (uint32_t)(C[0].off - out.src_off);
					nbytes = C[0].len;
				}
			}
			else
			{
				b.data_off = (uint32_t)out.pool.size();
				for(size_t ki = 0; ki < C.size(); ++ki)
				{
					const Chunk& k = C[ki];
					out.pool.insert(out.pool.end(), cs.data + k.off, cs.data + k.off + k.len);
					nbytes += k.len;
				}
			}
			b.data_len = nbytes;
			uint32_t ns = 0;
			const Segs& SS = bsegs[slot[bi]];
			for(size_t si2 = 0; si2 < SS.size(); ++si2)
			{
				const SegState& sg = SS[si2];
				if(sg.kpasses == 0)
					continue;
				out.segs.push_back(grkb_seg{sg.klen, sg.kpasses});
				++ns;
			}
			b.numsegs = (uint16_t)ns;
			if(nbytes == 0)
			{
				// included with zero bytes everywhere: nothing to decode (the reference skips blocks
				// without segment buffers, TileComponent.cpp:407-415)
				out.segs.resize(b.seg_first);
				b.numsegs = 0;
			}
		}
	if(!single)
		out.pool.resize(out.pool.size() + 16, 0);
	return true;
}
} // namespace grkb_cs
