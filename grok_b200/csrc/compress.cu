// compress.cu -- the encode side of the codestream-level C-ABI (include/grk_b200.h, section E):
// image planes in, a complete JPEG 2000 / HTJ2K codestream out, with T2 on our side.
// The device part of every tile is grkb_encoder_* (DC shift, MCT, forward wavelet, quantiser, block
// coders, packing); this file adds what the reference host does around it for a single quality
// layer that keeps every coding pass:
//   main header         CodeStreamCompress::start / write_siz / write_cap / write_cod / write_qcd /
//                       write_com   src/lib/jp2/codestream/CodeStreamCompress.cpp:640-1130
//   packets             T2Compress::compressPacket / compressHeader   t2/T2Compress.cpp:112-330,
//                       TagTree::compress t2/TagTree.h:160-200, BitIO::putbit / flush t2/BitIO.cpp:36-150
//   packet order        packet_order.h (t2/PacketIter.cpp)
//   tile parts          one tile part per tile: SOT, SOD, packets (SOTMarker.cpp:30-58)
// Without a rate target the reference keeps all passes too (makeSingleLosslessLayer,
// tile/TileProcessor.cpp:969-977), so for these parameters the codestream is byte-identical to the
// stock encoder's (tests/test_compress_gpu.py).  A single layer with a rate target is allocated as
// TileProcessor::pcrdBisectSimple does it (rate_allocate below).
#include "pipeline.h"
#include "codestream.h"
#include "geometry.h"
#include "packet_order.h"
#include "pcrd.h"
#include <chrono>
#include <math.h>
#include <memory>
#include <string.h>
#include <string>

namespace
{
// packet-header bit writer: MSB first, seven bits in the byte after 0xFF (BitIO::putbit / writeByte)
struct BitWriter
{
	std::vector<uint8_t>& out;
	uint32_t buf = 0, ct = 8;
	// The reference's rate allocation writes a packet header into a buffer of `limit` bytes (what is left
	// of the layer's budget) and gives the layer up when the buffer fills (BitIO::writeByte: the byte that
	// makes the count reach the size fails) -- but only where the caller looks at the result: the writes
	// of the number of passes and of the Lblock comma code go unchecked (T2Compress.cpp:208, :228), the
	// failed byte is written again with the next bit, the count runs past the limit for good and the
	// remaining budget wraps: from there on everything "fits".  event: 0 none, 1 the header does not fit,
	// 2 the limit was passed inside an unchecked field.  limit 0 never fails (the count starts at 1).
	uint64_t limit = 0, count = 0;
	bool unchecked = false;
	int event = 0;
	explicit BitWriter(std::vector<uint8_t>& o) : out(o) {}
	inline void byteout(bool from_flush = false)
	{
		if(++count == limit && !event)
			event = (from_flush || !unchecked) ? 1 : 2;
		out.push_back((uint8_t)buf);
		ct = buf == 0xFF ? 7 : 8;
		buf = 0;
	}
	inline void bit(uint32_t b)
	{
		if(ct == 0)
			byteout();
		--ct;
		buf |= (b & 1u) << ct;
	}
	inline void bits(uint32_t v, uint32_t n)
	{
		for(int i = (int)n - 1; i >= 0; --i)
			bit((v >> i) & 1u);
	}
	inline void comma(uint32_t n)
	{
		while(n--)
			bit(1);
		bit(0);
	}
	inline void numpasses(uint32_t n)
	{
		if(n == 1)
			bits(0, 1);
		else if(n == 2)
			bits(2, 2);
		else if(n <= 5)
			bits(0xC | (n - 3), 4);
		else if(n <= 36)
			bits(0x1E0 | (n - 6), 9);
		else
			bits(0xFF80 | (n - 37), 16);
	}
	inline void flush()
	{
		byteout(true);
		if(ct == 7)
			byteout(true);
	}
};

struct TNode
{
	uint32_t value, low;
	bool known;
};
// tag tree over a w x h grid, levels halving (rounding up) to 1 x 1; compress() as TagTree::compress
struct TagTreeEnc
{
	std::vector<TNode> nodes;
	std::vector<uint32_t> lvl_off, lvl_w;
	uint32_t w = 0, h = 0;
	void init(uint32_t W, uint32_t H)
	{
		w = W;
		h = H;
		lvl_off.clear();
		lvl_w.clear();
		uint32_t n = 0, lw = W, lh = H;
		while(true)
		{
			lvl_off.push_back(n);
			lvl_w.push_back(lw);
			n += lw * lh;
			if(lw <= 1 && lh <= 1)
				break;
			lw = (lw + 1) >> 1;
			lh = (lh + 1) >> 1;
		}
		nodes.assign(n, TNode{0xFFFFFFFFu, 0, false});
	}
	void setvalue(uint32_t leaf, uint32_t value)
	{
		uint32_t x = leaf % w, y = leaf / w;
		for(size_t l = 0; l < lvl_off.size(); ++l)
		{
			TNode& nd = nodes[lvl_off[l] + y * lvl_w[l] + x];
			if(nd.value <= value)
				break;
			nd.value = value;
			x >>= 1;
			y >>= 1;
		}
	}
	void compress(BitWriter& bw, uint32_t leaf, uint32_t threshold)
	{
		uint32_t path[34];
		int depth = 0;
		uint32_t x = leaf % w, y = leaf / w;
		for(size_t l = 0; l < lvl_off.size(); ++l)
		{
			path[depth++] = lvl_off[l] + y * lvl_w[l] + x;
			x >>= 1;
			y >>= 1;
		}
		uint32_t low = 0;
		for(int i = depth - 1; i >= 0; --i)
		{
			TNode& nd = nodes[path[i]];
			if(nd.low < low)
				nd.low = low;
			else
				low = nd.low;
			while(low < threshold)
			{
				if(low >= nd.value)
				{
					if(!nd.known)
					{
						bw.bit(1);
						nd.known = true;
					}
					break;
				}
				bw.bit(0);
				++low;
			}
			nd.low = low;
		}
	}
};

uint32_t floorlog2u(uint32_t v)
{
	uint32_t l = 0;
	while(v > 1)
	{
		v >>= 1;
		++l;
	}
	return l;
}

void put16(std::vector<uint8_t>& o, uint32_t v)
{
	o.push_back((uint8_t)(v >> 8));
	o.push_back((uint8_t)v);
}
void put32(std::vector<uint8_t>& o, uint32_t v)
{
	put16(o, v >> 16);
	put16(o, v & 0xFFFF);
}

// everything about one tile geometry that does not change from image to image
struct TileKit
{
	grkb_tile_params tp{};
	grkb_encoder* enc = nullptr;
	std::vector<grkb_cblk_geom> geom; // all blocks of one component, degenerate ones included
	grkb_layout_index idx;
	std::vector<int32_t> table_of; // geom index -> index in the encoder's block table of component 0, or -1
	std::vector<grkb_enc_cblk> blocks; // component-major, non-degenerate blocks only
	std::vector<uint32_t> band_numbps; // per (component, band in codestream order)
	uint32_t per_comp = 0; // blocks per component in `blocks`
	std::vector<uint32_t> npass; // coding passes of every block that go into the layer (rate allocation)
	// several layers: passes of every block in each layer, passes in the layers already final (the
	// reference's numPassesInPreviousPackets), and what a walk over the packets carries from layer to
	// layer: passes written so far, Lblock, the tag trees of every (component, band, precinct)
	std::vector<std::vector<uint32_t>> layer_np;
	std::vector<uint32_t> prev_np, in_packet;
	std::vector<uint8_t> lblock;
	std::vector<TagTreeEnc> incl_trees, imsb_trees;
	// results of the last encode (page-locked)
	grkb_coded* coded = nullptr;
	grkb_pass* passes = nullptr;
	uint8_t* bytes = nullptr;
	size_t bytes_cap = 0;
	// rate allocation on the device (pcrd.cu): tables of this geometry, results (page-locked)
	grkb_pcrd* pcrd = nullptr;
	uint32_t* npass_dev = nullptr;
	PcrdResult* pcrd_result = nullptr;
	~TileKit()
	{
		grkb_pcrd_destroy(pcrd);
		grkb_host_free(npass_dev);
		grkb_host_free(pcrd_result);
		if(enc)
			grkb_encoder_destroy(enc);
		grkb_host_free(coded);
		grkb_host_free(passes);
		grkb_host_free(bytes);
	}
};
} // namespace

struct grkb_compressor
{
	int device = 0;
	grkb_compress_params p{};
	std::string comment;
	std::vector<uint32_t> expn[GRKB_MAX_COMPS], mant[GRKB_MAX_COMPS];
	std::vector<std::unique_ptr<TileKit>> kits;
	std::vector<uint8_t> hdr; // packet headers of the packet being written
	double t2_ms = 0, device_ms = 0;
	uint32_t pcrd_iterations = 0; // bisection steps of the last tile allocated on the device
	uint32_t tiles_dev = 0, tiles_host = 0; // tiles of the last call allocated on the device / host
};

namespace
{
bool same_geometry(const grkb_tile_params& a, const grkb_tile_params& b)
{
	return memcmp(&a, &b, sizeof(a)) == 0;
}

// the packets of the tile's single layer in progression order, as the device allocator wants them
// (pcrd.h): the same walk as packet_header() / simulate() below
bool build_pcrd(grkb_compressor* z, TileKit* k)
{
	const grkb_compress_params& p = z->p;
	const uint32_t nb = 3 * p.numres - 2;
	PcrdLayout lay;
	std::vector<grkb_cs::Poc> progs;
	progs.push_back(grkb_cs::Poc{0, (uint8_t)p.numres, (uint8_t)p.prog_order, 0, (uint16_t)p.numcomps, 1});
	grkb_cs::for_each_packet(progs, p.numres, p.numcomps, 1, k->idx, k->tp.x0, k->tp.y0, k->tp.x1, k->tp.y1,
							 [&](uint32_t, uint32_t resno, uint32_t compno, uint64_t precinct) {
		const grkb_layout_res& R = k->idx.res[resno];
		PcrdPacket pk{};
		pk.band_first = (uint32_t)lay.bands.size();
		pk.leaf_first = (uint32_t)lay.leaf_block.size();
		pk.node_first = lay.nnodes;
		for(uint32_t band = 0; band < R.nbands; ++band)
		{
			const grkb_layout_pb& pb = k->idx.pb[(size_t)R.first_pb + (size_t)band * R.gw * R.gh + (size_t)precinct];
			const uint32_t ncb = pb.cgw * pb.cgh;
			if(ncb == 0 || pb.band_empty)
				continue;
			PcrdBand b{};
			b.leaf_first = (uint32_t)lay.leaf_block.size();
			b.cgw = pb.cgw;
			b.cgh = pb.cgh;
			b.band_bps = k->band_numbps[(size_t)compno * nb + (resno == 0 ? 0 : 3 * resno - 2) + band];
			b.node_first = lay.nnodes;
			b.packet = (uint32_t)lay.packets.size();
			uint32_t lw = pb.cgw, lh = pb.cgh;
			while(true)
			{
				lay.nnodes += lw * lh;
				if(lw <= 1 && lh <= 1)
					break;
				lw = (lw + 1) >> 1;
				lh = (lh + 1) >> 1;
			}
			for(uint32_t cb = 0; cb < ncb; ++cb)
			{
				const int32_t ti = k->table_of[pb.first + cb];
				lay.leaf_block.push_back(ti >= 0 ? (int32_t)(compno * k->per_comp + (uint32_t)ti) : -1);
				lay.leaf_band.push_back((uint32_t)lay.bands.size());
			}
			lay.bands.push_back(b);
		}
		pk.nbands = (uint32_t)lay.bands.size() - pk.band_first;
		pk.nleaves = (uint32_t)lay.leaf_block.size() - pk.leaf_first;
		pk.nnodes = lay.nnodes - pk.node_first;
		lay.packets.push_back(pk);
		return true;
	});
	if(grkb_pcrd_create(lay, (uint32_t)k->blocks.size(), &k->pcrd))
		return false;
	k->npass_dev = (uint32_t*)grkb_host_alloc(std::max<size_t>(k->blocks.size(), 1) * sizeof(uint32_t));
	k->pcrd_result = (PcrdResult*)grkb_host_alloc(sizeof(PcrdResult));
	return k->npass_dev && k->pcrd_result;
}

TileKit* kit_for(grkb_compressor* z, const grkb_tile_params& tp)
{
	for(auto& k : z->kits)
		if(same_geometry(k->tp, tp))
			return k.get();
	const grkb_compress_params& p = z->p;
	auto k = std::make_unique<TileKit>();
	k->tp = tp;
	grkb_layout_params lp{};
	lp.x0 = tp.x0;
	lp.y0 = tp.y0;
	lp.x1 = tp.x1;
	lp.y1 = tp.y1;
	lp.numres = p.numres;
	lp.cblkw_exp = p.cblkw_exp;
	lp.cblkh_exp = p.cblkh_exp;
	for(uint32_t r = 0; r < 33; ++r)
	{
		lp.prec_w_exp[r] = r < p.numres ? p.prec_w_exp[r] : 15;
		lp.prec_h_exp[r] = r < p.numres ? p.prec_h_exp[r] : 15;
	}
	uint32_t ng = 0;
	if(grkb_layout_walk(&lp, nullptr, 0, &ng, nullptr))
		return nullptr;
	k->geom.resize(ng);
	if(grkb_layout_walk(&lp, k->geom.data(), ng, &ng, &k->idx) && ng)
		return nullptr;
	const bool ht = (p.cblk_sty & GRKB_CBLKSTY_HT) != 0;
	const uint32_t nb = 3 * p.numres - 2;
	k->table_of.assign(ng, -1);
	uint32_t n0 = 0;
	for(uint32_t i = 0; i < ng; ++i)
		if(k->geom[i].w && k->geom[i].h)
			k->table_of[i] = (int32_t)n0++;
	k->per_comp = n0;
	k->blocks.assign((size_t)n0 * p.numcomps, grkb_enc_cblk{});
	k->band_numbps.assign((size_t)nb * p.numcomps, 0);
	for(uint32_t c = 0; c < p.numcomps; ++c)
	{
		for(uint32_t b = 0; b < nb; ++b)
		{
			float step;
			uint32_t bps;
			const uint32_t orient = b == 0 ? 0 : (b - 1) % 3 + 1;
			if(grkb_band_quant(z->expn[c][b], z->mant[c][b], orient, p.prec, p.reversible, 1, ht, p.numgbits, 0, &step,
							   &bps))
				return nullptr;
			k->band_numbps[(size_t)c * nb + b] = bps;
		}
		for(uint32_t i = 0; i < ng; ++i)
		{
			if(k->table_of[i] < 0)
				continue;
			const grkb_cblk_geom& g = k->geom[i];
			grkb_enc_cblk& e = k->blocks[(size_t)c * n0 + (uint32_t)k->table_of[i]];
			memset(&e, 0, sizeof(e));
			e.x = g.x;
			e.y = g.y;
			e.w = g.w;
			e.h = g.h;
			e.comp = (uint8_t)c;
			e.orient = g.orient;
			e.level = g.level;
			e.cblk_sty = (uint8_t)p.cblk_sty;
			const uint32_t qb = (g.res == 0 ? 0 : 3 * g.res - 2) + g.band;
			float step;
			uint32_t bps;
			grkb_band_quant(z->expn[c][qb], z->mant[c][qb], g.orient, p.prec, p.reversible, 1, ht, p.numgbits, 0, &step,
							&bps);
			e.stepsize = step;
			e.k_msbs = (uint8_t)bps;
		}
	}
	if(grkb_encoder_create(&tp, &k->enc))
		return nullptr;
	const size_t n = k->blocks.size();
	k->bytes_cap = (size_t)(tp.x1 - tp.x0) * (tp.y1 - tp.y0) * p.numcomps * 4 + 65536;
	k->coded = (grkb_coded*)grkb_host_alloc(std::max<size_t>(n, 1) * sizeof(grkb_coded));
	k->passes = (grkb_pass*)grkb_host_alloc(std::max<size_t>(n, 1) * GRKB_ENC_MAX_PASSES * sizeof(grkb_pass));
	k->bytes = (uint8_t*)grkb_host_alloc(k->bytes_cap);
	if(!k->coded || !k->passes || !k->bytes)
		return nullptr;
	if(p.rate > 0 && p.numlayers <= 1 && !build_pcrd(z, k.get()))
		return nullptr;
	z->kits.push_back(std::move(k));
	return z->kits.back().get();
}

// main header (CodeStreamCompress.cpp:640-1130 in the order the reference writes it)
void write_main_header(const grkb_compressor* z, std::vector<uint8_t>& o)
{
	const grkb_compress_params& p = z->p;
	const bool ht = (p.cblk_sty & GRKB_CBLKSTY_HT) != 0;
	put16(o, 0xFF4F);
	// SIZ
	put16(o, 0xFF51);
	put16(o, 38 + 3 * p.numcomps);
	put16(o, ht ? 0x4000 : 0); // Rsiz: GRK_JPH_RSIZ_FLAG for Part 15
	put32(o, p.x0 + p.width);
	put32(o, p.y0 + p.height);
	put32(o, p.x0);
	put32(o, p.y0);
	put32(o, p.tile_w ? p.tile_w : p.x0 + p.width - p.tile_x0);
	put32(o, p.tile_h ? p.tile_h : p.y0 + p.height - p.tile_y0);
	put32(o, p.tile_x0);
	put32(o, p.tile_y0);
	put16(o, p.numcomps);
	for(uint32_t c = 0; c < p.numcomps; ++c)
	{
		o.push_back((uint8_t)((p.prec - 1) | (p.sgnd ? 0x80 : 0)));
		o.push_back(1);
		o.push_back(1);
	}
	// CAP (Part 15): Pcap bit 15, Ccap15 = reversibility | magnitude bound (write_cap, :1064-1112)
	if(ht)
	{
		uint32_t B = 0;
		for(uint32_t b = 0; b < 3 * p.numres - 2; ++b)
			B = std::max<uint32_t>(B, p.reversible ? z->expn[0][b] + p.numgbits - 1
												   : z->expn[0][b] + p.numgbits - (p.numres - 1 - (b ? (b - 1) / 3 : 0)));
		const uint32_t Bp = B <= 8 ? 0 : B < 28 ? B - 8 : B < 48 ? 13 + (B >> 2) : 31;
		put16(o, 0xFF50);
		put16(o, 8);
		put32(o, 0x00020000);
		put16(o, (p.reversible ? 0 : 0x20) | Bp);
	}
	// COD
	bool prt = false;
	for(uint32_t r = 0; r < p.numres; ++r)
		prt = prt || p.prec_w_exp[r] != 15 || p.prec_h_exp[r] != 15;
	put16(o, 0xFF52);
	put16(o, 12 + (prt ? p.numres : 0));
	o.push_back((uint8_t)((prt ? 1 : 0) | (p.csty & 6)));
	o.push_back((uint8_t)p.prog_order);
	put16(o, p.numlayers > 1 ? p.numlayers : 1);
	o.push_back((uint8_t)(p.mct ? 1 : 0));
	o.push_back((uint8_t)(p.numres - 1));
	o.push_back((uint8_t)(p.cblkw_exp - 2));
	o.push_back((uint8_t)(p.cblkh_exp - 2));
	o.push_back((uint8_t)p.cblk_sty);
	o.push_back((uint8_t)(p.reversible ? 1 : 0));
	if(prt)
		for(uint32_t r = 0; r < p.numres; ++r)
			o.push_back((uint8_t)(p.prec_w_exp[r] | (p.prec_h_exp[r] << 4)));
	// QCD
	const uint32_t nb = 3 * p.numres - 2;
	put16(o, 0xFF5C);
	put16(o, 3 + (p.reversible ? nb : 2 * nb));
	o.push_back((uint8_t)((p.reversible ? 0 : 2) | (p.numgbits << 5)));
	for(uint32_t b = 0; b < nb; ++b)
	{
		if(p.reversible)
			o.push_back((uint8_t)(z->expn[0][b] << 3));
		else
			put16(o, (z->expn[0][b] << 11) | z->mant[0][b]);
	}
	// COM
	if(!z->comment.empty())
	{
		put16(o, 0xFF64);
		put16(o, 4 + (uint32_t)z->comment.size());
		put16(o, 1); // Latin text
		o.insert(o.end(), z->comment.begin(), z->comment.end());
	}
}

// One packet: header bits into z->hdr (returns false on an inconsistent table).  The passes of a block
// that go into the layer are k->npass[block].
// first and last byte (exclusive) of block bi's passes [from, from + np) in its stream
inline void pass_span(const TileKit* k, size_t bi, uint32_t from, uint32_t np, uint32_t* b0, uint32_t* b1)
{
	const grkb_pass* ps = k->passes + bi * GRKB_ENC_MAX_PASSES;
	const uint32_t dl = k->coded[bi].data_len;
	*b0 = from ? std::min<uint32_t>(ps[from - 1].rate, dl) : 0;
	*b1 = np ? std::min<uint32_t>(ps[from + np - 1].rate, dl) : *b0;
}

// T2Compress::compressHeader (t2/T2Compress.cpp:113-258) for the packet (layno, resno, compno, precinct)
bool packet_header(grkb_compressor* z, TileKit* k, uint32_t resno, uint32_t compno, uint64_t precinct,
				   TagTreeEnc& tmp_incl, TagTreeEnc& tmp_imsb, std::string& err, uint64_t limit = 0, int* event = nullptr,
				   uint32_t layno = 0)
{
	const grkb_compress_params& p = z->p;
	const uint32_t nb = 3 * p.numres - 2;
	const grkb_layout_res& R = k->idx.res[resno];
	const bool layered = !k->layer_np.empty();
	const std::vector<uint32_t>& npl = layered ? k->layer_np[layno] : k->npass;
	z->hdr.clear();
	BitWriter bw(z->hdr);
	bw.limit = limit;
	bw.bit(1); // the reference never writes the empty-packet form (T2Compress.cpp:150-153)
	for(uint32_t band = 0; band < R.nbands; ++band)
	{
		const size_t pbi = (size_t)R.first_pb + (size_t)band * R.gw * R.gh + (size_t)precinct;
		const grkb_layout_pb& pb = k->idx.pb[pbi];
		const uint32_t ncb = pb.cgw * pb.cgh;
		if(ncb == 0 || pb.band_empty)
			continue;
		const uint32_t band_bps = k->band_numbps[(size_t)compno * nb + (resno == 0 ? 0 : 3 * resno - 2) + band];
		// the trees live from the precinct's first layer to its last
		TagTreeEnc& incl = layered ? k->incl_trees[(size_t)compno * k->idx.pb.size() + pbi] : tmp_incl;
		TagTreeEnc& imsb = layered ? k->imsb_trees[(size_t)compno * k->idx.pb.size() + pbi] : tmp_imsb;
		if(layno == 0)
		{
			incl.init(pb.cgw, pb.cgh);
			imsb.init(pb.cgw, pb.cgh);
			for(uint32_t cb = 0; cb < ncb; ++cb)
			{
				const int32_t ti = k->table_of[pb.first + cb];
				const size_t bi = ti >= 0 ? (size_t)compno * k->per_comp + (uint32_t)ti : 0;
				const uint32_t bps = ti >= 0 ? k->coded[bi].numbps : 0;
				if(bps > band_bps)
				{
					err = "a code block has more bit planes than its band";
					return false;
				}
				imsb.setvalue(cb, band_bps - bps);
				if(layered && ti >= 0)
					k->in_packet[bi] = 0;
			}
		}
		for(uint32_t cb = 0; cb < ncb; ++cb)
		{
			const int32_t ti = k->table_of[pb.first + cb];
			const size_t bi = ti >= 0 ? (size_t)compno * k->per_comp + (uint32_t)ti : 0;
			if(ti >= 0 && npl[bi] && !(layered && k->in_packet[bi]))
				incl.setvalue(cb, layno);
		}
		for(uint32_t cb = 0; cb < ncb; ++cb)
		{
			const int32_t ti = k->table_of[pb.first + cb];
			const size_t bi = ti >= 0 ? (size_t)compno * k->per_comp + (uint32_t)ti : 0;
			const uint32_t np = ti >= 0 ? npl[bi] : 0;
			const uint32_t before = (layered && ti >= 0) ? k->in_packet[bi] : 0;
			if(!before)
				incl.compress(bw, cb, layno + 1);
			else
				bw.bit(np != 0);
			if(!np)
				continue;
			uint32_t numlenbits = 3;
			if(!before)
				imsb.compress(bw, cb, 0xFF);
			else
				numlenbits = k->lblock[bi];
			bw.unchecked = true;
			bw.numpasses(np);
			bw.unchecked = false;
			const grkb_pass* ps = k->passes + bi * GRKB_ENC_MAX_PASSES + before;
			// Lblock increment, then one length per codeword segment (T2Compress.cpp:213-256)
			uint32_t inc = 0, nump = 0, len = 0;
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
			bw.unchecked = true;
			bw.comma(inc);
			bw.unchecked = false;
			numlenbits += inc;
			if(layered)
				k->lblock[bi] = (uint8_t)numlenbits;
			for(uint32_t q = 0; q < np; ++q)
			{
				++nump;
				len += ps[q].len;
				if(ps[q].term || q == np - 1)
				{
					bw.bits(len, numlenbits + floorlog2u(nump));
					len = 0;
					nump = 0;
				}
			}
		}
	}
	bw.flush();
	if(event)
		*event = bw.event;
	return true;
}

// bytes of block bi that go into the layer: the rate of its last included pass
inline uint32_t layer_len(const TileKit* k, size_t bi, uint32_t layno = 0)
{
	if(!k->layer_np.empty())
	{
		uint32_t b0, b1;
		pass_span(k, bi, k->in_packet[bi], k->layer_np[layno][bi], &b0, &b1);
		return b1 - b0;
	}
	const uint32_t np = k->npass[bi];
	if(!np)
		return 0;
	return std::min<uint32_t>(k->passes[bi * GRKB_ENC_MAX_PASSES + np - 1].rate, k->coded[bi].data_len);
}

// T2Compress::compressPacketsSimulate (t2/T2Compress.cpp:58-110, :331-420): would the packets of the
// layer fit in max_bytes?  The reference's own accounting, packet by packet: a header must stay
// strictly below what is left (BitIO::writeByte fails when it fills its buffer), a body may use it up.
bool simulate(grkb_compressor* z, TileKit* k, uint64_t max_bytes, uint64_t* total, uint32_t layers = 1)
{
	const grkb_compress_params& p = z->p;
	const bool layered = !k->layer_np.empty();
	std::vector<grkb_cs::Poc> progs;
	progs.push_back(grkb_cs::Poc{0, (uint8_t)p.numres, (uint8_t)p.prog_order, 0, (uint16_t)p.numcomps, (uint16_t)layers});
	TagTreeEnc incl, imsb;
	bool fits = true;
	const bool limited = max_bytes != UINT32_MAX;
	uint64_t all = 0, left = max_bytes;
	std::string err;
	grkb_cs::for_each_packet(progs, p.numres, p.numcomps, layers, k->idx, k->tp.x0, k->tp.y0, k->tp.x1, k->tp.y1,
							 [&](uint32_t layno, uint32_t resno, uint32_t compno, uint64_t precinct) {
		uint64_t avail = left, bytes = 0;
		if(p.csty & 2)
		{
			if(limited)
				avail -= 6;
			bytes += 6;
		}
		int event = 0;
		if(!packet_header(z, k, resno, compno, precinct, incl, imsb, err, limited ? avail : 0, &event, layno) ||
		   event == 1)
		{
			fits = false;
			return false;
		}
		if(event == 2)
			return false; // the limit went by unnoticed: this and every later packet "fit"
		bytes += z->hdr.size();
		if(limited)
			avail -= z->hdr.size();
		if(p.csty & 4)
		{
			if(limited)
				avail -= 2;
			bytes += 2;
		}
		const grkb_layout_res& R = k->idx.res[resno];
		for(uint32_t band = 0; band < R.nbands && fits; ++band)
		{
			const grkb_layout_pb& pb = k->idx.pb[(size_t)R.first_pb + (size_t)band * R.gw * R.gh + (size_t)precinct];
			// (the reference's simulation walks the bodies of empty bands too: they hold no passes)
			for(uint32_t cb = 0; cb < pb.cgw * pb.cgh; ++cb)
			{
				const int32_t ti = k->table_of[pb.first + cb];
				if(ti < 0)
					continue;
				const size_t bi = (size_t)compno * k->per_comp + (uint32_t)ti;
				const uint32_t np = layered ? k->layer_np[layno][bi] : k->npass[bi];
				if(!np)
					continue;
				const uint32_t len = layer_len(k, bi, layno);
				if(limited && len > avail)
				{
					fits = false;
					break;
				}
				if(layered)
					k->in_packet[bi] += np;
				bytes += len;
				if(limited)
					avail -= len;
			}
		}
		if(!fits)
			return false;
		all += bytes;
		if(limited)
			left -= bytes;
		return true;
	});
	if(total)
		*total = all;
	return fits;
}

// TileProcessor::makeLayerSimple (tile/TileProcessor.cpp:1366-1440) for the single layer
void make_layer(TileKit* k, double thresh)
{
	for(size_t bi = 0; bi < k->blocks.size(); ++bi)
	{
		const uint32_t total = k->coded[bi].numpasses;
		const grkb_pass* ps = k->passes + bi * GRKB_ENC_MAX_PASSES;
		uint32_t inc = 0;
		if(thresh == 0)
			inc = total;
		else
			for(uint32_t q = 0; q < total; ++q)
			{
				uint32_t dr;
				double dd;
				if(inc == 0)
				{
					dr = ps[q].rate;
					dd = ps[q].distortion;
				}
				else
				{
					dr = ps[q].rate - ps[inc - 1].rate;
					dd = ps[q].distortion - ps[inc - 1].distortion;
				}
				if(!dr)
				{
					if(dd != 0)
						inc = q + 1;
					continue;
				}
				const double slope = dd / dr;
				if(thresh - slope < 2.2204460492503131e-16) // DBL_EPSILON
					inc = q + 1;
			}
		k->npass[bi] = inc;
	}
}

// makeLayerSimple for layer `layno` of several: passes beyond those of the layers already final
// (prev_np); with `final_attempt` the layer becomes final.  thresh 0 and makeLayerFinal: all that is left.
void make_layer_n(TileKit* k, uint32_t layno, double thresh, bool final_attempt)
{
	std::vector<uint32_t>& out = k->layer_np[layno];
	for(size_t bi = 0; bi < k->blocks.size(); ++bi)
	{
		const uint32_t total = k->coded[bi].numpasses, prev = k->prev_np[bi];
		const grkb_pass* ps = k->passes + bi * GRKB_ENC_MAX_PASSES;
		uint32_t inc = prev;
		if(thresh == 0)
			inc = total;
		else
			for(uint32_t q = prev; q < total; ++q)
			{
				uint32_t dr;
				double dd;
				if(inc == 0)
				{
					dr = ps[q].rate;
					dd = ps[q].distortion;
				}
				else
				{
					dr = ps[q].rate - ps[inc - 1].rate;
					dd = ps[q].distortion - ps[inc - 1].distortion;
				}
				if(!dr)
				{
					if(dd != 0)
						inc = q + 1;
					continue;
				}
				const double slope = dd / dr;
				if(thresh - slope < 2.2204460492503131e-16) // DBL_EPSILON
					inc = q + 1;
			}
		if(inc < prev) // (thresh 0 with nothing left)
			inc = prev;
		out[bi] = inc - prev;
		if(final_attempt)
			k->prev_np[bi] = inc;
	}
}

// pcrdBisectSimple over several layers (tile/TileProcessor.cpp:1277-1357): the layers are allocated
// in turn against the bytes of all layers so far; a layer's search for the slope threshold starts
// below the previous layer's lower bound (upperBound = lowerBound - 1); a layer without a rate takes
// every pass that is left (makeLayerFinal)
void rate_allocate_layers(grkb_compressor* z, TileKit* k, const std::vector<double>& layer_bytes)
{
	const uint32_t numlayers = (uint32_t)layer_bytes.size();
	double min_slope = 1.7976931348623157e308, max_slope = -1;
	for(size_t bi = 0; bi < k->blocks.size(); ++bi)
	{
		const grkb_pass* ps = k->passes + bi * GRKB_ENC_MAX_PASSES;
		for(uint32_t q = 0; q < k->coded[bi].numpasses; ++q)
		{
			const int32_t dr = q ? (int32_t)(ps[q].rate - ps[q - 1].rate) : (int32_t)ps[q].rate;
			const double dd = q ? ps[q].distortion - ps[q - 1].distortion : ps[q].distortion;
			if(dr == 0)
				continue;
			const double sl = dd / dr;
			if(sl < min_slope)
				min_slope = sl;
			if(sl > max_slope)
				max_slope = sl;
		}
	}
	k->layer_np.assign(numlayers, std::vector<uint32_t>(k->blocks.size(), 0));
	k->prev_np.assign(k->blocks.size(), 0);
	k->in_packet.assign(k->blocks.size(), 0);
	k->lblock.assign(k->blocks.size(), 3);
	k->incl_trees.resize((size_t)z->p.numcomps * k->idx.pb.size());
	k->imsb_trees.resize((size_t)z->p.numcomps * k->idx.pb.size());
	double upper = max_slope;
	for(uint32_t layno = 0; layno < numlayers; ++layno)
	{
		if(!(layer_bytes[layno] > 0))
		{
			make_layer_n(k, layno, 0, true);
			continue;
		}
		const uint64_t max_len = (uint64_t)(uint32_t)ceil(layer_bytes[layno]);
		double lower = min_slope, prev = -1, thresh = 0;
		for(uint32_t i = 0; i < 128; ++i)
		{
			thresh = (upper == -1) ? lower : (lower + upper) / 2;
			make_layer_n(k, layno, thresh, false);
			if(prev != -1 && fabs(prev - thresh) < 0.001)
				break;
			prev = thresh;
			if(!simulate(z, k, max_len, nullptr, layno + 1))
			{
				lower = thresh;
				continue;
			}
			upper = thresh;
		}
		make_layer_n(k, layno, (upper == -1) ? thresh : upper, true);
		upper = lower - 1;
	}
}

// TileProcessor::pcrdBisectSimple (tile/TileProcessor.cpp:1195-1357), one layer, rate target in bytes:
// the slope threshold is bisected on the reference's own packet accounting until the layer fits.
void rate_allocate(grkb_compressor* z, TileKit* k, double rate_bytes)
{
	double min_slope = 1.7976931348623157e308, max_slope = -1;
	for(size_t bi = 0; bi < k->blocks.size(); ++bi)
	{
		const grkb_pass* ps = k->passes + bi * GRKB_ENC_MAX_PASSES;
		for(uint32_t q = 0; q < k->coded[bi].numpasses; ++q)
		{
			const int32_t dr = q ? (int32_t)(ps[q].rate - ps[q - 1].rate) : (int32_t)ps[q].rate;
			const double dd = q ? ps[q].distortion - ps[q - 1].distortion : ps[q].distortion;
			if(dr == 0)
				continue;
			const double s = dd / dr;
			if(s < min_slope)
				min_slope = s;
			if(s > max_slope)
				max_slope = s;
		}
	}
	const uint64_t max_len = rate_bytes > 0 ? (uint64_t)(uint32_t)ceil(rate_bytes) : UINT32_MAX;
	double lower = min_slope, upper = max_slope, prev = -1, thresh = 0;
	for(uint32_t i = 0; i < 128; ++i)
	{
		thresh = (upper == -1) ? lower : (lower + upper) / 2;
		make_layer(k, thresh);
		if(prev != -1 && fabs(prev - thresh) < 0.001)
			break;
		prev = thresh;
		if(!simulate(z, k, max_len, nullptr))
		{
			lower = thresh;
			continue;
		}
		upper = thresh;
	}
	make_layer(k, (upper == -1) ? thresh : upper);
}

// T2 of one tile: SOT, SOD and every packet of the single layer, appended to o
bool write_tile(grkb_compressor* z, TileKit* k, uint32_t tileno, std::vector<uint8_t>& o, std::string& err)
{
	const grkb_compress_params& p = z->p;
	const uint32_t nc = p.numcomps, numres = p.numres;
	const size_t sot = o.size();
	put16(o, 0xFF90);
	put16(o, 10);
	put16(o, tileno);
	put32(o, 0); // Psot, patched below
	o.push_back(0);
	o.push_back(1);
	put16(o, 0xFF93);
	const bool layered = !k->layer_np.empty();
	const uint32_t numlayers = layered ? (uint32_t)k->layer_np.size() : 1;
	std::vector<grkb_cs::Poc> progs;
	progs.push_back(grkb_cs::Poc{0, (uint8_t)numres, (uint8_t)p.prog_order, 0, (uint16_t)nc, (uint16_t)numlayers});
	TagTreeEnc incl, imsb;
	uint32_t packet_no = 0;
	bool ok = true;
	grkb_cs::for_each_packet(progs, numres, nc, numlayers, k->idx, k->tp.x0, k->tp.y0, k->tp.x1, k->tp.y1,
							 [&](uint32_t layno, uint32_t resno, uint32_t compno, uint64_t precinct) {
		const grkb_layout_res& R = k->idx.res[resno];
		if(p.csty & 2) // SOP
		{
			put16(o, 0xFF91);
			put16(o, 4);
			put16(o, packet_no & 0xFFFF);
		}
		if(!packet_header(z, k, resno, compno, precinct, incl, imsb, err, 0, nullptr, layno))
		{
			ok = false;
			return false;
		}
		o.insert(o.end(), z->hdr.begin(), z->hdr.end());
		if(p.csty & 4) // EPH
			put16(o, 0xFF92);
		for(uint32_t band = 0; band < R.nbands; ++band)
		{
			const grkb_layout_pb& pb = k->idx.pb[(size_t)R.first_pb + (size_t)band * R.gw * R.gh + (size_t)precinct];
			const uint32_t ncb = pb.cgw * pb.cgh;
			if(ncb == 0 || pb.band_empty)
				continue;
			for(uint32_t cb = 0; cb < ncb; ++cb)
			{
				const int32_t ti = k->table_of[pb.first + cb];
				if(ti < 0)
					continue;
				const size_t bi = (size_t)compno * k->per_comp + (uint32_t)ti;
				if(layered)
				{
					const uint32_t np = k->layer_np[layno][bi];
					if(!np)
						continue;
					uint32_t b0, b1;
					pass_span(k, bi, k->in_packet[bi], np, &b0, &b1);
					o.insert(o.end(), k->bytes + k->coded[bi].data_off + b0, k->bytes + k->coded[bi].data_off + b1);
					k->in_packet[bi] += np;
					continue;
				}
				const uint32_t len = layer_len(k, bi);
				if(len)
					o.insert(o.end(), k->bytes + k->coded[bi].data_off, k->bytes + k->coded[bi].data_off + len);
			}
		}
		++packet_no;
		return true;
	});
	if(!ok)
		return false;
	const uint64_t psot = o.size() - sot;
	if(psot > 0xFFFFFFFFull)
	{
		err = "tile part longer than 4 GB";
		return false;
	}
	o[sot + 6] = (uint8_t)(psot >> 24);
	o[sot + 7] = (uint8_t)(psot >> 16);
	o[sot + 8] = (uint8_t)(psot >> 8);
	o[sot + 9] = (uint8_t)psot;
	return true;
}
} // namespace

extern "C" {

int grkb_compressor_create(int device, const grkb_compress_params* p, grkb_compressor** out)
{
	if(!p || !out || p->width == 0 || p->height == 0 || p->numcomps == 0 || p->numcomps > GRKB_MAX_COMPS ||
	   p->prec == 0 || p->prec > 16 || p->numres == 0 || p->numres > 33 || p->cblkw_exp < 2 || p->cblkh_exp < 2 ||
	   p->cblkw_exp > 10 || p->cblkh_exp > 10 || p->cblkw_exp + p->cblkh_exp > 12 || p->prog_order > 4 ||
	   p->tile_x0 > p->x0 || p->tile_y0 > p->y0)
	{
		grkb_set_error("grkb_compressor_create: bad parameters");
		return -1;
	}
	const bool ht = (p->cblk_sty & GRKB_CBLKSTY_HT) != 0;
	if(ht && (p->cblk_sty != GRKB_CBLKSTY_HT || !p->reversible))
	{
		grkb_set_error("grkb_compressor_create: HT is written with the reversible transform and no other mode switch");
		return -1;
	}
	if(p->mct && p->numcomps < 3)
	{
		grkb_set_error("grkb_compressor_create: the component transform needs three components");
		return -1;
	}
	if(p->numlayers > GRKB_MAX_LAYERS || (p->numlayers > 1 && ht))
	{
		grkb_set_error("grkb_compressor_create: %u layers (at most %u; HT blocks are written in one)", p->numlayers,
					   (unsigned)GRKB_MAX_LAYERS);
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(device));
	auto z = std::make_unique<grkb_compressor>();
	z->device = device;
	z->p = *p;
	if(z->p.numgbits == 0)
		z->p.numgbits = ht ? 1 : 2; // grk_compress defaults (src/bin/jp2/grk_compress.cpp:1123-1124)
	z->comment = p->comment[0] ? std::string(p->comment) : std::string("Created by grok_b200");
	const uint32_t nb = 3 * p->numres - 2;
	for(uint32_t c = 0; c < p->numcomps; ++c)
	{
		z->expn[c].resize(nb);
		z->mant[c].resize(nb);
		if(grkb_qcd_generate(p->numres, p->prec, p->reversible, ht, z->expn[c].data(), z->mant[c].data()))
			return -1;
	}
	*out = z.release();
	return 0;
}

void grkb_compressor_destroy(grkb_compressor* z)
{
	if(!z)
		return;
	cudaSetDevice(z->device);
	delete z;
}

int grkb_compress(grkb_compressor* z, const int32_t* const* planes, uint32_t stride, uint8_t* out, size_t cap,
				  size_t* used)
{
	if(!z || !planes || !out || !used || stride < z->p.width)
	{
		grkb_set_error("grkb_compress: bad arguments");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(z->device));
	const grkb_compress_params& p = z->p;
	const uint32_t ix0 = p.x0, iy0 = p.y0, ix1 = p.x0 + p.width, iy1 = p.y0 + p.height;
	const uint32_t tw = p.tile_w ? p.tile_w : ix1 - p.tile_x0, th = p.tile_h ? p.tile_h : iy1 - p.tile_y0;
	const uint32_t ntx = (uint32_t)(((uint64_t)ix1 - p.tile_x0 + tw - 1) / tw);
	const uint32_t nty = (uint32_t)(((uint64_t)iy1 - p.tile_y0 + th - 1) / th);
	if((uint64_t)ntx * nty > 65535)
	{
		grkb_set_error("grkb_compress: more than 65535 tiles");
		return -1;
	}
	std::vector<uint8_t> o;
	o.reserve((size_t)p.width * p.height * p.numcomps + 65536);
	write_main_header(z, o);
	const size_t header_size = o.size();
	z->t2_ms = z->device_ms = 0;
	z->tiles_dev = z->tiles_host = 0;
	for(uint32_t t = 0; t < ntx * nty; ++t)
	{
		const uint32_t tp = t % ntx, tq = t / ntx;
		grkb_tile_params g{};
		g.x0 = (uint32_t)std::max<uint64_t>((uint64_t)p.tile_x0 + (uint64_t)tp * tw, ix0);
		g.y0 = (uint32_t)std::max<uint64_t>((uint64_t)p.tile_y0 + (uint64_t)tq * th, iy0);
		g.x1 = (uint32_t)std::min<uint64_t>((uint64_t)p.tile_x0 + (uint64_t)(tp + 1) * tw, ix1);
		g.y1 = (uint32_t)std::min<uint64_t>((uint64_t)p.tile_y0 + (uint64_t)(tq + 1) * th, iy1);
		g.numcomps = p.numcomps;
		g.numres = p.numres;
		g.reversible = p.reversible ? 1 : 0;
		g.mct = p.mct ? 1 : 0;
		for(uint32_t c = 0; c < p.numcomps; ++c)
			g.shift[c] = p.sgnd ? 0 : 1 << (p.prec - 1);
		TileKit* k = kit_for(z, g);
		if(!k)
		{
			if(!grkb_last_error()[0])
				grkb_set_error("grkb_compress: tile %u: cannot set up the encoder", t);
			return -1;
		}
		const int32_t* src[GRKB_MAX_COMPS] = {};
		for(uint32_t c = 0; c < p.numcomps; ++c)
			src[c] = planes[c] + (size_t)(g.y0 - iy0) * stride + (g.x0 - ix0);
		const auto t0 = std::chrono::steady_clock::now();
		size_t nbytes = 0;
		static const double norms_rev[3] = {1.732, .8292, .8292}, norms_irrev[3] = {1.732, 1.805, 1.573};
		const uint32_t numlayers = p.numlayers > 1 ? p.numlayers : 1;
		bool rc = p.rate > 0;
		if(numlayers > 1)
		{
			rc = false;
			for(uint32_t l = 0; l < numlayers; ++l)
				rc = rc || p.layer_rates[l] > 0;
		}
		const double* norms = (rc && p.mct) ? (p.reversible ? norms_rev : norms_irrev) : nullptr;
		// the tile's byte budget as CodeStreamCompress::updateRates leaves it (:950-1018): its share of
		// the target size minus its share of the main header, at least 30 bytes
		double budget = 0;
		std::vector<double> layer_bytes(numlayers, 0.0);
		if(rc)
		{
			const double pixels = (double)(g.x1 - g.x0) * (double)(g.y1 - g.y0);
			const double sot_adjust = (pixels * (double)header_size) / ((double)p.width * (double)p.height);
			for(uint32_t l = 0; l < numlayers; ++l)
			{
				const double r = numlayers > 1 ? p.layer_rates[l] : p.rate;
				if(r > 0)
					layer_bytes[l] = ((double)(p.numcomps * p.prec) * pixels) / (r * 8.0);
			}
			// first layer, the layers between, the last of several (:998-1021)
			if(layer_bytes[0] > 0)
			{
				layer_bytes[0] -= sot_adjust;
				if(layer_bytes[0] < 30.0)
					layer_bytes[0] = 30.0;
			}
			for(uint32_t l = 1; l < numlayers; ++l)
				if(layer_bytes[l] > 0)
				{
					layer_bytes[l] -= l + 1 == numlayers ? sot_adjust + 2.0 : sot_adjust;
					if(layer_bytes[l] < layer_bytes[l - 1] + 10.0)
						layer_bytes[l] = layer_bytes[l - 1] + 20.0;
				}
			budget = layer_bytes[0];
		}
		// the allocation runs on the device right behind the block coders (pcrd.cu); GRKB_PCRD=host
		// keeps it on the host (rate_allocate below: same algorithm, same result)
		static const bool pcrd_host = [] {
			const char* e = getenv("GRKB_PCRD");
			return e && e[0] == 'h';
		}();
		const bool on_device = rc && numlayers == 1 && !pcrd_host && k->pcrd;
		grkb_encoder_set_pcrd(k->enc, on_device ? k->pcrd : nullptr, (uint64_t)(uint32_t)ceil(budget), p.csty, k->npass_dev,
							  k->pcrd_result);
		if(grkb_encoder_encode(k->enc, src, stride, k->blocks.data(), (uint32_t)k->blocks.size(), norms, norms ? 3 : 0,
							   rc ? 1 : 0, k->coded, k->passes, k->bytes, k->bytes_cap, &nbytes))
			return -1;
		for(size_t i = 0; i < k->blocks.size(); ++i)
			if(k->coded[i].failed)
			{
				grkb_set_error("grkb_compress: tile %u: block %zu overflowed its stream buffer", t, i);
				return -1;
			}
		const auto t1 = std::chrono::steady_clock::now();
		k->npass.resize(k->blocks.size());
		for(size_t i = 0; i < k->blocks.size(); ++i)
			k->npass[i] = k->coded[i].numpasses;
		k->layer_np.clear();
		if(numlayers > 1)
		{
			rate_allocate_layers(z, k, layer_bytes);
			++z->tiles_host;
		}
		else if(on_device && k->pcrd_result->error == 0)
		{
			for(size_t i = 0; i < k->blocks.size(); ++i)
				k->npass[i] = k->npass_dev[i];
			z->pcrd_iterations = k->pcrd_result->iterations;
			++z->tiles_dev;
		}
		else if(rc)
		{
			rate_allocate(z, k, budget);
			++z->tiles_host;
		}
		std::string err;
		if(!write_tile(z, k, t, o, err))
		{
			grkb_set_error("grkb_compress: tile %u: %s", t, err.c_str());
			return -1;
		}
		const auto t2 = std::chrono::steady_clock::now();
		z->device_ms += std::chrono::duration<double, std::milli>(t1 - t0).count();
		z->t2_ms += std::chrono::duration<double, std::milli>(t2 - t1).count();
	}
	put16(o, 0xFFD9);
	*used = o.size();
	if(o.size() > cap)
	{
		grkb_set_error("grkb_compress: the codestream needs %zu bytes, the buffer holds %zu", o.size(), cap);
		return -2;
	}
	memcpy(out, o.data(), o.size());
	return 0;
}

int grkb_compress_allocation(const grkb_compressor* z, uint32_t* tiles_on_device, uint32_t* tiles_on_host,
							 uint32_t* steps)
{
	if(!z)
		return -1;
	if(tiles_on_device)
		*tiles_on_device = z->tiles_dev;
	if(tiles_on_host)
		*tiles_on_host = z->tiles_host;
	if(steps)
		*steps = z->pcrd_iterations;
	return 0;
}

int grkb_compress_stats(const grkb_compressor* z, double* device_ms, double* t2_ms)
{
	if(!z)
		return -1;
	if(device_ms)
		*device_ms = z->device_ms;
	if(t2_ms)
		*t2_ms = z->t2_ms;
	return 0;
}

} // extern "C"
