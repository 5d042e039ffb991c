// decompress.cu -- codestream-level C-ABI (include/grk_b200.h, section E): the codestream index and
// per-tile T2 as exported calls, and the whole-image engine that feeds the tiles of an image
// through a ring of tile decoders and composites their rectangles into the caller's planes.
// The device side of grk_decompress (src/lib/jp2/grok.cpp:560-640 ->
// CodeStreamDecompress::decompressTiles, codestream/CodeStreamDecompress.cpp:412-640).
#include "pipeline.h"
#include "codestream.h"
#include <algorithm>
#include <chrono>
#include <memory>
#include <mutex>
#include <string.h>

struct grkb_codestream
{
	grkb_cs::Codestream cs;
	std::vector<std::unique_ptr<grkb_cs::TileTables>> tables; // one slot per tile, filled on demand
};

extern "C" {

int grkb_codestream_open(const uint8_t* data, size_t len, grkb_codestream** out)
{
	if(!data || !out)
	{
		grkb_set_error("grkb_codestream_open: null argument");
		return -1;
	}
	auto c = std::make_unique<grkb_codestream>();
	std::string err;
	if(!grkb_cs::open(data, len, c->cs, err))
	{
		grkb_set_error("grkb_codestream_open: %s", err.c_str());
		return -1;
	}
	c->tables.resize(c->cs.tiles.size());
	*out = c.release();
	return 0;
}

void grkb_codestream_close(grkb_codestream* cs)
{
	delete cs;
}

int grkb_codestream_get_info(const grkb_codestream* c, grkb_codestream_info* info)
{
	if(!c || !info)
	{
		grkb_set_error("grkb_codestream_get_info: null argument");
		return -1;
	}
	const grkb_cs::Codestream& cs = c->cs;
	memset(info, 0, sizeof(*info));
	info->x0 = cs.x0;
	info->y0 = cs.y0;
	info->x1 = cs.x1;
	info->y1 = cs.y1;
	info->tile_x0 = cs.tx0;
	info->tile_y0 = cs.ty0;
	info->tile_w = cs.tw;
	info->tile_h = cs.th;
	info->tiles_x = cs.ntx;
	info->tiles_y = cs.nty;
	info->numcomps = cs.numcomps;
	for(uint32_t i = 0; i < cs.numcomps; ++i)
	{
		info->prec[i] = cs.comps[i].prec;
		info->sgnd[i] = cs.comps[i].sgnd;
	}
	const grkb_cs::Coding& cod = cs.main.cod[0];
	info->numres = cod.numres;
	info->numlayers = cs.main.numlayers;
	info->prog_order = cs.main.prog;
	info->reversible = cod.reversible;
	info->mct = cs.main.mct;
	info->cblkw_exp = cod.cblkw_exp;
	info->cblkh_exp = cod.cblkh_exp;
	info->cblk_sty = cod.cblk_sty;
	info->csty = cs.main.csty;
	info->ht = (cod.cblk_sty & GRKB_CBLKSTY_HT) ? 1 : 0;
	info->truncated = cs.truncated ? 1 : 0;
	info->rsiz = cs.rsiz;
	for(const auto& t : cs.tiles)
		info->tile_parts += (uint32_t)t.parts.size();
	auto scan = [&](const grkb_cs::TileHeader& h, bool tile) {
		for(uint32_t i = 0; i < cs.numcomps; ++i)
		{
			if(h.roishift[i])
				info->features |= GRKB_CS_RGN;
			if(h.q[i].level == 2 || h.q[i].level == 4)
				info->features |= GRKB_CS_QCC;
			if(h.cod[i].level == 2 || h.cod[i].level == 4)
				info->features |= GRKB_CS_COC;
			if(tile && (h.q[i].level >= 3 || h.cod[i].level >= 3))
				info->features |= GRKB_CS_TILE_OVERRIDES;
			if(cs.comps[i].prec != cs.comps[0].prec || cs.comps[i].sgnd != cs.comps[0].sgnd)
				info->features |= GRKB_CS_MIXED_PRECISION;
		}
		if(!h.pocs.empty())
			info->features |= GRKB_CS_POC;
	};
	scan(cs.main, false);
	for(const auto& t : cs.tiles)
		if(t.seen)
			scan(t.hdr, true);
	return 0;
}

int grkb_codestream_parse_tile(grkb_codestream* c, uint32_t tileno, uint32_t reduce, uint32_t max_layers,
							   grkb_tile_tables* out)
{
	if(!c || !out || tileno >= c->tables.size())
	{
		grkb_set_error("grkb_codestream_parse_tile: bad arguments");
		return -1;
	}
	auto t = std::make_unique<grkb_cs::TileTables>();
	std::string err;
	if(!grkb_cs::parse_tile(c->cs, tileno, reduce, max_layers, *t, err))
	{
		grkb_set_error("grkb_codestream_parse_tile: tile %u: %s", tileno, err.c_str());
		return -1;
	}
	memset(out, 0, sizeof(*out));
	out->params = t->params;
	out->blocks = t->blocks.data();
	out->nblocks = (uint32_t)t->blocks.size();
	out->segs = t->segs.data();
	out->nsegs = (uint32_t)t->segs.size();
	if(t->zero_copy)
	{
		out->bytes = c->cs.data + t->src_off;
		out->nbytes = t->src_len;
	}
	else
	{
		out->bytes = t->pool.data();
		out->nbytes = t->pool.size();
	}
	out->zero_copy = t->zero_copy ? 1 : 0;
	out->packets = t->packets;
	out->truncated = t->truncated ? 1 : 0;
	c->tables[tileno] = std::move(t);
	return 0;
}

} // extern "C"

// ------------------------------------------------------------------------------------------------
// whole image

namespace
{
struct Slot
{
	std::vector<std::pair<grkb_tile_params, grkb_decoder*>> decs; // cached per (canonical) geometry
	grkb_decoder* busy = nullptr;
	grkb_cs::TileTables tables; // must outlive the submit that reads them
};

// A decoder depends on the tile origin only through the parities and sizes of its band
// rectangles; both survive a shift of the rectangle by a multiple of 2^(numres-1).  Tiles of a
// regular grid therefore share one decoder.
grkb_tile_params canonical(grkb_tile_params p)
{
	const uint32_t m = p.numres > 1 ? (p.numres - 1 >= 31 ? 0x7FFFFFFFu : (1u << (p.numres - 1)) - 1u) : 0u;
	const uint32_t w = p.x1 - p.x0, h = p.y1 - p.y0;
	p.x0 &= m;
	p.y0 &= m;
	p.x1 = p.x0 + w;
	p.y1 = p.y0 + h;
	return p;
}
} // namespace

struct grkb_decompressor
{
	int device = 0;
	std::vector<Slot> slots;
	uint32_t rejected = 0;
	bool failed = false;
	double t2_ms = 0;
	uint32_t tiles = 0;
	~grkb_decompressor()
	{
		for(auto& s : slots)
			for(auto& d : s.decs)
				grkb_decoder_destroy(d.second);
	}
};

namespace
{
int drain(grkb_decompressor* d, Slot& s)
{
	if(!s.busy)
		return 0;
	uint32_t rej = 0;
	const int rc = grkb_decoder_wait(s.busy, &rej);
	s.busy = nullptr;
	d->rejected += rej;
	if(rc)
		d->failed = true;
	return rc;
}

// win: x0, y0, x1, y1 relative to the image origin on the reference grid (grk_decompress_set_window),
// or null for the whole image
int submit_any(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce, uint32_t max_layers,
			   int32_t* const* planes, uint32_t stride, void* out, uint32_t format, size_t row_bytes,
			   const uint32_t* win = nullptr, uint32_t first_tile = 0, uint32_t tile_step = 1)
{
	if(!d || !data || (!planes && !out))
	{
		grkb_set_error("grkb_decompress_submit: null argument");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(d->device));
	for(auto& s : d->slots)
		drain(d, s);
	d->rejected = 0;
	d->failed = false;
	d->tiles = 0;
	const auto t0 = std::chrono::steady_clock::now();
	double t2 = 0;
	grkb_cs::Codestream cs;
	std::string err;
	if(!grkb_cs::open(data, len, cs, err))
	{
		grkb_set_error("grkb_decompress: %s", err.c_str());
		return -1;
	}
	t2 += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
	auto cdp2 = [](uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a + ((uint64_t)1 << b) - 1) >> b); };
	const uint32_t ix0 = cdp2(cs.x0, reduce), iy0 = cdp2(cs.y0, reduce);
	// the reference sizes the reduced image as ceil(width / 2^reduce) (codestream/markers/
	// SIZMarker.cpp:54-57), which can be one sample more than the ceil(x1 / 2^r) - ceil(x0 / 2^r) the
	// tiles cover; that last row / column is not written (the reference leaves zeros there)
	uint32_t IW = cdp2(cs.x1 - cs.x0, reduce), IH = cdp2(cs.y1 - cs.y0, reduce);
	const uint32_t nc = cs.numcomps;
	// the window on the reference grid, clipped to the image (CodeStreamDecompress::setDecompressWindow,
	// codestream/CodeStreamDecompress.cpp:295-398), and in reduced coordinates: the output image is
	// then [ceil(x0 / 2^r), ceil(x1 / 2^r)) as subsampleAndReduce sizes it (util/GrkImage.cpp:74-115)
	uint32_t wx0 = cs.x0, wy0 = cs.y0, wx1 = cs.x1, wy1 = cs.y1;
	uint32_t ox0 = ix0, oy0 = iy0; // origin of the output in reduced coordinates
	if(win && !(win[0] == 0 && win[1] == 0 && win[2] == 0 && win[3] == 0))
	{
		if(out)
		{
			grkb_set_error("grkb_decompress: a window is served as int32 planes only");
			return -1;
		}
		wx0 = (uint32_t)std::min<uint64_t>((uint64_t)cs.x0 + win[0], cs.x1);
		wy0 = (uint32_t)std::min<uint64_t>((uint64_t)cs.y0 + win[1], cs.y1);
		wx1 = (uint32_t)std::min<uint64_t>((uint64_t)cs.x0 + win[2], cs.x1);
		wy1 = (uint32_t)std::min<uint64_t>((uint64_t)cs.y0 + win[3], cs.y1);
		if(wx1 <= wx0 || wy1 <= wy0)
		{
			grkb_set_error("grkb_decompress: empty window");
			return -1;
		}
		ox0 = cdp2(wx0, reduce);
		oy0 = cdp2(wy0, reduce);
		IW = cdp2(wx1, reduce) - ox0;
		IH = cdp2(wy1, reduce) - oy0;
		if(IW == 0 || IH == 0)
		{
			grkb_set_error("grkb_decompress: the window holds no sample at this resolution");
			return -1;
		}
	}
	else
		win = nullptr;
	const uint32_t rwx1 = cdp2(wx1, reduce), rwy1 = cdp2(wy1, reduce); // window end, reduced coordinates
	size_t bps = 0, plane_bytes = 0;
	bool interleaved = false;
	if(out)
	{
		const uint32_t base = format & ~(uint32_t)GRKB_SAMPLES_BIG_ENDIAN;
		interleaved = base == GRKB_SAMPLES_16_INTERLEAVED || base == GRKB_SAMPLES_8_INTERLEAVED;
		bps = (base == GRKB_SAMPLES_16 || base == GRKB_SAMPLES_16_INTERLEAVED) ? 2 : 1;
		if(grkb_samples_min_row_bytes(format, IW, nc) == 0 || row_bytes < grkb_samples_min_row_bytes(format, IW, nc))
		{
			grkb_set_error("grkb_decompress_submit_packed: bad format %u or rows of %zu bytes", format, row_bytes);
			return -1;
		}
		plane_bytes = (size_t)IH * row_bytes;
	}
	else if(stride < IW)
	{
		grkb_set_error("grkb_decompress_submit: stride %u smaller than the image width %u", stride, IW);
		return -1;
	}
	uint32_t next = 0;
	for(uint32_t t = 0; t < cs.tiles.size(); ++t)
	{
		if(cs.tiles[t].parts.empty())
			continue; // no data for this tile: its rectangle is left as the caller gave it
		if(t < first_tile || (t - first_tile) % tile_step != 0)
			continue; // another rank's tile (tile t -> GPU t mod G)
		Slot& s = d->slots[next];
		next = (next + 1) % (uint32_t)d->slots.size();
		if(drain(d, s))
			return -1;
		uint32_t tw[4] = {ox0, oy0, rwx1, rwy1};
		if(win)
		{
			// tiles the window does not touch are not parsed at all
			const uint32_t tp = t % cs.ntx, tq = t / cs.ntx;
			const uint64_t tx0 = std::max<uint64_t>((uint64_t)cs.tx0 + (uint64_t)tp * cs.tw, cs.x0);
			const uint64_t ty0 = std::max<uint64_t>((uint64_t)cs.ty0 + (uint64_t)tq * cs.th, cs.y0);
			const uint64_t tx1 = std::min<uint64_t>((uint64_t)cs.tx0 + (uint64_t)(tp + 1) * cs.tw, cs.x1);
			const uint64_t ty1 = std::min<uint64_t>((uint64_t)cs.ty0 + (uint64_t)(tq + 1) * cs.th, cs.y1);
			if(tx1 <= wx0 || tx0 >= wx1 || ty1 <= wy0 || ty0 >= wy1)
				continue;
		}
		const auto t1 = std::chrono::steady_clock::now();
		if(!grkb_cs::parse_tile(cs, t, reduce, max_layers, s.tables, err, win ? tw : nullptr))
		{
			grkb_set_error("grkb_decompress: tile %u: %s", t, err.c_str());
			return -1;
		}
		t2 += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count();
		const grkb_tile_params& P = s.tables.params;
		if(P.x1 <= P.x0 || P.y1 <= P.y0)
			continue;
		const grkb_tile_params key = canonical(P);
		grkb_decoder* dec = nullptr;
		for(auto& k : s.decs)
			if(memcmp(&k.first, &key, sizeof(key)) == 0)
				dec = k.second;
		if(!dec)
		{
			if(s.decs.size() >= 6)
			{
				grkb_decoder_destroy(s.decs.front().second);
				s.decs.erase(s.decs.begin());
			}
			const size_t nb = s.tables.zero_copy ? s.tables.src_len : s.tables.pool.size();
			if(grkb_decoder_create(&key, (uint32_t)s.tables.blocks.size(), nb, &dec))
				return -1;
			s.decs.emplace_back(key, dec);
		}
		const uint8_t* bytes = s.tables.zero_copy ? data + s.tables.src_off : s.tables.pool.data();
		const size_t nbytes = s.tables.zero_copy ? s.tables.src_len : s.tables.pool.size();
		// the part of the tile that lands in the output
		const uint32_t cx0 = std::max(P.x0, ox0), cy0 = std::max(P.y0, oy0);
		const uint32_t cx1 = std::min(P.x1, win ? rwx1 : P.x1), cy1 = std::min(P.y1, win ? rwy1 : P.y1);
		if(cx1 <= cx0 || cy1 <= cy0)
			continue;
		const size_t ox = cx0 - ox0, oy = cy0 - oy0;
		int rc;
		if(out)
		{
			uint8_t* dst = (uint8_t*)out + oy * row_bytes + ox * bps * (interleaved ? nc : 1);
			rc = grkb_decoder_submit_packed_region(dec, s.tables.blocks.data(), (uint32_t)s.tables.blocks.size(),
												   s.tables.segs.data(), (uint32_t)s.tables.segs.size(), bytes, nbytes,
												   dst, format, row_bytes, plane_bytes);
		}
		else
		{
			int32_t* dst[GRKB_MAX_COMPS] = {};
			for(uint32_t c = 0; c < nc; ++c)
				dst[c] = planes[c] + oy * stride + ox;
			rc = grkb_decoder_submit_region(dec, s.tables.blocks.data(), (uint32_t)s.tables.blocks.size(),
											s.tables.segs.data(), (uint32_t)s.tables.segs.size(), bytes, nbytes, dst,
											stride, cx0 - P.x0, cy0 - P.y0, cx1 - cx0, cy1 - cy0);
		}
		if(rc)
			return -1;
		s.busy = dec;
		++d->tiles;
	}
	d->t2_ms = t2;
	return 0;
}
} // namespace

extern "C" {

int grkb_decompressor_create(int device, uint32_t slots, grkb_decompressor** out)
{
	if(!out || slots == 0 || slots > 64)
	{
		grkb_set_error("grkb_decompressor_create: bad arguments");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(device));
	auto d = new grkb_decompressor();
	d->device = device;
	d->slots.resize(slots);
	*out = d;
	return 0;
}

void grkb_decompressor_destroy(grkb_decompressor* d)
{
	if(!d)
		return;
	cudaSetDevice(d->device);
	for(auto& s : d->slots)
		drain(d, s);
	delete d;
}

int grkb_decompress_submit(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
						   uint32_t max_layers, int32_t* const* planes, uint32_t stride)
{
	return submit_any(d, data, len, reduce, max_layers, planes, stride, nullptr, 0, 0);
}

int grkb_decompress_submit_window(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
								  uint32_t max_layers, uint32_t wx0, uint32_t wy0, uint32_t wx1, uint32_t wy1,
								  int32_t* const* planes, uint32_t stride)
{
	const uint32_t win[4] = {wx0, wy0, wx1, wy1};
	return submit_any(d, data, len, reduce, max_layers, planes, stride, nullptr, 0, 0, win);
}

int grkb_decompress_submit_tiles(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
								 uint32_t max_layers, uint32_t first_tile, uint32_t tile_step, int32_t* const* planes,
								 uint32_t stride)
{
	if(tile_step == 0)
	{
		grkb_set_error("grkb_decompress_submit_tiles: tile_step must be at least 1");
		return -1;
	}
	return submit_any(d, data, len, reduce, max_layers, planes, stride, nullptr, 0, 0, nullptr, first_tile, tile_step);
}

int grkb_decompress_submit_packed(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
								  uint32_t max_layers, void* out, uint32_t format, size_t row_bytes)
{
	return submit_any(d, data, len, reduce, max_layers, nullptr, 0, out, format, row_bytes);
}

int grkb_decompress_wait(grkb_decompressor* d, uint32_t* rejected_out)
{
	if(!d)
	{
		grkb_set_error("grkb_decompress_wait: null decompressor");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(d->device));
	int rc = 0;
	for(auto& s : d->slots)
		if(drain(d, s))
			rc = -1;
	if(rejected_out)
		*rejected_out = d->rejected;
	return (rc || d->failed) ? -1 : 0;
}

int grkb_decompress(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
					uint32_t max_layers, int32_t* const* planes, uint32_t stride, uint32_t* rejected_out)
{
	if(grkb_decompress_submit(d, data, len, reduce, max_layers, planes, stride))
	{
		if(d)
			for(auto& s : d->slots)
				drain(d, s);
		return -1;
	}
	return grkb_decompress_wait(d, rejected_out);
}

int grkb_decompress_stats(const grkb_decompressor* d, double* t2_ms, uint32_t* tiles)
{
	if(!d)
		return -1;
	if(t2_ms)
		*t2_ms = d->t2_ms;
	if(tiles)
		*tiles = d->tiles;
	return 0;
}

} // extern "C"
