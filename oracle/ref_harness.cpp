/*
 * ref_harness.cpp -- thin extern "C" doors into the UNMODIFIED reference library.
 *
 * TEST INFRASTRUCTURE ONLY.  Compiled by oracle/Makefile together with the reference's own
 * sources (read in place from /root/reference) into oracle/_ref/libgrok_ref.so.  It contains
 * no codec arithmetic of its own: every function builds the reference's own objects
 * (TileComponent, T1, ...) and calls the reference's own entry points, so that the C
 * restatement in oracle/*.c and the CUDA kernels can be compared with what Grok computes.
 */
/* ojph headers first: grk_includes.h poisons malloc/free (same order as t1/t1_ht/T1HT.cpp:20-25) */
#include "ojph_block_decoder.h"
#include "ojph_block_encoder.h"
#include "ojph_mem.h"
#include "grk_includes.h"
#include "T1.h"
#include "WaveletFwd.h"
#include "mct.h"
#include "PostDecompressFilters.h"
#include <atomic>
#include <chrono>
#include <thread>

/* zeroed raw storage (the reference poisons malloc/calloc/free) */
template<typename T>
static T* zalloc(size_t n)
{
	void* p = ::operator new(n * sizeof(T));
	memset(p, 0, n * sizeof(T));
	return (T*)p;
}
static void zfree(void* p)
{
	::operator delete(p);
}

using namespace grk;

namespace
{
struct RefTileComp
{
	CodingParams cp;
	TileCodingParams tcp;
	TileComponentCodingParams tccp;
	TileComponent tilec;
	bool ok = false;

	RefTileComp(bool compress, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t numres,
				uint32_t qmfbid, uint32_t cblkw_exp = 6, uint32_t cblkh_exp = 6, uint8_t prec = 8)
	{
		memset(&cp.m_coding_params, 0, sizeof(cp.m_coding_params));
		tccp.numresolutions = (uint8_t)numres;
		tccp.cblkw = (uint8_t)cblkw_exp;
		tccp.cblkh = (uint8_t)cblkh_exp;
		tccp.qmfbid = (uint8_t)qmfbid;
		tccp.numgbits = 2;
		tccp.roishift = 0;
		for(uint32_t i = 0; i < GRK_J2K_MAXRLVLS; ++i)
		{
			tccp.precinctWidthExp[i] = 15;
			tccp.precinctHeightExp[i] = 15;
		}
		for(uint32_t i = 0; i < GRK_J2K_MAXBANDS; ++i)
		{
			tccp.stepsizes[i].expn = 10;
			tccp.stepsizes[i].mant = 0;
		}
		grkRectU32 r(x0, y0, x1, y1);
		if(!tilec.init(compress, true, r, prec, &cp, &tcp, &tccp, nullptr))
			return;
		if(!tilec.allocWindowBuffer(r))
			return;
		if(!tilec.getBuffer()->alloc())
			return;
		ok = true;
	}
	int32_t* data()
	{
		return tilec.getBuffer()->getHighestBufferResWindowREL()->getBuffer();
	}
	uint32_t stride()
	{
		return tilec.getBuffer()->getHighestBufferResWindowREL()->stride;
	}
};
} // namespace

extern "C" {

int ref_harness_version(void)
{
	return 2;
}

void ref_set_threads(uint32_t n)
{
	ThreadPool::release();
	ThreadPool::instance(n);
}

/* row stride the reference gives a tile-component buffer of this width */
uint32_t ref_stride(uint32_t width)
{
	return grkMakeAlignedWidth(width);
}

/* Whole-tile wavelet transform through the reference's own WaveletFwdImpl::compress /
 * WaveletReverse::decompress on a real TileComponent.  data: dense (x1-x0) x (y1-y0) plane,
 * transformed in place (int32 for 5/3, float bits for 9/7). */
int ref_dwt(int32_t* data, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t numres,
			uint32_t qmfbid, int forward)
{
	ThreadPool::get();
	RefTileComp tc(forward != 0, x0, y0, x1, y1, numres, qmfbid);
	if(!tc.ok)
		return -1;
	uint32_t w = x1 - x0, h = y1 - y0, s = tc.stride();
	auto buf = tc.data();
	for(uint32_t y = 0; y < h; ++y)
		memcpy(buf + (size_t)y * s, data + (size_t)y * w, sizeof(int32_t) * w);
	bool rc;
	if(forward)
	{
		WaveletFwdImpl f;
		rc = f.compress(&tc.tilec, (uint8_t)qmfbid);
	}
	else
	{
		/* WaveletReverse::decompress only reads p_tcd->wholeTileDecompress
		 * (transform/WaveletReverse.cpp:2232-2261) */
		auto tp = zalloc<TileProcessor>(1);
		tp->wholeTileDecompress = true;
		WaveletReverse r;
		rc = r.decompress(tp, &tc.tilec, 0, grkRectU32(x0, y0, x1, y1), (uint8_t)numres,
						  (uint8_t)qmfbid);
		zfree(tp);
	}
	for(uint32_t y = 0; y < h; ++y)
		memcpy(data + (size_t)y * w, buf + (size_t)y * s, sizeof(int32_t) * w);
	return rc ? 0 : -2;
}

/* forward MCT: mct::compress_rev / compress_irrev (point_transform/mct.cpp:583-596) */
void ref_mct_fwd(int32_t* c0, int32_t* c1, int32_t* c2, uint64_t n, int irrev)
{
	ThreadPool::get();
	/* the Highway kernels use aligned loads: stage through the reference's aligned allocator */
	int32_t* src[3] = {c0, c1, c2};
	int32_t* al[3];
	for(int i = 0; i < 3; ++i)
	{
		al[i] = (int32_t*)grkAlignedMalloc((n + 64) * sizeof(int32_t));
		memcpy(al[i], src[i], n * sizeof(int32_t));
	}
	if(irrev)
		mct::compress_irrev(al[0], al[1], al[2], n);
	else
		mct::compress_rev(al[0], al[1], al[2], n);
	for(int i = 0; i < 3; ++i)
	{
		memcpy(src[i], al[i], n * sizeof(int32_t));
		grkAlignedFree(al[i]);
	}
}

/* inverse MCT fused with DC shift + clamp: mct::decompress_rev / decompress_irrev on a real
 * Tile (point_transform/mct.cpp:480-513, :544-579). Planes are dense w x h.
 * mct == 0: per-component DC shift only (mct::decompress_dc_shift_*), numcomps planes. */
int ref_mct_inv(int32_t** planes, uint32_t numcomps, uint32_t w, uint32_t h, int irrev, int mct_on,
				const uint8_t* prec, const uint8_t* sgnd, const int32_t* dcshift)
{
	ThreadPool::get();
	std::vector<RefTileComp*> tcs;
	auto tccps = new TileComponentCodingParams[numcomps];
	auto tile = new Tile();
	tile->numcomps = (uint16_t)numcomps;
	tile->comps = nullptr;
	/* Tile owns a TileComponent array; we hand it our own, built one by one */
	auto comps = zalloc<TileComponent>(numcomps);
	for(uint32_t c = 0; c < numcomps; ++c)
	{
		auto t = new RefTileComp(false, 0, 0, w, h, 1, irrev ? 0 : 1);
		if(!t->ok)
			return -1;
		tcs.push_back(t);
		uint32_t s = t->stride();
		for(uint32_t y = 0; y < h; ++y)
			memcpy(t->data() + (size_t)y * s, planes[c] + (size_t)y * w, sizeof(int32_t) * w);
		memcpy((void*)(comps + c), (void*)&t->tilec, sizeof(TileComponent)); /* shallow alias */
		tccps[c].m_dc_level_shift = dcshift[c];
		tccps[c].qmfbid = irrev ? 0 : 1;
	}
	tile->comps = comps;
	auto image = zalloc<GrkImage>(1);
	auto icomps = zalloc<grk_image_comp>(numcomps);
	for(uint32_t c = 0; c < numcomps; ++c)
	{
		icomps[c].prec = prec[c];
		icomps[c].sgnd = sgnd[c];
	}
	image->comps = icomps;
	image->numcomps = (uint16_t)numcomps;
	if(mct_on)
	{
		if(irrev)
			mct::decompress_irrev(tile, image, tccps);
		else
			mct::decompress_rev(tile, image, tccps);
	}
	else
	{
		for(uint32_t c = 0; c < numcomps; ++c)
		{
			if(irrev)
				mct::decompress_dc_shift_irrev(tile, image, tccps, c);
			else
				mct::decompress_dc_shift_rev(tile, image, tccps, c);
		}
	}
	for(uint32_t c = 0; c < numcomps; ++c)
	{
		uint32_t s = tcs[c]->stride();
		for(uint32_t y = 0; y < h; ++y)
			memcpy(planes[c] + (size_t)y * w, tcs[c]->data() + (size_t)y * s, sizeof(int32_t) * w);
	}
	tile->comps = nullptr; /* aliases: not ours to destroy through Tile */
	tile->numcomps = 0;
	delete tile;
	zfree(comps);
	for(auto t : tcs)
		delete t;
	delete[] tccps;
	zfree(icomps);
	zfree(image);
	return 0;
}

} // extern "C"

/* ------------------------------------------------------------------------------------- */
/* T1 Part-1 (EBCOT + MQ): the reference's T1 object, called the way T1Part1::compress /
 * T1Part1::decompress call it (t1/t1_part1/T1Part1.cpp:89-161).                          */

extern "C" {

struct ref_pass
{
	uint32_t rate;
	uint32_t len;
	uint32_t term;
	double distortiondec;
};

/* samples: w*h sign-magnitude (as T1Part1::preCompress leaves them). out: >= w*h*4+2 bytes.
 * returns cumulative distortion; fills numbps, numpasses, passes[<=100]. */
double ref_t1_encode(const int32_t* samples, uint32_t w, uint32_t h, uint32_t max, uint32_t orient,
					 uint32_t compno, uint32_t level, uint32_t qmfbid, double stepsize,
					 uint32_t cblksty, const double* mct_norms, uint32_t mct_numcomps,
					 int doRateControl, uint8_t* out, uint32_t* numbps, uint32_t* numpasses,
					 ref_pass* passes, uint32_t* numbytes)
{
	T1 t1(true, w, h);
	t1.alloc(w, h);
	memcpy(t1.getUncompressedData(), samples, sizeof(int32_t) * w * h);
	std::vector<uint8_t> buf((size_t)w * h * 4 + 64 + 2, 0);
	cblk_enc cb;
	memset(&cb, 0, sizeof(cb));
	cb.x0 = 0;
	cb.y0 = 0;
	cb.x1 = w;
	cb.y1 = h;
	cb.data = buf.data() + 2; /* two zero bytes before the stream (t1/Codeblock.h:156-172) */
	double d = t1.compress_cblk(&cb, max, (uint8_t)orient, (uint16_t)compno, (uint8_t)level,
								(uint8_t)qmfbid, stepsize, cblksty, mct_norms,
								(uint16_t)mct_numcomps, doRateControl != 0);
	*numbps = cb.numbps;
	*numpasses = cb.numPassesTotal;
	uint32_t total = 0;
	for(uint32_t i = 0; i < cb.numPassesTotal; ++i)
	{
		passes[i].rate = cb.passes[i].rate;
		passes[i].len = cb.passes[i].len;
		passes[i].term = cb.passes[i].term;
		passes[i].distortiondec = cb.passes[i].distortiondec;
		total = cb.passes[i].rate;
	}
	*numbytes = total;
	memcpy(out, cb.data, total);
	t1.code_block_enc_deallocate(&cb);
	return d;
}

/* out: w*h int32 as T1::decompress_cblk leaves them (x2 fixed point, before dequantisation) */
int ref_t1_decode(const uint8_t* data, const uint32_t* seg_len, const uint32_t* seg_passes,
				  uint32_t numsegs, uint32_t numbps, uint32_t w, uint32_t h, uint32_t orient,
				  uint32_t cblksty, int32_t* out)
{
	T1 t1(false, w, h);
	DecompressCodeblock cblk;
	cblk.numbps = (uint8_t)numbps;
	size_t total = 0;
	for(uint32_t i = 0; i < numsegs; ++i)
	{
		auto seg = cblk.nextSegment();
		seg->len = seg_len[i];
		seg->numpasses = seg_passes[i];
		total += seg_len[i];
	}
	memset(out, 0, sizeof(int32_t) * w * h);
	t1.attachUncompressedData(out, w, h);
	t1.allocCompressedData(total + grk_cblk_dec_compressed_data_pad_right);
	auto cd = t1.getCompressedDataBuffer();
	memcpy(cd, data, total);
	cd[total] = 0;
	cd[total + 1] = 0;
	bool rc = t1.decompress_cblk(&cblk, cd, (uint8_t)orient, cblksty);
	t1.attachUncompressedData(nullptr, w, h); /* detach: `out` is the caller's */
	return rc ? 0 : -1;
}

/* ------------------------------------------------------------------------------------- */
/* HTJ2K block coder: ojph_encode_codeblock / ojph_decode_codeblock called the way
 * T1HT::compress / T1HT::decompress call them (t1/t1_ht/T1HT.cpp:109-187).               */

/* samples: w*h uint32 sign-magnitude with the magnitude MSB-aligned as T1HT::preCompress
 * leaves it. out >= 16384+3072+16 bytes. returns coded length, or -1. */
int ref_ht_encode(const uint32_t* samples, uint32_t missing_msbs, uint32_t w, uint32_t h,
				  uint8_t* out)
{
	try
	{
		ojph::mem_elastic_allocator elastic(1048576);
		ojph::coded_lists* coded = nullptr;
		uint32_t lengths[2] = {0, 0};
		std::vector<uint32_t> buf(samples, samples + (size_t)w * h);
		ojph::local::ojph_encode_codeblock(buf.data(), missing_msbs, 1, w, h, w, lengths, &elastic,
										   coded);
		memcpy(out, coded->buf, lengths[0]);
		return (int)lengths[0];
	}
	catch(...)
	{
		return -1;
	}
}

/* data: lengths1+lengths2 coded bytes; the harness adds the 8-byte zero pads both sides.
 * out: w*h uint32 sign-magnitude. returns 0 ok, -1 decoder reported failure. */
int ref_ht_decode(const uint8_t* data, uint32_t missing_msbs, uint32_t num_passes,
				  uint32_t lengths1, uint32_t lengths2, uint32_t w, uint32_t h, uint32_t* out)
{
	try
	{
		size_t total = (size_t)lengths1 + lengths2;
		std::vector<uint8_t> buf(total + 16, 0);
		memcpy(buf.data() + 8, data, total);
		/* the decoder also writes the phantom row / columns of a partial last quad row, as it does
		 * into T1HT's nominal-size buffer (T1HT.cpp:38-45): give it room, then copy w x h out */
		std::vector<uint32_t> scratch((size_t)w * (h + 2) + 16, 0);
		bool rc = ojph::local::ojph_decode_codeblock(buf.data() + 8, scratch.data(), missing_msbs,
													 num_passes, lengths1, lengths2, w, h, w);
		memcpy(out, scratch.data(), sizeof(uint32_t) * (size_t)w * h);
		return rc ? 0 : -1;
	}
	catch(...)
	{
		return -1;
	}
}

} // extern "C"

/* ------------------------------------------------------------------------------------- */
/* Geometry: the code-block list of a tile-component as TileComponent::init builds it
 * (tile/TileComponent.cpp:69-167), with each block's position in the tile buffer as
 * toRelativeCoordinates computes it (tile/TileComponentWindowBuffer.h:404-436).           */
extern "C" {
struct ref_cblk_geom
{
	uint32_t x, y, cx0, cy0, w, h, res, band, orient, precinct, cblkno;
};
int ref_layout(uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t numres, uint32_t cblkw,
			   uint32_t cblkh, const uint32_t* pw, const uint32_t* ph, ref_cblk_geom* out,
			   uint32_t cap)
{
	CodingParams cp;
	TileCodingParams tcp;
	TileComponentCodingParams tccp;
	TileComponent tilec;
	memset(&cp.m_coding_params, 0, sizeof(cp.m_coding_params));
	tccp.numresolutions = (uint8_t)numres;
	tccp.cblkw = (uint8_t)cblkw;
	tccp.cblkh = (uint8_t)cblkh;
	tccp.qmfbid = 1;
	tccp.numgbits = 2;
	tccp.roishift = 0;
	for(uint32_t i = 0; i < GRK_J2K_MAXRLVLS; ++i)
	{
		tccp.precinctWidthExp[i] = i < numres ? pw[i] : 15;
		tccp.precinctHeightExp[i] = i < numres ? ph[i] : 15;
	}
	for(uint32_t i = 0; i < GRK_J2K_MAXBANDS; ++i)
	{
		tccp.stepsizes[i].expn = 10;
		tccp.stepsizes[i].mant = 0;
	}
	grkRectU32 r(x0, y0, x1, y1);
	if(!tilec.init(true, true, r, 8, &cp, &tcp, &tccp, nullptr))
		return -1;
	if(!tilec.allocWindowBuffer(r))
		return -1;
	uint32_t n = 0;
	for(uint8_t resno = 0; resno < numres; ++resno)
	{
		auto res = tilec.tileCompResolution + resno;
		for(uint8_t bi = 0; bi < res->numTileBandWindows; ++bi)
		{
			auto band = res->tileBand + bi;
			for(auto prc : band->precincts)
			{
				for(uint64_t cb = 0; cb < prc->getNumCblks(); ++cb)
				{
					auto b = prc->getCodeBlockBounds(cb);
					if(n < cap)
					{
						uint32_t bx = b.x0, by = b.y0;
						tilec.getBuffer()->toRelativeCoordinates(resno, band->orientation, bx, by);
						out[n] = ref_cblk_geom{bx,	  by,	 b.x0,	b.y0, b.width(), b.height(), resno, bi,
											   (uint32_t)band->orientation, (uint32_t)prc->precinctIndex,
											   (uint32_t)cb};
					}
					++n;
				}
			}
		}
	}
	return (int)n;
}
}

/* ------------------------------------------------------------------------------------- */
/* Whole codec through the reference's PUBLIC C API (grok.h) on memory streams: the stock
 * CPU encoder and decoder, used as the end-to-end oracle and as bench.py's reference arm. */
extern "C" {

/* planes: numcomps dense w*h int32 planes. Returns codestream length, or -1.
 * Mirrors what grk_compress does after argument parsing (src/bin/jp2/grk_compress.cpp:
 * 820-828 lossless/all-pass default, :1978-1981 MCT default, :1123-1124 HT guard bits). */
/* compression ratio of the single quality layer for the next encodes (-r of grk_compress);
 * 0 = no rate limit: all passes, no rate control */
static double g_layer_rate = 0;
void ref_set_rate(double ratio)
{
	g_layer_rate = ratio;
}

/* Further encoder parameters for the next encodes, as a comma-free list of key=value pairs separated
 * by ';' -- the switches of grk_compress that the tests need, set on grk_cparameters the way the
 * CLI sets them (src/bin/jp2/grk_compress.cpp:820-1400):
 *   prog=LRCP|RLCP|RPCL|PCRL|CPRL     -p          rates=20:10:1         -r (layers)
 *   prc=256x256:128x128               -c          sop / eph             -S / -E
 *   tp=R|L|C                          -u          plt / tlm             -L / -X
 *   poc=0,0,1,3,3,CPRL/3,0,1,6,3,CPRL -P (resS,compS,layE,resE,compE,order per record)
 *   cinema=4k24|4k48|2k24|2k48        -x / -w     roi=c:s               -R
 *   sty=N   (code-block style bits)   -M          gbits=N               -N
 *   off=x0:y0 (image offset)          -d          torg=x0:y0 (tile origin) -T
 *   mct=0|1                           -Y                                                       */
static std::string g_enc_opts;
void ref_set_encode_options(const char* opts)
{
	g_enc_opts = opts ? opts : "";
}
static GRK_PROG_ORDER prog_by_name(const std::string& v)
{
	if(v == "RLCP")
		return GRK_RLCP;
	if(v == "RPCL")
		return GRK_RPCL;
	if(v == "PCRL")
		return GRK_PCRL;
	if(v == "CPRL")
		return GRK_CPRL;
	return GRK_LRCP;
}
static void apply_encode_options(grk_cparameters& p)
{
	size_t i = 0;
	while(i < g_enc_opts.size())
	{
		size_t j = g_enc_opts.find(';', i);
		if(j == std::string::npos)
			j = g_enc_opts.size();
		std::string kv = g_enc_opts.substr(i, j - i);
		i = j + 1;
		const size_t eq = kv.find('=');
		const std::string k = kv.substr(0, eq), v = eq == std::string::npos ? "" : kv.substr(eq + 1);
		if(k == "prog")
			p.prog_order = prog_by_name(v);
		else if(k == "rates")
		{
			p.numlayers = 0;
			size_t a = 0;
			while(a <= v.size() && p.numlayers < 100)
			{
				size_t b = v.find(':', a);
				if(b == std::string::npos)
					b = v.size();
				p.layer_rate[p.numlayers++] = atof(v.substr(a, b - a).c_str());
				a = b + 1;
			}
			p.allocationByRateDistoration = true;
		}
		else if(k == "prc")
		{
			p.res_spec = 0;
			size_t a = 0;
			while(a <= v.size() && p.res_spec < GRK_J2K_MAXRLVLS)
			{
				size_t b = v.find(':', a);
				if(b == std::string::npos)
					b = v.size();
				unsigned w = 0, h = 0;
				if(sscanf(v.substr(a, b - a).c_str(), "%ux%u", &w, &h) == 2)
				{
					p.prcw_init[p.res_spec] = w;
					p.prch_init[p.res_spec] = h;
					p.res_spec++;
				}
				a = b + 1;
			}
			p.csty |= 0x01;
		}
		else if(k == "sop")
			p.csty |= 0x02;
		else if(k == "eph")
			p.csty |= 0x04;
		else if(k == "tp")
		{
			p.newTilePartProgressionDivider = (uint8_t)v[0];
			p.enableTilePartGeneration = true;
		}
		else if(k == "plt")
			p.writePLT = true;
		else if(k == "tlm")
			p.writeTLM = true;
		else if(k == "poc")
		{
			uint32_t n = 0;
			size_t a = 0;
			while(a <= v.size() && n < GRK_J2K_MAXRLVLS)
			{
				size_t b = v.find('/', a);
				if(b == std::string::npos)
					b = v.size();
				unsigned rs, cs_, le, re, ce;
				char name[8] = {};
				if(sscanf(v.substr(a, b - a).c_str(), "%u,%u,%u,%u,%u,%4s", &rs, &cs_, &le, &re, &ce, name) == 6)
				{
					auto& g = p.progression[n++];
					g.tileno = 0;
					g.resS = (uint8_t)rs;
					g.compS = (uint16_t)cs_;
					g.layE = (uint16_t)le;
					g.resE = (uint8_t)re;
					g.compE = (uint16_t)ce;
					g.specifiedCompressionPocProg = prog_by_name(name);
				}
				a = b + 1;
			}
			p.numpocs = n ? n - 1 : 0;
		}
		else if(k == "cinema")
		{
			const bool k4 = v.substr(0, 2) == "4k", f48 = v.size() >= 4 && v.substr(2) == "48";
			p.rsiz = k4 ? GRK_PROFILE_CINEMA_4K : GRK_PROFILE_CINEMA_2K;
			p.max_comp_size = f48 ? GRK_CINEMA_48_COMP : GRK_CINEMA_24_COMP;
			p.max_cs_size = f48 ? GRK_CINEMA_48_CS : GRK_CINEMA_24_CS;
			p.framerate = f48 ? 48 : 24;
		}
		else if(k == "roi")
		{
			int c = 0;
			unsigned sh = 0;
			if(sscanf(v.c_str(), "%d:%u", &c, &sh) == 2)
			{
				p.roi_compno = c;
				p.roi_shift = sh;
			}
		}
		else if(k == "sty")
			p.cblk_sty = (uint8_t)atoi(v.c_str());
		else if(k == "gbits")
			p.numgbits = (uint8_t)atoi(v.c_str());
		else if(k == "off")
			sscanf(v.c_str(), "%u:%u", &p.image_offset_x0, &p.image_offset_y0);
		else if(k == "torg")
			sscanf(v.c_str(), "%u:%u", &p.tx0, &p.ty0);
		else if(k == "mct")
			p.mct = (uint8_t)atoi(v.c_str());
	}
}

long ref_codec_encode(int32_t** planes, uint32_t numcomps, uint32_t w, uint32_t h, uint32_t prec,
					  int irreversible, int ht, uint32_t numres, uint32_t cblkw, uint32_t cblkh,
					  uint32_t tile_w, uint32_t tile_h, uint32_t numthreads, uint8_t* out,
					  size_t out_cap)
{
	grk_initialize(nullptr, numthreads);
	grk_cparameters p;
	grk_compress_set_default_params(&p);
	p.cod_format = GRK_J2K_FMT;
	p.numresolution = (uint8_t)numres;
	p.cblockw_init = cblkw;
	p.cblockh_init = cblkh;
	p.irreversible = irreversible != 0;
	p.layer_rate[0] = g_layer_rate;
	p.numlayers = 1;
	p.allocationByRateDistoration = true;
	p.mct = numcomps >= 3 ? 1 : 0;
	p.numThreads = numthreads;
	if(ht)
	{
		p.isHT = true;
		p.numgbits = 1;
		p.cblk_sty = GRK_CBLKSTY_HT;
	}
	if(tile_w && tile_h)
	{
		p.tile_size_on = true;
		p.t_width = tile_w;
		p.t_height = tile_h;
	}
	apply_encode_options(p);
	std::vector<grk_image_cmptparm> cp(numcomps);
	for(auto& c : cp)
	{
		memset(&c, 0, sizeof(c));
		c.prec = (uint8_t)prec;
		c.sgnd = false;
		c.dx = c.dy = 1;
		c.w = w;
		c.h = h;
	}
	auto image = grk_image_new((uint16_t)numcomps, cp.data(),
							   numcomps >= 3 ? GRK_CLRSPC_SRGB : GRK_CLRSPC_GRAY, true);
	if(!image)
		return -1;
	image->x0 = p.image_offset_x0;
	image->y0 = p.image_offset_y0;
	image->x1 = image->x0 + w;
	image->y1 = image->y0 + h;
	for(uint32_t c = 0; c < numcomps; ++c)
	{
		auto comp = image->comps + c;
		comp->x0 = image->x0;
		comp->y0 = image->y0;
		for(uint32_t y = 0; y < h; ++y)
			memcpy(comp->data + (size_t)y * comp->stride, planes[c] + (size_t)y * w,
				   sizeof(int32_t) * w);
	}
	long rc = -1;
	auto stream = grk_stream_create_mem_stream(out, out_cap, false, false);
	auto codec = stream ? grk_compress_create(GRK_CODEC_J2K, stream) : nullptr;
	if(codec && grk_compress_init(codec, &p, image) && grk_compress_start(codec) &&
	   grk_compress(codec) && grk_compress_end(codec))
		rc = (long)grk_stream_get_write_mem_stream_length(stream);
	if(codec)
		grk_object_unref(codec);
	if(stream)
		grk_object_unref(stream);
	grk_object_unref(&image->obj);
	return rc;
}

/* Full decode of a codestream in memory with the stock multithreaded CPU decoder.
 * planes may be NULL (timing only). Returns 0 on success. */
/* resolutions to drop (-r of grk_decompress) and quality layers to keep (-l; 0 = all) for the next
 * ref_codec_decode calls; w and h are then those of the reduced image */
static uint32_t g_dec_reduce = 0, g_dec_layers = 0;
static uint32_t g_dec_win[4] = {0, 0, 0, 0};
void ref_set_decode_options(uint32_t reduce, uint32_t layers)
{
	g_dec_reduce = reduce;
	g_dec_layers = layers;
}
/* decode window of the next ref_codec_decode calls (-d of grk_decompress: grk_decompress_set_window),
 * relative to the image origin; 0,0,0,0 = whole image.  w and h are then those of the window. */
void ref_set_decode_window(uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1)
{
	g_dec_win[0] = x0;
	g_dec_win[1] = y0;
	g_dec_win[2] = x1;
	g_dec_win[3] = y1;
}

int ref_codec_decode(uint8_t* cs, size_t len, uint32_t numthreads, int32_t** planes,
					 uint32_t numcomps, uint32_t w, uint32_t h)
{
	grk_initialize(nullptr, numthreads);
	grk_decompress_parameters dp;
	memset(&dp, 0, sizeof(dp));
	grk_decompress_set_default_params(&dp.core);
	dp.core.cp_reduce = (uint8_t)g_dec_reduce;
	dp.core.cp_layer = (uint16_t)g_dec_layers;
	auto stream = grk_stream_create_mem_stream(cs, len, false, true);
	if(!stream)
		return -1;
	int rc = -1;
	auto codec = grk_decompress_create(GRK_CODEC_J2K, stream);
	grk_header_info hdr;
	memset(&hdr, 0, sizeof(hdr));
	if(codec && grk_decompress_init(codec, &dp.core) && grk_decompress_read_header(codec, &hdr) &&
	   grk_decompress_set_window(codec, g_dec_win[0], g_dec_win[1], g_dec_win[2], g_dec_win[3]) &&
	   grk_decompress(codec, nullptr))
	{
		auto image = grk_decompress_get_composited_image(codec);
		rc = image ? 0 : -2;
		if(image && planes)
		{
			if(image->numcomps < numcomps || image->comps[0].w != w || image->comps[0].h != h)
				rc = -3;
			else
				for(uint32_t c = 0; c < numcomps; ++c)
				{
					auto comp = image->comps + c;
					for(uint32_t y = 0; y < h; ++y)
						memcpy(planes[c] + (size_t)y * w, comp->data + (size_t)y * comp->stride,
							   sizeof(int32_t) * w);
				}
		}
	}
	if(codec)
		grk_object_unref(codec);
	grk_object_unref(stream);
	return rc;
}

/* dimensions and origin of component 0 of the image the stock decoder would deliver (after
 * cp_reduce): dims = {w, h, x0, y0} */
int ref_codec_decode_dims(uint8_t* cs, size_t len, uint32_t* dims)
{
	grk_decompress_parameters dp;
	memset(&dp, 0, sizeof(dp));
	grk_decompress_set_default_params(&dp.core);
	dp.core.cp_reduce = (uint8_t)g_dec_reduce;
	auto stream = grk_stream_create_mem_stream(cs, len, false, true);
	if(!stream)
		return -1;
	int rc = -1;
	auto codec = grk_decompress_create(GRK_CODEC_J2K, stream);
	grk_header_info hdr;
	memset(&hdr, 0, sizeof(hdr));
	if(codec && grk_decompress_init(codec, &dp.core) && grk_decompress_read_header(codec, &hdr))
	{
		auto image = grk_decompress_get_composited_image(codec);
		if(image)
		{
			dims[0] = image->comps[0].w;
			dims[1] = image->comps[0].h;
			dims[2] = image->comps[0].x0;
			dims[3] = image->comps[0].y0;
			rc = 0;
		}
	}
	if(codec)
		grk_object_unref(codec);
	grk_object_unref(stream);
	return rc;
}

/* ---- plugin drivers: what grk_compress / grk_decompress do around a loaded plugin -------------
 * (src/bin/jp2/grk_compress.cpp:1742-2160 pluginCompressCallback, grk_decompress.cpp:996-1300
 * decompress_callback), reduced to memory streams.  Only the public API of grok.h is used; the
 * stock grk_decompress binary cannot drive a plugin decode in this snapshot (user_data is not
 * forwarded, grok.cpp:740-753), hence this driver. */
static uint8_t* g_enc_out = nullptr;
static size_t g_enc_cap = 0;
static long g_enc_len = -1;

static bool plugin_enc_cb(grk_plugin_compress_user_callback_info* info)
{
	g_enc_len = -1;
	if(!info || !info->image || !info->tile)
		return false;
	auto stream = grk_stream_create_mem_stream(g_enc_out, g_enc_cap, false, false);
	auto codec = stream ? grk_compress_create(GRK_CODEC_J2K, stream) : nullptr;
	if(codec && grk_compress_init(codec, info->compressor_parameters, info->image) &&
	   grk_compress_start(codec) && grk_compress_with_plugin(codec, info->tile) &&
	   grk_compress_end(codec))
		g_enc_len = (long)grk_stream_get_write_mem_stream_length(stream);
	if(codec)
		grk_object_unref(codec);
	if(stream)
		grk_object_unref(stream);
	return g_enc_len >= 0;
}

static bool g_plugin_ready = false;
int ref_plugin_open(const char* plugin_dir, int device, uint32_t numthreads)
{
	if(g_plugin_ready)
		return 0;
	grk_initialize(nullptr, numthreads);
	if(getenv("GRKB_PLUGIN_VERBOSE"))
	{
		auto pr = [](const char* msg, void*) { fprintf(stderr, "[ref host] %s\n", msg); };
		grk_set_warning_handler(pr, nullptr);
		grk_set_error_handler(pr, nullptr);
	}
	grk_plugin_load_info li;
	li.pluginPath = plugin_dir;
	if(!grk_plugin_load(li))
		return -1;
	grk_plugin_init_info ii;
	ii.deviceId = device;
	ii.verbose = getenv("GRKB_PLUGIN_VERBOSE") != nullptr;
	if(!grk_plugin_init(ii))
		return -2;
	g_plugin_ready = true;
	return 0;
}

/* infile: binary PNM; parameters as ref_codec_encode.  Returns the codestream length or < 0. */
long ref_plugin_encode(const char* infile, int irreversible, int ht, uint32_t numres, uint32_t cblkw,
					   uint32_t cblkh, uint32_t numthreads, uint8_t* out, size_t out_cap)
{
	if(!g_plugin_ready)
		return -10;
	grk_cparameters p;
	grk_compress_set_default_params(&p);
	p.cod_format = GRK_J2K_FMT;
	p.decod_format = GRK_PXM_FMT;
	strncpy(p.infile, infile, sizeof(p.infile) - 1);
	p.numresolution = (uint8_t)numres;
	p.cblockw_init = cblkw;
	p.cblockh_init = cblkh;
	p.irreversible = irreversible != 0;
	p.layer_rate[0] = g_layer_rate;
	p.numlayers = 1;
	p.allocationByRateDistoration = true;
	p.mct = 255; /* decided from the image, like the CLI does */
	p.numThreads = numthreads;
	if(ht)
	{
		p.isHT = true;
		p.numgbits = 1;
		p.cblk_sty = GRK_CBLKSTY_HT;
	}
	apply_encode_options(p);
	g_enc_out = out;
	g_enc_cap = out_cap;
	g_enc_len = -1;
	/* the CLI resolves mct = 255 inside its callback, before grk_compress_init */
	struct Fix
	{
		static bool cb(grk_plugin_compress_user_callback_info* info)
		{
			if(info && info->image && info->compressor_parameters->mct == 255)
				info->compressor_parameters->mct = info->image->numcomps >= 3 ? 1 : 0;
			return plugin_enc_cb(info);
		}
	};
	if(grk_plugin_compress(&p, Fix::cb) != 0)
		return g_enc_len >= 0 ? g_enc_len : -11;
	return g_enc_len;
}

/* The batch door, as grk_compress drives it (src/bin/jp2/grk_compress.cpp:2216-2240): every
 * binary PNM of input_dir goes through the plugin and comes back to this callback, possibly on
 * several threads at once, which runs the stock rate allocation + T2 into output_file_name.
 * Returns the number of frames the callback completed, or < 0. */
static std::atomic<int> g_batch_ok{0}, g_batch_bad{0};
static bool plugin_batch_cb(grk_plugin_compress_user_callback_info* info)
{
	bool ok = false;
	if(info && info->image && info->tile && info->output_file_name)
	{
		if(info->compressor_parameters->mct == 255)
			info->compressor_parameters->mct = info->image->numcomps >= 3 ? 1 : 0;
		auto stream = grk_stream_create_file_stream(info->output_file_name, 1 << 20, false);
		auto codec = stream ? grk_compress_create(GRK_CODEC_J2K, stream) : nullptr;
		ok = codec && grk_compress_init(codec, info->compressor_parameters, info->image) &&
			 grk_compress_start(codec) && grk_compress_with_plugin(codec, info->tile) &&
			 grk_compress_end(codec);
		if(codec)
			grk_object_unref(codec);
		if(stream)
			grk_object_unref(stream);
	}
	(ok ? g_batch_ok : g_batch_bad)++;
	return ok;
}

int ref_plugin_batch_encode(const char* input_dir, const char* output_dir, int irreversible, int ht,
							uint32_t numres, uint32_t cblkw, uint32_t cblkh, uint32_t numthreads)
{
	if(!g_plugin_ready)
		return -10;
	grk_cparameters p;
	grk_compress_set_default_params(&p);
	p.cod_format = GRK_J2K_FMT;
	p.decod_format = GRK_PXM_FMT;
	p.numresolution = (uint8_t)numres;
	p.cblockw_init = cblkw;
	p.cblockh_init = cblkh;
	p.irreversible = irreversible != 0;
	p.layer_rate[0] = g_layer_rate;
	p.numlayers = 1;
	p.allocationByRateDistoration = true;
	p.mct = 255;
	p.numThreads = numthreads;
	if(ht)
	{
		p.isHT = true;
		p.numgbits = 1;
		p.cblk_sty = GRK_CBLKSTY_HT;
	}
	apply_encode_options(p);
	g_batch_ok = 0;
	g_batch_bad = 0;
	if(grk_plugin_batch_compress(input_dir, output_dir, &p, plugin_batch_cb) != 0)
		return -11;
	while(!grk_plugin_is_batch_complete())
		std::this_thread::sleep_for(std::chrono::milliseconds(2));
	grk_plugin_stop_batch_compress();
	return g_batch_bad ? -100 - g_batch_bad.load() : g_batch_ok.load();
}

static int32_t** g_dec_planes = nullptr;
static uint32_t g_dec_nc = 0, g_dec_w = 0, g_dec_h = 0;
static int g_dec_rc = -1;
static grk_decompress_parameters g_dec_params;

static int32_t plugin_dec_cb_impl(grk_plugin_decompress_callback_info* info);
static int32_t plugin_dec_cb(grk_plugin_decompress_callback_info* info)
{
	const int32_t rc = plugin_dec_cb_impl(info);
	if(getenv("GRKB_PLUGIN_VERBOSE"))
		fprintf(stderr, "[ref driver] decode callback flags=0x%x rc=%d\n",
				info ? info->decompress_flags : 0, rc);
	return rc;
}
static int32_t plugin_dec_cb_impl(grk_plugin_decompress_callback_info* info)
{
	if(!info)
		return -1;
	if(info->decompress_flags & GRK_PLUGIN_DECODE_CLEAN)
	{
		if(info->codec)
			grk_object_unref(info->codec);
		if(info->stream)
			grk_object_unref(info->stream);
		info->codec = nullptr;
		info->stream = nullptr;
		return 0;
	}
	if(info->decompress_flags & GRK_DECODE_HEADER)
	{
		static std::vector<uint8_t> file;
		file.clear();
		if(FILE* f = fopen(info->input_file_name, "rb"))
		{
			uint8_t buf[65536];
			size_t n;
			while((n = fread(buf, 1, sizeof(buf), f)) > 0)
				file.insert(file.end(), buf, buf + n);
			fclose(f);
		}
		info->stream = file.empty() ? nullptr
									: grk_stream_create_mem_stream(file.data(), file.size(), false, true);
		if(!info->stream)
			return -1;
		info->codec = grk_decompress_create(GRK_CODEC_J2K, info->stream);
		if(!info->codec || !grk_decompress_init(info->codec, &g_dec_params.core) ||
		   !grk_decompress_read_header(info->codec, &info->header_info))
			return -1;
		info->image = grk_decompress_get_composited_image(info->codec);
		if(!info->image || !info->init_decompressors_func)
			return -1;
		return info->init_decompressors_func(&info->header_info, info->image);
	}
	if(info->decompress_flags & GRK_DECODE_T2)
	{
		if(!info->tile || !info->codec)
			return -1;
		info->tile->decompress_flags = info->decompress_flags;
		return grk_decompress(info->codec, info->tile) ? 0 : -1;
	}
	if(info->decompress_flags & GRK_DECODE_POST_T1)
	{
		auto image = info->image;
		if(!image || image->numcomps < g_dec_nc || image->comps[0].w != g_dec_w ||
		   image->comps[0].h != g_dec_h)
			return -1;
		if(!g_dec_planes)
		{
			/* timing runs: the finished image is the host's from here on; look at it, do not copy it */
			g_dec_rc = image->comps[0].data ? 0 : -1;
			return g_dec_rc;
		}
		for(uint32_t c = 0; c < g_dec_nc; ++c)
			for(uint32_t y = 0; y < g_dec_h; ++y)
				memcpy(g_dec_planes[c] + (size_t)y * g_dec_w,
					   image->comps[c].data + (size_t)y * image->comps[c].stride,
					   sizeof(int32_t) * g_dec_w);
		g_dec_rc = 0;
		return 0;
	}
	return -1;
}

/* The batch door of the decoder (grk_decompress.cpp:996-1300 reduced to file streams): every
 * codestream of input_dir goes through host T2 + plugin, possibly on several threads at once;
 * the samples the plugin hands back are written to output_file_name as a flat dump
 * (uint32 numcomps, w, h, then the int32 planes).  All per-frame state rides in `info`. */
static std::atomic<int> g_bdec_ok{0}, g_bdec_bad{0};
static bool g_bdec_nowrite = false;
static int32_t plugin_batch_dec_cb(grk_plugin_decompress_callback_info* info)
{
	if(!info)
		return -1;
	if(info->decompress_flags & GRK_PLUGIN_DECODE_CLEAN)
	{
		if(info->codec)
			grk_object_unref(info->codec);
		if(info->stream)
			grk_object_unref(info->stream);
		info->codec = nullptr;
		info->stream = nullptr;
		return 0;
	}
	if(info->decompress_flags & GRK_DECODE_HEADER)
	{
		info->stream = grk_stream_create_file_stream(info->input_file_name, 1 << 20, true);
		if(!info->stream)
			return -1;
		info->codec = grk_decompress_create(GRK_CODEC_J2K, info->stream);
		if(!info->codec || !grk_decompress_init(info->codec, &g_dec_params.core) ||
		   !grk_decompress_read_header(info->codec, &info->header_info))
			return -1;
		info->image = grk_decompress_get_composited_image(info->codec);
		if(!info->image || !info->init_decompressors_func)
			return -1;
		return info->init_decompressors_func(&info->header_info, info->image);
	}
	if(info->decompress_flags & GRK_DECODE_T2)
	{
		if(!info->tile || !info->codec)
			return -1;
		info->tile->decompress_flags = info->decompress_flags;
		return grk_decompress(info->codec, info->tile) ? 0 : -1;
	}
	if(info->decompress_flags & GRK_DECODE_POST_T1)
	{
		auto image = info->image;
		FILE* f = image && info->output_file_name ? fopen(info->output_file_name, "wb") : nullptr;
		if(!f)
		{
			g_bdec_bad++;
			return -1;
		}
		if(g_bdec_nowrite)
		{
			/* timing runs: an empty output file, no 100 MB dump through the page cache */
			fclose(f);
			g_bdec_ok++;
			return 0;
		}
		const uint32_t hdr[3] = {image->numcomps, image->comps[0].w, image->comps[0].h};
		fwrite(hdr, sizeof(uint32_t), 3, f);
		for(uint32_t c = 0; c < image->numcomps; ++c)
			for(uint32_t y = 0; y < image->comps[c].h; ++y)
				fwrite(image->comps[c].data + (size_t)y * image->comps[c].stride, sizeof(int32_t),
					   image->comps[c].w, f);
		fclose(f);
		g_bdec_ok++;
		return 0;
	}
	return -1;
}

int ref_plugin_batch_decode(const char* input_dir, const char* output_dir)
{
	if(!g_plugin_ready)
		return -10;
	memset(&g_dec_params, 0, sizeof(g_dec_params));
	grk_decompress_set_default_params(&g_dec_params.core);
	g_dec_params.decod_format = GRK_J2K_FMT;
	g_bdec_ok = 0;
	g_bdec_bad = 0;
	if(grk_plugin_init_batch_decompress(input_dir, output_dir, &g_dec_params, plugin_batch_dec_cb) != 0)
		return -11;
	if(grk_plugin_batch_decompress() != 0)
		return -12;
	while(!grk_plugin_is_batch_complete())
		std::this_thread::sleep_for(std::chrono::milliseconds(2));
	grk_plugin_stop_batch_decompress();
	return g_bdec_bad ? -100 - g_bdec_bad.load() : g_bdec_ok.load();
}

/* Decode of a .j2k file with T2 on the host and everything after it in the plugin. */
/* timing doors for bench.py: the same calls, `iters` times, the host callback reduced to looking at
 * the image it is handed.  ms = mean wall time per grk_plugin_decompress call. */
int ref_plugin_decode_bench(const char* infile, uint32_t numcomps, uint32_t w, uint32_t h, uint32_t iters,
							uint32_t warmup, double* ms)
{
	if(!g_plugin_ready)
		return -10;
	memset(&g_dec_params, 0, sizeof(g_dec_params));
	grk_decompress_set_default_params(&g_dec_params.core);
	strncpy(g_dec_params.infile, infile, sizeof(g_dec_params.infile) - 1);
	g_dec_params.decod_format = GRK_J2K_FMT;
	g_dec_planes = nullptr;
	g_dec_nc = numcomps;
	g_dec_w = w;
	g_dec_h = h;
	std::chrono::steady_clock::time_point t0;
	for(uint32_t i = 0; i < warmup + iters; ++i)
	{
		if(i == warmup)
			t0 = std::chrono::steady_clock::now();
		g_dec_rc = -1;
		const int rc = grk_plugin_decompress(&g_dec_params, plugin_dec_cb);
		if(rc || g_dec_rc)
			return rc ? rc : g_dec_rc;
	}
	*ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / iters;
	return 0;
}

/* grk_plugin_batch_decompress over input_dir without the sample dump; ms = wall time of the batch */
int ref_plugin_batch_decode_bench(const char* input_dir, const char* output_dir, double* ms)
{
	g_bdec_nowrite = true;
	const auto t0 = std::chrono::steady_clock::now();
	const int n = ref_plugin_batch_decode(input_dir, output_dir);
	*ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
	g_bdec_nowrite = false;
	return n;
}

int ref_plugin_decode(const char* infile, int32_t** planes, uint32_t numcomps, uint32_t w, uint32_t h)
{
	if(!g_plugin_ready)
		return -10;
	memset(&g_dec_params, 0, sizeof(g_dec_params));
	grk_decompress_set_default_params(&g_dec_params.core);
	strncpy(g_dec_params.infile, infile, sizeof(g_dec_params.infile) - 1);
	g_dec_params.decod_format = GRK_J2K_FMT;
	g_dec_params.core.cp_reduce = (uint8_t)g_dec_reduce;
	g_dec_params.core.cp_layer = (uint16_t)g_dec_layers;
	g_dec_planes = planes;
	g_dec_nc = numcomps;
	g_dec_w = w;
	g_dec_h = h;
	g_dec_rc = -1;
	const int rc = grk_plugin_decompress(&g_dec_params, plugin_dec_cb);
	return rc ? rc : g_dec_rc;
}
}

/* ------------------------------------------------------------------------------------- */
/* The dequantising copy: the reference's own filter classes (filters/PostDecompressFilters.h)
 * instantiated on a DecompressBlockExec that carries roishift / k_msbs / stepsize, chosen as
 * TileComponent::postProcess / postProcessHT choose them (tile/TileComponent.cpp:333-386).    */
extern "C" int ref_dequant(int kind, int32_t* src, int32_t* dst, uint32_t n, uint32_t roishift,
						   uint32_t k_msbs, float stepsize)
{
	DecompressBlockExec block;
	block.roishift = (uint8_t)roishift;
	block.k_msbs = (uint8_t)k_msbs;
	block.stepsize = stepsize;
	switch(kind * 2 + (roishift ? 1 : 0))
	{
		case 0: { ShiftFilter<int32_t> f(&block); f.copy(dst, src, n); break; }
		case 1: { RoiShiftFilter<int32_t> f(&block); f.copy(dst, src, n); break; }
		case 2: { ScaleFilter<int32_t> f(&block); f.copy(dst, src, n); break; }
		case 3: { RoiScaleFilter<int32_t> f(&block); f.copy(dst, src, n); break; }
		case 4: { ShiftHTFilter<int32_t> f(&block); f.copy(dst, src, n); break; }
		case 5: { RoiShiftHTFilter<int32_t> f(&block); f.copy(dst, src, n); break; }
		case 6: { ScaleHTFilter<int32_t> f(&block); f.copy(dst, src, n); break; }
		case 7: { RoiScaleHTFilter<int32_t> f(&block); f.copy(dst, src, n); break; }
		default: return -1;
	}
	return 0;
}


/* array-based MCT: mct::compress_custom / decompress_custom (point_transform/mct.cpp:622-686) on the
 * caller's planes */
extern "C" int ref_mct_custom(void** planes, uint32_t nc, uint64_t n, float* matrix, int forward)
{
	std::vector<uint8_t*> p(nc);
	for(uint32_t c = 0; c < nc; ++c)
		p[c] = (uint8_t*)planes[c];
	const bool ok = forward ? mct::compress_custom((uint8_t*)matrix, n, p.data(), nc, 0)
							: mct::decompress_custom((uint8_t*)matrix, n, p.data(), nc, 0);
	return ok ? 0 : -1;
}
