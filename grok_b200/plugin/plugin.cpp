// plugin.cpp -- libgrokj2k_plugin.so: Grok's plugin ABI implemented over libgrok_b200.so.
//
// Built against the HOST's own headers (grok.h, plugin/plugin_interface.h, plugin/minpf_plugin.h
// from the Grok source tree the plugin is deployed with: `make -C grok_b200/plugin REF=<tree>`),
// exactly like any out-of-tree Grok plugin; nothing of those headers is copied here.
//
// What the host expects of each symbol is documented in SURVEY.md section 8b; the call sites are
//   loader        src/lib/jp2/plugin/minpf_plugin_manager.cpp:148-232 (dlsym "minpf_post_load_plugin")
//   entry points  src/lib/jp2/grok.cpp:606-825 (dlsym of the names in plugin/plugin_interface.h)
//   encode bridge src/lib/jp2/plugin/plugin_bridge.cpp:146-272 (compress_synch_with_plugin)
//   decode bridge src/lib/jp2/plugin/plugin_bridge.cpp:24-91  (decompress_synch_plugin_with_host)
// All compute goes through the C-ABI of include/grk_b200.h; there is no CPU path in here.
#include "grok.h"
#include "minpf_plugin.h"
#include "plugin_interface.h"

#include "../../include/grk_b200.h"

#include <atomic>
#include <chrono>
#include <algorithm>
#include <dirent.h>
#include <dlfcn.h>
#include <link.h>
#include <memory>
#include <mutex>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <sys/stat.h>
#include <thread>
#include <vector>

#define PLUGIN_EXPORT extern "C" __attribute__((visibility("default")))

namespace
{
bool g_verbose = false;
int g_device = 0; // -1: all devices, frames round-robin
int g_ndev = 0;
std::atomic<bool> g_stop{false};
std::atomic<bool> g_batch_done{true};
// joined when the library goes away (a host that never calls the plugin's exit function would
// otherwise end in std::terminate on a finished but unjoined thread)
struct BatchThread : std::thread
{
	using std::thread::operator=;
	~BatchThread()
	{
		g_stop = true;
		if(joinable())
			join();
	}
};
BatchThread g_batch_thread;

void logf(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
void logf(const char* fmt, ...)
{
	if(!g_verbose)
		return;
	va_list ap;
	va_start(ap, fmt);
	fprintf(stderr, "[grk_b200 plugin] ");
	vfprintf(stderr, fmt, ap);
	fprintf(stderr, "\n");
	va_end(ap);
}

// ---- host services the plugin needs (the image object is a host-side C++ object) -------------
typedef grk_image* (*image_new_fn)(uint16_t, grk_image_cmptparm*, GRK_COLOR_SPACE, bool);
typedef void (*unref_fn)(grk_object*);
// The host library is already in the process (it dlopen'ed this plugin); it may have been loaded
// RTLD_LOCAL (e.g. through ctypes), so every loaded object is asked, not just the global scope.
void* find_host_symbol(const char* name)
{
	if(void* p = dlsym(RTLD_DEFAULT, name))
		return p;
	struct Ctx
	{
		const char* name;
		void* sym;
	} ctx{name, nullptr};
	dl_iterate_phdr(
		[](dl_phdr_info* info, size_t, void* data) -> int {
			Ctx* c = (Ctx*)data;
			if(!info->dlpi_name || !*info->dlpi_name)
				return 0;
			if(void* h = dlopen(info->dlpi_name, RTLD_LAZY | RTLD_NOLOAD))
			{
				void* s = dlsym(h, c->name);
				dlclose(h);
				if(s)
				{
					c->sym = s;
					return 1;
				}
			}
			return 0;
		},
		&ctx);
	return ctx.sym;
}
image_new_fn host_image_new()
{
	static image_new_fn f = (image_new_fn)find_host_symbol("grk_image_new");
	return f;
}
unref_fn host_unref()
{
	static unref_fn f = (unref_fn)find_host_symbol("grk_object_unref");
	return f;
}

// ---- the grk_plugin_tile tree (grok.h:995-1077), owned by the plugin --------------------------
struct TileTree
{
	grk_plugin_tile tile{};
	std::vector<grk_plugin_tile_component> comps;
	std::vector<grk_plugin_tile_component*> comp_ptrs;
	std::vector<grk_plugin_resolution> ress;
	std::vector<grk_plugin_resolution*> res_ptrs;
	std::vector<grk_plugin_band> bands;
	std::vector<grk_plugin_band*> band_ptrs;
	std::vector<grk_plugin_precinct> precs;
	std::vector<grk_plugin_precinct*> prec_ptrs;
	std::vector<grk_plugin_code_block> blocks;
	std::vector<grk_plugin_code_block*> block_ptrs;
	std::vector<uint8_t> data; // decode: per-block byte buffers the host copies into
	// flat views, component-major in layout order (the order of the descriptor tables)
	std::vector<grkb_cblk_geom> geom; // one component's worth
	std::vector<grk_plugin_code_block*> flat; // numcomps * geom.size()
	std::vector<grk_plugin_band*> flat_band;
	uint32_t numcomps = 0, numres = 0;
};

struct TileGeom
{
	uint32_t x0, y0, x1, y1, numcomps, numres, cblkw_exp, cblkh_exp;
	uint32_t pw[33], ph[33];
};

uint32_t ceildivpow2(uint32_t a, uint32_t b)
{
	return (uint32_t)(((uint64_t)a + ((uint64_t)1 << b) - 1) >> b);
}

constexpr size_t TREE_GUARD_BYTES = 16u << 20;

// After the host's T2 filled the tree: no block may hold more than its buffer.  Returns false when
// one does (the frame is failed: the neighbouring buffers were overwritten); a block that ran past
// the guard region has corrupted memory that is not ours, and the process is stopped.
bool check_tree_capacity(const TileTree& t)
{
	bool ok = true;
	const uint8_t* end = t.data.data() + t.data.size();
	for(const grk_plugin_code_block& k : t.blocks)
	{
		if(!k.compressedData)
			continue;
		const size_t cap = (size_t)k.numPix * 4 + 64;
		if(k.compressedDataLength > cap)
		{
			ok = false;
			if(k.compressedData + k.compressedDataLength > end)
			{
				fprintf(stderr, "[grk_b200 plugin] fatal: the host wrote %u bytes for one code block, past the "
								"plugin's buffers; memory is corrupt\n", k.compressedDataLength);
				abort();
			}
		}
	}
	return ok;
}

// decode_cap: bytes reserved per block for the host's memcpy (0 on encode)
bool build_tree(const TileGeom& g, bool decode, TileTree& t)
{
	grkb_layout_params lp{};
	lp.x0 = g.x0;
	lp.y0 = g.y0;
	lp.x1 = g.x1;
	lp.y1 = g.y1;
	lp.numres = g.numres;
	lp.cblkw_exp = g.cblkw_exp;
	lp.cblkh_exp = g.cblkh_exp;
	for(int r = 0; r < 33; ++r)
	{
		lp.prec_w_exp[r] = g.pw[r];
		lp.prec_h_exp[r] = g.ph[r];
	}
	uint32_t n = 0;
	if(grkb_tile_layout(&lp, nullptr, 0, &n))
		return false;
	std::vector<grkb_cblk_geom> all(n);
	if(n && grkb_tile_layout(&lp, all.data(), n, &n))
		return false;
	t.numcomps = g.numcomps;
	t.numres = g.numres;
	// count bands / precincts per resolution the way the host does (Resolution::init)
	uint32_t nbands_total = 0;
	std::vector<uint64_t> nprec(g.numres);
	uint64_t nprec_total = 0;
	for(uint32_t r = 0; r < g.numres; ++r)
	{
		const uint32_t nd = g.numres - 1 - r;
		const uint32_t rx0 = ceildivpow2(g.x0, nd), ry0 = ceildivpow2(g.y0, nd);
		const uint32_t rx1 = ceildivpow2(g.x1, nd), ry1 = ceildivpow2(g.y1, nd);
		const uint32_t pwe = g.pw[r], phe = g.ph[r];
		const uint64_t gw = rx0 == rx1 ? 0 : (((uint64_t)ceildivpow2(rx1, pwe) << pwe) - ((uint64_t)(rx0 >> pwe) << pwe)) >> pwe;
		const uint64_t gh = ry0 == ry1 ? 0 : (((uint64_t)ceildivpow2(ry1, phe) << phe) - ((uint64_t)(ry0 >> phe) << phe)) >> phe;
		nprec[r] = gw * gh;
		const uint32_t nb = r ? 3 : 1;
		nbands_total += nb;
		nprec_total += nprec[r] * nb;
	}
	size_t nblk = 0;
	for(auto& b : all)
		nblk += (b.w && b.h) ? 1 : 0;
	const uint32_t C = g.numcomps;
	t.comps.resize(C);
	t.comp_ptrs.resize(C);
	t.ress.resize((size_t)C * g.numres);
	t.res_ptrs.resize(t.ress.size());
	t.bands.resize((size_t)C * nbands_total);
	t.band_ptrs.resize(t.bands.size());
	t.precs.resize((size_t)C * nprec_total);
	t.prec_ptrs.resize(t.precs.size());
	t.blocks.resize((size_t)C * all.size());
	t.block_ptrs.resize(t.blocks.size());
	size_t cap_total = 0;
	if(decode)
	{
		for(auto& b : all)
			cap_total += (size_t)b.w * b.h * 4 + 64; // the host's own bound (t1/T1.cpp:256)
		// The host copies every block's bytes into its buffer with no bound of its own
		// (plugin_bridge.cpp:84, Codeblock.h:245-259).  A block longer than the reference's own
		// T1 bound spills into the buffers that follow -- all inside this vector, whose tail is a
		// guard region -- and is caught by check_tree_capacity() before anything reads the data.
		t.data.assign(cap_total * C + TREE_GUARD_BYTES, 0);
	}
	t.geom.clear();
	for(auto& b : all)
		if(b.w && b.h)
			t.geom.push_back(b);
	t.flat.clear();
	t.flat_band.clear();
	size_t ri = 0, bi = 0, pi = 0, ki = 0, doff = 0;
	for(uint32_t c = 0; c < C; ++c)
	{
		t.comp_ptrs[c] = &t.comps[c];
		t.comps[c].numResolutions = g.numres;
		t.comps[c].resolutions = &t.res_ptrs[ri];
		size_t gi = 0; // cursor in `all` (sorted res, band, precinct, cblkno)
		for(uint32_t r = 0; r < g.numres; ++r, ++ri)
		{
			t.res_ptrs[ri] = &t.ress[ri];
			t.ress[ri].level = r;
			t.ress[ri].numBands = r ? 3 : 1;
			t.ress[ri].band = &t.band_ptrs[bi];
			for(uint32_t b = 0; b < (r ? 3u : 1u); ++b, ++bi)
			{
				t.band_ptrs[bi] = &t.bands[bi];
				grk_plugin_band& pb = t.bands[bi];
				pb.orientation = (uint8_t)(r ? b + 1 : 0);
				pb.numPrecincts = nprec[r];
				pb.precincts = &t.prec_ptrs[pi];
				pb.stepsize = 1.0f;
				for(uint64_t p = 0; p < nprec[r]; ++p, ++pi)
				{
					t.prec_ptrs[pi] = &t.precs[pi];
					grk_plugin_precinct& pp = t.precs[pi];
					pp.blocks = &t.block_ptrs[ki];
					uint64_t cnt = 0;
					while(gi < all.size() && all[gi].res == r && all[gi].band == b && all[gi].precinct == p)
					{
						const grkb_cblk_geom& q = all[gi++];
						grk_plugin_code_block& k = t.blocks[ki];
						t.block_ptrs[ki] = &k;
						memset(&k, 0, sizeof(k));
						k.x0 = q.cx0;
						k.y0 = q.cy0;
						k.x1 = q.cx0 + q.w;
						k.y1 = q.cy0 + q.h;
						k.numPix = (uint32_t)q.w * q.h;
						if(decode)
						{
							k.compressedData = t.data.data() + doff;
							doff += (size_t)q.w * q.h * 4 + 64;
						}
						if(q.w && q.h)
						{
							t.flat.push_back(&k);
							t.flat_band.push_back(&pb);
						}
						++ki;
						++cnt;
					}
					pp.numBlocks = cnt;
				}
			}
		}
	}
	t.tile.decompress_flags = 0;
	t.tile.numComponents = C;
	t.tile.tileComponents = t.comp_ptrs.data();
	return t.flat.size() == t.geom.size() * C && nblk == t.geom.size();
}

// ---- input images for plugin-driven encode: binary PNM (P5/P6, 8/16 bit) and raw ---------------
// page-locked host memory that only grows (kept by a batch worker from frame to frame)
struct PinnedBuf
{
	uint8_t* p = nullptr;
	size_t cap = 0;
	bool reserve(size_t n)
	{
		if(n <= cap)
			return true;
		if(p)
			grkb_host_free(p);
		p = (uint8_t*)grkb_host_alloc(n);
		cap = p ? n : 0;
		return p != nullptr;
	}
	~PinnedBuf()
	{
		if(p)
			grkb_host_free(p);
	}
};

struct HostImage
{
	uint32_t w = 0, h = 0, numcomps = 0, prec = 0;
	bool sgnd = false;
	std::vector<std::vector<int32_t>> planes; // raw input: one int32 plane per component
	// PNM input: the file's samples as they are (pixel-interleaved, 1 or 2 bytes, big-endian),
	// widened on the device (grkb_encoder_submit_packed) instead of here
	const uint8_t* packed = nullptr;
	uint32_t packed_bps = 0;
};

bool read_pnm(FILE* f, HostImage& im, PinnedBuf& keep)
{
	char magic[3] = {0, 0, 0};
	unsigned w = 0, h = 0, maxv = 0;
	auto token = [&](unsigned& v) -> bool {
		int c = fgetc(f);
		while(c == ' ' || c == '\n' || c == '\r' || c == '\t' || c == '#')
		{
			if(c == '#')
				while(c != '\n' && c != EOF)
					c = fgetc(f);
			c = fgetc(f);
		}
		if(c < '0' || c > '9')
			return false;
		v = 0;
		while(c >= '0' && c <= '9')
		{
			v = v * 10 + (unsigned)(c - '0');
			c = fgetc(f);
		}
		return true; // exactly one whitespace byte has been consumed after the number
	};
	if(fread(magic, 1, 2, f) != 2 || magic[0] != 'P' || (magic[1] != '5' && magic[1] != '6'))
		return false;
	if(!token(w) || !token(h) || !token(maxv) || !w || !h || !maxv || maxv > 65535)
		return false;
	im.w = w;
	im.h = h;
	im.numcomps = magic[1] == '6' ? 3 : 1;
	im.prec = 0;
	while((1u << im.prec) <= maxv)
		++im.prec;
	const size_t bps = maxv > 255 ? 2 : 1, n = (size_t)w * h * im.numcomps;
	if(!keep.reserve(n * bps + 16) || fread(keep.p, 1, n * bps, f) != n * bps)
		return false;
	im.packed = keep.p;
	im.packed_bps = (uint32_t)bps;
	return true;
}

// raw: component-planar, big-endian for > 8 bits (the reference's RAW_FMT; RAWL is little-endian)
bool read_raw(FILE* f, const grk_raw_cparameters& rc, bool little, HostImage& im)
{
	im.w = rc.width;
	im.h = rc.height;
	im.numcomps = rc.numcomps;
	im.prec = rc.prec;
	im.sgnd = rc.sgnd;
	if(!im.w || !im.h || !im.numcomps || !im.prec || im.prec > 16)
		return false;
	const size_t bps = im.prec > 8 ? 2 : 1, n = (size_t)im.w * im.h;
	std::vector<uint8_t> raw(n * bps);
	im.planes.assign(im.numcomps, std::vector<int32_t>(n));
	for(uint32_t c = 0; c < im.numcomps; ++c)
	{
		if(fread(raw.data(), 1, raw.size(), f) != raw.size())
			return false;
		for(size_t i = 0; i < n; ++i)
		{
			int32_t v;
			if(bps == 1)
				v = im.sgnd ? (int8_t)raw[i] : raw[i];
			else
			{
				const uint16_t u = little ? (uint16_t)(raw[2 * i] | (raw[2 * i + 1] << 8))
										  : (uint16_t)((raw[2 * i] << 8) | raw[2 * i + 1]);
				v = im.sgnd ? (int16_t)u : u;
			}
			im.planes[c][i] = v;
		}
	}
	return true;
}

bool read_image(const grk_cparameters* p, const char* path, HostImage& im, PinnedBuf& keep)
{
	FILE* f = fopen(path, "rb");
	if(!f)
		return false;
	bool ok;
	if(p->decod_format == GRK_RAW_FMT || p->decod_format == GRK_RAWL_FMT)
		ok = read_raw(f, p->raw_cp, p->decod_format == GRK_RAWL_FMT, im);
	else
		ok = read_pnm(f, im, keep);
	fclose(f);
	return ok;
}

// ---- one frame through the device encoder, then the host callback ------------------------------
struct EncodeJob
{
	std::string in, out;
	bool out_relative = false;
};

// A host image kept from frame to frame by a batch worker: a fresh 100 MB image costs more in
// page faults than everything the device does for the frame.  Reused only while the host left its
// shape alone; every sample is overwritten each frame.
struct CachedImage
{
	grk_image* image = nullptr;
	uint32_t w = 0, h = 0, nc = 0, prec = 0, x0 = 0, y0 = 0;
	bool sgnd = false;
	~CachedImage()
	{
		drop();
	}
	// The planes are page-locked memory of ours (the device writes them by DMA), lent to a host image
	// created without data; they are taken back before the host object is released, so the host's
	// allocator never sees them.  (Registering the host's own allocation instead is not safe: small
	// planes share pages with unrelated heap memory, and a copy that straddles registered and
	// unregistered pages is refused by the driver.)
	std::vector<void*> pinned;
	void drop()
	{
		if(image && !pinned.empty())
			for(uint32_t c = 0; c < image->numcomps && image->comps; ++c)
				image->comps[c].data = nullptr;
		for(void* p : pinned)
			grkb_host_free(p);
		pinned.clear();
		if(image && host_unref())
			host_unref()(&image->obj);
		image = nullptr;
	}
	// lend_planes: decode side (see above).  On the encode side the host takes the planes over
	// (its compressor moves the data pointers into its own image), so they are the host's allocation.
	grk_image* get(uint32_t W, uint32_t H, uint32_t numcomps, uint32_t precision, bool is_signed, uint32_t ox,
				   uint32_t oy, bool lend_planes = false)
	{
		bool ok = image && w == W && h == H && nc == numcomps && prec == precision && sgnd == is_signed &&
				  x0 == ox && y0 == oy && image->numcomps == numcomps && image->comps;
		for(uint32_t c = 0; ok && c < numcomps; ++c)
			ok = image->comps[c].data && image->comps[c].w == W && image->comps[c].h == H &&
				 image->comps[c].stride >= W && image->comps[c].prec == precision;
		if(ok)
			return image;
		drop();
		if(!host_image_new())
			return nullptr;
		std::vector<grk_image_cmptparm> cp(numcomps);
		for(auto& c : cp)
		{
			memset(&c, 0, sizeof(c));
			c.prec = (uint8_t)precision;
			c.sgnd = is_signed;
			c.dx = c.dy = 1;
			c.w = W;
			c.h = H;
			c.x0 = ox;
			c.y0 = oy;
		}
		image = host_image_new()((uint16_t)numcomps, cp.data(), numcomps >= 3 ? GRK_CLRSPC_SRGB : GRK_CLRSPC_GRAY,
								 !lend_planes);
		if(image && lend_planes)
		{
			// planes start from zeros: a reduced image may have a last row / column no tile covers,
			// which the reference leaves at zero
			const uint32_t stride = (W + 31u) & ~31u;
			for(uint32_t c = 0; c < numcomps; ++c)
			{
				const size_t bytes = (size_t)stride * H * sizeof(int32_t);
				void* p = grkb_host_alloc(bytes);
				if(!p)
				{
					drop();
					return nullptr;
				}
				memset(p, 0, bytes);
				pinned.push_back(p);
				image->comps[c].data = (int32_t*)p;
				image->comps[c].stride = stride;
			}
		}
		if(image)
		{
			image->x0 = ox;
			image->y0 = oy;
			image->x1 = ox + W;
			image->y1 = oy + H;
			w = W;
			h = H;
			nc = numcomps;
			prec = precision;
			sgnd = is_signed;
			x0 = ox;
			y0 = oy;
		}
		return image;
	}
};

// What a batch worker keeps from frame to frame: the device encoder (planes, pools and a stream
// of its own, so the copies of one worker's frame overlap the kernels of another's) and the
// page-locked buffers its results land in.
struct EncodeCtx
{
	grkb_encoder* enc = nullptr;
	grkb_tile_params tp{};
	int device = -1;
	uint8_t* bytes = nullptr; // grkb_host_alloc
	size_t bytes_cap = 0;
	grkb_coded* coded = nullptr; // grkb_host_alloc
	grkb_pass* passes = nullptr;
	size_t blocks_cap = 0;
	PinnedBuf raw; // the frame's file samples
	CachedImage himage; // the host-side image handed to the callback
	// tile tree and block table of the last frame, reused while the geometry stays the same
	std::unique_ptr<TileTree> tree;
	std::vector<grkb_enc_cblk> blocks;
	std::vector<float> band_step;
	TileGeom tg{};
	uint32_t t_prec = 0, t_numgbits = 0;
	bool t_reversible = false, t_ht = false;
	~EncodeCtx()
	{
		if(enc)
			grkb_encoder_destroy(enc);
		if(bytes)
			grkb_host_free(bytes);
		if(coded)
			grkb_host_free(coded);
		if(passes)
			grkb_host_free(passes);
	}
	bool reserve(size_t nblocks, size_t nbytes)
	{
		if(nbytes > bytes_cap)
		{
			if(bytes)
				grkb_host_free(bytes);
			bytes = (uint8_t*)grkb_host_alloc(nbytes);
			bytes_cap = bytes ? nbytes : 0;
		}
		if(nblocks > blocks_cap)
		{
			if(coded)
				grkb_host_free(coded);
			if(passes)
				grkb_host_free(passes);
			coded = (grkb_coded*)grkb_host_alloc(nblocks * sizeof(grkb_coded));
			passes = (grkb_pass*)grkb_host_alloc(nblocks * GRKB_ENC_MAX_PASSES * sizeof(grkb_pass));
			blocks_cap = coded && passes ? nblocks : 0;
		}
		return bytes_cap >= nbytes && blocks_cap >= nblocks;
	}
};

std::mutex g_cb_mutex; // GRKB_BATCH_SERIAL_CALLBACK=1: one host callback at a time
// The batch workers' contexts outlive a batch (creating a device encoder and page-locking its
// buffers costs as much as several frames).  Heap-allocated and released in exit_func only: no
// CUDA call is left to the static destructors.
std::vector<std::unique_ptr<EncodeCtx>>* g_ctx_pool = new std::vector<std::unique_ptr<EncodeCtx>>();
EncodeCtx* g_single_ctx = nullptr; // grk_plugin_compress (one frame per call)

// What the host's grk_compress_init will do to the parameters of a digital-cinema profile before its
// rate allocation and T2 run (codestream/CodeStreamCompress.cpp:219-235 -> Profile::is_cinema_compliant
// and Profile::set_cinema_parameters, codestream/Profile.cpp:1006-1215; `-x 24` / `-w 24` of
// grk_compress: BASELINE configs[4]).  The plugin codes its blocks before that call, so it has to
// partition the tile the way the host is going to: 32x32 blocks, 256x256 precincts (128 at the lowest
// resolution), 9/7, no mode switches, one or two guard bits.
void apply_profile(grk_cparameters& p, const HostImage& im)
{
	if(!GRK_IS_CINEMA(p.rsiz))
		return;
	const bool k4 = p.rsiz == GRK_PROFILE_CINEMA_4K;
	const bool compliant = (p.rsiz == GRK_PROFILE_CINEMA_2K || k4) && im.numcomps == 3 && im.prec == 12 && !im.sgnd &&
						   im.w <= (k4 ? 4096u : 2048u) && im.h <= (k4 ? 2160u : 1080u);
	if(!compliant)
	{
		p.rsiz = GRK_PROFILE_NONE; // the host falls back to the plain parameters
		return;
	}
	p.tile_size_on = false;
	p.tx0 = p.ty0 = 0;
	p.image_offset_x0 = p.image_offset_y0 = 0;
	p.cblockw_init = p.cblockh_init = 32;
	p.cblk_sty = 0;
	p.subsampling_dx = p.subsampling_dy = 1;
	p.irreversible = true;
	p.numlayers = 1;
	if(!k4 && p.numresolution > 6)
		p.numresolution = 6;
	if(k4 && p.numresolution < 2)
		p.numresolution = 1;
	if(k4 && p.numresolution > 7)
		p.numresolution = 7;
	p.csty |= 1;
	p.res_spec = (uint32_t)(p.numresolution - 1);
	for(uint32_t i = 0; i < p.res_spec; ++i)
		p.prcw_init[i] = p.prch_init[i] = 256;
	p.numgbits = k4 ? 2 : 1;
	// the byte budget of the profile makes the single layer rate controlled (Profile.cpp:1108-1150):
	// the host needs the passes' distortions
	p.allocationByRateDistoration = true;
	if(p.max_cs_size == 0 || p.max_cs_size > GRK_CINEMA_24_CS)
		p.max_cs_size = GRK_CINEMA_24_CS;
	p.layer_rate[0] = ((double)im.numcomps * im.w * im.h * im.prec) / ((double)p.max_cs_size * 8);
}

int encode_one(const grk_cparameters* params_in, const EncodeJob& job, int device,
			   grk::PLUGIN_ENCODE_USER_CALLBACK cb, EncodeCtx* ctx = nullptr, bool serial_cb = false)
{
	EncodeCtx local;
	if(!ctx)
		ctx = &local;
	const grk_cparameters* params = params_in; // for reading the input; replaced below
	using clk = std::chrono::steady_clock;
	const auto t_start = clk::now();
	auto ms_since = [](clk::time_point a) { return std::chrono::duration<double, std::milli>(clk::now() - a).count(); };
	HostImage im;
	if(!read_image(params, job.in.c_str(), im, ctx->raw))
	{
		logf("cannot read %s (binary PNM and raw input are supported)", job.in.c_str());
		return -1;
	}
	grk_cparameters eff = *params_in; // the parameters as the host's grk_compress_init will leave them
	apply_profile(eff, im);
	params = &eff;
	if(params->tile_size_on && (params->t_width < im.w || params->t_height < im.h))
	{
		logf("multi-tile encode is not routed through the plugin tile (one tile per grk_plugin_tile)");
		return -1;
	}
	if(params->subsampling_dx > 1 || params->subsampling_dy > 1 || im.prec > 16)
		return -1;
	const bool ht = params->isHT || (params->cblk_sty & GRK_CBLKSTY_HT);
	const uint8_t sty = (uint8_t)(params->cblk_sty & 0x3F);
	if(!ht && sty != 0)
	{
		logf("mode switches other than HT cannot cross the plugin bridge (pass->term is never set, "
			 "plugin_bridge.cpp:229-262)");
		return -1;
	}
	const bool reversible = !params->irreversible;
	const uint32_t numres = params->numresolution;
	const bool mct = params->mct == 255 ? im.numcomps >= 3 : params->mct == 1;
	if(grkb_set_device(device))
		return -1;
	// geometry
	TileGeom g{};
	g.x0 = params->image_offset_x0;
	g.y0 = params->image_offset_y0;
	g.x1 = g.x0 + im.w;
	g.y1 = g.y0 + im.h;
	g.numcomps = im.numcomps;
	g.numres = numres;
	auto log2u = [](uint32_t v) {
		uint32_t l = 0;
		while((1u << l) < v)
			++l;
		return l;
	};
	g.cblkw_exp = log2u(params->cblockw_init ? params->cblockw_init : 64);
	g.cblkh_exp = log2u(params->cblockh_init ? params->cblockh_init : 64);
	for(uint32_t r = 0; r < 33; ++r)
		g.pw[r] = g.ph[r] = 15;
	if((params->csty & 1) && params->res_spec)
	{
		// precinct sizes, highest resolution first; the last pair is halved for every further
		// resolution (CodeStreamCompress.cpp:542-596)
		auto flog2 = [](uint32_t v) {
			uint32_t l = 0;
			while(v > 1)
			{
				v >>= 1;
				++l;
			}
			return l;
		};
		for(uint32_t r = 0; r < numres; ++r)
		{
			const uint32_t p = numres - 1 - r, last = params->res_spec - 1;
			uint32_t pw, ph;
			if(p < params->res_spec)
			{
				pw = params->prcw_init[p];
				ph = params->prch_init[p];
			}
			else
			{
				pw = (p - last) < 32 ? params->prcw_init[last] >> (p - last) : 0;
				ph = (p - last) < 32 ? params->prch_init[last] >> (p - last) : 0;
			}
			g.pw[r] = pw < 1 ? 1 : flog2(pw);
			g.ph[r] = ph < 1 ? 1 : flog2(ph);
		}
	}
	const double ms_read = ms_since(t_start);
	const uint32_t numgbits = params->numgbits;
	const bool cached = ctx->tree && memcmp(&ctx->tg, &g, sizeof(g)) == 0 && ctx->t_prec == im.prec &&
						ctx->t_numgbits == numgbits && ctx->t_reversible == reversible && ctx->t_ht == ht;
	if(!cached)
	{
		ctx->tree = std::make_unique<TileTree>();
		if(!build_tree(g, false, *ctx->tree))
		{
			ctx->tree.reset();
			return -1;
		}
		// block table: quantisation parameters as the host will signal them in QCD
		const uint32_t nb = 3 * numres - 2;
		std::vector<uint32_t> expn(nb), mant(nb);
		if(grkb_qcd_generate(numres, im.prec, reversible, ht, expn.data(), mant.data()))
		{
			logf("%s", grkb_last_error());
			ctx->tree.reset();
			return -1;
		}
		const size_t ng0 = ctx->tree->geom.size();
		ctx->blocks.assign(ng0 * im.numcomps, grkb_enc_cblk{});
		ctx->band_step.assign(ng0, 1.0f);
		for(size_t i = 0; i < ng0; ++i)
		{
			const grkb_cblk_geom& q = ctx->tree->geom[i];
			float step = 1.0f;
			uint32_t bps = 0;
			const uint32_t qi = (q.res == 0 ? 0 : 3 * q.res - 2) + q.band;
			if(grkb_band_quant(expn[qi], mant[qi], q.orient, im.prec, reversible, 1, ht, numgbits, 0, &step,
							   &bps))
			{
				ctx->tree.reset();
				return -1;
			}
			ctx->band_step[i] = step;
			for(uint32_t c = 0; c < im.numcomps; ++c)
			{
				grkb_enc_cblk& b = ctx->blocks[c * ng0 + i];
				memset(&b, 0, sizeof(b));
				b.x = q.x;
				b.y = q.y;
				b.w = q.w;
				b.h = q.h;
				b.comp = (uint8_t)c;
				b.orient = q.orient;
				b.level = q.level;
				b.cblk_sty = ht ? GRKB_CBLKSTY_HT : 0;
				b.k_msbs = (uint8_t)bps;
				b.stepsize = step;
			}
		}
		ctx->tg = g;
		ctx->t_prec = im.prec;
		ctx->t_numgbits = numgbits;
		ctx->t_reversible = reversible;
		ctx->t_ht = ht;
	}
	TileTree* tree = ctx->tree.get();
	const std::vector<grkb_enc_cblk>& blocks = ctx->blocks;
	const std::vector<float>& band_step = ctx->band_step;
	const size_t ng = tree->geom.size(), n = ng * im.numcomps;
	grkb_tile_params tp{};
	tp.x0 = g.x0;
	tp.y0 = g.y0;
	tp.x1 = g.x1;
	tp.y1 = g.y1;
	tp.numcomps = im.numcomps;
	tp.numres = numres;
	tp.reversible = reversible;
	tp.mct = mct;
	for(uint32_t c = 0; c < im.numcomps; ++c)
		tp.shift[c] = im.sgnd ? 0 : 1 << (im.prec - 1);
	if(ctx->enc && (ctx->device != device || memcmp(&ctx->tp, &tp, sizeof(tp)) != 0))
	{
		grkb_encoder_destroy(ctx->enc);
		ctx->enc = nullptr;
	}
	if(!ctx->enc)
	{
		if(grkb_encoder_create(&tp, &ctx->enc))
		{
			ctx->enc = nullptr;
			logf("%s", grkb_last_error());
			return -1;
		}
		ctx->tp = tp;
		ctx->device = device;
	}
	grkb_encoder* enc = ctx->enc;
	const size_t bytes_cap = (size_t)im.w * im.h * im.numcomps * 4 + 65536;
	if(!ctx->reserve(n, bytes_cap))
	{
		logf("cannot allocate page-locked result buffers");
		return -1;
	}
	grkb_coded* coded = ctx->coded;
	grkb_pass* passes = ctx->passes;
	uint8_t* bytes = ctx->bytes;
	// the host runs rate control whenever a rate or quality target is set
	const bool rc = (params->allocationByRateDistoration && params->layer_rate[0] != 0) ||
					params->allocationByQuality || params->numlayers > 1;
	static const double norms_rev[3] = {1.732, .8292, .8292}, norms_irrev[3] = {1.732, 1.805, 1.573};
	const double* norms = mct ? (reversible ? norms_rev : norms_irrev) : nullptr;
	size_t used = 0;
	const double ms_prep = ms_since(t_start);
	int err;
	if(im.packed)
	{
		// PNM samples go to the device as the file holds them
		const uint32_t fmt = im.packed_bps == 2 ? (GRKB_SAMPLES_16_INTERLEAVED | GRKB_SAMPLES_BIG_ENDIAN)
												: GRKB_SAMPLES_8_INTERLEAVED;
		err = grkb_encoder_submit_packed(enc, im.packed, fmt, 0, (size_t)im.w * im.numcomps * im.packed_bps,
										 blocks.data(), (uint32_t)n, norms, mct ? 3 : 0, rc ? 1 : 0, coded, passes,
										 bytes, bytes_cap);
		if(!err)
			err = grkb_encoder_wait(enc, &used);
	}
	else
	{
		std::vector<const int32_t*> planes(im.numcomps);
		for(uint32_t c = 0; c < im.numcomps; ++c)
			planes[c] = im.planes[c].data();
		err = grkb_encoder_encode(enc, planes.data(), im.w, blocks.data(), (uint32_t)n, norms, mct ? 3 : 0,
								  rc ? 1 : 0, coded, passes, bytes, bytes_cap, &used);
	}
	if(err)
	{
		logf("%s", grkb_last_error());
		return -1;
	}
	const double ms_gpu = ms_since(t_start);
	// fill the tree the way compress_synch_with_plugin reads it (plugin_bridge.cpp:146-272)
	for(size_t i = 0; i < n; ++i)
	{
		grk_plugin_code_block& k = *tree->flat[i];
		const grkb_coded& c = coded[i];
		if(c.failed || c.numpasses > 67)
			return -1;
		tree->flat_band[i]->stepsize = band_step[i % ng];
		k.compressedData = bytes + c.data_off;
		k.compressedDataLength = c.data_len;
		k.numBitPlanes = c.numbps;
		k.numPasses = c.numpasses;
		for(uint32_t p = 0; p < c.numpasses; ++p)
		{
			const grkb_pass& ps = passes[i * GRKB_ENC_MAX_PASSES + p];
			k.passes[p].distortionDecrease = ps.distortion;
			k.passes[p].rate = (size_t)ps.rate - 1; // the bridge adds the 1 back (:249-257)
			k.passes[p].length = ps.len;
		}
	}
	// the image object is the host's (a C++ GrkImage behind the C struct)
	grk_image* image = ctx->himage.get(im.w, im.h, im.numcomps, im.prec, im.sgnd, g.x0, g.y0);
	{
		if(image)
		{
			if(im.packed)
			{
				// de-interleave the file's samples into the host's int32 planes (what the host's own
				// PNM reader does, src/bin/image_format/PNMFormat.cpp)
				const uint32_t nc = im.numcomps;
				for(uint32_t y = 0; y < im.h; ++y)
				{
					const uint8_t* src = im.packed + (size_t)y * im.w * nc * im.packed_bps;
					for(uint32_t c = 0; c < nc; ++c)
					{
						int32_t* dst = image->comps[c].data + (size_t)y * image->comps[c].stride;
						if(im.packed_bps == 2)
							for(uint32_t x = 0; x < im.w; ++x)
							{
								const uint8_t* q = src + ((size_t)x * nc + c) * 2;
								dst[x] = (int32_t)((q[0] << 8) | q[1]);
							}
						else
							for(uint32_t x = 0; x < im.w; ++x)
								dst[x] = src[(size_t)x * nc + c];
					}
				}
			}
			else
				for(uint32_t c = 0; c < im.numcomps; ++c)
					for(uint32_t y = 0; y < im.h; ++y)
						memcpy(image->comps[c].data + (size_t)y * image->comps[c].stride,
							   im.planes[c].data() + (size_t)y * im.w, sizeof(int32_t) * im.w);
		}
	}
	const double ms_fill = ms_since(t_start);
	grk::plugin_encode_user_callback_info info;
	memset(&info, 0, sizeof(info));
	info.input_file_name = job.in.c_str();
	info.outputFileNameIsRelative = job.out_relative;
	info.output_file_name = job.out.c_str();
	info.compressor_parameters = const_cast<grk_cparameters*>(params_in);
	info.image = image;
	info.tile = &tree->tile;
	info.error_code = 0;
	// synchronous: everything it reads stays alive until it returns
	if(serial_cb)
	{
		std::lock_guard<std::mutex> lk(g_cb_mutex);
		cb(&info);
	}
	else
		cb(&info);
	logf("%s: read %.1f ms, tables %.1f, device encode %.1f, tree + host image %.1f, host callback %.1f",
		 job.in.c_str(), ms_read, ms_prep - ms_read, ms_gpu - ms_prep, ms_fill - ms_gpu, ms_since(t_start) - ms_fill);
	return info.error_code ? -1 : 0; // the image stays with the worker's context
}

int next_device()
{
	static std::atomic<uint32_t> rr{0};
	if(g_device >= 0)
		return g_device;
	return g_ndev > 0 ? (int)(rr++ % (uint32_t)g_ndev) : 0;
}

// ---- decode ------------------------------------------------------------------------------------

// Read-only peek at the main header's QCD: the plugin tile tree carries each block's bit-plane
// count but not the band's, and the HT dequantiser needs their difference (BlockExec.h:40).
struct QcdPeek
{
	bool ok = false;
	uint32_t numgbits = 0, style = 0;
	std::vector<uint32_t> expn, mant;
};

QcdPeek peek_qcd(const char* path, uint32_t numres)
{
	QcdPeek q;
	FILE* f = path ? fopen(path, "rb") : nullptr;
	if(!f)
		return q;
	std::vector<uint8_t> b(1 << 16);
	b.resize(fread(b.data(), 1, b.size(), f));
	fclose(f);
	size_t i = 0;
	while(i + 3 < b.size() && !(b[i] == 0xFF && b[i + 1] == 0x4F && b[i + 2] == 0xFF && b[i + 3] == 0x51))
		++i;
	i += 2;
	while(i + 4 <= b.size() && b[i] == 0xFF && b[i + 1] != 0x90)
	{
		const uint32_t len = (b[i + 2] << 8) | b[i + 3];
		if(b[i + 1] == 0x5C && i + 2 + len <= b.size() && len >= 4)
		{
			const uint8_t sq = b[i + 4];
			q.numgbits = sq >> 5;
			q.style = sq & 0x1F;
			const uint8_t* p = &b[i + 5];
			const uint32_t nbody = len - 3, nb = 3 * numres - 2;
			q.expn.assign(nb, 0);
			q.mant.assign(nb, 0);
			for(uint32_t k = 0; k < nb; ++k)
			{
				const uint32_t res = k == 0 ? 0 : (k - 1) / 3 + 1;
				if(q.style == 0 && k < nbody)
					q.expn[k] = p[k] >> 3;
				else if(q.style == 2 && 2 * k + 1 < nbody)
				{
					const uint32_t v = (p[2 * k] << 8) | p[2 * k + 1];
					q.expn[k] = v >> 11;
					q.mant[k] = v & 0x7FF;
				}
				else if(q.style == 1 && nbody >= 2)
				{
					const uint32_t v = (p[0] << 8) | p[1], e0 = v >> 11;
					q.expn[k] = res == 0 ? e0 : (e0 > res - 1 ? e0 - (res - 1) : 0);
					q.mant[k] = v & 0x7FF;
				}
			}
			q.ok = true;
			return q;
		}
		i += 2 + len;
	}
	return q;
}

// What a batch worker keeps from frame to frame on the decode side: the device decoder, the tile
// tree of its geometry (with the per-block byte buffers the host copies into), and page-locked
// staging for the coded bytes on the way in and the samples on the way out.
struct DecoderSlot
{
	grkb_decompressor* dz = nullptr; // own T2 (the default): whole codestreams, any number of tiles
	int dz_device = -1;
	PinnedBuf file; // the codestream as read from disk, page-locked: its bytes are the byte pool
	grkb_decoder* dec = nullptr;
	grkb_tile_params tp{};
	int device = -1;
	std::unique_ptr<TileTree> tree;
	TileGeom tg{};
	PinnedBuf pool, out;
	CachedImage himage; // handed to the host's output callback with plugin_owns_image = true
	~DecoderSlot()
	{
		if(dz)
			grkb_decompressor_destroy(dz);
		if(dec)
			grkb_decoder_destroy(dec);
	}
};

struct DecodeCtx
{
	TileGeom geom{};
	DecoderSlot* slot = nullptr;
	TileTree* tree = nullptr; // owned by the slot
	uint32_t prec = 0;
	bool sgnd = false, reversible = true, mct = false, ht = false;
	uint8_t cblk_sty = 0;
	bool ready = false;
};
thread_local DecodeCtx* t_ctx = nullptr;

// GROK_INIT_DECOMPRESSORS: called by the host's callback once the main header is read
int init_decompressors(grk_header_info* h, grk_image* image)
{
	DecodeCtx* c = t_ctx;
	if(!c || !h || !image || image->numcomps == 0 || image->numcomps > GRKB_MAX_COMPS)
		return -1;
	if(h->t_grid_width * h->t_grid_height != 1)
	{
		logf("multi-tile images are decoded by the host (one tile per grk_plugin_tile)");
		return -1;
	}
	for(uint16_t i = 0; i < image->numcomps; ++i)
		if(image->comps[i].dx != 1 || image->comps[i].dy != 1 || image->comps[i].prec != image->comps[0].prec)
			return -1;
	TileGeom& g = c->geom;
	g.x0 = image->x0;
	g.y0 = image->y0;
	g.x1 = image->x1;
	g.y1 = image->y1;
	g.numcomps = image->numcomps;
	g.numres = h->numresolutions;
	auto flog2 = [](uint32_t v) {
		uint32_t l = 0;
		while(v > 1)
		{
			v >>= 1;
			++l;
		}
		return l;
	};
	g.cblkw_exp = flog2(h->cblockw_init);
	g.cblkh_exp = flog2(h->cblockh_init);
	for(uint32_t r = 0; r < 33; ++r)
	{
		g.pw[r] = r < g.numres ? flog2(h->prcw_init[r]) : 15;
		g.ph[r] = r < g.numres ? flog2(h->prch_init[r]) : 15;
	}
	c->prec = image->comps[0].prec;
	c->sgnd = image->comps[0].sgnd;
	c->reversible = !h->irreversible;
	c->mct = h->mct == 1 && image->numcomps >= 3;
	c->cblk_sty = h->cblk_sty;
	c->ht = (h->cblk_sty & GRK_CBLKSTY_HT) != 0;
	DecoderSlot* sl = c->slot;
	if(!sl)
		return -1;
	if(sl->tree && memcmp(&sl->tg, &g, sizeof(g)) == 0)
	{
		// same geometry as the last frame: the tree stands, only what the host fills in is reset
		for(auto& k : sl->tree->blocks)
		{
			k.compressedDataLength = 0;
			k.numPasses = 0;
			k.numBitPlanes = 0;
		}
		c->ready = true;
	}
	else
	{
		sl->tree = std::make_unique<TileTree>();
		c->ready = build_tree(g, true, *sl->tree);
		if(c->ready)
			sl->tg = g;
		else
			sl->tree.reset();
	}
	c->tree = sl->tree.get();
	return c->ready ? 0 : -1;
}

std::vector<std::unique_ptr<DecoderSlot>>* g_dec_pool = new std::vector<std::unique_ptr<DecoderSlot>>();
DecoderSlot* g_single_slot = nullptr; // grk_plugin_decompress (one frame per call)

// The host's callback runs library code that throws for streams it will not hand to a plugin
// (PluginDecodeUnsupportedException, plugin/plugin_bridge.cpp:55-72: several segments in a block, or
// more passes than its BIBO bound) and nothing above catches it: an exception must not unwind
// through this library or end a batch worker's thread.
int32_t call_host(grk::PLUGIN_DECODE_USER_CALLBACK cb, grk::PluginDecodeCallbackInfo* info)
{
	try
	{
		return cb(info);
	}
	catch(const std::exception& e)
	{
		logf("host callback threw (%s): the frame is left to the host's CPU path", e.what());
	}
	catch(...)
	{
		logf("host callback threw: the frame is left to the host's CPU path");
	}
	return -1;
}


// GROK_INIT_DECOMPRESSORS for the own-T2 path: the header the host read is not needed (the plugin
// parses the codestream itself); returning 0 tells the host to go on
int init_decompressors_own(grk_header_info*, grk_image*)
{
	return 0;
}

// uint16 / int16 rows -> the host's int32 planes, on a few threads (the 16-bit bus format, used
// when several GPUs share one host bridge)
void widen16(const uint8_t* src, size_t row16, grk_image* image, uint32_t numcomps, uint32_t W, uint32_t H, bool sgnd)
{
	const unsigned nt = std::max(1u, std::min(4u, std::thread::hardware_concurrency()));
	auto work = [&](unsigned t) {
		for(uint32_t c = 0; c < numcomps; ++c)
			for(uint32_t y = t; y < H; y += nt)
			{
				const uint16_t* s = (const uint16_t*)(src + ((size_t)c * H + y) * row16);
				int32_t* d = image->comps[c].data + (size_t)y * image->comps[c].stride;
				if(sgnd)
					for(uint32_t x = 0; x < W; ++x)
						d[x] = (int16_t)s[x];
				else
					for(uint32_t x = 0; x < W; ++x)
						d[x] = s[x];
			}
	};
	std::vector<std::thread> th;
	for(unsigned t = 1; t < nt; ++t)
		th.emplace_back(work, t);
	work(0);
	for(auto& t : th)
		t.join();
}

// Decode with T2 on our side: the plugin reads the file, parses markers and packets
// itself (grkb_decompress_*), decodes every tile on the device and hands the host a finished
// grk_image for its POST_T1 stage (colour management, file output).  The host's callback is asked
// for GRK_DECODE_HEADER (it opens its codec; the stock callback needs that for the later stages)
// and GRK_DECODE_POST_T1 only: its T2 and the grk_plugin_tile bridge -- one tile, one segment per
// block, 3*(prec+3) passes at most, one memcpy per block -- are not on the path.
int decode_one_own_t2(grk_decompress_parameters* params, const std::string& in, const std::string& out,
					  int device, grk::PLUGIN_DECODE_USER_CALLBACK cb, DecoderSlot* slot)
{
	using clk = std::chrono::steady_clock;
	const auto t_start = clk::now();
	auto ms_since = [](clk::time_point a) { return std::chrono::duration<double, std::milli>(clk::now() - a).count(); };
	// ---- the file, into page-locked memory
	FILE* f = fopen(in.c_str(), "rb");
	if(!f)
	{
		logf("cannot open %s", in.c_str());
		return -1;
	}
	struct stat sb;
	if(fstat(fileno(f), &sb) != 0 || sb.st_size <= 0 || !slot->file.reserve((size_t)sb.st_size + 16))
	{
		fclose(f);
		return -1;
	}
	const size_t len = fread(slot->file.p, 1, (size_t)sb.st_size, f);
	fclose(f);
	if(len != (size_t)sb.st_size)
		return -1;
	const double ms_read = ms_since(t_start);
	grk::PluginDecodeCallbackInfo info(in, out, params, params ? params->decod_format : GRK_UNK_FMT,
									   GRK_DECODE_HEADER);
	if(params)
		info.cod_format = params->cod_format;
	info.deviceId = (size_t)device;
	info.init_decompressors_func = init_decompressors_own;
	auto finish = [&](int rc) {
		info.decompress_flags = GRK_PLUGIN_DECODE_CLEAN;
		call_host(cb, &info);
		return rc;
	};
	if(call_host(cb, &info) != 0)
		return finish(-1);
	const double ms_header = ms_since(t_start);
	grkb_codestream* cs = nullptr;
	grkb_codestream_info ci;
	if(grkb_codestream_open(slot->file.p, len, &cs) || grkb_codestream_get_info(cs, &ci))
	{
		logf("%s: %s", in.c_str(), grkb_last_error());
		if(cs)
			grkb_codestream_close(cs);
		return finish(-1);
	}
	grkb_codestream_close(cs);
	const uint32_t reduce = params ? params->core.cp_reduce : 0, layers = params ? params->core.cp_layer : 0;
	if(reduce >= ci.numres)
	{
		logf("%s: cannot drop %u of %u resolutions", in.c_str(), reduce, ci.numres);
		return finish(-1);
	}
	if(ci.features & GRKB_CS_MIXED_PRECISION)
	{
		logf("%s: components of different precision or signedness are left to the host", in.c_str());
		return finish(-1);
	}
	uint32_t W = ceildivpow2(ci.x1 - ci.x0, reduce), H = ceildivpow2(ci.y1 - ci.y0, reduce);
	// decode window (-d of grk_decompress, relative to the image origin): the host image is the window
	uint32_t win[4] = {0, 0, 0, 0};
	uint32_t img_x0 = ceildivpow2(ci.x0, reduce), img_y0 = ceildivpow2(ci.y0, reduce);
	const bool windowed = params && (params->DA_x1 > params->DA_x0) && (params->DA_y1 > params->DA_y0);
	if(windowed)
	{
		win[0] = params->DA_x0;
		win[1] = params->DA_y0;
		win[2] = params->DA_x1;
		win[3] = params->DA_y1;
		const uint32_t ax0 = (uint32_t)std::min<uint64_t>((uint64_t)ci.x0 + win[0], ci.x1);
		const uint32_t ay0 = (uint32_t)std::min<uint64_t>((uint64_t)ci.y0 + win[1], ci.y1);
		const uint32_t ax1 = (uint32_t)std::min<uint64_t>((uint64_t)ci.x0 + win[2], ci.x1);
		const uint32_t ay1 = (uint32_t)std::min<uint64_t>((uint64_t)ci.y0 + win[3], ci.y1);
		img_x0 = ceildivpow2(ax0, reduce);
		img_y0 = ceildivpow2(ay0, reduce);
		W = ceildivpow2(ax1, reduce) - img_x0;
		H = ceildivpow2(ay1, reduce) - img_y0;
		if(ax1 <= ax0 || ay1 <= ay0 || W == 0 || H == 0)
		{
			logf("%s: empty decode window", in.c_str());
			return finish(-1);
		}
	}
	const uint32_t prec = ci.prec[0];
	const bool sgnd = ci.sgnd[0] != 0;
	if(!host_image_new() || !host_unref())
	{
		logf("host library symbols grk_image_new / grk_object_unref are not visible");
		return finish(-1);
	}
	if(grkb_set_device(device))
		return finish(-1);
	grk_image* image = slot->himage.get(W, H, ci.numcomps, prec, sgnd, img_x0, img_y0, true);
	if(!image)
		return finish(-1);
	if(slot->dz && slot->dz_device != device)
	{
		grkb_decompressor_destroy(slot->dz);
		slot->dz = nullptr;
	}
	if(!slot->dz)
	{
		if(grkb_decompressor_create(device, 3, &slot->dz))
		{
			slot->dz = nullptr;
			logf("%s", grkb_last_error());
			return finish(-1);
		}
		slot->dz_device = device;
	}
	const double ms_setup = ms_since(t_start);
	uint32_t rejected = 0;
	int rc;
	// one GPU: int32 planes by DMA straight into the host's (page-locked) image.  Several GPUs behind
	// one host bridge (deviceId = -1): 16-bit samples across the bus, widened here.
	const bool bus16 = !windowed && prec <= 16 && (getenv("GRKB_PLUGIN_BUS16") ? atoi(getenv("GRKB_PLUGIN_BUS16")) != 0
																  : (g_device < 0 && g_ndev > 1));
	const size_t row16 = (size_t)W * 2;
	if(bus16 && slot->out.reserve(row16 * H * ci.numcomps + 16))
	{
		memset(slot->out.p, 0, row16 * H * ci.numcomps);
		rc = grkb_decompress_submit_packed(slot->dz, slot->file.p, len, reduce, layers, slot->out.p, GRKB_SAMPLES_16,
										   row16);
		if(!rc)
			rc = grkb_decompress_wait(slot->dz, &rejected);
		if(!rc)
			widen16(slot->out.p, row16, image, ci.numcomps, W, H, sgnd);
	}
	else
	{
		int32_t* planes[GRKB_MAX_COMPS] = {};
		for(uint32_t c = 0; c < ci.numcomps; ++c)
			planes[c] = image->comps[c].data;
		rc = grkb_decompress_submit_window(slot->dz, slot->file.p, len, reduce, layers, win[0], win[1], win[2], win[3],
										   planes, image->comps[0].stride);
		if(!rc)
			rc = grkb_decompress_wait(slot->dz, &rejected);
	}
	if(rc || rejected)
	{
		logf("%s: decode failed: %s (%u blocks rejected)", in.c_str(), grkb_last_error(), rejected);
		return finish(-1);
	}
	const double ms_dev = ms_since(t_start);
	double t2_ms = 0;
	uint32_t tiles = 0;
	grkb_decompress_stats(slot->dz, &t2_ms, &tiles);
	info.image = image;
	info.plugin_owns_image = true;
	info.decompress_flags = GRK_DECODE_POST_T1;
	rc = call_host(cb, &info); // host: colour management, file output
	const double ms_out = ms_since(t_start);
	finish(rc);
	logf("%s: read %.1f ms, host header %.1f, setup %.1f, decode %.1f (own T2 %.1f, %u tiles), host output %.1f, "
		 "clean %.1f", in.c_str(), ms_read, ms_header - ms_read, ms_setup - ms_header, ms_dev - ms_setup, t2_ms, tiles,
		 ms_out - ms_dev, ms_since(t_start) - ms_out);
	return rc;
}

int decode_one(grk_decompress_parameters* params, const std::string& in, const std::string& out,
			   int device, grk::PLUGIN_DECODE_USER_CALLBACK cb, DecoderSlot* slot = nullptr)
{
	DecoderSlot local;
	if(!slot)
		slot = &local;
	// GRKB_PLUGIN_HOST_T2=1 selects the bridge path: the host's T2 fills a grk_plugin_tile
	const bool host_t2 = getenv("GRKB_PLUGIN_HOST_T2") && atoi(getenv("GRKB_PLUGIN_HOST_T2")) != 0;
	if(!host_t2)
		return decode_one_own_t2(params, in, out, device, cb, slot);
	// The bridge carries one tile, one QCD, no ROI shift and neither cp_reduce nor cp_layer: what it
	// cannot express is refused here instead of being decoded wrongly.
	{
		if(params && (params->core.cp_reduce || params->core.cp_layer))
		{
			fprintf(stderr, "[grk_b200 plugin] %s: bridge path: cp_reduce / cp_layer are not supported\n", in.c_str());
			return -1;
		}
		std::vector<uint8_t> head(1 << 20);
		FILE* f = fopen(in.c_str(), "rb");
		if(!f)
			return -1;
		head.resize(fread(head.data(), 1, head.size(), f));
		fclose(f);
		grkb_codestream* cs = nullptr;
		grkb_codestream_info ci;
		memset(&ci, 0, sizeof(ci));
		const bool ok = grkb_codestream_open(head.data(), head.size(), &cs) == 0 && grkb_codestream_get_info(cs, &ci) == 0;
		if(cs)
			grkb_codestream_close(cs);
		if(ok && (ci.features & (GRKB_CS_RGN | GRKB_CS_QCC | GRKB_CS_COC | GRKB_CS_TILE_OVERRIDES |
								 GRKB_CS_MIXED_PRECISION)))
		{
			fprintf(stderr, "[grk_b200 plugin] %s: bridge path: RGN / QCC / COC / tile-header overrides / mixed "
							"precision cannot cross grk_plugin_tile\n", in.c_str());
			return -1;
		}
	}
	using clk = std::chrono::steady_clock;
	const auto t_start = clk::now();
	auto ms_since = [](clk::time_point a) { return std::chrono::duration<double, std::milli>(clk::now() - a).count(); };
	DecodeCtx ctx;
	ctx.slot = slot;
	t_ctx = &ctx;
	grk::PluginDecodeCallbackInfo info(in, out, params, params ? params->decod_format : GRK_UNK_FMT,
									   GRK_DECODE_HEADER);
	if(params)
		info.cod_format = params->cod_format;
	info.deviceId = (size_t)device;
	info.init_decompressors_func = init_decompressors;
	auto finish = [&](int rc) {
		info.decompress_flags = GRK_PLUGIN_DECODE_CLEAN;
		call_host(cb, &info);
		t_ctx = nullptr;
		return rc;
	};
	if(call_host(cb, &info) != 0 || !ctx.ready)
		return finish(-1);
	const double ms_header = ms_since(t_start);
	TileTree& t = *ctx.tree;
	// T2 on the host.  GRK_DECODE_POST_T1 has to ride along: without it the host never advances
	// to the next tile part (CodeStreamDecompress::findNextTile only does so when the post stage
	// is enabled, codestream/CodeStreamDecompress.cpp:1048-1070) and then fails with "no SOT
	// marker found".  The host's own post stage then runs over its empty tile buffer; the samples
	// that count are the ones this plugin hands over below.
	info.decompress_flags = GRK_DECODE_T2 | GRK_DECODE_POST_T1;
	info.tile = &t.tile;
	if(call_host(cb, &info) != 0) // host: T2, then decompress_synch_plugin_with_host fills the tree
		return finish(-1);
	if(!check_tree_capacity(t))
	{
		fprintf(stderr, "[grk_b200 plugin] %s: a code block is longer than the buffer the plugin ABI lets us "
						"offer; frame refused\n", in.c_str());
		return finish(-1);
	}
	const double ms_t2 = ms_since(t_start);
	// descriptor table from the filled tree
	const TileGeom& g = ctx.geom;
	const size_t ng = t.geom.size(), n = ng * g.numcomps;
	QcdPeek qcd;
	if(ctx.ht)
	{
		qcd = peek_qcd(in.c_str(), g.numres);
		if(!qcd.ok)
		{
			logf("cannot read QCD of %s", in.c_str());
			return finish(-1);
		}
	}
	std::vector<grkb_cblk> blocks(n);
	std::vector<grkb_seg> segs;
	segs.reserve(n);
	size_t pool_bytes = 16;
	for(size_t i = 0; i < n; ++i)
		pool_bytes += t.flat[i]->numPasses ? t.flat[i]->compressedDataLength : 0;
	if(!slot->pool.reserve(pool_bytes))
		return finish(-1);
	uint8_t* pool = slot->pool.p;
	size_t pool_size = 0;
	for(size_t i = 0; i < n; ++i)
	{
		const grkb_cblk_geom& q = t.geom[i % ng];
		const grk_plugin_code_block& k = *t.flat[i];
		grkb_cblk& b = blocks[i];
		memset(&b, 0, sizeof(b));
		b.x = q.x;
		b.y = q.y;
		b.w = q.w;
		b.h = q.h;
		b.comp = (uint8_t)(i / ng);
		b.orient = q.orient;
		b.cblk_sty = ctx.cblk_sty;
		b.numbps = k.numBitPlanes;
		const float half_step = t.flat_band[i]->stepsize; // band->stepsize / 2 (bridge :46)
		if(ctx.ht)
		{
			float s;
			uint32_t band_bps = 0;
			const uint32_t qi = (q.res == 0 ? 0 : 3 * q.res - 2) + q.band;
			if(grkb_band_quant(qcd.expn[qi], qcd.mant[qi], q.orient, ctx.prec, ctx.reversible, 0, 0,
							   qcd.numgbits, 0, &s, &band_bps))
				return finish(-1);
			b.k_msbs = (uint8_t)(band_bps > k.numBitPlanes ? band_bps - k.numBitPlanes : 0);
			b.scale = half_step * 2.0f;
		}
		else
			b.scale = half_step;
		if(k.compressedDataLength && k.numPasses)
		{
			b.numsegs = 1;
			b.seg_first = (uint32_t)segs.size();
			b.data_off = (uint32_t)pool_size;
			b.data_len = k.compressedDataLength;
			segs.push_back(grkb_seg{k.compressedDataLength, (uint32_t)k.numPasses});
			memcpy(pool + pool_size, k.compressedData, k.compressedDataLength);
			pool_size += k.compressedDataLength;
		}
	}
	memset(pool + pool_size, 0, 16);
	pool_size += 16;
	if(grkb_set_device(device))
		return finish(-1);
	grkb_tile_params tp{};
	tp.x0 = g.x0;
	tp.y0 = g.y0;
	tp.x1 = g.x1;
	tp.y1 = g.y1;
	tp.numcomps = g.numcomps;
	tp.numres = g.numres;
	tp.reversible = ctx.reversible;
	tp.mct = ctx.mct;
	for(uint32_t c = 0; c < g.numcomps; ++c)
	{
		tp.shift[c] = ctx.sgnd ? 0 : 1 << (ctx.prec - 1);
		tp.lo[c] = ctx.sgnd ? -(1 << (ctx.prec - 1)) : 0;
		tp.hi[c] = ctx.sgnd ? (1 << (ctx.prec - 1)) - 1 : (1 << ctx.prec) - 1;
	}
	// the output image is a host object; it is handed over with plugin_owns_image = true
	if(!host_image_new() || !host_unref())
	{
		logf("host library symbols grk_image_new / grk_object_unref are not visible");
		return finish(-1);
	}
	const uint32_t W = g.x1 - g.x0, H = g.y1 - g.y0;
	grk_image* image = slot->himage.get(W, H, g.numcomps, ctx.prec, ctx.sgnd, g.x0, g.y0);
	if(!image)
		return finish(-1);
	const double ms_table = ms_since(t_start);
	if(slot->dec && (slot->device != device || memcmp(&slot->tp, &tp, sizeof(tp)) != 0))
	{
		grkb_decoder_destroy(slot->dec);
		slot->dec = nullptr;
	}
	int rc = 0;
	if(!slot->dec)
	{
		rc = grkb_decoder_create(&tp, (uint32_t)n, pool_size, &slot->dec);
		if(rc)
			slot->dec = nullptr;
		slot->tp = tp;
		slot->device = device;
	}
	grkb_decoder* dec = slot->dec;
	uint32_t rejected = 0;
	if(!rc)
	{
		const size_t row16 = (size_t)W * 2;
		if(ctx.prec <= 16 && slot->out.reserve(row16 * H * g.numcomps + 16))
		{
			// the samples cross the bus as 16-bit planes into page-locked memory and are
			// widened into the host's int32 image here
			rc = grkb_decoder_submit_packed(dec, blocks.data(), (uint32_t)n, segs.data(), (uint32_t)segs.size(),
											pool, pool_size, slot->out.p, GRKB_SAMPLES_16, row16);
			if(!rc)
				rc = grkb_decoder_wait(dec, &rejected);
			if(!rc)
				for(uint32_t c = 0; c < g.numcomps; ++c)
					for(uint32_t y = 0; y < H; ++y)
					{
						const uint16_t* src = (const uint16_t*)(slot->out.p + ((size_t)c * H + y) * row16);
						int32_t* dst = image->comps[c].data + (size_t)y * image->comps[c].stride;
						if(ctx.sgnd)
							for(uint32_t x = 0; x < W; ++x)
								dst[x] = (int16_t)src[x];
						else
							for(uint32_t x = 0; x < W; ++x)
								dst[x] = src[x];
					}
		}
		else
		{
			std::vector<int32_t*> planes(g.numcomps);
			for(uint32_t c = 0; c < g.numcomps; ++c)
				planes[c] = image->comps[c].data;
			rc = grkb_decoder_decode(dec, blocks.data(), (uint32_t)n, segs.data(), (uint32_t)segs.size(), pool,
									 pool_size, planes.data(), image->comps[0].stride, &rejected);
		}
	}
	if(rc || rejected)
	{
		logf("decode failed: %s (%u blocks rejected)", grkb_last_error(), rejected);
		return finish(-1);
	}
	const double ms_dev = ms_since(t_start);
	info.image = image;
	info.plugin_owns_image = true;
	info.decompress_flags = GRK_DECODE_POST_T1;
	rc = call_host(cb, &info); // host: colour management, file output
	const double ms_out = ms_since(t_start);
	finish(rc);
	logf("%s: host header %.1f ms, host T2 %.1f, table + host image %.1f, device decode %.1f, host output %.1f, "
		 "clean %.1f", in.c_str(), ms_header, ms_t2 - ms_header, ms_table - ms_t2, ms_dev - ms_table,
		 ms_out - ms_dev, ms_since(t_start) - ms_out);
	return rc; // the image stays with the slot
}

std::vector<std::string> list_dir(const char* dir)
{
	std::vector<std::string> v;
	if(DIR* d = dir ? opendir(dir) : nullptr)
	{
		while(dirent* e = readdir(d))
		{
			std::string p = std::string(dir) + "/" + e->d_name;
			struct stat st;
			if(e->d_name[0] != '.' && stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode))
				v.push_back(e->d_name);
		}
		closedir(d);
	}
	std::sort(v.begin(), v.end());
	return v;
}

std::string replace_ext(const std::string& name, const char* ext)
{
	const size_t dot = name.rfind('.');
	return (dot == std::string::npos ? name : name.substr(0, dot)) + "." + ext;
}

void join_batch()
{
	g_stop = true;
	if(g_batch_thread.joinable())
		g_batch_thread.join();
	g_stop = false;
}
} // namespace

// ---- minpf registration (plugin/minpf_plugin_manager.cpp:148-232) -------------------------------

PLUGIN_EXPORT void* create(grk::minpf_object_params*)
{
	return nullptr;
}
PLUGIN_EXPORT int32_t destroy(void*)
{
	return 0;
}
PLUGIN_EXPORT int32_t exit_func()
{
	join_batch();
	g_ctx_pool->clear();
	g_dec_pool->clear();
	delete g_single_slot;
	g_single_slot = nullptr;
	delete g_single_ctx;
	g_single_ctx = nullptr;
	return 0;
}
PLUGIN_EXPORT grk::minpf_exit_func minpf_post_load_plugin(const char*,
														   const grk::minpf_platform_services* params)
{
	if(!params || !params->registerObject)
		return nullptr;
	grk::minpf_register_params rp;
	rp.version.major = 1;
	rp.version.minor = 0;
	rp.createFunc = create;
	rp.destroyFunc = destroy;
	if(params->registerObject("grk_b200", &rp) < 0)
		return nullptr;
	return exit_func;
}

PLUGIN_EXPORT bool plugin_init(grk_plugin_init_info info)
{
	g_verbose = info.verbose;
	g_ndev = grkb_device_count();
	if(info.deviceId == -2 || g_ndev < 1)
	{
		logf("no CUDA device: the plugin stays disabled (there is no CPU path in it)");
		return false;
	}
	g_device = info.deviceId;
	if(g_device >= g_ndev)
		return false;
	return grkb_set_device(g_device < 0 ? 0 : g_device) == 0;
}

PLUGIN_EXPORT uint32_t plugin_get_debug_state(void)
{
	return GRK_PLUGIN_STATE_NO_DEBUG;
}
PLUGIN_EXPORT void plugin_debug_mqc_next_cxd(grk::grk_plugin_debug_mqc*, uint32_t) {}
PLUGIN_EXPORT void plugin_debug_mqc_next_plane(grk::grk_plugin_debug_mqc*) {}

// ---- encode ------------------------------------------------------------------------------------

PLUGIN_EXPORT int32_t plugin_encode(grk_cparameters* params, grk::PLUGIN_ENCODE_USER_CALLBACK cb)
{
	if(!params || !cb)
		return -1;
	EncodeJob job;
	job.in = params->infile;
	job.out = params->outfile;
	// the context (device encoder, page-locked buffers, host image) is kept from call to call;
	// concurrent callers take turns
	static std::mutex mu;
	std::lock_guard<std::mutex> lk(mu);
	if(!g_single_ctx)
		g_single_ctx = new EncodeCtx();
	return encode_one(params, job, next_device(), cb, g_single_ctx);
}

PLUGIN_EXPORT int32_t plugin_batch_encode(const char* input_dir, const char* output_dir,
										   grk_cparameters* params,
										   grk::PLUGIN_ENCODE_USER_CALLBACK cb)
{
	if(!input_dir || !output_dir || !params || !cb)
		return -1;
	join_batch();
	const std::string in = input_dir, out = output_dir;
	const grk_cparameters copy = *params;
	g_batch_done = false;
	// Frames are independent: a pool of workers takes them in directory order, each with a device
	// encoder of its own, so the H2D of one frame, the kernels of another and the host's rate
	// allocation + T2 + file write (the callback) of a third run at the same time.  The host's
	// callback works on its own codec and stream per call (grk_compress.cpp:1742-2160).
	g_batch_thread = std::thread([in, out, copy, cb]() {
		const char* ext = copy.cod_format == GRK_JP2_FMT ? "jp2" : "j2k";
		const std::vector<std::string> names = list_dir(in.c_str());
		// the host's callback (rate allocation + T2 + file write, ~150 ms per DCI-4K frame on one
		// core) bounds the batch: one worker per core, 16 at most (each holds a device encoder)
		unsigned nthreads = std::max(2u, std::min(16u, std::thread::hardware_concurrency()));
		if(const char* v = getenv("GRKB_BATCH_THREADS"))
			nthreads = std::max(1, atoi(v));
		nthreads = (unsigned)std::min<size_t>(nthreads, std::max<size_t>(names.size(), 1));
		const char* sv = getenv("GRKB_BATCH_SERIAL_CALLBACK");
		const bool serial_cb = sv && *sv && *sv != '0';
		std::atomic<size_t> next{0};
		std::atomic<uint32_t> failed{0};
		while(g_ctx_pool->size() < nthreads)
			g_ctx_pool->emplace_back(new EncodeCtx());
		auto worker = [&](unsigned slot) {
			EncodeCtx& ctx = *(*g_ctx_pool)[slot];
			for(;;)
			{
				const size_t i = next++;
				if(i >= names.size() || g_stop)
					break;
				grk_cparameters p = copy; // the callback may write to its parameters
				EncodeJob job;
				job.in = in + "/" + names[i];
				job.out = out + "/" + replace_ext(names[i], ext);
				strncpy(p.infile, job.in.c_str(), sizeof(p.infile) - 1);
				strncpy(p.outfile, job.out.c_str(), sizeof(p.outfile) - 1);
				// a worker stays on one device (deviceId = -1: workers spread over all of them)
				const int dev = g_device >= 0 ? g_device : (g_ndev > 0 ? (int)(slot % (unsigned)g_ndev) : 0);
				if(encode_one(&p, job, dev, cb, &ctx, serial_cb))
					failed++;
			}
		};
		std::vector<std::thread> pool;
		for(unsigned t = 1; t < nthreads; ++t)
			pool.emplace_back(worker, t);
		worker(0);
		for(auto& t : pool)
			t.join();
		if(failed)
			logf("batch encode: %u of %zu frames failed", failed.load(), names.size());
		g_batch_done = true;
	});
	return 0;
}

PLUGIN_EXPORT bool plugin_is_batch_complete(void)
{
	return g_batch_done;
}
PLUGIN_EXPORT void plugin_stop_batch_encode(void)
{
	g_stop = true; // async-signal tolerant: a flag only (grk_compress.cpp:79-82)
}

// ---- decode ------------------------------------------------------------------------------------

PLUGIN_EXPORT int32_t plugin_decompress(grk_decompress_parameters* params,
										 grk::PLUGIN_DECODE_USER_CALLBACK cb)
{
	if(!params || !cb)
		return -1;
	// one slot kept for this door too: the decoders, the page-locked file buffer and the host image
	// (whose planes are page-locked once) survive from call to call; concurrent callers take turns
	static std::mutex mu;
	std::lock_guard<std::mutex> lk(mu);
	if(!g_single_slot)
		g_single_slot = new DecoderSlot();
	return decode_one(params, params->infile, params->outfile, next_device(), cb, g_single_slot);
}

namespace
{
std::string g_dec_in, g_dec_out;
grk_decompress_parameters* g_dec_params = nullptr;
grk::PLUGIN_DECODE_USER_CALLBACK g_dec_cb = nullptr;
std::atomic<int> g_dec_failed{0}, g_dec_ok{0};

// extension of the output files, from the format the host asked for (params->cod_format)
const char* output_ext(const grk_decompress_parameters* p)
{
	switch(p ? p->cod_format : GRK_PXM_FMT)
	{
		case GRK_PGX_FMT:
			return "pgx";
		case GRK_BMP_FMT:
			return "bmp";
		case GRK_TIF_FMT:
			return "tif";
		case GRK_RAW_FMT:
			return "raw";
		case GRK_RAWL_FMT:
			return "rawl";
		case GRK_PNG_FMT:
			return "png";
		case GRK_JPG_FMT:
			return "jpg";
		default:
			return "ppm";
	}
}
} // namespace

PLUGIN_EXPORT int32_t plugin_init_batch_decompress(const char* input_dir, const char* output_dir,
													grk_decompress_parameters* params,
													grk::PLUGIN_DECODE_USER_CALLBACK cb)
{
	if(!input_dir || !output_dir || !params || !cb)
		return -1;
	g_dec_in = input_dir;
	g_dec_out = output_dir;
	g_dec_params = params;
	g_dec_cb = cb;
	g_dec_failed = 0;
	g_dec_ok = 0;
	return 0;
}

PLUGIN_EXPORT int32_t plugin_batch_decompress(void)
{
	if(!g_dec_cb)
		return -1;
	join_batch();
	g_batch_done = false;
	// frames are independent: workers take them in directory order, each with a device decoder of
	// its own; the host's T2 (first callback) and output stage (last callback) of different frames
	// run side by side, as do the copies and kernels on the device
	g_batch_thread = std::thread([]() {
		const std::vector<std::string> names = list_dir(g_dec_in.c_str());
		unsigned nthreads = std::max(2u, std::min(8u, std::thread::hardware_concurrency()));
		if(const char* v = getenv("GRKB_BATCH_THREADS"))
			nthreads = std::max(1, atoi(v));
		nthreads = (unsigned)std::min<size_t>(nthreads, std::max<size_t>(names.size(), 1));
		while(g_dec_pool->size() < nthreads)
			g_dec_pool->emplace_back(new DecoderSlot());
		std::atomic<size_t> next{0};
		auto worker = [&](unsigned slot) {
			for(;;)
			{
				const size_t i = next++;
				if(i >= names.size() || g_stop)
					break;
				const int rc = decode_one(g_dec_params, g_dec_in + "/" + names[i],
										  g_dec_out + "/" + replace_ext(names[i], output_ext(g_dec_params)),
										  g_device >= 0 ? g_device : (g_ndev > 0 ? (int)(slot % (unsigned)g_ndev) : 0),
										  g_dec_cb, (*g_dec_pool)[slot].get());
				if(rc)
				{
					// never silent: the frame has no output file
					++g_dec_failed;
					fprintf(stderr, "[grk_b200 plugin] batch decompress: %s was NOT decoded (%s)\n", names[i].c_str(),
							grkb_last_error());
				}
				else
					++g_dec_ok;
			}
		};
		std::vector<std::thread> pool;
		for(unsigned t = 1; t < nthreads; ++t)
			pool.emplace_back(worker, t);
		worker(0);
		for(auto& t : pool)
			t.join();
		if(g_dec_failed)
			fprintf(stderr, "[grk_b200 plugin] batch decompress: %d of %d frames failed\n", g_dec_failed.load(),
					g_dec_failed.load() + g_dec_ok.load());
		g_batch_done = true;
	});
	return 0;
}

PLUGIN_EXPORT void plugin_stop_batch_decompress(void)
{
	g_stop = true;
}
