// tile.cu -- whole-tile decode on the device and the tile-level C-ABI (include/grk_b200.h,
// section B).
//
// One grkb_decoder owns, for one tile geometry, everything resident in HBM: per component a
// coefficient plane (reference Mallat layout, stride = width rounded up to 32 samples), a
// full-resolution image plane and two quarter-size ping-pong planes for the intermediate
// resolutions; the descriptor table, segment table and compressed-byte pool of the tile in
// flight; one stream.  A decode is: H2D(descriptors, bytes) -> T1 kernels (dequantisation fused
// into their store; every block writes its whole rectangle, so no clear pass is needed) ->
// one wavelet launch per level for all components -> inverse MCT + DC shift + clamp fused
// into the last level -> D2H(planes).  Mirrors the device half of
// TileProcessor::decompressT1 + mctDecompress (src/lib/jp2/tile/TileProcessor.cpp:334-383,
// :457-504).
#include "pipeline.h"
#include "t1.cuh"
#include <vector>
#include <string.h>
#include <mutex>
#include <new>

struct grkb_decoder
{
	grkb_tile_params p;
	grkb_rect tc;
	uint32_t W, H, stride, sstride;
	int device;
	cudaStream_t stream = nullptr;
	void* coeff[GRKB_MAX_COMPS] = {};
	void* image[GRKB_MAX_COMPS] = {};
	void* scratch[GRKB_MAX_COMPS][2] = {};
	grkb_cblk* d_blocks = nullptr;
	grkb_seg* d_segs = nullptr;
	uint8_t* d_bytes = nullptr;
	uint32_t* d_order = nullptr;
	uint32_t* d_counters = nullptr; // [0] MQ work counter, [1] HT work counter, [2] status
	uint32_t cap_blocks = 0, cap_segs = 0;
	size_t cap_bytes = 0;
	uint32_t* h_order = nullptr; // pinned
	uint32_t* h_status = nullptr; // pinned
	// last tile
	uint32_t nblocks = 0, nsegs = 0;
	uint32_t max_area = 0, max_flag_words = 0, max_ht_w = 0;
	bool any_mq = false, any_ht = false, ht_refine = false;
	cudaEvent_t ev[7] = {};
	cudaEvent_t busy = nullptr; // recorded behind the block decoders of every tile
	void* packed = nullptr; // narrow output (row N3), allocated on first use
	size_t packed_cap = 0;
};

namespace
{
int ensure(grkb_decoder* d, uint32_t nblocks, uint32_t nsegs, size_t nbytes)
{
	if(nblocks > d->cap_blocks)
	{
		const uint32_t n = nblocks + nblocks / 4 + 64;
		if(d->d_blocks)
			cudaFree(d->d_blocks);
		if(d->d_order)
			cudaFree(d->d_order);
		if(d->h_order)
			cudaFreeHost(d->h_order);
		d->d_blocks = nullptr;
		d->d_order = nullptr;
		d->h_order = nullptr;
		d->cap_blocks = 0;
		GRKB_CUDA_CHECK(cudaMalloc(&d->d_blocks, (size_t)n * sizeof(grkb_cblk)));
		GRKB_CUDA_CHECK(cudaMalloc(&d->d_order, (size_t)n * sizeof(uint32_t)));
		GRKB_CUDA_CHECK(cudaMallocHost(&d->h_order, (size_t)n * sizeof(uint32_t)));
		d->cap_blocks = n;
	}
	if(nsegs > d->cap_segs)
	{
		const uint32_t n = nsegs + nsegs / 4 + 64;
		if(d->d_segs)
			cudaFree(d->d_segs);
		d->d_segs = nullptr;
		d->cap_segs = 0;
		GRKB_CUDA_CHECK(cudaMalloc(&d->d_segs, (size_t)n * sizeof(grkb_seg)));
		d->cap_segs = n;
	}
	if(nbytes + 16 > d->cap_bytes)
	{
		const size_t n = nbytes + nbytes / 4 + 4096;
		if(d->d_bytes)
			cudaFree(d->d_bytes);
		d->d_bytes = nullptr;
		d->cap_bytes = 0;
		GRKB_CUDA_CHECK(cudaMalloc(&d->d_bytes, n));
		d->cap_bytes = n;
	}
	return 0;
}

// Validates the descriptor table against the geometry (nothing the kernels index may leave the
// buffers) and builds the longest-first processing order with a counting sort on the length.
int scan_blocks(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks, const grkb_seg* segs,
				uint32_t nsegs, size_t nbytes, uint32_t plane_w, uint32_t plane_h, uint32_t ncomp)
{
	d->max_area = 1;
	d->max_flag_words = 9;
	d->max_ht_w = 4;
	d->any_mq = d->any_ht = d->ht_refine = false;
	constexpr uint32_t NB = 1024;
	uint32_t hist[NB + 1];
	memset(hist, 0, sizeof(hist));
	for(uint32_t i = 0; i < nblocks; ++i)
	{
		const grkb_cblk& b = blocks[i];
		if(b.w == 0 || b.h == 0 || b.w > 1024 || b.h > 1024 || (uint32_t)b.w * b.h > 4096 ||
		   (uint64_t)b.x + b.w > plane_w || (uint64_t)b.y + b.h > plane_h || b.comp >= ncomp ||
		   b.orient > 3)
		{
			grkb_set_error("block %u: bad geometry (x %u y %u w %u h %u comp %u)", i, b.x, b.y, b.w,
						   b.h, b.comp);
			return -1;
		}
		if((uint64_t)b.data_off + b.data_len > nbytes ||
		   (uint64_t)b.seg_first + b.numsegs > nsegs)
		{
			grkb_set_error("block %u: bytes or segments out of range", i);
			return -1;
		}
		uint64_t total = 0;
		for(uint32_t s = 0; s < b.numsegs; ++s)
			total += segs[b.seg_first + s].len;
		if(total > b.data_len)
		{
			grkb_set_error("block %u: segment lengths exceed the block's bytes", i);
			return -1;
		}
		if(b.cblk_sty & GRKB_CBLKSTY_HT)
		{
			d->any_ht = true;
			if(b.numsegs > 1)
				d->ht_refine = true;
			if(b.w > d->max_ht_w)
				d->max_ht_w = b.w;
		}
		else
		{
			d->any_mq = true;
			const uint32_t area = (uint32_t)b.w * b.h;
			const uint32_t fw = ((uint32_t)b.w + 2) * ((((uint32_t)b.h + 3) >> 2) + 2);
			if(area > d->max_area)
				d->max_area = area;
			if(fw > d->max_flag_words)
				d->max_flag_words = fw;
		}
		uint32_t key = b.data_len >> 4;
		if(key >= NB)
			key = NB - 1;
		hist[NB - 1 - key]++; // descending length
	}
	uint32_t acc = 0;
	for(uint32_t k = 0; k < NB; ++k)
	{
		const uint32_t c = hist[k];
		hist[k] = acc;
		acc += c;
	}
	for(uint32_t i = 0; i < nblocks; ++i)
	{
		uint32_t key = blocks[i].data_len >> 4;
		if(key >= NB)
			key = NB - 1;
		d->h_order[hist[NB - 1 - key]++] = i;
	}
	return 0;
}

// The decoders of the process, so that one can tell whether others have block decoding in flight on
// its device: with several frames in flight the EBCOT kernel trades a frame's latency for throughput
// (t1_mq_dec.cu: two code blocks per warp).
std::mutex g_registry_mu;
std::vector<grkb_decoder*> g_registry;

uint32_t others_in_flight(const grkb_decoder* me)
{
	std::lock_guard<std::mutex> lk(g_registry_mu);
	uint32_t n = 0;
	for(const grkb_decoder* o : g_registry)
		if(o != me && o->device == me->device && o->busy && cudaEventQuery(o->busy) == cudaErrorNotReady)
			++n;
	(void)cudaGetLastError(); // (cudaErrorNotReady is not an error)
	return n;
}

int run_device(grkb_decoder* d, void* const* coeff_planes, uint32_t stride, bool wavelet,
			   cudaEvent_t* ev, uint32_t* launches)
{
	T1Launch l{};
	l.blocks = d->d_blocks;
	l.segs = d->d_segs;
	l.bytes = d->d_bytes;
	l.order = d->d_order;
	l.nblocks = d->nblocks;
	for(uint32_t c = 0; c < d->p.numcomps; ++c)
		l.planes[c] = coeff_planes[c];
	l.stride = stride;
	l.reversible = d->p.reversible;
	l.status = d->d_counters + 2;
	l.max_area = d->max_area;
	l.max_flag_words = d->max_flag_words;
	l.max_ht_w = d->max_ht_w;
	l.ht_refine = d->ht_refine ? 1u : 0u;
	l.concurrent = d->any_mq ? others_in_flight(d) : 0u;
	if(ev)
		cudaEventRecord(ev[0], d->stream);
	if(d->any_mq)
	{
		l.counter = d->d_counters;
		if(grkb_launch_t1_mq_decode(l, d->stream))
			return -1;
		if(launches)
			++*launches;
	}
	if(d->any_ht)
	{
		l.counter = d->d_counters + 1;
		if(grkb_launch_t1_ht_decode(l, d->stream))
			return -1;
		if(launches)
			++*launches;
	}
	if(ev)
		cudaEventRecord(ev[1], d->stream);
	if(d->busy)
		cudaEventRecord(d->busy, d->stream);
	if(!wavelet)
		return 0;
	// wavelet + component transform: components 0..2 as one group when the MCT couples them,
	// every other component on its own
	const grkb_tile_params& p = d->p;
	uint32_t c = 0;
	while(c < p.numcomps)
	{
		DwtGroup g{};
		const bool grp3 = (c == 0 && p.numcomps >= 3 && p.mct);
		g.ncomp = grp3 ? 3 : 1;
		g.coeff_stride = d->stride;
		g.image_stride = d->stride;
		g.scratch_stride = d->sstride;
		for(int i = 0; i < g.ncomp; ++i)
		{
			g.coeff[i] = d->coeff[c + i];
			g.image[i] = d->image[c + i];
			g.scratch[i][0] = d->scratch[c + i][0];
			g.scratch[i][1] = d->scratch[c + i][1];
		}
		int32_t sh[3], lo[3], hi[3];
		for(int i = 0; i < 3; ++i)
		{
			const uint32_t cc = c + (i < g.ncomp ? i : 0);
			sh[i] = p.shift[cc];
			lo[i] = p.lo[cc];
			hi[i] = p.hi[cc];
		}
		if(grkb_idwt_device(g, d->tc, p.numres, p.reversible, grp3 ? 2 : 1, sh, lo, hi, d->stream,
							(ev && c == 0) ? ev + 2 : nullptr, launches))
			return -1;
		c += g.ncomp;
	}
	if(ev)
		cudaEventRecord(ev[4], d->stream);
	return 0;
}
} // namespace

extern "C" {

int grkb_decoder_create(const grkb_tile_params* p, uint32_t max_blocks, size_t max_bytes,
						grkb_decoder** out)
{
	if(!p || !out || p->numcomps == 0 || p->numcomps > GRKB_MAX_COMPS || p->x1 <= p->x0 ||
	   p->y1 <= p->y0 || p->numres == 0 || p->numres > 33)
	{
		grkb_set_error("grkb_decoder_create: bad parameters");
		return -1;
	}
	grkb_decoder* d = new(std::nothrow) grkb_decoder();
	if(!d)
		return -1;
	d->p = *p;
	d->tc = grkb_rect{p->x0, p->y0, p->x1, p->y1};
	d->W = p->x1 - p->x0;
	d->H = p->y1 - p->y0;
	d->stride = grkb_stride(d->W);
	if(cudaGetDevice(&d->device) != cudaSuccess)
	{
		grkb_set_error("grkb_decoder_create: no CUDA device (there is no CPU fallback)");
		delete d;
		return -1;
	}
	const size_t plane_bytes = (size_t)d->stride * d->H * 4;
	const size_t sbytes = grkb_dwt_scratch_bytes(d->tc, p->numres, &d->sstride);
	auto fail = [&]() {
		grkb_decoder_destroy(d);
		return -1;
	};
#define TRY(expr)                                                                        \
	if((expr) != cudaSuccess)                                                            \
	{                                                                                    \
		grkb_set_error("grkb_decoder_create: %s failed: %s", #expr,                      \
					   cudaGetErrorString(cudaGetLastError()));                          \
		return fail();                                                                   \
	}
	TRY(cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking));
	for(uint32_t c = 0; c < p->numcomps; ++c)
	{
		TRY(cudaMalloc(&d->coeff[c], plane_bytes));
		TRY(cudaMalloc(&d->image[c], plane_bytes));
		TRY(cudaMalloc(&d->scratch[c][0], sbytes ? sbytes : 16));
		TRY(cudaMalloc(&d->scratch[c][1], sbytes ? sbytes : 16));
		TRY(cudaMemsetAsync(d->coeff[c], 0, plane_bytes, d->stream));
	}
	TRY(cudaMalloc(&d->d_counters, 4 * sizeof(uint32_t)));
	TRY(cudaMemsetAsync(d->d_counters, 0, 4 * sizeof(uint32_t), d->stream));
	TRY(cudaMallocHost(&d->h_status, sizeof(uint32_t)));
	for(auto& e : d->ev)
		TRY(cudaEventCreate(&e));
	TRY(cudaEventCreateWithFlags(&d->busy, cudaEventDisableTiming));
#undef TRY
	if(ensure(d, max_blocks ? max_blocks : 1, max_blocks ? max_blocks : 1, max_bytes))
		return fail();
	if(cudaStreamSynchronize(d->stream) != cudaSuccess)
		return fail();
	{
		std::lock_guard<std::mutex> lk(g_registry_mu);
		g_registry.push_back(d);
	}
	*out = d;
	return 0;
}

void grkb_decoder_destroy(grkb_decoder* d)
{
	if(!d)
		return;
	{
		std::lock_guard<std::mutex> lk(g_registry_mu);
		for(size_t i = 0; i < g_registry.size(); ++i)
			if(g_registry[i] == d)
			{
				g_registry.erase(g_registry.begin() + (long)i);
				break;
			}
	}
	if(d->stream)
		cudaStreamSynchronize(d->stream);
	if(d->busy)
		cudaEventDestroy(d->busy);
	for(uint32_t c = 0; c < GRKB_MAX_COMPS; ++c)
	{
		cudaFree(d->coeff[c]);
		cudaFree(d->image[c]);
		cudaFree(d->scratch[c][0]);
		cudaFree(d->scratch[c][1]);
	}
	cudaFree(d->d_blocks);
	cudaFree(d->d_segs);
	cudaFree(d->d_bytes);
	cudaFree(d->d_order);
	cudaFree(d->d_counters);
	cudaFree(d->packed);
	if(d->h_order)
		cudaFreeHost(d->h_order);
	if(d->h_status)
		cudaFreeHost(d->h_status);
	for(auto& e : d->ev)
		if(e)
			cudaEventDestroy(e);
	if(d->stream)
		cudaStreamDestroy(d->stream);
	delete d;
}

// copies in, all kernels, status out: everything of a submit but the results
static int submit_front(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks, const grkb_seg* segs,
						uint32_t nsegs, const uint8_t* bytes, size_t nbytes)
{
	if(!d || (nblocks && !blocks) || (nsegs && !segs) || (nbytes && !bytes))
	{
		grkb_set_error("grkb_decoder_submit: null argument");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(d->device));
	if(ensure(d, nblocks, nsegs, nbytes))
		return -1;
	if(scan_blocks(d, blocks, nblocks, segs, nsegs, nbytes, d->W, d->H, d->p.numcomps))
		return -1;
	d->nblocks = nblocks;
	d->nsegs = nsegs;
	cudaStream_t s = d->stream;
	if(nblocks)
	{
		GRKB_CUDA_CHECK(cudaMemcpyAsync(d->d_blocks, blocks, (size_t)nblocks * sizeof(grkb_cblk),
										cudaMemcpyHostToDevice, s));
		GRKB_CUDA_CHECK(cudaMemcpyAsync(d->d_order, d->h_order, (size_t)nblocks * sizeof(uint32_t),
										cudaMemcpyHostToDevice, s));
	}
	if(nsegs)
		GRKB_CUDA_CHECK(cudaMemcpyAsync(d->d_segs, segs, (size_t)nsegs * sizeof(grkb_seg),
										cudaMemcpyHostToDevice, s));
	if(nbytes)
		GRKB_CUDA_CHECK(cudaMemcpyAsync(d->d_bytes, bytes, nbytes, cudaMemcpyHostToDevice, s));
	GRKB_CUDA_CHECK(cudaMemsetAsync(d->d_counters + 2, 0, sizeof(uint32_t), s));
	if(run_device(d, d->coeff, d->stride, true, nullptr, nullptr))
		return -1;
	GRKB_CUDA_CHECK(cudaMemcpyAsync(d->h_status, d->d_counters + 2, sizeof(uint32_t),
									cudaMemcpyDeviceToHost, s));
	return 0;
}

int grkb_decoder_submit(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks,
						const grkb_seg* segs, uint32_t nsegs, const uint8_t* bytes, size_t nbytes,
						int32_t* const* planes, uint32_t stride)
{
	if(!d)
	{
		grkb_set_error("grkb_decoder_submit: null decoder");
		return -1;
	}
	return grkb_decoder_submit_region(d, blocks, nblocks, segs, nsegs, bytes, nbytes, planes, stride, 0, 0, d->W, d->H);
}

} // extern "C"

// grkb_decoder_submit that copies back only the rectangle (rx, ry, rw, rh) of the tile: planes[c]
// points at where that rectangle's first sample goes (windowed decode)
int grkb_decoder_submit_region(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks, const grkb_seg* segs,
							   uint32_t nsegs, const uint8_t* bytes, size_t nbytes, int32_t* const* planes,
							   uint32_t stride, uint32_t rx, uint32_t ry, uint32_t rw, uint32_t rh)
{
	if(submit_front(d, blocks, nblocks, segs, nsegs, bytes, nbytes))
		return -1;
	cudaStream_t s = d->stream;
	if(planes && rw && rh)
	{
		if(stride < rw || (uint64_t)rx + rw > d->W || (uint64_t)ry + rh > d->H)
		{
			grkb_set_error("grkb_decoder_submit: stride smaller than the rectangle, or rectangle outside the tile");
			return -1;
		}
		for(uint32_t c = 0; c < d->p.numcomps; ++c)
			GRKB_CUDA_CHECK(cudaMemcpy2DAsync(planes[c], (size_t)stride * 4,
											  (const int32_t*)d->image[c] + (size_t)ry * d->stride + rx,
											  (size_t)d->stride * 4, (size_t)rw * 4, rh, cudaMemcpyDeviceToHost, s));
	}
	return 0;
}

extern "C" {

size_t grkb_samples_size(uint32_t format, uint32_t H, uint32_t numcomps, size_t row_bytes)
{
	if(format == GRKB_SAMPLES_32)
		return row_bytes * H * numcomps;
	return grkb_samples_bytes(format, H, numcomps, row_bytes);
}

int grkb_decoder_submit_packed(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks,
							   const grkb_seg* segs, uint32_t nsegs, const uint8_t* bytes, size_t nbytes,
							   void* out, uint32_t format, size_t row_bytes)
{
	return grkb_decoder_submit_packed_region(d, blocks, nblocks, segs, nsegs, bytes, nbytes, out, format, row_bytes,
											 0);
}

} // extern "C"

// plane_bytes: distance between the component planes of a planar `out` (0: H rows, the tile is the
// whole buffer).  With it a tile lands in its rectangle of a larger image buffer.
int grkb_decoder_submit_packed_region(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks,
									  const grkb_seg* segs, uint32_t nsegs, const uint8_t* bytes, size_t nbytes,
									  void* out, uint32_t format, size_t row_bytes, size_t plane_bytes)
{
	if(!d || !out)
	{
		grkb_set_error("grkb_decoder_submit_packed: null argument");
		return -1;
	}
	const uint32_t base = format & ~(uint32_t)GRKB_SAMPLES_BIG_ENDIAN;
	const uint32_t width = (base == GRKB_SAMPLES_16 || base == GRKB_SAMPLES_16_INTERLEAVED) ? 16
						   : (base == GRKB_SAMPLES_8 || base == GRKB_SAMPLES_8_INTERLEAVED)   ? 8
																							   : 0;
	const bool interleaved = base == GRKB_SAMPLES_16_INTERLEAVED || base == GRKB_SAMPLES_8_INTERLEAVED;
	if(width == 0 || row_bytes < grkb_samples_min_row_bytes(format, d->W, d->p.numcomps))
	{
		grkb_set_error("grkb_decoder_submit_packed: bad format %u or rows of %zu bytes", format, row_bytes);
		return -1;
	}
	// the clamp of the last stage must already fit the format: [lo, hi] inside its range
	for(uint32_t c = 0; c < d->p.numcomps; ++c)
	{
		const int64_t lo = d->p.lo[c], hi = d->p.hi[c];
		const bool fits = lo >= 0 ? hi < ((int64_t)1 << width)
								  : (lo >= -((int64_t)1 << (width - 1)) && hi < ((int64_t)1 << (width - 1)));
		if(!fits)
		{
			grkb_set_error("grkb_decoder_submit_packed: component %u ranges over [%lld, %lld], wider than %u bits", c,
						   (long long)lo, (long long)hi, width);
			return -1;
		}
	}
	GRKB_CUDA_CHECK(cudaSetDevice(d->device));
	// on the device the tile is packed densely (rows of `need` bytes rounded up to 16)
	const size_t need = grkb_samples_min_row_bytes(format, d->W, d->p.numcomps);
	const size_t drow = plane_bytes ? ((need + 15) & ~(size_t)15) : row_bytes;
	const size_t total = grkb_samples_bytes(format, d->H, d->p.numcomps, drow);
	if(total > d->packed_cap)
	{
		if(d->packed)
			cudaFree(d->packed);
		d->packed = nullptr;
		d->packed_cap = 0;
		GRKB_CUDA_CHECK(cudaMalloc(&d->packed, total));
		d->packed_cap = total;
	}
	if(submit_front(d, blocks, nblocks, segs, nsegs, bytes, nbytes))
		return -1;
	cudaStream_t s = d->stream;
	if(grkb_launch_pack_samples(d->image, d->p.numcomps, d->stride, d->W, d->H, format, d->packed, drow, s))
		return -1;
	if(plane_bytes)
	{
		const uint32_t nplanes = interleaved ? 1 : d->p.numcomps;
		for(uint32_t c = 0; c < nplanes; ++c)
			GRKB_CUDA_CHECK(cudaMemcpy2DAsync((uint8_t*)out + (size_t)c * plane_bytes, row_bytes,
											  (const uint8_t*)d->packed + (size_t)c * d->H * drow, drow, need, d->H,
											  cudaMemcpyDeviceToHost, s));
		return 0;
	}
	// padding between the caller's rows stays untouched
	if(need == row_bytes)
		GRKB_CUDA_CHECK(cudaMemcpyAsync(out, d->packed, total, cudaMemcpyDeviceToHost, s));
	else
		GRKB_CUDA_CHECK(cudaMemcpy2DAsync(out, row_bytes, d->packed, row_bytes, need, total / row_bytes,
										  cudaMemcpyDeviceToHost, s));
	return 0;
}

extern "C" {

int grkb_decoder_wait(grkb_decoder* d, uint32_t* status_out)
{
	if(!d)
	{
		grkb_set_error("grkb_decoder_wait: null decoder");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(d->device));
	GRKB_CUDA_CHECK(cudaStreamSynchronize(d->stream));
	if(status_out)
		*status_out = *d->h_status;
	return 0;
}

int grkb_decoder_decode(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks,
						const grkb_seg* segs, uint32_t nsegs, const uint8_t* bytes, size_t nbytes,
						int32_t* const* planes, uint32_t stride, uint32_t* status_out)
{
	if(grkb_decoder_submit(d, blocks, nblocks, segs, nsegs, bytes, nbytes, planes, stride))
		return -1;
	return grkb_decoder_wait(d, status_out);
}

int grkb_decoder_download(grkb_decoder* d, int32_t* const* planes, uint32_t stride)
{
	if(!d || !planes || stride < d->W)
	{
		grkb_set_error("grkb_decoder_download: bad arguments");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(d->device));
	for(uint32_t c = 0; c < d->p.numcomps; ++c)
		GRKB_CUDA_CHECK(cudaMemcpy2DAsync(planes[c], (size_t)stride * 4, d->image[c],
										  (size_t)d->stride * 4, (size_t)d->W * 4, d->H,
										  cudaMemcpyDeviceToHost, d->stream));
	GRKB_CUDA_CHECK(cudaStreamSynchronize(d->stream));
	return 0;
}

int grkb_decoder_time_device(grkb_decoder* d, uint32_t iters, uint32_t warmup, float ms[4],
							 uint32_t* launches_out)
{
	if(!d || iters == 0 || !ms)
	{
		grkb_set_error("grkb_decoder_time_device: bad arguments");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(d->device));
	cudaStream_t s = d->stream;
	uint32_t launches = 0;
	for(uint32_t i = 0; i < warmup; ++i)
		if(run_device(d, d->coeff, d->stride, true, nullptr, nullptr))
			return -1;
	// whole passes back to back between two events
	cudaEvent_t t0 = d->ev[5], t1 = d->ev[6];
	GRKB_CUDA_CHECK(cudaEventRecord(t0, s));
	for(uint32_t i = 0; i < iters; ++i)
	{
		launches = 0;
		if(run_device(d, d->coeff, d->stride, true, nullptr, &launches))
			return -1;
	}
	GRKB_CUDA_CHECK(cudaEventRecord(t1, s));
	GRKB_CUDA_CHECK(cudaEventSynchronize(t1));
	float total = 0;
	GRKB_CUDA_CHECK(cudaEventElapsedTime(&total, t0, t1));
	ms[0] = total / (float)iters;
	// stage shares: the same passes again, each bracketed by its own events
	double t_t1 = 0, t_dwt = 0, t_last = 0;
	for(uint32_t i = 0; i < iters; ++i)
	{
		if(run_device(d, d->coeff, d->stride, true, d->ev, nullptr))
			return -1;
		GRKB_CUDA_CHECK(cudaEventSynchronize(d->ev[4]));
		float a = 0, b = 0, c = 0;
		GRKB_CUDA_CHECK(cudaEventElapsedTime(&a, d->ev[0], d->ev[1]));
		GRKB_CUDA_CHECK(cudaEventElapsedTime(&b, d->ev[1], d->ev[4]));
		if(d->p.numres > 1)
			GRKB_CUDA_CHECK(cudaEventElapsedTime(&c, d->ev[2], d->ev[3]));
		t_t1 += a;
		t_dwt += b;
		t_last += c;
	}
	ms[1] = (float)(t_t1 / iters);
	ms[2] = (float)(t_dwt / iters);
	ms[3] = (float)(t_last / iters);
	if(launches_out)
		*launches_out = launches;
	return 0;
}

int grkb_t1_decode_host(const grkb_cblk* blocks, uint32_t nblocks, const grkb_seg* segs,
						uint32_t nsegs, const uint8_t* bytes, size_t nbytes, void* const* planes,
						uint32_t numcomps, uint32_t stride, uint32_t plane_h, int reversible,
						uint32_t* status_out)
{
	if(numcomps == 0 || numcomps > GRKB_MAX_COMPS || !planes || stride == 0 || plane_h == 0)
	{
		grkb_set_error("grkb_t1_decode_host: bad arguments");
		return -1;
	}
	grkb_tile_params p{};
	p.x0 = p.y0 = 0;
	p.x1 = stride;
	p.y1 = plane_h;
	p.numcomps = numcomps;
	p.numres = 1;
	p.reversible = reversible;
	grkb_decoder* d = nullptr;
	if(grkb_decoder_create(&p, nblocks, nbytes, &d))
		return -1;
	int rc = -1;
	do
	{
		// the dense host planes become the device coefficient planes (device stride may be wider)
		for(uint32_t c = 0; c < numcomps; ++c)
			if(cudaMemcpy2DAsync(d->coeff[c], (size_t)d->stride * 4, planes[c], (size_t)stride * 4,
								 (size_t)stride * 4, plane_h, cudaMemcpyHostToDevice,
								 d->stream) != cudaSuccess)
			{
				grkb_set_error("grkb_t1_decode_host: H2D failed");
				break;
			}
		if(ensure(d, nblocks, nsegs, nbytes) ||
		   scan_blocks(d, blocks, nblocks, segs, nsegs, nbytes, stride, plane_h, numcomps))
			break;
		d->nblocks = nblocks;
		d->nsegs = nsegs;
		cudaStream_t s = d->stream;
		if(nblocks)
		{
			cudaMemcpyAsync(d->d_blocks, blocks, (size_t)nblocks * sizeof(grkb_cblk),
							cudaMemcpyHostToDevice, s);
			cudaMemcpyAsync(d->d_order, d->h_order, (size_t)nblocks * sizeof(uint32_t),
							cudaMemcpyHostToDevice, s);
		}
		if(nsegs)
			cudaMemcpyAsync(d->d_segs, segs, (size_t)nsegs * sizeof(grkb_seg), cudaMemcpyHostToDevice, s);
		if(nbytes)
			cudaMemcpyAsync(d->d_bytes, bytes, nbytes, cudaMemcpyHostToDevice, s);
		cudaMemsetAsync(d->d_counters + 2, 0, sizeof(uint32_t), s);
		if(run_device(d, d->coeff, d->stride, false, nullptr, nullptr))
			break;
		cudaMemcpyAsync(d->h_status, d->d_counters + 2, sizeof(uint32_t), cudaMemcpyDeviceToHost, s);
		for(uint32_t c = 0; c < numcomps; ++c)
			cudaMemcpy2DAsync(planes[c], (size_t)stride * 4, d->coeff[c], (size_t)d->stride * 4,
							  (size_t)stride * 4, plane_h, cudaMemcpyDeviceToHost, s);
		cudaError_t e = cudaStreamSynchronize(s);
		if(e != cudaSuccess)
		{
			grkb_set_error("grkb_t1_decode_host: %s", cudaGetErrorString(e));
			break;
		}
		if(status_out)
			*status_out = *d->h_status;
		rc = 0;
	} while(0);
	grkb_decoder_destroy(d);
	return rc;
}

void* grkb_host_alloc(size_t bytes)
{
	void* p = nullptr;
	if(cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess)
	{
		grkb_set_error("grkb_host_alloc(%zu) failed", bytes);
		cudaGetLastError();
		return nullptr;
	}
	return p;
}

void grkb_host_free(void* p)
{
	if(p)
		cudaFreeHost(p);
}

int grkb_host_register(void* p, size_t bytes)
{
	if(!p || !bytes)
		return -1;
	// may fail harmlessly (small buffers sharing a page with registered memory): the copies then go
	// through the driver's staging buffer; the error must not linger for the next launch check
	const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
	if(e != cudaSuccess)
	{
		cudaGetLastError();
		grkb_set_error("grkb_host_register: %s", cudaGetErrorString(e));
		return -1;
	}
	return 0;
}

int grkb_host_unregister(void* p)
{
	if(!p)
		return -1;
	if(cudaHostUnregister(p) != cudaSuccess)
	{
		cudaGetLastError();
		return -1;
	}
	return 0;
}

} // extern "C"
