// encode.cu -- host side of the encode path: device buffers, block table validation, launch of
// the block encoders (t1_mq_enc.cu, t1_ht_enc.cu) and of the stream packer, C-ABI entry points.
//
// Mirrors the reference's T1CompressScheduler (src/lib/jp2/t1/T1CompressScheduler.cpp:31-154):
// one flat table of blocks for all components of a tile, one launch per coder.
#include "pcrd.h"
#include "pipeline.h"
#include "t1.cuh"
#include <new>
#include <stdlib.h>
#include <vector>

namespace
{
struct DevBuf
{
	void* p = nullptr;
	size_t cap = 0;
	int reserve(size_t bytes)
	{
		if(bytes <= cap)
			return 0;
		if(p)
			cudaFree(p);
		p = nullptr;
		cap = 0;
		const size_t want = bytes + bytes / 4 + 256;
		if(cudaMalloc(&p, want) != cudaSuccess)
		{
			grkb_set_error("encode: cudaMalloc of %zu bytes failed", want);
			return -1;
		}
		cap = want;
		return 0;
	}
	void release()
	{
		if(p)
			cudaFree(p);
		p = nullptr;
		cap = 0;
	}
};

struct EncState
{
	int device = -1;
	cudaStream_t stream = nullptr;
	DevBuf blocks, coded, passes, slots, packed, misc, sym, planes[GRKB_MAX_COMPS];
	uint32_t sym_bytes_per_sample = 8; // symbol pool sizing, doubled when a tile runs out
	void release()
	{
		blocks.release();
		coded.release();
		passes.release();
		slots.release();
		packed.release();
		misc.release();
		sym.release();
		for(auto& p : planes)
			p.release();
		if(stream)
			cudaStreamDestroy(stream);
		stream = nullptr;
	}
	// the per-thread state of grkb_t1_encode_host goes with its thread (release() is idempotent; the
	// encoder objects release theirs in grkb_encoder_destroy)
	~EncState()
	{
		if(device >= 0 && (stream || blocks.p))
		{
			int cur = -1;
			if(cudaGetDevice(&cur) == cudaSuccess && cudaSetDevice(device) == cudaSuccess)
			{
				release();
				if(cur >= 0)
					cudaSetDevice(cur);
			}
			cudaGetLastError();
		}
	}
};

constexpr uint32_t SLOT_LEAD = 16;

int scan_enc_blocks(const grkb_enc_cblk* b, uint32_t n, uint32_t ncomp, uint32_t stride,
					uint32_t plane_h, int reversible, T1EncLaunch& l, bool& any_mq, bool& any_ht)
{
	l.max_area = l.max_flag_words = l.max_w = 0;
	l.total_area = 0;
	any_mq = any_ht = false;
	for(uint32_t i = 0; i < n; ++i)
	{
		const grkb_enc_cblk& c = b[i];
		if(c.w == 0 || c.h == 0 || (uint32_t)c.w * c.h > 4096 || c.w > 1024 || c.h > 1024 ||
		   c.comp >= ncomp || c.orient > 3 || (uint64_t)c.x + c.w > stride ||
		   (uint64_t)c.y + c.h > plane_h)
		{
			grkb_set_error("encode: block %u has a bad geometry", i);
			return -1;
		}
		if(c.cblk_sty & GRKB_CBLKSTY_HT)
		{
			if(!reversible)
			{
				grkb_set_error("encode: HT with the irreversible transform is not supported "
							   "(the reference's quantiser for it is broken, T1HT.cpp:92-104)");
				return -1;
			}
			any_ht = true;
		}
		else
		{
			if(!reversible && !(c.stepsize > 0.0f))
			{
				grkb_set_error("encode: block %u has no step size", i);
				return -1;
			}
			any_mq = true;
			l.total_area += (uint64_t)c.w * c.h;
			const uint32_t fw = ((uint32_t)c.w + 2) * ((((uint32_t)c.h + 3) >> 2) + 2);
			if(fw > l.max_flag_words)
				l.max_flag_words = fw;
		}
		if((uint32_t)c.w * c.h > l.max_area)
			l.max_area = (uint32_t)c.w * c.h;
		if(c.w > l.max_w)
			l.max_w = c.w;
	}
	uint32_t slot = SLOT_LEAD + l.max_area * 4 + 256;
	if(any_ht && slot < SLOT_LEAD + 16384 + 192 + 3072 + 64)
		slot = SLOT_LEAD + 16384 + 192 + 3072 + 64;
	l.slot_bytes = (slot + 15) & ~15u;
	return 0;
}

// blocks already validated; planes on the device
int run_encoders(EncState& st, T1EncLaunch& l, bool any_mq, bool any_ht)
{
	cudaStream_t s = st.stream;
	if(st.coded.reserve((size_t)l.nblocks * sizeof(grkb_coded)) ||
	   st.passes.reserve((size_t)l.nblocks * GRKB_ENC_MAX_PASSES * sizeof(grkb_pass)) ||
	   st.slots.reserve((size_t)l.nblocks * l.slot_bytes) || st.misc.reserve(64 + (size_t)l.nblocks * 8))
		return -1;
	l.coded = (grkb_coded*)st.coded.p;
	l.passes = (grkb_pass*)st.passes.p;
	l.slots = (uint8_t*)st.slots.p;
	l.packed_used = (uint64_t*)st.misc.p;
	l.counter = (uint32_t*)((uint8_t*)st.misc.p + 16);
	l.sym_ctrl = (uint32_t*)((uint8_t*)st.misc.p + 8); // read back together with packed_used
	l.code_ctrl = (uint32_t*)((uint8_t*)st.misc.p + 24);
	l.sym_total = (uint32_t*)((uint8_t*)st.misc.p + 64);
	l.code_order = l.sym_total + l.nblocks;
	if(any_mq)
	{
		// a symbol is one byte; a block hands its symbols over in 1 KB chunks
		const size_t want = (size_t)l.total_area * st.sym_bytes_per_sample + (size_t)l.nblocks * 2048 + 4096;
		if(st.sym.reserve(want))
			return -1;
		l.sym_pool = (uint8_t*)st.sym.p;
		l.sym_cap = st.sym.cap;
	}
	GRKB_CUDA_CHECK(cudaMemsetAsync(l.sym_ctrl, 0, 8, s));
	// every slot starts with zero bytes: the MQ coder reads the byte before the stream
	GRKB_CUDA_CHECK(cudaMemset2DAsync(l.slots, l.slot_bytes, 0, SLOT_LEAD, l.nblocks, s));
	if(any_mq && grkb_launch_t1_mq_encode(l, s))
		return -1;
	if(any_ht && grkb_launch_t1_ht_encode(l, s))
		return -1;
	return grkb_launch_t1_pack(l, s);
}

EncState& tls_state()
{
	static thread_local EncState st;
	return st;
}
} // namespace

extern "C" {

int grkb_t1_encode_host(const grkb_enc_cblk* blocks, uint32_t nblocks, const void* const* planes,
						uint32_t numcomps, uint32_t stride, uint32_t plane_h, int reversible,
						const double* mct_norms, uint32_t mct_numcomps, int rate_control,
						grkb_coded* coded, grkb_pass* passes, uint8_t* bytes, size_t cap,
						size_t* used)
{
	if((nblocks && (!blocks || !coded || !passes)) || !planes || numcomps == 0 ||
	   numcomps > GRKB_MAX_COMPS || (cap && !bytes) || mct_numcomps > GRKB_MAX_COMPS)
	{
		grkb_set_error("grkb_t1_encode_host: bad arguments");
		return -1;
	}
	if(used)
		*used = 0;
	if(nblocks == 0)
		return 0;
	EncState& st = tls_state();
	int dev = 0;
	GRKB_CUDA_CHECK(cudaGetDevice(&dev));
	if(st.device != dev)
	{
		st.release();
		st.device = dev;
	}
	if(!st.stream)
		GRKB_CUDA_CHECK(cudaStreamCreateWithFlags(&st.stream, cudaStreamNonBlocking));
	T1EncLaunch l{};
	bool any_mq, any_ht;
	if(scan_enc_blocks(blocks, nblocks, numcomps, stride, plane_h, reversible, l, any_mq, any_ht))
		return -1;
	cudaStream_t s = st.stream;
	const size_t plane_bytes = (size_t)stride * plane_h * 4;
	if(st.blocks.reserve((size_t)nblocks * sizeof(grkb_enc_cblk)) || st.packed.reserve(cap ? cap : 16))
		return -1;
	for(uint32_t c = 0; c < numcomps; ++c)
	{
		if(!planes[c])
		{
			grkb_set_error("grkb_t1_encode_host: plane %u is null", c);
			return -1;
		}
		if(st.planes[c].reserve(plane_bytes))
			return -1;
		GRKB_CUDA_CHECK(cudaMemcpyAsync(st.planes[c].p, planes[c], plane_bytes, cudaMemcpyHostToDevice, s));
		l.planes[c] = st.planes[c].p;
	}
	GRKB_CUDA_CHECK(cudaMemcpyAsync(st.blocks.p, blocks, (size_t)nblocks * sizeof(grkb_enc_cblk),
									cudaMemcpyHostToDevice, s));
	l.blocks = (const grkb_enc_cblk*)st.blocks.p;
	l.nblocks = nblocks;
	l.stride = stride;
	l.reversible = reversible;
	l.mct_numcomps = mct_norms ? mct_numcomps : 0;
	for(uint32_t c = 0; c < l.mct_numcomps; ++c)
		l.mct_norms[c] = mct_norms[c];
	l.rate_control = rate_control;
	l.packed = (uint8_t*)st.packed.p;
	l.packed_cap = cap;
	uint64_t ctrl[2] = {0, 0}; // packed bytes | symbol pool: chunks used, overflow flag
	for(int attempt = 0;; ++attempt)
	{
		if(run_encoders(st, l, any_mq, any_ht))
			return -1;
		GRKB_CUDA_CHECK(cudaMemcpyAsync(ctrl, l.packed_used, sizeof(ctrl), cudaMemcpyDeviceToHost, s));
		GRKB_CUDA_CHECK(cudaStreamSynchronize(s));
		if((ctrl[1] >> 32) == 0 || attempt == 4)
			break;
		st.sym_bytes_per_sample *= 2; // the symbol pool ran out: code the tile again with a larger one
	}
	if(ctrl[1] >> 32)
	{
		grkb_set_error("grkb_t1_encode_host: symbol pool exhausted");
		return -1;
	}
	const uint64_t total = ctrl[0];
	GRKB_CUDA_CHECK(cudaMemcpyAsync(coded, l.coded, (size_t)nblocks * sizeof(grkb_coded),
									cudaMemcpyDeviceToHost, s));
	GRKB_CUDA_CHECK(cudaMemcpyAsync(passes, l.passes,
									(size_t)nblocks * GRKB_ENC_MAX_PASSES * sizeof(grkb_pass),
									cudaMemcpyDeviceToHost, s));
	GRKB_CUDA_CHECK(cudaStreamSynchronize(s));
	if(total == ~(uint64_t)0)
	{
		grkb_set_error("grkb_t1_encode_host: output pool of %zu bytes is too small", cap);
		return -2;
	}
	if(total)
		GRKB_CUDA_CHECK(cudaMemcpy(bytes, l.packed, (size_t)total, cudaMemcpyDeviceToHost));
	if(used)
		*used = (size_t)total;
	return 0;
}

// ---- whole-tile encoder ------------------------------------------------------------------------

struct grkb_encoder
{
	grkb_tile_params p;
	grkb_rect tc;
	uint32_t W, H, stride, sstride;
	EncState st;
	void* coeff[GRKB_MAX_COMPS] = {};
	void* image[GRKB_MAX_COMPS] = {};
	void* scratch[GRKB_MAX_COMPS][2] = {};
	cudaEvent_t ev[4] = {};
	// last tile, for grkb_encoder_time_device
	T1EncLaunch last{};
	bool any_mq = false, any_ht = false, have_last = false;
	// tile in flight (grkb_encoder_submit .. grkb_encoder_wait)
	bool pending = false;
	uint32_t p_nblocks = 0;
	grkb_coded* p_coded = nullptr;
	grkb_pass* p_passes = nullptr;
	uint8_t* p_bytes = nullptr;
	size_t p_cap = 0;
	uint64_t* h_total = nullptr; // pinned
	void* packed_in = nullptr; // narrow input (row N3), allocated on first use
	size_t packed_in_cap = 0;
	// rate allocation to run behind the block coders of the next tile (pcrd.h), null: none
	grkb_pcrd* pcrd = nullptr;
	uint64_t pcrd_max_bytes = 0;
	uint32_t pcrd_csty = 0;
	uint32_t* pcrd_npass = nullptr;
	PcrdResult* pcrd_result = nullptr;
};

extern "C++" void grkb_encoder_set_pcrd(grkb_encoder* e, grkb_pcrd* p, uint64_t max_bytes, uint32_t csty,
									   uint32_t* npass_host, PcrdResult* result_host)
{
	e->pcrd = p;
	e->pcrd_max_bytes = max_bytes;
	e->pcrd_csty = csty;
	e->pcrd_npass = npass_host;
	e->pcrd_result = result_host;
}

static int encoder_front(grkb_encoder* e, cudaStream_t s)
{
	// DC shift + forward MCT + forward wavelet: components 0..2 as one group when the MCT couples
	// them, every other component on its own (mirror of the decoder's back end)
	const grkb_tile_params& p = e->p;
	uint32_t c = 0;
	while(c < p.numcomps)
	{
		DwtGroup g{};
		const bool grp3 = (c == 0 && p.numcomps >= 3 && p.mct);
		g.ncomp = grp3 ? 3 : 1;
		g.coeff_stride = e->stride;
		g.image_stride = e->stride;
		g.scratch_stride = e->sstride;
		int32_t sh[3];
		for(int i = 0; i < g.ncomp; ++i)
		{
			g.coeff[i] = e->coeff[c + i];
			g.image[i] = e->image[c + i];
			g.scratch[i][0] = e->scratch[c + i][0];
			g.scratch[i][1] = e->scratch[c + i][1];
		}
		for(int i = 0; i < 3; ++i)
			sh[i] = p.shift[c + (i < g.ncomp ? i : 0)];
		if(grkb_fdwt_device(g, e->tc, p.numres, p.reversible, grp3 ? 2 : 1, sh, s))
			return -1;
		c += g.ncomp;
	}
	return 0;
}

int grkb_encoder_create(const grkb_tile_params* p, grkb_encoder** out)
{
	if(!p || !out || p->numcomps == 0 || p->numcomps > GRKB_MAX_COMPS || p->x1 <= p->x0 ||
	   p->y1 <= p->y0 || p->numres == 0 || p->numres > 33)
	{
		grkb_set_error("grkb_encoder_create: bad parameters");
		return -1;
	}
	grkb_encoder* e = new(std::nothrow) grkb_encoder();
	if(!e)
		return -1;
	e->p = *p;
	e->tc = grkb_rect{p->x0, p->y0, p->x1, p->y1};
	e->W = p->x1 - p->x0;
	e->H = p->y1 - p->y0;
	e->stride = grkb_stride(e->W);
	if(cudaGetDevice(&e->st.device) != cudaSuccess)
	{
		grkb_set_error("grkb_encoder_create: no CUDA device (there is no CPU fallback)");
		delete e;
		return -1;
	}
	const size_t plane_bytes = (size_t)e->stride * e->H * 4;
	const size_t sbytes = grkb_dwt_scratch_bytes(e->tc, p->numres, &e->sstride);
	bool ok = cudaStreamCreateWithFlags(&e->st.stream, cudaStreamNonBlocking) == cudaSuccess;
	for(uint32_t c = 0; ok && c < p->numcomps; ++c)
	{
		ok = cudaMalloc(&e->coeff[c], plane_bytes) == cudaSuccess &&
			 cudaMalloc(&e->image[c], plane_bytes) == cudaSuccess &&
			 cudaMalloc(&e->scratch[c][0], sbytes ? sbytes : 16) == cudaSuccess &&
			 cudaMalloc(&e->scratch[c][1], sbytes ? sbytes : 16) == cudaSuccess &&
			 cudaMemsetAsync(e->image[c], 0, plane_bytes, e->st.stream) == cudaSuccess;
	}
	for(auto& v : e->ev)
		ok = ok && cudaEventCreate(&v) == cudaSuccess;
	ok = ok && cudaMallocHost(&e->h_total, 2 * sizeof(uint64_t)) == cudaSuccess;
	if(!ok || cudaStreamSynchronize(e->st.stream) != cudaSuccess)
	{
		grkb_set_error("grkb_encoder_create: device allocation failed: %s",
					   cudaGetErrorString(cudaGetLastError()));
		grkb_encoder_destroy(e);
		return -1;
	}
	*out = e;
	return 0;
}

void grkb_encoder_destroy(grkb_encoder* e)
{
	if(!e)
		return;
	cudaSetDevice(e->st.device);
	if(e->st.stream)
		cudaStreamSynchronize(e->st.stream);
	for(uint32_t c = 0; c < GRKB_MAX_COMPS; ++c)
	{
		cudaFree(e->coeff[c]);
		cudaFree(e->image[c]);
		cudaFree(e->scratch[c][0]);
		cudaFree(e->scratch[c][1]);
	}
	for(auto& v : e->ev)
		if(v)
			cudaEventDestroy(v);
	if(e->h_total)
		cudaFreeHost(e->h_total);
	if(e->packed_in)
		cudaFree(e->packed_in);
	e->st.release();
	delete e;
}

// the image arrives as int32 planes (image, stride) or in a narrow format (packed, format, ...)
static int submit_any(grkb_encoder* e, const int32_t* const* image, uint32_t stride, const void* packed,
					  uint32_t format, int sgnd, size_t row_bytes, const grkb_enc_cblk* blocks, uint32_t nblocks,
					  const double* mct_norms, uint32_t mct_numcomps, int rate_control, grkb_coded* coded,
					  grkb_pass* passes, uint8_t* bytes, size_t cap)
{
	if(!e || (!image && !packed) || (image && stride < e->W) || (nblocks && (!blocks || !coded || !passes)) ||
	   (cap && !bytes) || mct_numcomps > GRKB_MAX_COMPS)
	{
		grkb_set_error("grkb_encoder_submit: bad arguments");
		return -1;
	}
	if(e->pending)
	{
		grkb_set_error("grkb_encoder_submit: a tile is already in flight (call grkb_encoder_wait)");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(e->st.device));
	cudaStream_t s = e->st.stream;
	T1EncLaunch l{};
	bool any_mq = false, any_ht = false;
	if(nblocks && scan_enc_blocks(blocks, nblocks, e->p.numcomps, e->stride, e->H, e->p.reversible, l,
								  any_mq, any_ht))
		return -1;
	if(image)
	{
		for(uint32_t c = 0; c < e->p.numcomps; ++c)
		{
			if(!image[c])
			{
				grkb_set_error("grkb_encoder_encode: plane %u is null", c);
				return -1;
			}
			GRKB_CUDA_CHECK(cudaMemcpy2DAsync(e->image[c], (size_t)e->stride * 4, image[c],
											  (size_t)stride * 4, (size_t)e->W * 4, e->H,
											  cudaMemcpyHostToDevice, s));
		}
	}
	else
	{
		if(format == GRKB_SAMPLES_32 || row_bytes < grkb_samples_min_row_bytes(format, e->W, e->p.numcomps) ||
		   grkb_samples_min_row_bytes(format, e->W, e->p.numcomps) == 0)
		{
			grkb_set_error("grkb_encoder_submit_packed: bad format %u or rows of %zu bytes", format, row_bytes);
			return -1;
		}
		const size_t total = grkb_samples_bytes(format, e->H, e->p.numcomps, row_bytes);
		if(total > e->packed_in_cap)
		{
			if(e->packed_in)
				cudaFree(e->packed_in);
			e->packed_in = nullptr;
			e->packed_in_cap = 0;
			GRKB_CUDA_CHECK(cudaMalloc(&e->packed_in, total));
			e->packed_in_cap = total;
		}
		GRKB_CUDA_CHECK(cudaMemcpyAsync(e->packed_in, packed, total, cudaMemcpyHostToDevice, s));
		if(grkb_launch_unpack_samples(e->packed_in, row_bytes, format, sgnd, e->image, e->p.numcomps, e->stride,
									  e->W, e->H, s))
			return -1;
	}
	if(encoder_front(e, s))
		return -1;
	e->have_last = false;
	e->pending = true;
	// a tile counts as in flight only once everything is enqueued: any early return below gives the
	// encoder back
	struct Guard
	{
		grkb_encoder* e;
		bool ok = false;
		~Guard()
		{
			if(!ok)
				e->pending = false;
		}
	} guard{e};
	e->p_nblocks = nblocks;
	e->p_coded = coded;
	e->p_passes = passes;
	e->p_bytes = bytes;
	e->p_cap = cap;
	e->h_total[0] = e->h_total[1] = 0;
	if(nblocks == 0)
	{
		guard.ok = true;
		return 0;
	}
	if(e->st.blocks.reserve((size_t)nblocks * sizeof(grkb_enc_cblk)) ||
	   e->st.packed.reserve(cap ? cap : 16))
	{
		e->pending = false;
		return -1;
	}
	GRKB_CUDA_CHECK(cudaMemcpyAsync(e->st.blocks.p, blocks, (size_t)nblocks * sizeof(grkb_enc_cblk),
									cudaMemcpyHostToDevice, s));
	l.blocks = (const grkb_enc_cblk*)e->st.blocks.p;
	l.nblocks = nblocks;
	for(uint32_t c = 0; c < e->p.numcomps; ++c)
		l.planes[c] = e->coeff[c];
	l.stride = e->stride;
	l.reversible = e->p.reversible;
	l.mct_numcomps = mct_norms ? mct_numcomps : 0;
	for(uint32_t c = 0; c < l.mct_numcomps; ++c)
		l.mct_norms[c] = mct_norms[c];
	l.rate_control = rate_control;
	l.packed = (uint8_t*)e->st.packed.p;
	l.packed_cap = cap;
	if(run_encoders(e->st, l, any_mq, any_ht))
	{
		e->pending = false;
		return -1;
	}
	e->last = l;
	e->any_mq = any_mq;
	e->any_ht = any_ht;
	e->have_last = true;
	// the block table and the total come back with the kernels; the pass table and the bytes
	// are sized from them in grkb_encoder_wait
	GRKB_CUDA_CHECK(cudaMemcpyAsync(coded, l.coded, (size_t)nblocks * sizeof(grkb_coded),
									cudaMemcpyDeviceToHost, s));
	GRKB_CUDA_CHECK(cudaMemcpyAsync(e->h_total, l.packed_used, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
	if(e->pcrd && grkb_pcrd_enqueue(e->pcrd, l.coded, l.passes, e->pcrd_max_bytes, e->pcrd_csty, e->pcrd_npass,
									e->pcrd_result, s))
		return -1;
	guard.ok = true;
	return 0;
}

int grkb_encoder_submit(grkb_encoder* e, const int32_t* const* image, uint32_t stride,
						const grkb_enc_cblk* blocks, uint32_t nblocks, const double* mct_norms,
						uint32_t mct_numcomps, int rate_control, grkb_coded* coded,
						grkb_pass* passes, uint8_t* bytes, size_t cap)
{
	if(!image)
	{
		grkb_set_error("grkb_encoder_submit: bad arguments");
		return -1;
	}
	return submit_any(e, image, stride, nullptr, 0, 0, 0, blocks, nblocks, mct_norms, mct_numcomps, rate_control,
					  coded, passes, bytes, cap);
}

int grkb_encoder_submit_packed(grkb_encoder* e, const void* image, uint32_t format, int sgnd,
							   size_t row_bytes, const grkb_enc_cblk* blocks, uint32_t nblocks,
							   const double* mct_norms, uint32_t mct_numcomps, int rate_control,
							   grkb_coded* coded, grkb_pass* passes, uint8_t* bytes, size_t cap)
{
	if(!image)
	{
		grkb_set_error("grkb_encoder_submit_packed: bad arguments");
		return -1;
	}
	return submit_any(e, nullptr, 0, image, format, sgnd, row_bytes, blocks, nblocks, mct_norms, mct_numcomps,
					  rate_control, coded, passes, bytes, cap);
}

int grkb_encoder_wait(grkb_encoder* e, size_t* used)
{
	if(used)
		*used = 0;
	if(!e || !e->pending)
	{
		grkb_set_error("grkb_encoder_wait: no tile in flight");
		return -1;
	}
	e->pending = false;
	GRKB_CUDA_CHECK(cudaSetDevice(e->st.device));
	cudaStream_t s = e->st.stream;
	GRKB_CUDA_CHECK(cudaStreamSynchronize(s));
	if(e->p_nblocks == 0)
		return 0;
	for(int attempt = 0; (e->h_total[1] >> 32) != 0; ++attempt)
	{
		// the symbol pool of the MQ coder ran out: code the tile again with a larger one
		if(attempt == 4)
		{
			grkb_set_error("grkb_encoder_wait: symbol pool exhausted");
			return -1;
		}
		e->st.sym_bytes_per_sample *= 2;
		T1EncLaunch l = e->last;
		if(run_encoders(e->st, l, e->any_mq, e->any_ht))
			return -1;
		e->last = l;
		GRKB_CUDA_CHECK(cudaMemcpyAsync(e->p_coded, l.coded, (size_t)e->p_nblocks * sizeof(grkb_coded),
										cudaMemcpyDeviceToHost, s));
		GRKB_CUDA_CHECK(cudaMemcpyAsync(e->h_total, l.packed_used, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
		if(e->pcrd && grkb_pcrd_enqueue(e->pcrd, l.coded, l.passes, e->pcrd_max_bytes, e->pcrd_csty, e->pcrd_npass,
										e->pcrd_result, s))
			return -1;
		GRKB_CUDA_CHECK(cudaStreamSynchronize(s));
	}
	e->pcrd = nullptr;
	const uint64_t total = e->h_total[0];
	if(total == ~(uint64_t)0)
	{
		grkb_set_error("grkb_encoder_encode: output pool of %zu bytes is too small", e->p_cap);
		return -2;
	}
	// only the pass records that exist cross the bus: the first max(numpasses) of every row
	uint32_t maxp = 1;
	for(uint32_t i = 0; i < e->p_nblocks; ++i)
		if(e->p_coded[i].numpasses > maxp)
			maxp = e->p_coded[i].numpasses;
	if(maxp > GRKB_ENC_MAX_PASSES)
		maxp = GRKB_ENC_MAX_PASSES;
	const size_t pitch = (size_t)GRKB_ENC_MAX_PASSES * sizeof(grkb_pass);
	GRKB_CUDA_CHECK(cudaMemcpy2DAsync(e->p_passes, pitch, e->last.passes, pitch, (size_t)maxp * sizeof(grkb_pass),
									  e->p_nblocks, cudaMemcpyDeviceToHost, s));
	if(total)
		GRKB_CUDA_CHECK(cudaMemcpyAsync(e->p_bytes, e->last.packed, (size_t)total, cudaMemcpyDeviceToHost, s));
	GRKB_CUDA_CHECK(cudaStreamSynchronize(s));
	if(used)
		*used = (size_t)total;
	return 0;
}

int grkb_encoder_encode(grkb_encoder* e, const int32_t* const* image, uint32_t stride,
						const grkb_enc_cblk* blocks, uint32_t nblocks, const double* mct_norms,
						uint32_t mct_numcomps, int rate_control, grkb_coded* coded,
						grkb_pass* passes, uint8_t* bytes, size_t cap, size_t* used)
{
	if(used)
		*used = 0;
	if(grkb_encoder_submit(e, image, stride, blocks, nblocks, mct_norms, mct_numcomps, rate_control, coded,
						   passes, bytes, cap))
		return -1;
	return grkb_encoder_wait(e, used);
}

int grkb_encoder_time_device(grkb_encoder* e, uint32_t iters, uint32_t warmup, float ms[3])
{
	if(!e || !e->have_last || iters == 0 || !ms)
	{
		grkb_set_error("grkb_encoder_time_device: nothing to time");
		return -1;
	}
	GRKB_CUDA_CHECK(cudaSetDevice(e->st.device));
	cudaStream_t s = e->st.stream;
	double acc[3] = {0, 0, 0};
	for(uint32_t i = 0; i < warmup + iters; ++i)
	{
		// the image planes are still resident: the forward stages read them, never write them
		cudaEventRecord(e->ev[0], s);
		if(encoder_front(e, s))
			return -1;
		cudaEventRecord(e->ev[1], s);
		T1EncLaunch l = e->last;
		if(run_encoders(e->st, l, e->any_mq, e->any_ht))
			return -1;
		cudaEventRecord(e->ev[2], s);
		GRKB_CUDA_CHECK(cudaStreamSynchronize(s));
		if(i >= warmup)
		{
			float a = 0, b = 0;
			cudaEventElapsedTime(&a, e->ev[0], e->ev[1]);
			cudaEventElapsedTime(&b, e->ev[1], e->ev[2]);
			acc[0] += a + b;
			acc[1] += a;
			acc[2] += b;
		}
	}
	for(int k = 0; k < 3; ++k)
		ms[k] = (float)(acc[k] / iters);
	return 0;
}

} // extern "C"
