// encode.cu -- host side of the encode path: device buffers, block table validation, launch of
// the block encoders (t1_mq_enc.cu, t1_ht_enc.cu) and of the stream packer, C-ABI entry points.
//
// Mirrors the reference's T1CompressScheduler (src/lib/jp2/t1/T1CompressScheduler.cpp:31-154):
// one flat table of blocks for all components of a tile, one launch per coder.
#include "pipeline.h"
#include "t1.cuh"
#include <new>
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
	DevBuf blocks, coded, passes, slots, packed, misc, planes[GRKB_MAX_COMPS];
	void release()
	{
		blocks.release();
		coded.release();
		passes.release();
		slots.release();
		packed.release();
		misc.release();
		for(auto& p : planes)
			p.release();
		if(stream)
			cudaStreamDestroy(stream);
		stream = nullptr;
	}
};

constexpr uint32_t SLOT_LEAD = 16;

int scan_enc_blocks(const grkb_enc_cblk* b, uint32_t n, uint32_t ncomp, uint32_t stride,
					uint32_t plane_h, int reversible, T1EncLaunch& l, bool& any_mq, bool& any_ht)
{
	l.max_area = l.max_flag_words = l.max_w = 0;
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
	   st.slots.reserve((size_t)l.nblocks * l.slot_bytes) || st.misc.reserve(64))
		return -1;
	l.coded = (grkb_coded*)st.coded.p;
	l.passes = (grkb_pass*)st.passes.p;
	l.slots = (uint8_t*)st.slots.p;
	l.packed_used = (uint64_t*)st.misc.p;
	l.counter = (uint32_t*)((uint8_t*)st.misc.p + 16);
	l.order = nullptr;
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
	if(run_encoders(st, l, any_mq, any_ht))
		return -1;
	uint64_t total = 0;
	GRKB_CUDA_CHECK(cudaMemcpyAsync(coded, l.coded, (size_t)nblocks * sizeof(grkb_coded),
									cudaMemcpyDeviceToHost, s));
	GRKB_CUDA_CHECK(cudaMemcpyAsync(passes, l.passes,
									(size_t)nblocks * GRKB_ENC_MAX_PASSES * sizeof(grkb_pass),
									cudaMemcpyDeviceToHost, s));
	GRKB_CUDA_CHECK(cudaMemcpyAsync(&total, l.packed_used, sizeof(total), cudaMemcpyDeviceToHost, s));
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

} // extern "C"
