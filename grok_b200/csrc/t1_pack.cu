// t1_pack.cu -- packs the per-block streams the block encoders left in their slots back to back
// (block order) into one byte pool, so that the D2H copy moves coded bytes only.  The host side
// of the reference keeps one buffer per block (CompressCodeblock::paddedCompressedStream,
// src/lib/jp2/t1/Codeblock.h:156-172); the pool + data_off is this library's ABI for the same.
#include "t1.cuh"

namespace
{
constexpr int SCAN_THREADS = 1024;
constexpr uint32_t SLOT_LEAD = 16;

// exclusive prefix sum of coded[].data_len -> coded[].data_off; one CTA
__global__ void __launch_bounds__(SCAN_THREADS) pack_scan_kernel(const T1EncLaunch l)
{
	__shared__ uint64_t warp_sum[SCAN_THREADS / 32];
	__shared__ uint64_t base;
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if(threadIdx.x == 0)
		base = 0;
	__syncthreads();
	for(uint32_t start = 0; start < l.nblocks; start += SCAN_THREADS)
	{
		const uint32_t i = start + threadIdx.x;
		const uint64_t len = i < l.nblocks ? l.coded[i].data_len : 0;
		uint64_t incl = len;
		for(int o = 1; o < 32; o <<= 1)
		{
			const uint64_t t = __shfl_up_sync(0xffffffffu, incl, o);
			if((int)lane >= o)
				incl += t;
		}
		if(lane == 31)
			warp_sum[warp] = incl;
		__syncthreads();
		if(warp == 0)
		{
			uint64_t v = warp_sum[lane];
			for(int o = 1; o < 32; o <<= 1)
			{
				const uint64_t t = __shfl_up_sync(0xffffffffu, v, o);
				if((int)lane >= o)
					v += t;
			}
			warp_sum[lane] = v; // inclusive over warps
		}
		__syncthreads();
		const uint64_t off = base + (warp ? warp_sum[warp - 1] : 0) + incl - len;
		if(i < l.nblocks)
			l.coded[i].data_off = off;
		__syncthreads();
		if(threadIdx.x == 0)
			base += warp_sum[SCAN_THREADS / 32 - 1];
		__syncthreads();
	}
	if(threadIdx.x == 0)
		*l.packed_used = base <= l.packed_cap ? base : ~(uint64_t)0;
}

// one warp per block
__global__ void __launch_bounds__(256) pack_copy_kernel(const T1EncLaunch l)
{
	if(*l.packed_used == ~(uint64_t)0)
		return; // pool too small: nothing is copied, the host reports it
	const uint32_t bi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if(bi >= l.nblocks)
		return;
	const uint32_t lane = threadIdx.x & 31;
	const grkb_coded c = l.coded[bi];
	const uint8_t* src = l.slots + (size_t)bi * l.slot_bytes + SLOT_LEAD;
	uint8_t* dst = l.packed + c.data_off;
	const uint32_t len = c.data_len;
	// aligned 32-bit stores: the bytes up to the first aligned destination word one by one, then
	// every destination word from the two source words it straddles (the slot is 16-byte aligned,
	// so the straddle is the same for all words), then the tail
	const uint32_t head = min(len, (4u - ((uint32_t)reinterpret_cast<uintptr_t>(dst) & 3u)) & 3u);
	if(lane < head)
		dst[lane] = src[lane];
	const uint32_t nwords = (len - head) >> 2;
	const uint32_t* sw = reinterpret_cast<const uint32_t*>(src);
	uint32_t* dw = reinterpret_cast<uint32_t*>(dst + head);
	const uint32_t sh = 8u * head;
	for(uint32_t j = lane; j < nwords; j += 32)
		dw[j] = __funnelshift_r(sw[j], sw[j + 1], sh);
	const uint32_t done = head + 4u * nwords;
	if(done + lane < len)
		dst[done + lane] = src[done + lane];
}
} // namespace

int grkb_launch_t1_pack(const T1EncLaunch& l, cudaStream_t s)
{
	if(l.nblocks == 0)
	{
		GRKB_CUDA_CHECK(cudaMemsetAsync(l.packed_used, 0, sizeof(uint64_t), s));
		return 0;
	}
	pack_scan_kernel<<<1, SCAN_THREADS, 0, s>>>(l);
	GRKB_CUDA_CHECK(cudaGetLastError());
	pack_copy_kernel<<<(l.nblocks + 7) / 8, 256, 0, s>>>(l);
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
