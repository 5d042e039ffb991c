// bulk.cuh -- mbarrier + bulk asynchronous copy (cp.async.bulk, SASS UBLKCP) helpers shared by
// the strip kernels of the inverse and forward wavelet transforms (sm_100a).
#pragma once
#include <stdint.h>

static __device__ __forceinline__ uint32_t smem_u32(const void* p)
{
	return (uint32_t)__cvta_generic_to_shared(p);
}
static __device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
static __device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
static __device__ __forceinline__ void mbar_arrive_tx(uint32_t bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
				 : "memory");
}
static __device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity)
{
	uint32_t ok;
	asm volatile("{\n\t.reg .pred p;\n\t"
				 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
				 "selp.u32 %0, 1, 0, p;\n\t}"
				 : "=r"(ok)
				 : "r"(bar), "r"(parity)
				 : "memory");
	return ok != 0;
}
static __device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
	while(!mbar_try(bar, parity))
	{
	}
}
static __device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
				 "l"(src), "r"(bytes), "r"(bar)
				 : "memory");
}
