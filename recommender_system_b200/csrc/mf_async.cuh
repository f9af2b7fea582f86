/*
 * csrc/mf_async.cuh -- mbarrier + bulk-copy (TMA 1-D) primitives for sm_100a.
 *
 * The loop kernels move whole factor rows (F doubles, 16-byte aligned, a
 * multiple of 16 bytes) from global to shared memory with one
 * cp.async.bulk per row; completion is tracked by an mbarrier's transaction
 * count, so no thread holds the data in registers while it is in flight.
 * Waits are bounded: a barrier that never completes traps instead of hanging
 * the device.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mf {

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
	return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

/* make the initialised barriers visible to the async proxy before first use */
__device__ __forceinline__ void mbar_fence_init()
{
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
	             : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
	const uint32_t a = smem_u32(bar);
	uint32_t done = 0;
	for (int spin = 0; spin < (1 << 24); spin++) {
		asm volatile(
		    "{\n\t.reg .pred p;\n\t"
		    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
		    "selp.u32 %0, 1, 0, p;\n\t}"
		    : "=r"(done)
		    : "r"(a), "r"(parity)
		    : "memory");
		if (done)
			return;
	}
	__trap();   /* a copy that never lands is a bug: fail the launch, do not hang the GPU */
}

/* global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned */
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
	asm volatile(
	    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
	        smem_u32(dst)),
	    "l"(src), "r"(bytes), "r"(smem_u32(bar))
	    : "memory");
}

} // namespace mf
