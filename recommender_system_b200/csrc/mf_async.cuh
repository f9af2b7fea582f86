/*
 * csrc/mf_async.cuh -- asynchronous global -> shared copies for sm_100a.
 *
 * Factor rows are gathered (one row per nonzero, 240 B .. 8 KB each).  Measured
 * on B200 (profiles/r01_v2_*): one cp.async.bulk (TMA) per row sustains only
 * about one copy per ~100 cycles per SM -- 16 B/clk for 1600-byte rows -- so
 * gathers of rows this small go through the LSU instead: 16-byte cp.async
 * (LDGSTS), which needs no registers for the data in flight and reaches the
 * L2 -> SM bandwidth.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mf {

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
	return (uint32_t)__cvta_generic_to_shared(p);
}

/* 16-byte asynchronous copy, L2 only (the rows are not re-read through L1) */
__device__ __forceinline__ void cp_async16(void *dst, const void *src)
{
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}

/* the same with the destination as a shared-window address and a predicate (no branch) */
__device__ __forceinline__ void cp_async16_if(uint32_t dst, const void *src, bool pred)
{
	asm volatile(
	    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t"
	    "@p cp.async.cg.shared.global [%0], [%1], 16;\n\t}"
	    ::"r"(dst), "l"(src), "r"((int)pred)
	    : "memory");
}

__device__ __forceinline__ void cp_async_commit()
{
	asm volatile("cp.async.commit_group;" ::: "memory");
}

template <int N>
__device__ __forceinline__ void cp_async_wait()
{
	asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

/* 8-byte asynchronous copy (for operands that are only 8-byte aligned) */
__device__ __forceinline__ void cp_async8(void *dst, const void *src)
{
	asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}

__device__ __forceinline__ void cp_async4(void *dst, const void *src)
{
	asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}

/* ---- mbarriers in shared memory (CTA scope): producer / consumer hand-over of ring stages ---- */

__device__ __forceinline__ void mbar_init(uint32_t bar, int count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}

/* release: everything this thread read or wrote before is ordered before the arrival */
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

/* one (pre-counted) arrival that happens when every cp.async this thread has issued so far has
 * landed: the copies of a stage signal their own completion, nobody waits for them */
__device__ __forceinline__ void mbar_arrive_on_copies(uint32_t bar)
{
	asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
	uint32_t ok;
	asm volatile(
	    "{\n\t.reg .pred p;\n\t"
	    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
	    "selp.u32 %0, 1, 0, p;\n\t}"
	    : "=r"(ok)
	    : "r"(bar), "r"(parity)
	    : "memory");
	return ok != 0;
}

/* wait for the phase with the given parity to complete; a wait that cannot end (a protocol
 * error) traps after ~10^8 polls instead of hanging the device */
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
	unsigned spins = 0;
	while (!mbar_try_wait(bar, parity))
		if (++spins > (1u << 27))
			__trap();
}


/* ---- waiting for another GPU ------------------------------------------------------------
 * A wait for a peer is bounded by TIME (%globaltimer, nanoseconds), not by a poll count: under a
 * profiler's replay, a debugger or a peer delayed by a page fault a poll count runs out long
 * before anything is wrong.  A wait that does run out sets the error word of the rank's sync
 * block and gives up (the kernel finishes with garbage in the affected elements); the host finds
 * the word after its next sync and reports MFGPU_ECOMM -- no trap, no poisoned context. */
__device__ __forceinline__ unsigned long long global_ns()
{
	unsigned long long t;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
	return t;
}

struct WaitCtl {
	unsigned long long *err;           /* error word of this rank's sync block */
	unsigned long long timeout_ns;
};

/* true: keep polling; false: out of time (the error word is set) */
__device__ __forceinline__ bool wait_more(const WaitCtl &w, unsigned &spins, unsigned long long &t0, int who)
{
	if ((++spins & 1023u) != 0)
		return true;
	if (*(volatile unsigned long long *)w.err)         /* somebody here gave up already: do not wait again */
		return false;
	const unsigned long long now = global_ns();
	if (t0 == 0)
		t0 = now;
	if (now - t0 <= w.timeout_ns)
		return true;
	atomicCAS(w.err, 0ull, (unsigned long long)(who + 1));
	return false;
}

/* (value, tag) pairs of the low-latency exchange (mf_iter.cu): a pair travels as ONE 16-byte
 * store and is read by ONE 16-byte load, so a reader that sees the tag it waits for also has the
 * value -- no fence, no flag of its own (the idea of NCCL's LL protocol, with 8-byte payloads).
 * "One" in the sense of the PTX memory model: the accesses use the SCALAR type .b128 (ld/st
 * .b128, PTX ISA 8.3, sm_70+) with .relaxed.sys semantics -- one strong memory operation each
 * (SASS: STG.E.128.STRONG.SYS / LDG.E.128.STRONG.SYS) -- and not the vector form .v2.u64 of
 * round 1, which the model defines as two independent 8-byte operations that a reader may see
 * one without the other.  tests/rank_mode_worker.py soaks the exchange on top of that (replicas
 * hashed against each other every few iterations). */
__device__ __forceinline__ void st_pair(double *p, double v, unsigned long long tag)
{
	asm volatile("{\n\t.reg .b128 q;\n\tmov.b128 q, {%1, %2};\n\tst.relaxed.sys.global.b128 [%0], q;\n\t}"
	             ::"l"(p), "l"(__double_as_longlong(v)), "l"(tag) : "memory");
}

/* one look at a pair (the callers issue several before they use the first) */
__device__ __forceinline__ void ld_pair_try(const double *p, unsigned long long &v, unsigned long long &t)
{
	asm volatile("{\n\t.reg .b128 q;\n\tld.relaxed.sys.global.b128 q, [%2];\n\tmov.b128 {%0, %1}, q;\n\t}"
	             : "=l"(v), "=l"(t) : "l"(p) : "memory");
}

__device__ __forceinline__ double ld_pair_when(const double *p, unsigned long long tag, const WaitCtl &w, int who)
{
	unsigned long long v, t;
	unsigned spins = 0;
	unsigned long long t0 = 0;
	for (;;) {
		asm volatile("{\n\t.reg .b128 q;\n\tld.relaxed.sys.global.b128 q, [%2];\n\tmov.b128 {%0, %1}, q;\n\t}"
		             : "=l"(v), "=l"(t) : "l"(p) : "memory");
		if (t == tag)
			return __longlong_as_double((long long)v);
		if (!wait_more(w, spins, t0, who))
			return 0.0;
	}
}

} // namespace mf
