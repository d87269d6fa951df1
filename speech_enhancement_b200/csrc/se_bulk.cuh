// 1-D bulk-async copies (the TMA unit's cp.async.bulk, global -> shared memory, completion on an mbarrier) for the
// int16 input stream of the NoiseSup kernels: one elected lane posts a whole 160- / 320-byte block per instruction
// several blocks ahead; nothing waits for DRAM and no register holds prefetched samples.  sm_90+ PTX.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace se_bulk {

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(void *bar, int arrivals)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(arrivals) : "memory");
}

// makes the initialised barriers visible to the async proxy (the copy engine completes transactions on them)
__device__ __forceinline__ void mbar_init_fence()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// one arrival + `bytes` expected, then the copy: src and dst 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, void *bar)
{
    const uint32_t b = smem_addr(bar);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(b)
                 : "memory");
}

// plain arrival (no transaction bytes): completes a phase whose data was written with ordinary stores
__device__ __forceinline__ void mbar_arrive(void *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}

// orders this thread's ordinary shared-memory accesses before later async-proxy (bulk copy) accesses to the same bytes
__device__ __forceinline__ void proxy_fence()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    return done != 0;
}

// Every lane polls and the warp votes, so the loop condition is warp-uniform: a lane-0-only (or per-lane) spin makes
// the compiler treat everything after it as possibly diverged and wrap each later shuffle in WARPSYNC / ENDCOLLECTIVE
// (+25 % instructions in the NoiseSup kernels).
__device__ __forceinline__ void mbar_wait_warp(void *bar, uint32_t parity)
{
    const uint32_t b = smem_addr(bar);
    while (!__all_sync(0xffffffffu, mbar_try_wait(b, parity))) {}
}

}  // namespace se_bulk
