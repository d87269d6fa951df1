// Tensor memory (TMEM) as a per-lane constant store.  Blackwell's 256 KB of tensor memory per SM is reached through
// its own datapath (tcgen05.ld / tcgen05.st), not through the shared-memory pipe, which is what bounds the NoiseSup
// kernels (75 % of its wavefront peak, a quarter of it the dense 25 x 128 Gamma bank each lane re-reads every frame).
// A warp reads the 32 TMEM lanes of its quadrant (warp id % 4), lane l its own row, N consecutive 32-bit columns per
// instruction -- exactly the access pattern of a per-lane coefficient table.  sm_100a PTX.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace se_tmem {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// one whole warp: allocate `cols` columns (a power of two >= 32) and leave the base address in shared memory
__device__ __forceinline__ void alloc(uint32_t *smem_dst, uint32_t cols)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void dealloc(uint32_t taddr, uint32_t cols)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// address of column `col` in the lane quadrant of the calling warp
__device__ __forceinline__ uint32_t warp_addr(uint32_t base, int warp, int col)
{
    return base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)col;
}

__device__ __forceinline__ void st4(uint32_t taddr, float a, float b, float c, float d)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(__float_as_uint(a)),
                 "r"(__float_as_uint(b)), "r"(__float_as_uint(c)), "r"(__float_as_uint(d))
                 : "memory");
}
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// loads are asynchronous: the registers are valid after wait_ld, which takes them as operands so that no use of them
// can be scheduled above it
__device__ __forceinline__ void ld8(uint32_t taddr, uint32_t (&r)[8])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void wait_ld8(uint32_t (&r)[8])
{
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7])
                 :
                 : "memory");
}
__device__ __forceinline__ void ld4(uint32_t taddr, uint32_t (&r)[4])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void wait_ld4(uint32_t (&r)[4])
{
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]) : : "memory");
}

}  // namespace se_tmem
