// bulk_copy.cuh -- 1-D bulk copies global -> shared memory through the TMA engine (cp.async.bulk, SASS UBLKCP) completing
// on an mbarrier: one thread issues the copy of a whole table, no thread spends registers or issue slots on moving it, and
// the CTA does other staging work until it needs the bytes.  Source, destination and size must be multiples of 16 bytes.
#pragma once
#include <cstdint>

namespace b200jpeg {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t arrivals)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); // visible to the async proxy before a copy names it
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) // one arrival + the bytes the copies will deliver
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void bulk_copy_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "B2J_WAIT_%=:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra B2J_DONE_%=;\n"
                 "bra B2J_WAIT_%=;\n"
                 "B2J_DONE_%=:\n"
                 "}" ::"r"(smem_u32(bar)),
                 "r"(parity)
                 : "memory");
}

} // namespace b200jpeg
