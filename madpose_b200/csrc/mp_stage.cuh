// Staging helpers: 1-D bulk asynchronous copies (TMA engine, UBLKCP in SASS) from
// global memory into shared memory, completion tracked by an mbarrier.
#pragma once
#include <cstdint>

namespace mp {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra.uni WAIT_DONE;\n"
        "bra.uni WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// bytes must be a multiple of 16; src and dst 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Stage `bytesP` + `bytesQ` (two arrays) with one elected thread; everyone waits.
// `phase` is the caller-tracked parity of the barrier (flip after every use).
__device__ __forceinline__ void stage_two(void *dP, const void *gP, uint32_t bytesP, void *dQ, const void *gQ,
                                          uint32_t bytesQ, uint64_t *bar, uint32_t &phase) {
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, bytesP + bytesQ);
        // the bulk engine takes at most a few hundred KB per request; split in 64 KB pieces
        for (uint32_t o = 0; o < bytesP; o += 65536u) {
            uint32_t b = bytesP - o < 65536u ? bytesP - o : 65536u;
            bulk_g2s((char *)dP + o, (const char *)gP + o, b, bar);
        }
        for (uint32_t o = 0; o < bytesQ; o += 65536u) {
            uint32_t b = bytesQ - o < 65536u ? bytesQ - o : 65536u;
            bulk_g2s((char *)dQ + o, (const char *)gQ + o, b, bar);
        }
    }
    mbar_wait(bar, phase);
    phase ^= 1u;
}

} // namespace mp
