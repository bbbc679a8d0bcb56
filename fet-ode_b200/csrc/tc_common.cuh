// tc_common.cuh -- PTX wrappers shared by the tcgen05 kernels (kanlinear_tc.cu forward, kanlinear_tc_bwd.cu backward):
// mbarriers, the TMA bulk-copy engine, tcgen05.mma / commit / fences, UMMA shared-memory descriptors.
#pragma once
#include <cstdio>
#include "kanlinear.cuh"

namespace fetode {
namespace tc {

constexpr int BM = 128, BK = 32, NSTAGE = 3;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Spins on try_wait.  A wait that does not complete within ~2^28 polls (seconds; every legitimate wait in these
// kernels is microseconds) reports itself and traps: a protocol bug then surfaces as a CUDA error instead of a hung GPU.
static __device__ __noinline__ void mbar_timeout(uint32_t bar, uint32_t parity) {
    printf("fetode: mbarrier wait timed out: smem 0x%x parity %u block (%d,%d,%d) thread %d\n", bar, parity,
           (int)blockIdx.x, (int)blockIdx.y, (int)blockIdx.z, (int)threadIdx.x);
    __trap();
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0, polls = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (!done && ++polls == (1u << 28)) mbar_timeout(bar, parity);
    }
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// SWIZZLE_128B operand tile made of 1024-byte atoms (8 x 128 B).  K-major: atoms sbo apart along M/N.
// MN-major tf32 operands must use SWIZZLE_128B_BASE32B (layout 1): atoms of 32 M/N elements (128 B) x 4 K rows,
// 32-byte chunks XOR-ed with the row index; atoms lbo apart along M/N and sbo apart along K.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes = 16, uint32_t sbo_bytes = 1024, uint32_t layout = 2) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);            // start address
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;   // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;   // stride byte offset
    d |= (uint64_t)1 << 46;                             // descriptor version (Blackwell)
    d |= (uint64_t)layout << 61;                        // 2 = SWIZZLE_128B, 1 = SWIZZLE_128B_BASE32B (MN-major tf32)
    return d;
}
// float offset of element (row, col) inside a [rows x 32] swizzled tile
__host__ __device__ __forceinline__ int swz(int row, int col) {
    return (row >> 3) * 256 + (row & 7) * 32 + ((((col >> 2) ^ (row & 7)) << 2) | (col & 3));
}

}  // namespace tc
}  // namespace fetode
