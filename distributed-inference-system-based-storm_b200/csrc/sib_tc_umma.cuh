// sib_tc_umma.cuh -- tcgen05 / TMEM primitives for the conv1 stage of the whole-network kernel (sm_100a).
// Included by sib_kernels_tc.cu inside namespace sib::{anonymous}. Conventions were pinned with
// tools/tcgen05_probe.cu on the B200 (profiles/tcgen05_probe_r01.jsonl): A operand in TMEM = one matrix row per
// lane, 32-bit column j holds K elements (2j | 2j+1 << 16); B in shared memory, K-major, no swizzle:
// [n/8][k/8][n%8][k%8] with leading byte offset 128 (next K chunk) and stride byte offset 256 (next 8 rows).
#pragma once

__device__ __forceinline__ void tmem_alloc_all(uint32_t* slot) {  // one converged warp; 512 columns = the whole TMEM
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free_all(uint32_t addr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(512u) : "memory");
}
__device__ __forceinline__ void tc5_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc5_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
                 "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] . B[smem descriptor]^T ; issued by ONE elected lane of a converged warp
__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(d), "r"(a), "l"(bdesc),
        "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {  // arrives on the mbarrier when every prior mma has retired
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ uint64_t umma_bdesc(uint32_t saddr) {  // 16x16 halves, K-major, no swizzle, version 1
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)(128u >> 4) << 16) | ((uint64_t)(256u >> 4) << 32) | (1ull << 46);
}
template <typename T> __device__ __forceinline__ constexpr uint32_t umma_idesc_m128_n16() {
    // c_format f32 (bit 4), a/b format (bits 7..12: 0 = f16, 1 = bf16), K-major both, N>>3 at bit 17, M>>4 at bit 24
    constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
}
