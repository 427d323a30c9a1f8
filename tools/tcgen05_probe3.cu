// tcgen05_probe3.cu -- can tcgen05.cp build conv1's A operand in TMEM straight from a staged half-precision row?
//
// Layout under test (tools/tcgen05_probe2.cu proved it exact for an SS mma): an input row stored as [pixel][4 halves]
// (8 B a pixel); a no-swizzle K-major descriptor with 16 B between K chunks and 128 B between 8-row groups makes
// matrix row m the 32-byte window that starts at byte 16*m, i.e. at pixel 2m. Here the same descriptor feeds
// tcgen05.cp.128x256b, which should leave row m's 16 halves in TMEM lane m, 8 columns -- the form a TS mma reads its A
// operand in -- with no register pass.
//   part 1: contents of TMEM after the copy vs the expected windows
//   part 2: clocks per tcgen05.cp back to back; per TS mma (N = 32) back to back; and per "row pair" of the conv1
//           schedule (4 copies + 12 mma) issued by one thread, to see how far copies and mma overlap
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tcgen05_probe3 tools/tcgen05_probe3.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void cp_128x256b(uint32_t taddr, uint64_t sdesc) {
    asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(d), "r"(a), "l"(bdesc),
        "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    long long t0 = clock64();
    while (!ok) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (!ok && clock64() - t0 > 2000000000ll) __trap();
    }
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
           (1ull << 46);
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(pred));
    return pred != 0;
}
__host__ __device__ constexpr uint32_t idesc_for(uint32_t n) { return (1u << 4) | ((n >> 3) << 17) | ((128u >> 4) << 24); }

constexpr int kABuf = 8 * 128 + 64;  // halves: row m starts at half 8*m, K = 32 reaches 8*127 + 31

// part 1: two copies (K chunks 0-1 and 2-3 of every row) -> TMEM columns 0..15; read back
__global__ void __launch_bounds__(128) probe_cp(const __half* abuf, uint32_t* out) {
    __shared__ __align__(128) __half sA[kABuf];
    __shared__ uint32_t s_taddr;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < kABuf; i += 128) sA[i] = abuf[i];
    if (tid == 0) mbar_init(&bar, 1);
    if (warp == 0) tmem_alloc(&s_taddr, 32);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tbase = s_taddr;
    if (warp == 0) {
        if (elect_one()) {
            cp_128x256b(tbase, make_desc(smem_u32(sA), 16, 128));
            cp_128x256b(tbase + 8, make_desc(smem_u32(sA) + 32, 16, 128));
            mma_commit(&bar);
        }
        __syncwarp();
    }
    mbar_wait(&bar, 0);
    fence_after();
    uint32_t r[8];
    const uint32_t lane_base = tbase + ((uint32_t)(warp * 32) << 16);
    tmem_ld8(lane_base, r);
    tmem_wait_ld();
    for (int c = 0; c < 8; c++) out[tid * 16 + c] = r[c];
    tmem_ld8(lane_base + 8, r);
    tmem_wait_ld();
    for (int c = 0; c < 8; c++) out[tid * 16 + 8 + c] = r[c];
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 32);
}

// part 2: mode 0 = copies only, 1 = TS mma only (N = 32), 2 = the conv1 row-pair schedule (4 copies then 12 mma)
__global__ void __launch_bounds__(128) probe_rate(int iters, long long* cycles, int mode) {
    __shared__ __align__(128) __half sA[8][kABuf];   // 8 staged rows
    __shared__ __align__(128) __half sB[2][96 * 16];  // stacked B per K step
    __shared__ uint32_t s_taddr;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 8 * kABuf; i += 128) (&sA[0][0])[i] = __float2half(0.001f * (i & 127));
    for (int i = tid; i < 2 * 96 * 16; i += 128) (&sB[0][0])[i] = __float2half(0.001f * (i & 255));
    if (tid == 0) mbar_init(&bar, 1);
    if (warp == 0) tmem_alloc(&s_taddr, 512);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tbase = s_taddr;
    const long long t0 = clock64();
    if (warp == 0) {
        const uint32_t i32 = idesc_for(32), i16 = idesc_for(16);
        if (elect_one()) {
            for (int it = 0; it < iters; it++) {
                const uint32_t a = tbase + (it & 3) * 32;          // A slots: 2 rows x 16 columns
                const uint32_t d = tbase + 128 + (it & 3) * 96;    // three D pairs of 32 columns
                if (mode != 1) {
                    for (int r = 0; r < 2; r++)
                        for (int ks = 0; ks < 2; ks++)
                            cp_128x256b(a + r * 16 + ks * 8, make_desc(smem_u32(sA[(2 * it + r) & 7]) + ks * 32, 16, 128));
                }
                if (mode != 0) {
                    for (int i = 0; i < 3; i++)
                        for (int ks = 0; ks < 2; ks++) {
                            mma_ts(d + i * 32, a + ks * 8, make_desc(smem_u32(sB[ks]) + (4 - 2 * i) * 512, 128, 256), i32, 1u);
                            mma_ts(i < 2 ? d + i * 32 : d + i * 32 + 16, a + 16 + ks * 8, make_desc(smem_u32(sB[ks]) + (i < 2 ? 3 - 2 * i : 0) * 512, 128, 256),
                                   i < 2 ? i32 : i16, 1u);
                        }
                }
            }
            mma_commit(&bar);
        }
        __syncwarp();
    }
    mbar_wait(&bar, 0);
    __syncthreads();
    const long long t1 = clock64();
    if (tid == 0) cycles[0] = t1 - t0;
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 512);
}

int main() {
    std::vector<__half> hA(kABuf);
    srand(11);
    for (int i = 0; i < kABuf; i++) hA[i] = __float2half((float)((rand() % 2001) - 1000) / 64.f);
    __half* dA;
    uint32_t* dO;
    long long* dC;
    CK(cudaMalloc(&dA, hA.size() * 2)); CK(cudaMalloc(&dO, 128 * 16 * 4)); CK(cudaMalloc(&dC, 16));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemset(dO, 0xff, 128 * 16 * 4));
    probe_cp<<<1, 128>>>(dA, dO);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("{\"probe\": \"tcgen05.cp 128x256b, overlapped descriptor\", \"error\": \"%s\"}\n", cudaGetErrorString(e)); return 1; }
    std::vector<uint32_t> out(128 * 16);
    CK(cudaMemcpy(out.data(), dO, out.size() * 4, cudaMemcpyDeviceToHost));
    int bad = 0, first_bad = -1;
    for (int m = 0; m < 128; m++)
        for (int c = 0; c < 16; c++) {
            uint16_t lo, hi;
            memcpy(&lo, &hA[8 * m + 2 * c], 2);
            memcpy(&hi, &hA[8 * m + 2 * c + 1], 2);
            const uint32_t want = (uint32_t)lo | ((uint32_t)hi << 16);
            if (out[m * 16 + c] != want) { bad++; if (first_bad < 0) first_bad = m * 16 + c; }
        }
    printf("{\"probe\": \"tcgen05.cp 128x256b x2, A = sliding window (LBO 16 B, SBO 128 B): TMEM lane m, columns 0..15 = halves [8m, 8m+32)\", "
           "\"words_checked\": %d, \"mismatches\": %d, \"first_mismatch_lane\": %d, \"first_mismatch_column\": %d}\n", 128 * 16, bad,
           first_bad < 0 ? -1 : first_bad / 16, first_bad < 0 ? -1 : first_bad % 16);
    const char* names[3] = {"tcgen05.cp 128x256b back to back (4 per iteration)", "tcgen05.mma TS N=32/16 conv schedule (12 per iteration)",
                            "4 x tcgen05.cp + 12 x tcgen05.mma per iteration (conv1 row pair)"};
    for (int mode = 0; mode < 3; mode++) {
        const int iters = 4096;
        probe_rate<<<1, 128>>>(iters, dC, mode);
        probe_rate<<<1, 128>>>(iters, dC, mode);
        e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("{\"probe\": \"%s\", \"error\": \"%s\"}\n", names[mode], cudaGetErrorString(e)); return 1; }
        long long c = 0;
        CK(cudaMemcpy(&c, dC, 8, cudaMemcpyDeviceToHost));
        printf("{\"probe\": \"%s\", \"clk_per_iteration\": %.1f}\n", names[mode], (double)c / iters);
    }
    return 0;
}
