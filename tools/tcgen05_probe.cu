// tcgen05_probe.cu -- de-risking probe for moving conv1 onto the 5th-gen tensor cores (round-2 design input).
//
// Question: can the conv1 im2col operand live in TMEM (written from registers with tcgen05.st, one output column
// per TMEM lane), be multiplied against a 16x16 filter slice in shared memory with tcgen05.mma (A from TMEM), and
// be read back with tcgen05.ld fast enough to beat mma.sync's 8.5 clk per m16n8k16 per sub-partition?
//   part 1: correctness of the operand conventions (A packing in TMEM, B descriptor LBO/SBO, D layout)
//   part 2: rates -- tcgen05.mma M=128 N=16 K=16 (A in TMEM), tcgen05.st 32x32b.x8, tcgen05.ld 32x32b.x16
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tcgen05_probe tools/tcgen05_probe.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
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

__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
                 "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] . B[smem desc]; issued by ONE thread
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

__device__ __forceinline__ uint64_t make_bdesc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
           (1ull << 46);  // version 1 (Blackwell), no swizzle, base offset 0
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(pred));
    return pred != 0;
}
__host__ __device__ constexpr uint32_t idesc_for(uint32_t n) { return (1u << 4) | ((n >> 3) << 17) | ((128u >> 4) << 24); }
constexpr uint32_t kIdesc = (1u << 4) | ((16u >> 3) << 17) | ((128u >> 4) << 24);  // f16 x f16 -> f32, M=128, N=16, K-major

// part 1: out[m][n] = sum_k A[m][k] * B[n][k]; A handed in as [128][16] halves, B as [16][16] halves (n-major rows)
__global__ void __launch_bounds__(128) probe_correct(const __half* A, const __half* B, float* out, int swap_lbo_sbo) {
    __shared__ __align__(128) __half sB[16 * 16];
    __shared__ uint32_t s_taddr;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5;
    // B in the canonical K-major no-swizzle layout: [n/8][k/8][n%8][k%8]
    for (int i = tid; i < 256; i += 128) {
        const int n = i / 16, k = i % 16;
        sB[(n / 8) * 128 + (k / 8) * 64 + (n % 8) * 8 + (k % 8)] = B[n * 16 + k];
    }
    if (tid == 0) mbar_init(&bar, 1);
    if (warp == 0) tmem_alloc(&s_taddr, 64);
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tbase = s_taddr;
    // A: this thread owns row m = tid, 16 halves = 8 words, word j = (k=2j | k=2j+1 << 16)
    uint32_t a[8];
    for (int j = 0; j < 8; j++) {
        const uint32_t lo = __half_as_ushort(A[tid * 16 + 2 * j]), hi = __half_as_ushort(A[tid * 16 + 2 * j + 1]);
        a[j] = lo | (hi << 16);
    }
    const uint32_t lane_base = tbase + ((uint32_t)(warp * 32) << 16);
    tmem_st8(lane_base + 0, a);  // A at columns 0..7
    tmem_wait_st();
    fence_before();
    __syncthreads();
    fence_after();
    if (tid == 0) {
        const uint32_t lbo = swap_lbo_sbo ? 256u : 128u, sbo = swap_lbo_sbo ? 128u : 256u;
        mma_ts(tbase + 16, tbase + 0, make_bdesc(smem_u32(sB), lbo, sbo), kIdesc, 0u);  // D at columns 16..31
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    fence_after();
    uint32_t d[16];
    tmem_ld16(lane_base + 16, d);
    tmem_wait_ld();
    for (int n = 0; n < 16; n++) out[tid * 16 + n] = __uint_as_float(d[n]);
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 64);
}

// part 2: rates. mode 0: back-to-back MMAs (one issuing thread); mode 1: tcgen05.st x8 by 4 warps; mode 2: tcgen05.ld x16
__global__ void __launch_bounds__(128) probe_rate(int mode, int iters, long long* cycles, int nn) {
    __shared__ __align__(128) __half sB[256 * 16];
    __shared__ uint32_t s_taddr;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 256 * 16; i += 128) sB[i] = __float2half(0.001f * (i & 255));
    if (tid == 0) mbar_init(&bar, 1);
    if (warp == 0) tmem_alloc(&s_taddr, 512);
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tbase = s_taddr, lane_base = tbase + ((uint32_t)(warp * 32) << 16);
    uint32_t a[8] = {1, 2, 3, 4, 5, 6, 7, 8};
    for (int c = 0; c < 256; c += 8) tmem_st8(lane_base + c, a);
    tmem_wait_st();
    fence_before();
    __syncthreads();
    fence_after();
    const long long t0 = clock64();
    if (mode == 0) {
        if (warp == 0) {  // whole warp runs the loop, one elected lane issues: no divergence-handling code per mma
            const uint64_t bd = make_bdesc(smem_u32(sB), 128, 256);
            const uint32_t id = idesc_for((uint32_t)nn);
            if (elect_one()) {  // one lane stays elected for the whole burst (CUTLASS's issue pattern)
#pragma unroll 8
                for (int i = 0; i < iters; i++)
                    mma_ts(tbase + 256 + ((nn <= 128) ? nn * (i & 1) : 0), tbase + 8 * (i & 7), bd, id, 1u);
                mma_commit(&bar);
            }
            __syncwarp();
        }
        mbar_wait(&bar, 0);
    } else if (mode == 1) {
        for (int i = 0; i < iters; i++) tmem_st8(lane_base + 8 * (i & 15), a);
        tmem_wait_st();
    } else {
        uint32_t d[16], acc = 0;
        for (int i = 0; i < iters; i++) {
            tmem_ld16(lane_base + 16 * (i & 7), d);
            tmem_wait_ld();
            acc += d[i & 15];
        }
        if (acc == 0x12345678u) cycles[1] = acc;
    }
    __syncthreads();
    const long long t1 = clock64();
    if (tid == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 512);
}

int main() {
    std::vector<__half> hA(128 * 16), hB(16 * 16);
    std::vector<float> fA(128 * 16), fB(16 * 16);
    srand(1);
    for (int i = 0; i < 128 * 16; i++) { fA[i] = (float)((rand() % 17) - 8) / 8.f; hA[i] = __float2half(fA[i]); }
    for (int i = 0; i < 256; i++) { fB[i] = (float)((rand() % 13) - 6) / 4.f; hB[i] = __float2half(fB[i]); }
    __half *dA, *dB;
    float* dO;
    long long* dC;
    CK(cudaMalloc(&dA, hA.size() * 2)); CK(cudaMalloc(&dB, hB.size() * 2)); CK(cudaMalloc(&dO, 128 * 16 * 4)); CK(cudaMalloc(&dC, 16));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    std::vector<float> ref(128 * 16);
    for (int m = 0; m < 128; m++)
        for (int n = 0; n < 16; n++) {
            float s = 0;
            for (int k = 0; k < 16; k++) s += fA[m * 16 + k] * fB[n * 16 + k];
            ref[m * 16 + n] = s;
        }
    for (int swap = 0; swap < 2; swap++) {
        CK(cudaMemset(dO, 0, 128 * 16 * 4));
        probe_correct<<<1, 128>>>(dA, dB, dO, swap);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("{\"probe\": \"correct\", \"swap_lbo_sbo\": %d, \"error\": \"%s\"}\n", swap, cudaGetErrorString(e)); return 1; }
        std::vector<float> out(128 * 16);
        CK(cudaMemcpy(out.data(), dO, out.size() * 4, cudaMemcpyDeviceToHost));
        float maxerr = 0;
        for (size_t i = 0; i < out.size(); i++) maxerr = fmaxf(maxerr, fabsf(out[i] - ref[i]));
        printf("{\"probe\": \"tcgen05.mma A-in-TMEM M128 N16 K16\", \"lbo\": %d, \"sbo\": %d, \"max_abs_err\": %g, \"out00\": %g, \"ref00\": %g}\n",
               swap ? 256 : 128, swap ? 128 : 256, maxerr, out[0], ref[0]);
    }
    const char* names[3] = {"tcgen05.mma M128 N16 K16 (A in TMEM), back to back, 1 SM", "tcgen05.st 32x32b.x8 by 4 warps", "tcgen05.ld 32x32b.x16 by 4 warps (+wait each)"};
    for (int nn : {16, 32, 64, 128, 256}) {
        const int iters = 8192;
        probe_rate<<<1, 128>>>(0, iters, dC, nn);
        probe_rate<<<1, 128>>>(0, iters, dC, nn);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("{\"probe\": \"mma N=%d\", \"error\": \"%s\"}\n", nn, cudaGetErrorString(e)); return 1; }
        long long c = 0;
        CK(cudaMemcpy(&c, dC, 8, cudaMemcpyDeviceToHost));
        printf("{\"probe\": \"tcgen05.mma TS M128 N%d K16 back to back (elect_one issue)\", \"clk_per_mma\": %.2f, \"macs_per_clk\": %.0f}\n",
               nn, (double)c / iters, 128.0 * nn * 16 / ((double)c / iters));
    }
    for (int mode = 1; mode < 3; mode++) {
        const int iters = 4096;
        probe_rate<<<1, 128>>>(mode, iters, dC, 16);
        probe_rate<<<1, 128>>>(mode, iters, dC, 16);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("{\"probe\": \"%s\", \"error\": \"%s\"}\n", names[mode], cudaGetErrorString(e)); return 1; }
        long long c = 0;
        CK(cudaMemcpy(&c, dC, 8, cudaMemcpyDeviceToHost));
        double per = (double)c / iters;
        if (mode == 0) printf("{\"probe\": \"%s\", \"clk_per_mma\": %.2f, \"macs_per_clk\": %.0f}\n", names[mode], per, 128.0 * 16 * 16 / per);
        else printf("{\"probe\": \"%s\", \"clk_per_instr_per_warp\": %.2f, \"bytes_per_clk_sm\": %.1f}\n", names[mode], per,
                    (mode == 1 ? 4 * 32 * 8 * 4 : 4 * 32 * 16 * 4) / per);
    }
    return 0;
}
