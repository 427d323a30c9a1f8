// tcgen05_probe2.cu -- second de-risking probe: conv1 with the im2col operand read STRAIGHT from shared memory.
//
// Idea under test: store an input row as [pixel][4 halves] (8 B a pixel). A no-swizzle K-major A descriptor with
// stride-between-8-row-groups SBO = 128 B and K-chunk stride LBO = 16 B makes operand row m start at byte 16*m and run
// contiguously along K, i.e. row m IS the sliding window that starts at pixel 2m -- overlapping "core matrices", no
// im2col copy, no register traffic. One mma with N = 5 kh x 6 channels (padded to 32) then adds input row h into the
// accumulators of output rows h-4..h at once, provided D may start at a TMEM column that is a multiple of 6.
//   part 1: numerics of the overlapped A descriptor; D at column offsets 0/6/8/12/22 with accumulate onto known values
//   part 2: tcgen05.ld.x8 / tcgen05.st.x2 at a column offset that is not a multiple of their width
//   part 3: issue rate of SS mmas, M=128, N = 16/32/48/64, K=16
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tcgen05_probe2 tools/tcgen05_probe2.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cmath>
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
__device__ __forceinline__ void tmem_st1(uint32_t taddr, uint32_t v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(v) : "memory");
}
__device__ __forceinline__ void tmem_st2(uint32_t taddr, uint32_t v0, uint32_t v1) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1,%2};" ::"r"(taddr), "r"(v0), "r"(v1) : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t tmem_ld1(uint32_t taddr) {
    uint32_t v;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr) : "memory");
    return v;
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(d), "l"(adesc), "l"(bdesc),
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

constexpr int kABuf = 8 * 128 + 64;  // halves: row m starts at 8*m, K = 32 reaches 8*127 + 31

// part 1: D[m][dcol + n] += sum_{k<32} abuf[8m + k] * B[n][k]  (two K=16 mmas), N = 32, D pre-set to init(m, col)
__global__ void __launch_bounds__(128) probe_ss(const __half* abuf, const __half* B, float* out, int dcol, int a_lbo, int a_sbo) {
    __shared__ __align__(128) __half sA[kABuf];
    __shared__ __align__(128) __half sB[2][32 * 16];  // two K steps, each [n/8][k/8][n%8][k%8]
    __shared__ uint32_t s_taddr;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < kABuf; i += 128) sA[i] = abuf[i];
    for (int i = tid; i < 32 * 32; i += 128) {
        const int n = i / 32, k = i % 32, ks = k / 16, kk = k % 16;
        sB[ks][(n / 8) * 128 + (kk / 8) * 64 + (n % 8) * 8 + (kk % 8)] = B[n * 32 + k];
    }
    if (tid == 0) mbar_init(&bar, 1);
    if (warp == 0) tmem_alloc(&s_taddr, 64);
    fence_async_smem();  // generic-proxy writes of sA/sB -> visible to the tensor core's async-proxy reads
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tbase = s_taddr, lane_base = tbase + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < 64; c++) tmem_st1(lane_base + c, __float_as_uint((float)tid + 0.25f * c));
    tmem_wait_st();
    fence_before();
    __syncthreads();
    fence_after();
    if (warp == 0) {
        if (elect_one()) {
            for (int ks = 0; ks < 2; ks++)
                mma_ss(tbase + dcol, make_desc(smem_u32(sA) + ks * 32, a_lbo, a_sbo), make_desc(smem_u32(sB[ks]), 128, 256),
                       idesc_for(32), 1u);
            mma_commit(&bar);
        }
        __syncwarp();
    }
    mbar_wait(&bar, 0);
    fence_after();
    for (int c = 0; c < 64; c++) {
        const uint32_t v = tmem_ld1(lane_base + c);
        tmem_wait_ld();
        out[tid * 64 + c] = __uint_as_float(v);
    }
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 64);
}

// part 2: ld.x8 from column `col`, st.x2 to column `col` -- neither a multiple of the access width
__global__ void __launch_bounds__(128) probe_ldst(float* out, int col) {
    __shared__ uint32_t s_taddr;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tmem_alloc(&s_taddr, 32);
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t lane_base = s_taddr + ((uint32_t)(warp * 32) << 16);
    for (int c = 0; c < 32; c++) tmem_st1(lane_base + c, __float_as_uint((float)tid + 0.25f * c));
    tmem_wait_st();
    tmem_st2(lane_base + col, __float_as_uint(-1.f), __float_as_uint(-2.f));
    tmem_wait_st();
    uint32_t r[8];
    tmem_ld8(lane_base + col, r);
    tmem_wait_ld();
    for (int c = 0; c < 8; c++) out[tid * 8 + c] = __uint_as_float(r[c]);
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(s_taddr, 32);
}

// part 3: SS mma issue rate (A: overlapped descriptor in smem, B: smem), D alternating between two regions
__global__ void __launch_bounds__(128) probe_rate_ss(int iters, long long* cycles, int nn, int dstep) {
    __shared__ __align__(128) __half sA[kABuf];
    __shared__ __align__(128) __half sB[64 * 16];
    __shared__ uint32_t s_taddr;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < kABuf; i += 128) sA[i] = __float2half(0.001f * (i & 127));
    for (int i = tid; i < 64 * 16; i += 128) sB[i] = __float2half(0.001f * (i & 255));
    if (tid == 0) mbar_init(&bar, 1);
    if (warp == 0) tmem_alloc(&s_taddr, 512);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tbase = s_taddr;
    const long long t0 = clock64();
    if (warp == 0) {
        const uint64_t ad = make_desc(smem_u32(sA), 16, 128), bd = make_desc(smem_u32(sB), 128, 256);
        const uint32_t id = idesc_for((uint32_t)nn);
        if (elect_one()) {
#pragma unroll 8
            for (int i = 0; i < iters; i++) mma_ss(tbase + dstep * (i & 31), ad + 2 * (i & 1), bd, id, 1u);
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

int main(int argc, char** argv) {
    // usage: tcgen05_probe2 [ss <dcol> | ldst <col> | rate]   (one risky case per process: a trap kills the context)
    const char* what = argc > 1 ? argv[1] : "all";
    const int arg = argc > 2 ? atoi(argv[2]) : -1;
    auto want = [&](const char* w, int v) { return !strcmp(what, "all") || (!strcmp(what, w) && (arg < 0 || arg == v)); };
    std::vector<__half> hA(kABuf), hB(32 * 32);
    std::vector<float> fA(kABuf), fB(32 * 32);
    srand(7);
    for (int i = 0; i < kABuf; i++) { fA[i] = (float)((rand() % 17) - 8) / 8.f; hA[i] = __float2half(fA[i]); }
    for (int i = 0; i < 32 * 32; i++) { fB[i] = (float)((rand() % 13) - 6) / 4.f; hB[i] = __float2half(fB[i]); }
    __half *dA, *dB;
    float* dO;
    long long* dC;
    CK(cudaMalloc(&dA, hA.size() * 2)); CK(cudaMalloc(&dB, hB.size() * 2)); CK(cudaMalloc(&dO, 128 * 64 * 4)); CK(cudaMalloc(&dC, 16));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    for (int dcol : {0, 8, 6, 12, 22, 2}) {
        if (!want("ss", dcol)) continue;
        CK(cudaMemset(dO, 0, 128 * 64 * 4));
        probe_ss<<<1, 128>>>(dA, dB, dO, dcol, 16, 128);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("{\"probe\": \"ss overlapped A, D at column %d\", \"error\": \"%s\"}\n", dcol, cudaGetErrorString(e)); return 1; }
        std::vector<float> out(128 * 64);
        CK(cudaMemcpy(out.data(), dO, out.size() * 4, cudaMemcpyDeviceToHost));
        float err_in = 0, err_out = 0;
        for (int m = 0; m < 128; m++)
            for (int c = 0; c < 64; c++) {
                float want = (float)m + 0.25f * c;
                if (c >= dcol && c < dcol + 32) {
                    const int n = c - dcol;
                    for (int k = 0; k < 32; k++) want += fA[8 * m + k] * fB[n * 32 + k];
                    err_in = fmaxf(err_in, fabsf(out[m * 64 + c] - want));
                } else {
                    err_out = fmaxf(err_out, fabsf(out[m * 64 + c] - want));
                }
            }
        printf("{\"probe\": \"tcgen05.mma SS, A = sliding window (LBO 16 B, SBO 128 B), M128 N32 K32, accumulate\", \"d_column\": %d, "
               "\"max_abs_err_window\": %g, \"max_abs_err_untouched_columns\": %g}\n", dcol, err_in, err_out);
    }
    for (int col : {6, 3}) {
        if (!want("ldst", col)) continue;
        probe_ldst<<<1, 128>>>(dO, col);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("{\"probe\": \"ld.x8/st.x2 at column %d\", \"error\": \"%s\"}\n", col, cudaGetErrorString(e)); return 1; }
        std::vector<float> out(128 * 8);
        CK(cudaMemcpy(out.data(), dO, out.size() * 4, cudaMemcpyDeviceToHost));
        float err = 0;
        for (int m = 0; m < 128; m++)
            for (int c = 0; c < 8; c++) {
                const float want = c == 0 ? -1.f : c == 1 ? -2.f : (float)m + 0.25f * (col + c);
                err = fmaxf(err, fabsf(out[m * 8 + c] - want));
            }
        printf("{\"probe\": \"tcgen05.st.x2 then tcgen05.ld.x8 at an unaligned column\", \"column\": %d, \"max_abs_err\": %g}\n", col, err);
    }
    for (int nn : {16, 32, 48, 64}) {
        for (int dstep : {0, 6}) {
            if (!want("rate", dstep)) continue;
            const int iters = 8192;
            probe_rate_ss<<<1, 128>>>(iters, dC, nn, dstep);
            probe_rate_ss<<<1, 128>>>(iters, dC, nn, dstep);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("{\"probe\": \"ss rate N=%d\", \"error\": \"%s\"}\n", nn, cudaGetErrorString(e)); return 1; }
            long long c = 0;
            CK(cudaMemcpy(&c, dC, 8, cudaMemcpyDeviceToHost));
            printf("{\"probe\": \"tcgen05.mma SS M128 N%d K16 back to back, D column step %d\", \"clk_per_mma\": %.2f, \"macs_per_clk\": %.0f}\n",
                   nn, dstep, (double)c / iters, 128.0 * nn * 16 / ((double)c / iters));
        }
    }
    return 0;
}
