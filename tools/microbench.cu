// microbench.cu -- design-decision probes for the B200 (run once under gpurun; results land in profiles/).
//   1. legacy mma.sync rates (m16n8k16 f16/bf16, m16n8k8 tf32): is the register-fragment tensor path fast enough
//      for N=8/16-wide conv GEMMs, or is tcgen05 required?  (SURVEY.md 7 "decide by ncu, and say so")
//   2. FFMA issue rate (ceiling of the fp32 guard path)
//   3. shared-memory LDS.32 bandwidth (the im2col fragment loads are LDS-bound)
//   4. pinned H2D / D2H bandwidth (the T1 engine roof)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("{\"error\": \"%s at %s:%d\"}\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

template <int KIND>  // 0 f16, 1 bf16, 2 tf32
__global__ void mma_rate(float* out, int iters) {
    unsigned a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    float c[8][4];
#pragma unroll
    for (int j = 0; j < 8; j++) c[j][0] = c[j][1] = c[j][2] = c[j][3] = 0.f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {  // 8 independent accumulators: hides mma latency
            if (KIND == 0)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 1)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
        }
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    if (s == 12345.678f) out[0] = s;
}

__global__ void ffma_rate(float* out, int iters) {
    float a[16];
#pragma unroll
    for (int j = 0; j < 16; j++) a[j] = threadIdx.x * 0.001f + j;
    float x = 1.0001f, y = 0.9999f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) a[j] = fmaf(a[j], x, y);
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) s += a[j];
    if (s == 12345.678f) out[0] = s;
}

// packed fp32 FMA (fma.rn.f32x2, sm_100): 16 independent float2 chains, and a 50/50 mix with shared-memory loads
__global__ void ffma2_rate(float* out, int iters) {
    float2 a[16];
#pragma unroll
    for (int j = 0; j < 16; j++) a[j] = make_float2(threadIdx.x * 0.001f + j, threadIdx.x * 0.002f + j);
    const float2 x = make_float2(1.0001f, 0.9998f), y = make_float2(0.9999f, 1.0002f);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) a[j] = __ffma2_rn(a[j], x, y);
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) s += a[j].x + a[j].y;
    if (s == 12345.678f) out[0] = s;
}

__global__ void lds_rate(float* out, int iters) {
    __shared__ float buf[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) buf[i] = i;
    __syncthreads();
    float s = 0;
    int idx = threadIdx.x;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) s += buf[(idx + j * 67) & 4095];
        idx = (idx + 1) & 4095;
    }
    if (s == 12345.678f) out[0] = s;
}

template <typename F>
float time_ms(F f) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    f();  // warm-up
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    f();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    return ms;
}

int main() {
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    float* d;
    CK(cudaMalloc(&d, 1024));
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", p.name, sms, clk_khz);
    const int iters = 20000;
    for (int warps : {4, 8, 16, 32}) {
        int threads = warps * 32;
        int blocks = sms * (warps <= 16 ? 2 : 1);
        double n_mma = (double)blocks * warps * iters * 8;
        float ms0 = time_ms([&] { mma_rate<0><<<blocks, threads>>>(d, iters); });
        float ms1 = time_ms([&] { mma_rate<1><<<blocks, threads>>>(d, iters); });
        float ms2 = time_ms([&] { mma_rate<2><<<blocks, threads>>>(d, iters); });
        printf("{\"probe\": \"mma.sync\", \"warps_per_cta\": %d, \"ctas\": %d, \"f16_m16n8k16_tflops\": %.1f, \"bf16_m16n8k16_tflops\": %.1f, \"tf32_m16n8k8_tflops\": %.1f, \"f16_mma_per_ns_per_sm\": %.3f}\n",
               warps, blocks, n_mma * 4096 / ms0 * 1e-9, n_mma * 4096 / ms1 * 1e-9, n_mma * 2048 / ms2 * 1e-9, n_mma / (ms0 * 1e6) / sms);
    }
    {
        int blocks = sms * 4, threads = 512;
        float ms = time_ms([&] { ffma_rate<<<blocks, threads>>>(d, iters); });
        printf("{\"probe\": \"ffma\", \"tflops\": %.1f}\n", (double)blocks * threads * iters * 16 * 2 / ms * 1e-9);
        ms = time_ms([&] { ffma2_rate<<<blocks, threads>>>(d, iters); });
        printf("{\"probe\": \"ffma2 (fma.rn.f32x2)\", \"tflops\": %.1f}\n", (double)blocks * threads * iters * 16 * 4 / ms * 1e-9);
        ms = time_ms([&] { lds_rate<<<blocks, threads>>>(d, iters); });
        printf("{\"probe\": \"lds32\", \"TBps\": %.2f, \"bytes_per_clk_per_sm_at_max_clock\": %.1f}\n", (double)blocks * threads * iters * 16 * 4 / ms * 1e-9,
               (double)blocks * threads * iters * 16 * 4 / (ms * 1e-3) / sms / (clk_khz * 1e3));
    }
    for (size_t mb : {1, 16, 64, 256}) {
        size_t bytes = mb << 20;
        void *h, *dd;
        CK(cudaHostAlloc(&h, bytes, cudaHostAllocDefault));
        CK(cudaMalloc(&dd, bytes));
        float h2d = time_ms([&] { for (int i = 0; i < 8; i++) cudaMemcpyAsync(dd, h, bytes, cudaMemcpyHostToDevice); });
        float d2h = time_ms([&] { for (int i = 0; i < 8; i++) cudaMemcpyAsync(h, dd, bytes, cudaMemcpyDeviceToHost); });
        printf("{\"probe\": \"pcie_pinned\", \"MiB\": %zu, \"h2d_GBps\": %.1f, \"d2h_GBps\": %.1f}\n", mb, 8.0 * bytes / h2d * 1e-6, 8.0 * bytes / d2h * 1e-6);
        cudaFreeHost(h); cudaFree(dd);
    }
    return 0;
}
