// sib_kernels_guard.cu -- the margin guard's recompute path: fp32 CUDA-core LeNet-5 forward, up to five listed images per CTA pass.
//
// The tensor-core kernel lists the images whose top1-top2 logit margin is below tau (sib_kernels_tc.cu); this
// kernel re-runs exactly those images in fp32 and overwrites their outputs, so that the emitted argmax equals the
// fp32 path's (BASELINE.json: "argmax predictions must agree exactly"). The list is short (~2% of a batch), so what
// matters is the latency of one image: a CTA owns an image, every stage is register-tiled so that FFMA, not
// shared-memory traffic, is the limiter, and the dense weights are streamed from L2 with 128-bit loads.
// Op semantics as in sib_kernels_fp32.cu (TF 1.x Conv2D/BiasAdd/Relu/MaxPool/MatMul/Softmax; InferenceBolt.java:81-85).
#include <cuda_runtime.h>

#include <atomic>

#include "sib_internal.hpp"

namespace sib {
namespace {

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// One CTA per SM, 24 warps, up to 10 listed images per pass. A 65 536-image batch lists ~1 000-1 400 images, i.e.
// 6-10 per SM: that is 1 250-1 850 conv1 work items, 2-3 rounds of 768 threads. (256-thread CTAs, two per SM, with
// 3-5 images each needed 3-4 conv1 rounds and 2 conv2 rounds for the same images: r01 guard profiles.)
#ifndef SIB_GUARD_THREADS
#define SIB_GUARD_THREADS 512   // 16 warps x 128 registers: the 2x4 / 2x2x8 register tiles below need ~100 (768 threads cap at 85)
#endif
#ifndef SIB_GUARD_IMGS
#define SIB_GUARD_IMGS 10
#endif
constexpr int kGT = SIB_GUARD_THREADS;
constexpr int kGI = SIB_GUARD_IMGS;  // images per CTA pass (shared-memory bound: 10 cifar images + scratch = 206 KB)

template <int H_, int W_, int C_>
struct GNet {
    static constexpr int H = H_, W = W_, C = C_;
    static constexpr int PH1 = (H - 4) / 2, PW1 = (W - 4) / 2, PH2 = (PH1 - 4) / 2, PW2 = (PW1 - 4) / 2;
    static constexpr int C1 = 6, C2 = 16, F1 = 120, F2 = 84, F3 = 10;
    static constexpr int IMG = H * W * C, P1 = PH1 * PW1 * C1, FEAT = PH2 * PW2 * C2;
    static constexpr int NW1 = 25 * C * C1, NW2 = 25 * C1 * C2;
    // dense stages: as many k slices as the threads allow, so that the per-thread chain of L2 weight loads is short
    static constexpr int KSL = kGT >= 600 ? FEAT / (FEAT % 20 == 0 ? 20 : 16) : kGT >= 480 ? 16 : 8;  // k slices of fc1 (x 30 float4 columns)
    static constexpr int S2 = kGT >= 8 * F2 ? 8 : 3, K2 = F1 / S2;                   // k slices of fc2 (x 84 outputs)
    static_assert(PW1 % 2 == 0 && FEAT % KSL == 0 && IMG % 4 == 0 && KSL * (F1 / 4) <= kGT && S2 * F2 <= kGT && kGI * 32 <= kGT && F2 % 3 == 0 && F1 % S2 == 0, "mapping");
    // shared memory (floats). The images are dead once conv1 has run: the dense stages' scratch overlays them.
    static constexpr int DENSE = kGI * KSL * F1 + kGI * F1 + kGI * S2 * F2;
    static constexpr int S_IMG = 0, S_PART = 0, S_H1 = S_PART + kGI * KSL * F1, S_H2 = S_H1 + kGI * F1,
                         S_W1 = ((kGI * IMG > DENSE ? kGI * IMG : DENSE) + 3) / 4 * 4, S_P1 = S_W1 + ((NW1 + C1 + 3) / 4) * 4,
                         S_W2 = S_P1 + kGI * P1, S_P2 = S_W2 + NW2 + C2, S_W3 = S_P2 + kGI * FEAT,
                         S_TOTAL = S_W3 + F2 * F3 + F3 + 2;
};

template <typename G>
__global__ void __launch_bounds__(kGT) guard_fp32_kernel(DeviceModel m, const float* __restrict__ in,
                                                           const int32_t* __restrict__ list,
                                                           int32_t* count, int32_t* out_count, int self_reset, int cap,
                                                           float* __restrict__ probs, float* __restrict__ logits,
                                                           int32_t* __restrict__ argmax) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x;
    const int n = min(*reinterpret_cast<volatile int32_t*>(count), cap);
    // self_reset: every CTA has read the length above before it takes a ticket; the last one to finish hands the
    // length to the host's result block and clears counter and ticket for the next launch on this scratch
    auto finish = [&]() {
        if (!self_reset) return;
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            if (atomicAdd(count + 1, 1) == (int)gridDim.x - 1) {
                if (out_count) *out_count = n;
                count[0] = 0;
                count[1] = 0;
            }
        }
    };
    // The list is dealt out in contiguous, near-equal shares (they differ by at most one image), so every CTA is done
    // after the same number of passes; dealing whole 4-image passes round-robin left a second, mostly empty wave.
    const int per = n / (int)gridDim.x, rem = n % (int)gridDim.x;
    const int first = (int)blockIdx.x * per + min((int)blockIdx.x, rem), last = first + per + ((int)blockIdx.x < rem ? 1 : 0);
    if (first >= last) { finish(); return; }
    const float* W = m.weights;
    const DeviceLayer L0 = m.layer[0], L1 = m.layer[1], L2 = m.layer[2], L3 = m.layer[3], L4 = m.layer[4];
    auto load_images = [&](int it, int imgs) {  // asynchronous copies: nothing waits until cp_async_wait_all()
        for (int g = 0; g < imgs; g++) {
            const int img = list[it + g];
            const float* src = in + (size_t)img * G::IMG;
            float* dst = sm + G::S_IMG + g * G::IMG;
            for (int i = tid; i < G::IMG / 4; i += kGT) cp_async16(dst + 4 * i, src + 4 * i);
        }
    };
    load_images(first, min(kGI, last - first));
    // conv filters, fc3 and their biases stay in shared memory for every image this CTA handles (their load latency
    // runs under the image copies above)
    for (int i = tid; i < G::NW1; i += kGT) sm[G::S_W1 + i] = __ldg(W + L0.w_off + i);
    for (int i = tid; i < G::C1; i += kGT) sm[G::S_W1 + G::NW1 + i] = __ldg(W + L0.b_off + i);
    for (int i = tid; i < G::NW2; i += kGT) sm[G::S_W2 + i] = __ldg(W + L1.w_off + i);
    for (int i = tid; i < G::C2; i += kGT) sm[G::S_W2 + G::NW2 + i] = __ldg(W + L1.b_off + i);
    for (int i = tid; i < G::F2 * G::F3; i += kGT) sm[G::S_W3 + i] = __ldg(W + L4.w_off + i);
    for (int i = tid; i < G::F3; i += kGT) sm[G::S_W3 + G::F2 * G::F3 + i] = __ldg(W + L4.b_off + i);

    for (int it = first; it < last; it += kGI) {
        const int imgs = min(kGI, last - it);
        if (it != first) {
            __syncthreads();  // the dense scratch of the previous pass overlays the image area
            load_images(it, imgs);
        }
        cp_async_wait_all();
        __syncthreads();
        // ---- conv1 + bias + ReLU + pool: item = (image, pooled row, PAIR of pooled columns): 2 x 4 conv outputs x 6 channels in
        // registers as channel pairs (packed fp32 FMA, fma.rn.f32x2). The stage is bound by the broadcast weight loads from
        // shared memory (stall_mio on them, short_scoreboard on the FMA behind them: profiles/ncu_guard_r02.txt); a 2x4 tile
        // does twice the FMA per weight load of the 2x2 tile it replaces.
        for (int item = tid; item < imgs * G::PH1 * (G::PW1 / 2); item += kGT) {
            const int g = item / (G::PH1 * (G::PW1 / 2)), pp = item % (G::PH1 * (G::PW1 / 2));
            const int py = pp / (G::PW1 / 2), px = 2 * (pp % (G::PW1 / 2));
            float2 acc[2][4][G::C1 / 2];
#pragma unroll
            for (int a = 0; a < 2; a++)
#pragma unroll
                for (int b = 0; b < 4; b++)
#pragma unroll
                    for (int c = 0; c < G::C1 / 2; c++) acc[a][b][c] = make_float2(0.f, 0.f);
            const float* ip = sm + G::S_IMG + g * G::IMG + ((2 * py) * G::W + 2 * px) * G::C;
#pragma unroll 1
            for (int r = 0; r < 6; r++) {  // patch row r feeds output row dy with kernel row r-dy
                float row[8 * G::C];
#pragma unroll
                for (int i = 0; i < 8 * G::C; i++) row[i] = ip[r * G::W * G::C + i];
#pragma unroll
                for (int dy = 0; dy < 2; dy++) {
                    const int kh = r - dy;
                    if (kh < 0 || kh > 4) continue;
#pragma unroll
                    for (int kw = 0; kw < 5; kw++)
#pragma unroll
                        for (int c = 0; c < G::C; c++) {
                            const float2* wp = reinterpret_cast<const float2*>(sm + G::S_W1 + ((kh * 5 + kw) * G::C + c) * G::C1);  // broadcast reads
                            float2 w[G::C1 / 2];
#pragma unroll
                            for (int co = 0; co < G::C1 / 2; co++) w[co] = wp[co];
#pragma unroll
                            for (int dx = 0; dx < 4; dx++) {
                                const float x = row[(kw + dx) * G::C + c];
                                const float2 v = make_float2(x, x);
#pragma unroll
                                for (int co = 0; co < G::C1 / 2; co++) acc[dy][dx][co] = __ffma2_rn(v, w[co], acc[dy][dx][co]);
                            }
                        }
                }
            }
#pragma unroll
            for (int h = 0; h < 2; h++) {  // the two pooled columns
                float* op = sm + G::S_P1 + g * G::P1 + (py * G::PW1 + px + h) * G::C1;
#pragma unroll
                for (int co = 0; co < G::C1 / 2; co++) {
                    const float mx = fmaxf(fmaxf(acc[0][2 * h][co].x, acc[0][2 * h + 1][co].x), fmaxf(acc[1][2 * h][co].x, acc[1][2 * h + 1][co].x));
                    const float my = fmaxf(fmaxf(acc[0][2 * h][co].y, acc[0][2 * h + 1][co].y), fmaxf(acc[1][2 * h][co].y, acc[1][2 * h + 1][co].y));
                    op[2 * co] = fmaxf(mx + sm[G::S_W1 + G::NW1 + 2 * co], 0.f);
                    op[2 * co + 1] = fmaxf(my + sm[G::S_W1 + G::NW1 + 2 * co + 1], 0.f);
                }
            }
        }
        __syncthreads();
        // ---- conv2 + bias + ReLU + pool: item = (image, pooled position, HALF of the 16 output channels): 2 x 2 conv outputs x 8
        // channels, two 128-bit weight loads per (kw, c) feeding 8 packed FMA
        for (int item = tid; item < imgs * G::PH2 * G::PW2 * 2; item += kGT) {
            const int q = item & 1, t2 = item >> 1;
            const int g = t2 / (G::PH2 * G::PW2), pos = t2 % (G::PH2 * G::PW2);
            const int py = pos / G::PW2, px = pos % G::PW2;
            float2 acc[2][2][4];
#pragma unroll
            for (int a = 0; a < 2; a++)
#pragma unroll
                for (int b = 0; b < 2; b++)
#pragma unroll
                    for (int c = 0; c < 4; c++) acc[a][b][c] = make_float2(0.f, 0.f);
            const float* ip = sm + G::S_P1 + g * G::P1 + ((2 * py) * G::PW1 + 2 * px) * G::C1;
#pragma unroll 1
            for (int r = 0; r < 6; r++) {
                float row[6 * G::C1];
#pragma unroll
                for (int i = 0; i < 6 * G::C1; i++) row[i] = ip[r * G::PW1 * G::C1 + i];
#pragma unroll
                for (int dy = 0; dy < 2; dy++) {
                    const int kh = r - dy;
                    if (kh < 0 || kh > 4) continue;
#pragma unroll
                    for (int kw = 0; kw < 5; kw++)
#pragma unroll
                        for (int c = 0; c < G::C1; c++) {
                            const float4* wp = reinterpret_cast<const float4*>(sm + G::S_W2 + ((kh * 5 + kw) * G::C1 + c) * G::C2 + 8 * q);
                            const float4 wa = wp[0], wb = wp[1];
                            const float2 w0 = make_float2(wa.x, wa.y), w1 = make_float2(wa.z, wa.w), w2 = make_float2(wb.x, wb.y), w3 = make_float2(wb.z, wb.w);
                            const float x0 = row[kw * G::C1 + c], x1 = row[(kw + 1) * G::C1 + c];
                            const float2 v0 = make_float2(x0, x0), v1 = make_float2(x1, x1);
                            acc[dy][0][0] = __ffma2_rn(v0, w0, acc[dy][0][0]); acc[dy][0][1] = __ffma2_rn(v0, w1, acc[dy][0][1]);
                            acc[dy][0][2] = __ffma2_rn(v0, w2, acc[dy][0][2]); acc[dy][0][3] = __ffma2_rn(v0, w3, acc[dy][0][3]);
                            acc[dy][1][0] = __ffma2_rn(v1, w0, acc[dy][1][0]); acc[dy][1][1] = __ffma2_rn(v1, w1, acc[dy][1][1]);
                            acc[dy][1][2] = __ffma2_rn(v1, w2, acc[dy][1][2]); acc[dy][1][3] = __ffma2_rn(v1, w3, acc[dy][1][3]);
                        }
                }
            }
#pragma unroll
            for (int c = 0; c < 8; c++) {
                auto at = [&](int a, int b) { return (c & 1) ? acc[a][b][c >> 1].y : acc[a][b][c >> 1].x; };
                const float mx = fmaxf(fmaxf(at(0, 0), at(0, 1)), fmaxf(at(1, 0), at(1, 1)));
                sm[G::S_P2 + g * G::FEAT + pos * G::C2 + 8 * q + c] = fmaxf(mx + sm[G::S_W2 + G::NW2 + 8 * q + c], 0.f);  // (h,w,c) flatten
            }
        }
        __syncthreads();
        // ---- fc1: KSL k-slices x 30 float4 columns; each 128-bit weight load from L2 serves all the images of the pass
        // (rows of images past `imgs` hold stale data; their results are never read)
        if (tid < G::KSL * (G::F1 / 4)) {
            const int n4 = tid % (G::F1 / 4), ks = tid / (G::F1 / 4);
            constexpr int KL = G::FEAT / G::KSL;
            const float4* wp = reinterpret_cast<const float4*>(W + L2.w_off) + (size_t)(ks * KL) * (G::F1 / 4) + n4;
            const float* xp = sm + G::S_P2 + ks * KL;
            float2 alo[kGI], ahi[kGI];
#pragma unroll
            for (int g = 0; g < kGI; g++) alo[g] = ahi[g] = make_float2(0.f, 0.f);
#pragma unroll 5
            for (int k = 0; k < KL; k++) {
                const float4 w = __ldg(wp + (size_t)k * (G::F1 / 4));
                const float2 wlo = make_float2(w.x, w.y), whi = make_float2(w.z, w.w);
#pragma unroll
                for (int g = 0; g < kGI; g++) {
                    const float x = xp[g * G::FEAT + k];
                    const float2 xx = make_float2(x, x);
                    alo[g] = __ffma2_rn(xx, wlo, alo[g]);
                    ahi[g] = __ffma2_rn(xx, whi, ahi[g]);
                }
            }
#pragma unroll
            for (int g = 0; g < kGI; g++)
                *reinterpret_cast<float4*>(sm + G::S_PART + (g * G::KSL + ks) * G::F1 + 4 * n4) = make_float4(alo[g].x, alo[g].y, ahi[g].x, ahi[g].y);
        }
        __syncthreads();
        for (int o = tid; o < kGI * G::F1; o += kGT) {
            const int g = o / G::F1, j = o % G::F1;
            float v = __ldg(W + L2.b_off + j);
#pragma unroll
            for (int ks = 0; ks < G::KSL; ks++) v += sm[G::S_PART + (g * G::KSL + ks) * G::F1 + j];
            sm[G::S_H1 + g * G::F1 + j] = fmaxf(v, 0.f);
        }
        __syncthreads();
        // ---- fc2: S2 k-slices x 84 outputs, one weight load per k for all the images of the pass
        if (tid < G::S2 * G::F2) {
            const int j = tid % G::F2, ks = tid / G::F2;
            const float* wp = W + L3.w_off + (size_t)(ks * G::K2) * G::F2 + j;
            float a[kGI];
#pragma unroll
            for (int g = 0; g < kGI; g++) a[g] = 0.f;
#pragma unroll 5
            for (int k = 0; k < G::K2; k++) {
                const float w = __ldg(wp + (size_t)k * G::F2);
#pragma unroll
                for (int g = 0; g < kGI; g++) a[g] = fmaf(sm[G::S_H1 + g * G::F1 + ks * G::K2 + k], w, a[g]);
            }
#pragma unroll
            for (int g = 0; g < kGI; g++) sm[G::S_H2 + (g * G::S2 + ks) * G::F2 + j] = a[g];
        }
        __syncthreads();
        for (int o = tid; o < kGI * G::F2; o += kGT) {
            const int g = o / G::F2, j = o % G::F2;
            float v = __ldg(W + L3.b_off + j);
#pragma unroll
            for (int ks = 0; ks < G::S2; ks++) v += sm[G::S_H2 + (g * G::S2 + ks) * G::F2 + j];
            sm[G::S_H1 + g * G::F1 + j] = fmaxf(v, 0.f);  // h2 reuses the h1 area (fc2 has finished reading it)
        }
        __syncthreads();
        // ---- fc3 + softmax / argmax: warp g handles image g; lane = (k third, class), thirds summed by shuffles
        const int g = tid >> 5, lane = tid & 31;
        if (g < imgs) {
            const int img = list[it + g];
            constexpr int KT = G::F2 / 3;
            const int cls = lane % G::F3, kt = min(lane / G::F3, 2);
            float a = 0.f;
            {
                const float* hp = sm + G::S_H1 + g * G::F1 + kt * KT;
                const float* wp = sm + G::S_W3 + (kt * KT) * G::F3 + cls;
#pragma unroll 7
                for (int k = 0; k < KT; k++) a = fmaf(hp[k], wp[k * G::F3], a);
            }
            const float a1 = __shfl_down_sync(0xffffffffu, a, G::F3), a2 = __shfl_down_sync(0xffffffffu, a, 2 * G::F3);
            const float x = lane < G::F3 ? (a + a1) + a2 + sm[G::S_W3 + G::F2 * G::F3 + lane] : -INFINITY;
            float mx = x;
            int am = lane;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ox = __shfl_xor_sync(0xffffffffu, mx, o);
                const int oa = __shfl_xor_sync(0xffffffffu, am, o);
                if (ox > mx || (ox == mx && oa < am)) { mx = ox; am = oa; }
            }
            const float e = lane < G::F3 ? expf(x - mx) : 0.f;
            float s = e;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane < G::F3) {
                probs[(size_t)img * G::F3 + lane] = e / s;
                if (logits) logits[(size_t)img * G::F3 + lane] = x;
            }
            if (lane == 0 && argmax) argmax[img] = am;
        }
    }
    finish();
}

template <typename G>
int launch_guard(const DeviceModel& m, const float* d_in, int cap, int expected, const GuardScratch& g, float* d_probs,
                 float* d_logits, int32_t* d_argmax, cudaStream_t st) {
    static SmemLimit limit;                   // function attributes are per device
    static std::atomic<int> ctas_per_sm[64];  // per device; zero = not asked yet (threads may race, to the same value)
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    const int smem = G::S_TOTAL * (int)sizeof(float);
    if (int rc = limit.ensure(dev, smem, [](int b) {
            return (int)cudaFuncSetAttribute(guard_fp32_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, b);
        }))
        return rc;
    if (!ctas_per_sm[dev & 63].load()) {
        int occ = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, guard_fp32_kernel<G>, kGT, smem);
        if (e != cudaSuccess) return (int)e;
        ctas_per_sm[dev & 63].store(occ > 0 ? occ : 1);
    }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = max(1, min(expected, sms * ctas_per_sm[dev & 63].load()));  // one wave; a CTA's share may take several passes
    guard_fp32_kernel<G><<<grid, kGT, smem, st>>>(m, d_in, g.d_list, g.d_count, g.d_out_count, g.self_reset ? 1 : 0, cap, d_probs, d_logits, d_argmax);
    return (int)cudaGetLastError();
}

}  // namespace

// Re-runs the images d_list[0 .. *d_count) in fp32 and overwrites their rows. preset: 0 mnist-lenet5, 1 cifar-lenet5.
int launch_guard_fp32(int preset, const DeviceModel& m, const float* d_in, int n, const GuardScratch& g, float* d_probs,
                      float* d_logits, int32_t* d_argmax, void* stream) {
    if (n <= 0) return 0;
    if (preset != 0 && preset != 1) return (int)cudaErrorInvalidValue;  // other shapes: launch_forward_fp32 with the list
    const int expected = min(n, max(256, n / 16));  // grid sizing only; the kernel grid-strides over the real count
    if (preset == 0)
        return launch_guard<GNet<28, 28, 1>>(m, d_in, n, expected, g, d_probs, d_logits, d_argmax, (cudaStream_t)stream);
    return launch_guard<GNet<32, 32, 3>>(m, d_in, n, expected, g, d_probs, d_logits, d_argmax, (cudaStream_t)stream);
}

}  // namespace sib
