// sib_kernels_fp32.cu -- fp32 CUDA-core forward pass for any model the blob can describe.
//
// Replaces Session.run("input:0" -> "output/Softmax:0") (InferenceBolt.java:81-85) with TF 1.x op semantics:
// Conv2D NHWC/HWIO VALID s1 + BiasAdd + Relu + MaxPool 2x2/2, flatten (h,w,c), MatMul+BiasAdd(+Relu), Softmax.
// Used (1) as the SIB_PREC_FP32 engine path, (2) as the margin guard's recompute path for the BF16 tensor-core
// kernel (d_idx != null: only the listed images are re-run, results overwritten in place).
//
// One CTA runs IMGS images end to end; activations never leave shared memory; conv, bias, ReLU and the 2x2 pool
// are one fused stage (each thread owns a pooled position x CO_T output channels = 4*CO_T accumulators).
#include <cuda_runtime.h>

#include "sib_internal.hpp"

namespace sib {
namespace {

constexpr int kThreads = 256;
// Images per CTA. 4 for whole batches: dense-layer weights are read once per CTA and reused across the images.
// 1 for the margin guard's short list (a few hundred images): there the latency of one CTA is what counts, so
// each image gets a CTA of its own and the conv stages are cut into finer work items (CO_T below).
constexpr int kImgsBatch = 4;

template <int CO_T>
__device__ __forceinline__ void conv_relu_pool(const float* __restrict__ in, int in_stride, float* __restrict__ out,
                                               int out_stride, int H, int W, int Cin, int KH, int KW, int Cout,
                                               const float* __restrict__ wt, const float* __restrict__ bias, int imgs) {
    const int OH = H - KH + 1, OW = W - KW + 1, PH = OH >> 1, PW = OW >> 1;
    const int nchunk = (Cout + CO_T - 1) / CO_T;
    const int total = imgs * PH * PW * nchunk;
    const int rowC = W * Cin;
    for (int it = threadIdx.x; it < total; it += blockDim.x) {
        int chunk = it % nchunk;
        int t = it / nchunk;
        int px = t % PW;
        t /= PW;
        int py = t % PH;
        int img = t / PH;
        const int co0 = chunk * CO_T;
        float acc[4][CO_T];
#pragma unroll
        for (int q = 0; q < 4; q++)
#pragma unroll
            for (int c = 0; c < CO_T; c++) acc[q][c] = 0.0f;
        const float* ip = in + img * in_stride + ((2 * py) * W + 2 * px) * Cin;
        for (int kh = 0; kh < KH; kh++) {
            for (int kw = 0; kw < KW; kw++) {
                const float* q = ip + kh * rowC + kw * Cin;
                const float* wp = wt + ((kh * KW + kw) * Cin) * Cout + co0;
                for (int ci = 0; ci < Cin; ci++) {
                    float v00 = q[ci], v01 = q[ci + Cin], v10 = q[ci + rowC], v11 = q[ci + rowC + Cin];
#pragma unroll
                    for (int c = 0; c < CO_T; c++) {
                        float w = (co0 + c < Cout) ? __ldg(wp + ci * Cout + c) : 0.0f;
                        acc[0][c] = fmaf(v00, w, acc[0][c]);
                        acc[1][c] = fmaf(v01, w, acc[1][c]);
                        acc[2][c] = fmaf(v10, w, acc[2][c]);
                        acc[3][c] = fmaf(v11, w, acc[3][c]);
                    }
                }
            }
        }
        float* op = out + img * out_stride + (py * PW + px) * Cout + co0;
#pragma unroll
        for (int c = 0; c < CO_T; c++) {
            if (co0 + c < Cout) {
                float b = __ldg(bias + co0 + c);
                // max over the window of relu(x+b) == relu(max(x)+b): both are monotone
                float m = fmaxf(fmaxf(acc[0][c], acc[1][c]), fmaxf(acc[2][c], acc[3][c])) + b;
                op[c] = fmaxf(m, 0.0f);
            }
        }
    }
}

template <int kImgs>
__device__ __forceinline__ void dense_layer(const float* __restrict__ in, int in_stride, float* __restrict__ out,
                                            int out_stride, int K, int M, const float* __restrict__ wt,
                                            const float* __restrict__ bias, bool relu, int imgs) {
    // thread = (output neuron, k-slice); 4 k-slices per neuron keep more of the CTA busy on 120/84/10-wide layers
    constexpr int KS = 2;
    const int total = M * KS;
    __shared__ float part[kImgs][KS][128];
    for (int base = 0; base < M; base += 128) {
        int j = base + (threadIdx.x >> 1), ks = threadIdx.x & 1;
        float acc[kImgs];
#pragma unroll
        for (int g = 0; g < kImgs; g++) acc[g] = 0.0f;
        if (j < M && j < base + 128) {
            for (int i = ks; i < K; i += KS) {
                float w = __ldg(wt + i * M + j);
#pragma unroll
                for (int g = 0; g < kImgs; g++) acc[g] = fmaf(in[g * in_stride + i], w, acc[g]);
            }
#pragma unroll
            for (int g = 0; g < kImgs; g++) part[g][ks][j - base] = acc[g];
        }
        __syncthreads();
        if (j < M && j < base + 128 && ks == 0) {
            float b = __ldg(bias + j);
#pragma unroll
            for (int g = 0; g < kImgs; g++) {
                if (g < imgs) {
                    float v = part[g][0][j - base] + part[g][1][j - base] + b;
                    out[g * out_stride + j] = relu ? fmaxf(v, 0.0f) : v;
                }
            }
        }
        __syncthreads();
    }
    (void)total;
}

template <int kImgs>
__global__ void __launch_bounds__(kThreads) forward_fp32_kernel(DeviceModel m, const float* __restrict__ in, int n,
                                                               const int32_t* __restrict__ idx,
                                                               const int32_t* __restrict__ count,
                                                               float* __restrict__ probs, float* __restrict__ logits,
                                                               int32_t* __restrict__ argmax, float* __restrict__ margin) {
    extern __shared__ __align__(16) float smem[];
    if (count) n = min(n, *count);
    const int stride = (int)m.max_store_act;
    float* bufA = smem;
    float* bufB = smem + kImgs * stride;
    for (int g0 = blockIdx.x * kImgs; g0 < n; g0 += gridDim.x * kImgs) {
        const int imgs = min(kImgs, n - g0);
        // ---- stage images (coalesced; float4 when the image size allows)
        for (int g = 0; g < kImgs; g++) {
            float* dst = bufA + g * stride;
            if (g < imgs) {
                const int src_img = idx ? idx[g0 + g] : g0 + g;
                const float* src = in + (size_t)src_img * m.img_elems;
                if ((m.img_elems & 3u) == 0) {
                    const float4* s4 = reinterpret_cast<const float4*>(src);
                    float4* d4 = reinterpret_cast<float4*>(dst);
                    for (int i = threadIdx.x; i < (int)(m.img_elems >> 2); i += blockDim.x) d4[i] = __ldg(s4 + i);
                } else {
                    for (int i = threadIdx.x; i < (int)m.img_elems; i += blockDim.x) dst[i] = __ldg(src + i);
                }
            } else {
                for (int i = threadIdx.x; i < (int)m.img_elems; i += blockDim.x) dst[i] = 0.0f;
            }
        }
        __syncthreads();
        float* cur = bufA;
        float* nxt = bufB;
        int H = m.in_h, W = m.in_w, Cc = m.in_c;
        for (uint32_t li = 0; li < m.n_layers; li++) {
            const DeviceLayer L = m.layer[li];
            const float* wt = m.weights + L.w_off;
            const float* bs = m.weights + L.b_off;
            if (L.kind == kKindConv) {
                // output channels per work item: wide for batches (register reuse), narrow for the 1-image guard CTA
                if (kImgs == 1 && L.d > 6) conv_relu_pool<2>(cur, stride, nxt, stride, H, W, Cc, L.a, L.b, L.d, wt, bs, kImgs);
                else if (kImgs == 1) conv_relu_pool<3>(cur, stride, nxt, stride, H, W, Cc, L.a, L.b, L.d, wt, bs, kImgs);
                else if (L.d <= 6) conv_relu_pool<6>(cur, stride, nxt, stride, H, W, Cc, L.a, L.b, L.d, wt, bs, kImgs);
                else conv_relu_pool<8>(cur, stride, nxt, stride, H, W, Cc, L.a, L.b, L.d, wt, bs, kImgs);
                H = (H - L.a + 1) >> 1;
                W = (W - L.b + 1) >> 1;
                Cc = L.d;
                __syncthreads();
            } else {
                dense_layer<kImgs>(cur, stride, nxt, stride, L.a, L.b, wt, bs, (L.flags & 1u) != 0, kImgs);
            }
            float* t = cur; cur = nxt; nxt = t;
        }
        // ---- softmax + argmax + top1-top2 margin: one warp per image, lanes hold the logits
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        if (warp < imgs) {
            const int nc = (int)m.n_classes;
            const int dst_img = idx ? idx[g0 + warp] : g0 + warp;
            float x = lane < nc ? cur[warp * stride + lane] : -INFINITY;
            float mx = x;
            int am = lane;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                float ox = __shfl_xor_sync(0xffffffffu, mx, o);
                int oa = __shfl_xor_sync(0xffffffffu, am, o);
                if (ox > mx || (ox == mx && oa < am)) { mx = ox; am = oa; }
            }
            float second = (lane == am) ? -INFINITY : x;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) second = fmaxf(second, __shfl_xor_sync(0xffffffffu, second, o));
            float e = lane < nc ? expf(x - mx) : 0.0f;
            float s = e;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane < nc) {
                probs[(size_t)dst_img * nc + lane] = e / s;
                if (logits) logits[(size_t)dst_img * nc + lane] = x;
            }
            if (lane == 0) {
                if (argmax) argmax[dst_img] = am;
                if (margin) margin[dst_img] = mx - second;
            }
        }
        __syncthreads();
    }
}

}  // namespace

int fp32_smem_bytes(const DeviceModel& m) { return (int)(2u * kImgsBatch * m.max_store_act * sizeof(float)); }

template <int kImgs>
static int launch_inst(const DeviceModel& m, const float* d_in, int n, int expected, const int32_t* d_idx,
                       const int32_t* d_count, float* d_probs, float* d_logits, int32_t* d_argmax, float* d_margin,
                       cudaStream_t st) {
    static SmemLimit limit;  // function attributes are per device
    const int smem = (int)(2u * kImgs * m.max_store_act * sizeof(float));
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    if (int rc = limit.ensure(dev, smem, [](int b) {
            return (int)cudaFuncSetAttribute(forward_fp32_kernel<kImgs>, cudaFuncAttributeMaxDynamicSharedMemorySize, b);
        }))
        return rc;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int ctas = (expected + kImgs - 1) / kImgs;  // the kernel grid-strides, so fewer CTAs than items is fine
    const int per_sm = max(1, min(8, (227 * 1024) / (smem + 1024)));
    const int grid = min(ctas, sms * per_sm);
    forward_fp32_kernel<kImgs><<<grid, kThreads, smem, st>>>(m, d_in, n, d_idx, d_count, d_probs, d_logits, d_argmax, d_margin);
    return (int)cudaGetLastError();
}

int launch_forward_fp32(const DeviceModel& m, const float* d_in, int n, const int32_t* d_idx, const int32_t* d_count,
                        float* d_probs, float* d_logits, int32_t* d_argmax, float* d_margin, void* stream) {
    if (n <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    // d_idx != null is the margin guard: a short list whose length is only known on the device. Size the grid for
    // a generous share of the batch (the list is ~2% of it); CTAs past the real count exit at once.
    if (d_idx) return launch_inst<1>(m, d_in, n, min(n, max(1024, n / 8)), d_idx, d_count, d_probs, d_logits, d_argmax, d_margin, st);
    return launch_inst<kImgsBatch>(m, d_in, n, n, d_idx, d_count, d_probs, d_logits, d_argmax, d_margin, st);
}

}  // namespace sib
