// sib_kernels_tcgen.cu -- tensor-core forward pass for ANY model of the blob's family (sm_100a), half or bfloat16 operands.
//
// The reference runs whatever graph model/saved_model.pb holds (InferenceBolt.java:57, :81-85); the persistent
// whole-network kernel (sib_kernels_tc.cu) is built around the two LeNet-5 presets' shapes. Every other model of the family
//     [conv KxK VALID s1 + bias + ReLU + maxpool 2x2/2]* -> flatten(h,w,c) -> [dense (+ReLU)]* -> softmax
// -- the wide presets of SURVEY.md D1, a user's own SavedModel -- takes this kernel instead of the 20x slower fp32 one:
//   * one CTA runs 8 images end to end, activations in shared memory in half precision, never in HBM;
//   * convolution = implicit GEMM on mma.sync.m16n8k16: a tile is two output rows x 8 columns of one image (fragment rows g
//     and g+8 are a vertical pool pair, lanes g and g^1 a horizontal one: the 2x2 pool is one max and one shuffle), K runs
//     over one kernel row at a time (its KW x Cin halves are contiguous in the NHWC map), the layer's packed B fragments sit
//     in shared memory, bias + ReLU + pool are the epilogue;
//   * dense layers: the 8 images are the (half-empty) 16-row M tile, B fragments stream from L2 once per CTA pass;
//   * softmax / argmax / top1-top2 margin by warp shuffles; low-margin and non-finite images go on the guard's list and are
//     re-run by forward_fp32_kernel (the generic fp32 path), exactly like the fused kernel's list.
// It is a layer-by-layer kernel on the legacy tensor path: the shapes are run-time values, so there are no register-resident
// sliding windows and no TMEM pipeline here; that specialisation is what sib_kernels_tc.cu is.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>
#include <vector>

#include "sib_internal.hpp"

namespace sib {
namespace {

constexpr int kGI = 8;        // images per CTA pass
constexpr int kGT = 512;      // threads (16 warps: 256 ran the wide presets 1.6x slower)
constexpr int kSlack = 32;    // halves behind every image's activation: K padding may read past a row's end (x zero weights)

template <typename T> struct GOps;
template <> struct GOps<__half> {
    __device__ static __forceinline__ uint16_t bits(float x) { return __half_as_ushort(__float2half_rn(x)); }
    __device__ static __forceinline__ void mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
    static uint16_t host_bits(float x) { return __half_as_ushort(__float2half_rn(x)); }
};
template <> struct GOps<__nv_bfloat16> {
    __device__ static __forceinline__ uint16_t bits(float x) { return __bfloat16_as_ushort(__float2bfloat16_rn(x)); }
    __device__ static __forceinline__ void mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
    static uint16_t host_bits(float x) { return __bfloat16_as_ushort(__float2bfloat16_rn(x)); }
};

__device__ __forceinline__ float fmax_nan_g(float a, float b) {  // hands a NaN on (an overflowed activation must reach the logits)
    float d;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}

struct GParams {
    TcgPlan plan;
    const uint32_t* wpack;  // packed B fragments of every layer (+ fp32 biases), device
    const float* in;
    int n;
    float* probs;
    float* logits;
    int32_t* argmax;
    float* margin;
    int32_t* list;
    int32_t* count;
    float tau;
};

// two halves at an address that may be odd (Cin odd): always correct, used by the first conv layer of 1- and 3-channel inputs
__device__ __forceinline__ uint32_t ld_pair(const uint16_t* p, bool aligned) {
    if (aligned) return *reinterpret_cast<const uint32_t*>(p);
    return (uint32_t)p[0] | ((uint32_t)p[1] << 16);
}

template <typename T>
__global__ void __launch_bounds__(kGT) forward_tcgen_kernel(const __grid_constant__ GParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const TcgPlan& pl = P.plan;
    const int stride = pl.act_stride;                                   // halves per image and buffer
    uint16_t* bufA = reinterpret_cast<uint16_t*>(smem_raw);
    uint16_t* bufB = bufA + kGI * stride;
    uint32_t* wsm = reinterpret_cast<uint32_t*>(bufB + kGI * stride);   // the current conv layer's B fragments
    float* lg = reinterpret_cast<float*>(wsm + pl.wsm_words);           // [kGI][n_classes] fp32 logits
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int nc = pl.n_classes;
    for (int i = tid; i < 2 * kGI * stride / 2; i += kGT) reinterpret_cast<uint32_t*>(bufA)[i] = 0u;  // slack and tails are finite
    __syncthreads();
    for (int g0 = blockIdx.x * kGI; g0 < P.n; g0 += gridDim.x * kGI) {
        const int imgs = min(kGI, P.n - g0);
        // ---- stage: fp32 NHWC -> half
        for (int im = 0; im < kGI; im++) {
            uint16_t* dst = bufA + im * stride;
            const float* src = P.in + (size_t)(g0 + min(im, imgs - 1)) * pl.img_elems;  // tail slots repeat the last image
            for (int i = tid; i < pl.img_elems; i += kGT) dst[i] = GOps<T>::bits(__ldg(src + i));
        }
        uint16_t* cur = bufA;
        uint16_t* nxt = bufB;
        int H = pl.in_h, W = pl.in_w, C = pl.in_c;
        for (int li = 0; li < pl.n_layers; li++) {
            const TcgLayer L = pl.layer[li];
            const float* bias = reinterpret_cast<const float*>(P.wpack + L.b_off);
            if (L.kind == kKindConv) {
                // ---- this layer's fragments into shared memory: [kh][ks][nt][lane][2]
                const int KS = L.ks, NT = L.nt, KH = L.a, KW = L.b, Cout = L.d;
                __syncthreads();  // the previous layer (and the staging) is done with cur / wsm
                for (int i = tid; i < KH * KS * NT * 64; i += kGT) wsm[i] = __ldg(P.wpack + L.w_off + i);
                __syncthreads();
                const int OH = H - KH + 1, OW = W - KW + 1, PH = OH >> 1, PW = OW >> 1;
                const int JT = (2 * PW + 7) >> 3;                    // column tiles of 8 output columns
                const int tiles = kGI * PH * JT;
                const bool aligned = (C & 1) == 0;
                for (int tile = warp; tile < tiles; tile += kGT / 32) {
                    const int j = tile % JT, ph = (tile / JT) % PH, im = tile / (JT * PH);
                    const int ow = min(8 * j + g, OW - 1);           // clamped: columns past the map are computed and dropped
                    const uint16_t* r0 = cur + im * stride + ((2 * ph) * W + ow) * C;  // row g: output row 2ph; row g+8: 2ph+1
                    float acc[8][4];
#pragma unroll
                    for (int n8 = 0; n8 < 8; n8++) acc[n8][0] = acc[n8][1] = acc[n8][2] = acc[n8][3] = 0.f;
                    // (kernel row, K step) pairs in one flat loop, the next step's A fragment loaded before this step's mma:
                    // the loads then have NT mma to land under (+3% on the wide presets: the loop is not bound by them)
                    const uint16_t* a_lo = r0 + 2 * t;
                    const int steps = KH * KS, wrap = W * C - KS * 16;
                    uint32_t a[4], an[4];
                    a[0] = ld_pair(a_lo, aligned); a[1] = ld_pair(a_lo + W * C, aligned);
                    a[2] = ld_pair(a_lo + 8, aligned); a[3] = ld_pair(a_lo + W * C + 8, aligned);
                    int ks = 0;
                    for (int st = 0; st < steps; st++) {
                        a_lo += 16;
                        if (++ks == KS) { ks = 0; a_lo += wrap; }
                        if (st + 1 < steps) {
                            an[0] = ld_pair(a_lo, aligned); an[1] = ld_pair(a_lo + W * C, aligned);
                            an[2] = ld_pair(a_lo + 8, aligned); an[3] = ld_pair(a_lo + W * C + 8, aligned);
                        }
                        const uint32_t* bp = wsm + (st * NT) * 64 + lane * 2;
#pragma unroll
                        for (int n8 = 0; n8 < 8; n8++)
                            if (n8 < NT) {
                                const uint2 b = *reinterpret_cast<const uint2*>(bp + n8 * 64);
                                GOps<T>::mma(acc[n8], a, b.x, b.y);
                            }
#pragma unroll
                        for (int i = 0; i < 4; i++) a[i] = an[i];
                    }
                    // epilogue: vertical max in the thread, horizontal max with the lane of the next column, + bias, ReLU
                    const int col = 8 * j + g;
                    const bool writer = (g & 1) == 0 && col + 1 < 2 * PW;
                    uint16_t* op = nxt + im * stride + (ph * PW + (col >> 1)) * Cout;
#pragma unroll
                    for (int n8 = 0; n8 < 8; n8++)
                        if (n8 < NT) {
                            float v0 = fmax_nan_g(acc[n8][0], acc[n8][2]), v1 = fmax_nan_g(acc[n8][1], acc[n8][3]);
                            v0 = fmax_nan_g(v0, __shfl_xor_sync(0xffffffffu, v0, 4));
                            v1 = fmax_nan_g(v1, __shfl_xor_sync(0xffffffffu, v1, 4));
                            const int c0 = n8 * 8 + 2 * t;
                            if (writer && c0 < Cout) {
                                op[c0] = GOps<T>::bits(fmax_nan_g(v0 + bias[c0], 0.f));
                                if (c0 + 1 < Cout) op[c0 + 1] = GOps<T>::bits(fmax_nan_g(v1 + bias[c0 + 1], 0.f));
                            }
                        }
                }
                H = PH; W = PW; C = Cout;
            } else {
                // ---- dense: rows 0..7 of the M tile are the images, rows 8..15 stay zero; fragments stream from L2
                __syncthreads();
                const int K = L.a, M = L.b, KS = L.ks, NT = L.nt;
                const bool last = li == pl.n_layers - 1, relu = (L.flags & 1u) != 0;
                const uint16_t* ap = cur + g * stride + 2 * t;
                for (int n8 = warp; n8 < NT; n8 += kGT / 32) {
                    float acc[4] = {0.f, 0.f, 0.f, 0.f};
                    const uint32_t* bp = P.wpack + L.w_off + (size_t)n8 * 64 + lane * 2;
#pragma unroll 4
                    for (int ks = 0; ks < KS; ks++) {
                        uint32_t a[4];
                        a[0] = *reinterpret_cast<const uint32_t*>(ap + ks * 16);
                        a[2] = *reinterpret_cast<const uint32_t*>(ap + ks * 16 + 8);
                        a[1] = a[3] = 0u;
                        const uint2 b = __ldg(reinterpret_cast<const uint2*>(bp + (size_t)ks * NT * 64));
                        GOps<T>::mma(acc, a, b.x, b.y);
                    }
                    const int c0 = n8 * 8 + 2 * t;
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const int c = c0 + e;
                        if (c < M) {
                            float v = acc[e] + bias[c];
                            if (relu) v = fmax_nan_g(v, 0.f);
                            if (last) lg[g * nc + c] = v;
                            else nxt[g * stride + c] = GOps<T>::bits(v);
                        }
                    }
                }
                (void)K;
                C = M; H = W = 1;
            }
            uint16_t* sw = cur; cur = nxt; nxt = sw;
        }
        __syncthreads();
        // ---- softmax, argmax (first maximum), margin, guard list: one warp per image
        if (warp < imgs) {
            const int img = g0 + warp;
            const float x = lane < nc ? lg[warp * nc + lane] : -INFINITY;
            float mx = x;
            int am = lane;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ox = __shfl_xor_sync(0xffffffffu, mx, o);
                const int oa = __shfl_xor_sync(0xffffffffu, am, o);
                if (ox > mx || (ox == mx && oa < am)) { mx = ox; am = oa; }
            }
            float second = (lane == am) ? -INFINITY : x;
            bool finite = lane >= nc || fabsf(x) <= 3.0e38f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) second = fmaxf(second, __shfl_xor_sync(0xffffffffu, second, o));
            finite = __all_sync(0xffffffffu, finite);
            const float e = lane < nc ? expf(x - mx) : 0.f;
            float s = e;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane < nc) {
                P.probs[(size_t)img * nc + lane] = e / s;
                if (P.logits) P.logits[(size_t)img * nc + lane] = x;
            }
            if (lane == 0) {
                if (P.argmax) P.argmax[img] = am;
                const float mg = mx - second;
                if (P.margin) P.margin[img] = mg;
                if (P.list && (!(mg >= P.tau * fmaxf(1.f, fabsf(mx) * (1.f / 12.f))) || !finite)) {
                    const int pos = atomicAdd(P.count, 1);
                    if (pos < P.n) P.list[pos] = img;
                }
            }
        }
        __syncthreads();
    }
}

// B fragment of mma.m16n8k16 (col layout): lane (g, t) holds {B[2t][g], B[2t+1][g]} and {B[2t+8][g], B[2t+9][g]}
template <typename T, typename F>
void pack_b(uint32_t* dst, F at) {
    for (int lane = 0; lane < 32; lane++) {
        const int gg = lane >> 2, tt = lane & 3;
        for (int r = 0; r < 2; r++) {
            const int k = 2 * tt + 8 * r;
            dst[lane * 2 + r] = (uint32_t)GOps<T>::host_bits(at(k, gg)) | ((uint32_t)GOps<T>::host_bits(at(k + 1, gg)) << 16);
        }
    }
}

template <typename T>
std::vector<uint32_t> pack_generic(const ModelDesc& m, TcgPlan* pl) {
    std::vector<uint32_t> w;
    const float* wt = m.weights.data();
    for (uint32_t li = 0; li < m.n_layers; li++) {
        const LayerDesc& L = m.layer[li];
        TcgLayer& G = pl->layer[li];
        G.w_off = (uint32_t)w.size();
        if (L.kind == kKindConv) {
            const int KH = (int)L.a, KW = (int)L.b, Cin = (int)L.c, Cout = (int)L.d, KR = KW * Cin;
            const float* f = wt + L.w_off;  // [kh][kw][cin][cout]
            w.resize(w.size() + (size_t)KH * G.ks * G.nt * 64);
            for (int kh = 0; kh < KH; kh++)
                for (int ks = 0; ks < G.ks; ks++)
                    for (int nt = 0; nt < G.nt; nt++)
                        pack_b<T>(&w[G.w_off + ((size_t)(kh * G.ks + ks) * G.nt + nt) * 64], [&](int k, int n) -> float {
                            const int kk = ks * 16 + k, co = nt * 8 + n;
                            return (kk < KR && co < Cout) ? f[(size_t)(kh * KR + kk) * Cout + co] : 0.f;
                        });
        } else {
            const int K = (int)L.a, M = (int)L.b;
            const float* f = wt + L.w_off;  // [in][out]
            w.resize(w.size() + (size_t)G.ks * G.nt * 64);
            for (int ks = 0; ks < G.ks; ks++)
                for (int nt = 0; nt < G.nt; nt++)
                    pack_b<T>(&w[G.w_off + ((size_t)ks * G.nt + nt) * 64], [&](int k, int n) -> float {
                        const int i = ks * 16 + k, j = nt * 8 + n;
                        return (i < K && j < M) ? f[(size_t)i * M + j] : 0.f;
                    });
        }
        G.b_off = (uint32_t)w.size();
        const int nb = (int)(L.kind == kKindConv ? L.d : L.b);
        w.resize(w.size() + (size_t)((nb + 3) & ~3));
        std::memcpy(&w[G.b_off], wt + L.b_off, (size_t)nb * sizeof(float));
    }
    return w;
}

}  // namespace

// Fills the plan; false (with a reason) if the shapes do not fit this kernel.
bool tcg_plan(const ModelDesc& m, TcgPlan* pl, std::string* why) {
    std::memset(pl, 0, sizeof *pl);
    auto no = [&](const char* s) { if (why) *why = s; return false; };
    if (m.n_layers == 0 || m.n_layers > (uint32_t)kMaxLayers || m.n_classes > 32) return no("more than 32 classes");
    pl->in_h = (int)m.in_h; pl->in_w = (int)m.in_w; pl->in_c = (int)m.in_c; pl->n_layers = (int)m.n_layers;
    pl->n_classes = (int)m.n_classes; pl->img_elems = (int)m.img_elems;
    int H = pl->in_h, W = pl->in_w, C = pl->in_c, wsm = 0;
    size_t act = m.img_elems;
    bool dense_seen = false;
    for (uint32_t li = 0; li < m.n_layers; li++) {
        const LayerDesc& L = m.layer[li];
        TcgLayer& G = pl->layer[li];
        G.kind = L.kind; G.a = L.a; G.b = L.b; G.c = L.c; G.d = L.d; G.flags = L.flags;
        if (L.kind == kKindConv) {
            if (dense_seen) return no("conv after dense");
            if ((int)L.c != C || (int)L.a > H || (int)L.b > W) return no("conv shape");
            if (L.d > 64 || (L.d & 1)) return no("conv output channels must be even and <= 64");
            G.ks = ((int)(L.b * L.c) + 15) / 16;
            G.nt = ((int)L.d + 7) / 8;
            wsm = std::max(wsm, (int)L.a * G.ks * G.nt * 64);
            H = (H - (int)L.a + 1) / 2; W = (W - (int)L.b + 1) / 2; C = (int)L.d;
            if (H < 1 || W < 1) return no("map vanishes");
            act = std::max(act, (size_t)H * W * C);
        } else {
            if (!dense_seen && (size_t)H * W * C != L.a) return no("flatten size");
            dense_seen = true;
            if (L.a & 1) return no("dense input width must be even");
            G.ks = ((int)L.a + 15) / 16;
            G.nt = ((int)L.b + 7) / 8;
            act = std::max(act, (size_t)L.b);
        }
    }
    if (!dense_seen || m.layer[m.n_layers - 1].b != m.n_classes) return no("last layer");
    pl->act_stride = (int)((act + kSlack + 15) & ~(size_t)15);
    pl->wsm_words = wsm;
    pl->smem_bytes = 2 * kGI * pl->act_stride * 2 + wsm * 4 + kGI * pl->n_classes * 4;
    if (pl->smem_bytes > 227 * 1024) return no("activations of 8 images exceed shared memory");
    return true;
}

std::vector<uint32_t> tcg_pack(const ModelDesc& m, int precision, TcgPlan* pl) {
    return precision == SIB_PREC_BF16 ? pack_generic<__nv_bfloat16>(m, pl) : pack_generic<__half>(m, pl);
}

int launch_forward_tcgen(const TcWeights& w, const float* d_in, int n, float* d_probs, float* d_logits, int32_t* d_argmax,
                         GuardScratch* guard, float tau, void* stream) {
    if (n <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    GParams p{};
    p.plan = w.plan;
    p.wpack = static_cast<const uint32_t*>(w.d_packed);
    p.in = d_in; p.n = n; p.probs = d_probs; p.logits = d_logits; p.argmax = d_argmax; p.tau = tau;
    if (guard) {
        p.margin = guard->d_margin; p.list = guard->d_list; p.count = guard->d_count;
        if (!guard->self_reset) {
            cudaError_t e = cudaMemsetAsync(guard->d_count, 0, sizeof(int32_t), st);
            if (e != cudaSuccess) return (int)e;
        }
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const bool bf = w.precision == SIB_PREC_BF16;
    static SmemLimit lim_h, lim_b;
    const int smem = w.plan.smem_bytes;
    if (int rc = (bf ? lim_b : lim_h).ensure(dev, smem, [bf](int b) {
            return bf ? (int)cudaFuncSetAttribute(forward_tcgen_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, b)
                      : (int)cudaFuncSetAttribute(forward_tcgen_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, b);
        }))
        return rc;
    const int per_sm = std::max(1, std::min(4, (227 * 1024) / std::max(smem, 1)));
    const int passes = (n + kGI - 1) / kGI;
    const int grid = std::max(1, std::min(passes, sms * per_sm));
    if (bf) forward_tcgen_kernel<__nv_bfloat16><<<grid, kGT, smem, st>>>(p);
    else forward_tcgen_kernel<__half><<<grid, kGT, smem, st>>>(p);
    return (int)cudaGetLastError();
}

}  // namespace sib
