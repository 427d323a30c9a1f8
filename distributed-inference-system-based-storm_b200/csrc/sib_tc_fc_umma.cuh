// sib_tc_fc_umma.cuh -- the dense head (FEAT -> 120 -> 84 -> 10 -> softmax) on tcgen05.mma, 16 images (one pair) at a time.
// Included by sib_kernels_tc.cu (namespace sib::{anonymous}) with SIB_CONV2_UMMA (the default).
//
// The mma.sync head was one warp running 483 dependent-ish HMMA + 640 fragment loads per pair: ~1400 clk per image,
// the slowest stage once conv2 had left the legacy tensor path (profiles/ncu_tc_r02_notes.md). An M = 128 tile of
// images would need 100 KB of pooled features in shared memory next to 121 KB of weights, so the product is
// transposed instead: the WEIGHTS are the M = 128 operand (row = output feature), the pair's 16 images are N:
//     D1[120 x 16] = W1^T[120 x FEAT] . p2^T      D2[84 x 16] = W2^T . h1^T      D3[10 x 16] = W3^T . h2^T
// All operands are read from shared memory, K-major without swizzle; conv2's epilogue already writes p2 as 16-byte
// (8 channels of one image) chunks, which is exactly that layout's unit, and h1 / h2 are written the same way by the
// epilogues here. Weight rows that do not exist (120..127, 88..127, 16..127) are whatever follows in shared memory:
// they only produce accumulator rows nobody reads. 39 mma per pair (25 + 8 + 6), ~39 clk each (shared-memory operand).
// The accumulator's 128 rows span all four TMEM lane quadrants, so the epilogues are spread over the warps that own
// them: the three conv2 workers (quadrants 0..2, between two of their rows) and this head warp (quadrant 3), which also
// issues the mma. fc3's weight rows are placed at rows 96.. so the logits land in the head warp's own quadrant.
#pragma once

__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {  // K-major, no swizzle
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// D[tmem] (+)= A[smem] . B[smem]^T ; issued by ONE elected lane of a converged warp
__device__ __forceinline__ void umma_ss(uint32_t d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    if (SIB_DIAG_SKIP & 8) return;  // diagnostic: everything but the tensor work
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(d), "l"(adesc), "l"(bdesc),
        "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void sts16(uint32_t addr, uint16_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(v) : "memory"); }

template <typename T> __device__ __forceinline__ uint16_t to_bits(float x);
template <> __device__ __forceinline__ uint16_t to_bits<__half>(float x) { return __half_as_ushort(__float2half_rn(x)); }
template <> __device__ __forceinline__ uint16_t to_bits<__nv_bfloat16>(float x) { return __bfloat16_as_ushort(__float2bfloat16_rn(x)); }

template <typename N>
struct FcBars {
    uint32_t f_full, h1_full, h2_full, fcw;
    __device__ __forceinline__ explicit FcBars(unsigned char* smem) {
        const uint32_t b = smem_u32(smem + N::S_BAR);
        f_full = b + N::B_FC * 8; h1_full = f_full + 24; h2_full = f_full + 32; fcw = b + N::B_FCW * 8;
    }
};
constexpr int kTmF = 480;  // TMEM columns of the head's accumulator (16): D1, then D2, then D3 of a pair in turn

// Epilogue of fc1 (stage 1: h1 = relu(D1 + b1), K padded to 128 with zeros) or fc2 (stage 2: h2, K padded to 96) for
// the accumulator rows of lane quadrant q: row f = 32 q + lane is one feature of all 16 images.
template <typename N, typename T, bool DBG, int STAGE>
__device__ __forceinline__ void fcu_epilogue(const TcParams& P, unsigned char* smem, int q, int lane, uint32_t tmem, int Qi) {
    const FcBars<N> F(smem);
    constexpr int REAL = STAGE == 1 ? 120 : 84, KPAD = STAGE == 1 ? 128 : 96, SBO = STAGE == 1 ? N::H1_SBO : N::H2_SBO;
    const int f = q * 32 + lane;
    const float b = f < REAL ? reinterpret_cast<const float*>(smem + N::S_BIAS)[(STAGE == 1 ? 24 : 152) + f] : 0.f;
    mbar_wait_s(F.f_full + (STAGE - 1) * 8, Qi & 1);
    tc5_fence_after();
    float v[16];
    tmem_ld16(tmem + ((uint32_t)(q * 32) << 16) + kTmF, v);
    tmem_wait_ld();
    tc5_fence_before();
    const uint32_t dst = smem_u32(smem + (STAGE == 1 ? N::S_H1 : N::S_H2)) + (f >> 3) * N::BK + (f & 7) * 2;
    if (f < KPAD) {  // features past the padded K do not exist in the next layer's operand
#pragma unroll
        for (int n = 0; n < 16; n++) {
            const float x = f < REAL ? fmax_nan(v[n] + b, 0.f) : 0.f;
            sts16(dst + (n >> 3) * SBO + (n & 7) * 16, to_bits<T>(x));
            if (DBG && STAGE == 1 && P.dbg && blockIdx.x == 0 && Qi == 0 && f < 120)
                P.dbg[2 * kG * N::PH1 * N::P1_ROW + N::DBG_P2 + n * 128 + f] = x;
        }
    }
    fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's (async proxy) operand reads
    mbar_arrive_s(STAGE == 1 ? F.h1_full : F.h2_full);
}

// ---- head warp (hardware warp 4k + 3): issues the pair's 39 mma, owns quadrant 3 of fc1 and all of fc3 + softmax ------
template <typename N, typename T, bool DBG>
__device__ __forceinline__ void fcu_head(const TcParams& P, unsigned char* smem, int lane, uint32_t tmem) {
    const U2Bars<N> B(smem);
    const FcBars<N> F(smem);
    const uint32_t a1 = smem_u32(smem + N::S_FCW), a2 = a1 + N::F1_GROUPS * N::F1_SBO;
    const uint32_t a3 = a2 + N::F2_GROUPS * N::F2_SBO - 12 * N::F3_SBO;  // fc3's rows sit at M rows 96..: quadrant 3
    const uint32_t b_p2 = smem_u32(smem + N::S_P2), b_h1 = smem_u32(smem + N::S_H1), b_h2 = smem_u32(smem + N::S_H2);
    const uint32_t idesc = umma_idesc_m128_n16<T>();
    const uint32_t d = tmem + kTmF;
    float* scratch = reinterpret_cast<float*>(smem + N::S_FCS);  // [16 images][10 classes]
    const float b3 = lane < 10 ? reinterpret_cast<const float*>(smem + N::S_BIAS)[248 + lane] : 0.f;
    mbar_wait_s(F.fcw, 0);  // the weight image is in
    if (SIB_DIAG_SKIP & 4) return;
    for (int Qi = 0;; Qi++) {
        int base, lim;
        if (!pair_images(P, Qi, base, lim)) break;
        const int buf = Qi % N::NP2, use = Qi / N::NP2;
        const uint32_t ph = Qi & 1;
        // ---- fc1
        mbar_wait_s(B.p2_full + buf * 8, use & 1);
        tc5_fence_after();
        if (elect_one()) {
#pragma unroll 1
            for (int ks = 0; ks < N::KS1; ks++)
                umma_ss(d, umma_desc(a1 + ks * 256, 128, N::F1_SBO), umma_desc(b_p2 + buf * N::P2_BUF * 2 + ks * 2 * N::BK, N::BK, N::P2_SBO), idesc, ks > 0);
            umma_commit_s(F.f_full);
            umma_commit_s(B.p2_empty + buf * 8);  // the pair's features have been read
        }
        __syncwarp();
        fcu_epilogue<N, T, DBG, 1>(P, smem, 3, lane, tmem, Qi);
        // ---- fc2
        mbar_wait_s(F.h1_full, ph);
        tc5_fence_after();
        if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 8; ks++)
                umma_ss(d, umma_desc(a2 + ks * 256, 128, N::F2_SBO), umma_desc(b_h1 + ks * 2 * N::BK, N::BK, N::H1_SBO), idesc, ks > 0);
            umma_commit_s(F.f_full + 8);
        }
        __syncwarp();
        // ---- fc3
        mbar_wait_s(F.h2_full, ph);
        tc5_fence_after();
        if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 6; ks++)
                umma_ss(d, umma_desc(a3 + ks * 256, 128, N::F3_SBO), umma_desc(b_h2 + ks * 2 * N::BK, N::BK, N::H2_SBO), idesc, ks > 0);
            umma_commit_s(F.f_full + 16);
        }
        __syncwarp();
        mbar_wait_s(F.f_full + 16, ph);
        tc5_fence_after();
        float x[16];
        tmem_ld16(tmem + ((uint32_t)96 << 16) + kTmF, x);  // lane c < 10: class c of the pair's 16 images
        tmem_wait_ld();
        tc5_fence_before();
        if (lane < 10) {
#pragma unroll
            for (int n = 0; n < 16; n++) scratch[n * 10 + lane] = x[n] + b3;
        }
        __syncwarp();
        const int img = base + lane;
        if (lane < 16 && img < lim) {  // one image per lane: softmax, argmax (first maximum), top1 - top2 margin
            float lg[10];
#pragma unroll
            for (int c = 0; c < 10; c++) lg[c] = scratch[lane * 10 + c];
            float mx = lg[0], second = -INFINITY;
            int am = 0;
            bool finite = fabsf(lg[0]) <= 3.0e38f;
#pragma unroll
            for (int c = 1; c < 10; c++) finite = finite && fabsf(lg[c]) <= 3.0e38f;  // false for NaN and Inf
#pragma unroll
            for (int c = 1; c < 10; c++) {
                if (lg[c] > mx) { second = mx; mx = lg[c]; am = c; }
                else second = fmaxf(second, lg[c]);
            }
            float e[10], s = 0.f;
#pragma unroll
            for (int c = 0; c < 10; c++) { e[c] = expf(lg[c] - mx); s += e[c]; }
            float* pr = P.probs + (size_t)img * 10;
#pragma unroll
            for (int c = 0; c < 10; c += 2) *reinterpret_cast<float2*>(pr + c) = make_float2(e[c] / s, e[c + 1] / s);
            if (P.logits) {
                float* lo = P.logits + (size_t)img * 10;
#pragma unroll
                for (int c = 0; c < 10; c += 2) *reinterpret_cast<float2*>(lo + c) = make_float2(lg[c], lg[c + 1]);
            }
            if (P.argmax) P.argmax[img] = am;
            const float mg = mx - second;
            if (P.margin) P.margin[img] = mg;
            // guard: recompute in fp32. The half-precision error grows with the size of the logits, so the threshold does
            // too (tau is calibrated for |logit| ~ 12); a NaN margin (overflowed activations) is listed as well.
            // A logit that is not finite means an activation overflowed half precision: same treatment.
            if (P.list && (!(mg >= P.tau * fmaxf(1.f, fabsf(mx) * (1.f / 12.f))) || !finite)) {
                const int pos = atomicAdd(P.count, 1);
                if (pos < P.n) P.list[pos] = img;  // a stale counter can never push the list past its capacity
            }
        }
        __syncwarp();  // the scratch is rewritten by the next pair
    }
}
