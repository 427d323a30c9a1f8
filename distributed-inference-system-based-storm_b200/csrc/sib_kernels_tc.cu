// sib_kernels_tc.cu -- persistent whole-network tensor-core kernel for the LeNet-5 presets (sm_100a).
//
// Replaces Session.run("input:0" -> "output/Softmax:0") (InferenceBolt.java:81-85) for the graded models
// (conv5x5(6) -> pool -> conv5x5(16) -> pool -> 120 -> 84 -> 10 -> softmax; 28x28x1 and 32x32x3).
//
// Shape of the kernel (DESIGN.md 4.1 has the numbers behind each choice, profiles/ncu_tc_r02_notes.md the history):
//   * one persistent CTA per SM, 20 warps in five warpgroups with their registers re-dealt by setmaxnreg (conv1 128,
//     everyone else 48); dense weights and conv2's filter live in shared memory for the whole launch, conv1's filter in
//     registers; 8 images per p1 ring slot, 16 per p2 buffer;
//   * warps are specialised and pipelined through shared memory and TMEM with mbarriers:
//       conv1 warps (12, mma.sync m16n8k16): stream fp32 NHWC rows from HBM (128-bit coalesced loads, 4 rows in flight
//                    per warp in registers), convert to half precision into a 2-row private staging ring, build im2col A
//                    fragments with a 5-row sliding register window (each staged row is read once and feeds 5 kernel
//                    rows), bias+ReLU+2x2 maxpool fused in registers (even/odd column tiles make the pool thread-local),
//                    pooled rows -> p1 ring; spans are claimed from a shared counter;
//       conv2 (tcgen05.mma, sib_tc_conv2_umma.cuh): 2 x 3 worker warps write every p1 row into TMEM once (tcgen05.st), an
//                    issuer warp adds it into the accumulators of the five output rows it touches (M128 x N32 x K16,
//                    A and D in TMEM, the stacked filter rows in shared memory), the workers pool the accumulators out
//                    of TMEM (tcgen05.ld) straight into p2 in the dense head's operand layout;
//       dense head (tcgen05.mma, sib_tc_fc_umma.cuh): transposed -- the weights are the M = 128 operand, a pair's 16 images
//                    N; epilogues on the workers and the head warp; softmax / argmax / top1-top2 margin and the margin
//                    guard's compaction list on the head warp.
//   * the dense-layer operands (119 KB) arrive by cp.async.bulk on an mbarrier; only the dense head waits for them;
//   * whole 16-image pairs are dealt to the CTAs round-robin, what is left after the last full round in equal runs of
//     images to all of them (pair_images): a small launch uses one SM per image, a large one ends on all SMs at once;
//   * conv1 stays on mma.sync: its N = 6, K = 75 contraction has no cheap operand path to tcgen05 -- from shared memory an
//     M128 x K16 operand costs 39-48 clk per mma, through registers (tcgen05.st: SIB_CONV1_UMMA, 208 M img/s) the im2col
//     rows cost more instructions than the mma save, and tcgen05.cp builds them exactly but at 64 clk per copy, unoverlapped
//     (tools/tcgen05_probe2/3.cu). -DSIB_CONV2_UMMA=0 builds the round-1 kernel (everything on mma.sync, 229 M img/s).
//
// Operand type is a template parameter: __half (default: 11-bit significand, max logit error 2e-2 on the presets)
// or __nv_bfloat16 (8-bit significand, ~1e-1). Accumulation is always fp32.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>
#include <type_traits>
#include <vector>

#include "sib_internal.hpp"

namespace sib {
namespace {

// Warp layout. The register file is split per SM sub-partition (warp % 4): 3 warps there may hold 168 registers
// each, 4 warps only 128. Two layouts were measured (tools/sweep_tc.sh, profiles/ncu_tc_r01_notes.md):
//   8 conv1 + 3 conv2 + 1 fc, 168 regs, two 4-row prefetch batches, split accumulator chains: 205 M img/s
//  12 conv1 + 3 conv2 + 1 fc, 128 regs, two 2-row prefetch batches, single chains (fits with 8 B spill): 216 M img/s
//  (230 M after the later trims: no L2 prefetch, no dead adds in the conv1 epilogue, bulk-copied dense weights)
// The second is the default: a 4th warp per sub-partition hides more of the mma / shared-memory latency than the
// extra registers bought. Each sub-partition hosts 3 conv1 warps + (a conv2 warp | the fc warp).
#ifndef SIB_CONV1_UMMA
#define SIB_CONV1_UMMA 0   // 1: conv1 on tcgen05.mma with operands in TMEM (sib_tc_conv1_umma.cuh)
#endif
#ifndef SIB_CONV2_UMMA
#define SIB_CONV2_UMMA (SIB_CONV1_UMMA ? 0 : 1)   // 1: conv2 on tcgen05.mma, A rows and accumulators in TMEM (sib_tc_conv2_umma.cuh)
#endif
#if SIB_CONV1_UMMA && SIB_CONV2_UMMA
#error "the tcgen05 variants of conv1 and conv2 each use the whole TMEM: pick one"
#endif
#ifndef SIB_CONV1_WARPS
#define SIB_CONV1_WARPS (SIB_CONV1_UMMA ? 8 : 12)
#endif
#ifndef SIB_BATCH_ROWS
#define SIB_BATCH_ROWS 2
#endif
#ifndef SIB_L2_PREFETCH
#define SIB_L2_PREFETCH 0   // conv1 (mma.sync path): prefetch.global.L2 of the row 2*PD rows ahead (measured: 218 M img/s with, 223 M without)
#endif
#ifndef SIB_L2_BULK_PREFETCH
#define SIB_L2_BULK_PREFETCH 0   // conv1: one cp.async.bulk.prefetch.L2 per span for the images the warp's next span reads
                                 // (measured: 65 536 images 264.2 us with, 262.1 without; mnist 189.9 / 180.2: off)
#endif
#ifndef SIB_CONV1_NANMAX
#define SIB_CONV1_NANMAX 1      // conv1's pool and ReLU hand a NaN on (FMNMX.NAN / FMNMX3.NAN, the same instruction count as fmaxf:
                                // 260.2 us per 65 536 images against 262.0) like the later stages' do: a NaN pixel, or Inf - Inf of
                                // two pixels beyond the half range, reaches the logits and the guard
#endif
#ifndef SIB_SPLIT_ACC
#define SIB_SPLIT_ACC 0
#endif
constexpr int kConv1Warps = SIB_CONV1_WARPS;
constexpr int kBR = SIB_BATCH_ROWS;  // rows per register prefetch batch (two batches in flight)
#ifndef SIB_CONV2_WARPS
#define SIB_CONV2_WARPS (SIB_CONV1_UMMA ? 5 : SIB_CONV2_UMMA ? 6 : 3)   // tcgen05 conv2: two sets of three workers (TMEM lane quadrants 0..2)
#endif
constexpr int kConv2Warps = SIB_CONV2_WARPS;
#ifndef SIB_CONV1_PACK2
#define SIB_CONV1_PACK2 1
#endif
#ifndef SIB_CONV2_SPLIT
#define SIB_CONV2_SPLIT 0
#endif
#ifndef SIB_TMA_WEIGHTS
#define SIB_TMA_WEIGHTS 1     // dense-layer fragments enter shared memory by cp.async.bulk (0: copied by all threads before the start)
#endif
#ifndef SIB_SKIP_EMPTY
#define SIB_SKIP_EMPTY 1      // conv1 spans / conv2 tiles that hold no real image are not computed (small launches)
#endif
#ifndef SIB_REG_STAGE
#define SIB_REG_STAGE 1       // 0 forces the cp.async staging ring (kept for heights that are not a multiple of 2 batches)
#endif
#ifndef SIB_DYN_SPANS
#define SIB_DYN_SPANS SIB_CONV2_UMMA  // 1: conv1 warps claim their spans from a shared counter. mma.sync build: 219 vs 228 M img/s dealt
                                     // round-robin; tcgen05 build: 242 vs 238 (the sub-partitions carry unequal side loads: two conv2
                                     // workers on three of them, the light issuer and head warps on the fourth)
#endif
#ifndef SIB_BIAS_IN_ACC1
#define SIB_BIAS_IN_ACC1 0     // conv1 / conv2 accumulators start at the bias instead of adding it after the pool
#endif
#ifndef SIB_BIAS_IN_ACC2
#define SIB_BIAS_IN_ACC2 0
#endif
#ifndef SIB_TAIL_FIRST
#define SIB_TAIL_FIRST 0     // 1: conv2 and fc warps get the lowest warp ids (they are the oldest warps of their sub-partition)
#endif
#ifndef SIB_DIAG_SKIP
#define SIB_DIAG_SKIP 0      // diagnostic builds only (results are wrong): 1 = conv1 warps only keep the slot protocol going, 2 = the same for the dense head, 3 = both
#endif
#ifndef SIB_FC_IN_CONV2
#define SIB_FC_IN_CONV2 0   // 1: no fc warp; conv2 warp (pair % kConv2Warps) runs the dense head of a pair after its tiles
#endif
constexpr int kFcWarps = SIB_FC_IN_CONV2 ? 0 : 1;
constexpr int kIssueWarps = SIB_CONV1_UMMA ? 2 : SIB_CONV2_UMMA ? 1 : 0;  // tcgen05.mma issuers
constexpr int kWarps = kConv1Warps + kConv2Warps + kFcWarps + kIssueWarps;
constexpr int kThreadsTc = kWarps * 32;
constexpr int kG = 8;     // images per ring slot (conv stages)
constexpr int kPair = 16; // images per fc M tile = 2 slots

template <int H_, int W_, int C_>
struct Net {
    static constexpr int H = H_, W = W_, C = C_;
    static constexpr int OH1 = H - 4, OW1 = W - 4, PH1 = OH1 / 2, PW1 = OW1 / 2;
    static constexpr int OH2 = PH1 - 4, OW2 = PW1 - 4, PH2 = OH2 / 2, PW2 = OW2 / 2;
    static constexpr int C1 = 6, C2 = 16, F1 = 120, F2 = 84, F3 = 10;
    static constexpr int IMG = H * W * C;
    static constexpr int ROWF = W * C;              // floats per image row
    static constexpr int K1ROW = 5 * C;             // real K per kernel row of conv1 (<= 15)
    // One input channel: a kernel row needs 5 (+1 for the odd-column shift) of the 16 K slots, so two kernel rows share one
    // mma (k 0..7 = row 2s, k 8..15 = row 2s+1): 3 mma per tile and output row instead of 5, and half the fragment loads.
    static constexpr bool PACK2 = (C == 1) && SIB_CONV1_PACK2;
    static constexpr int SPANS = kG * OW1 / 32;     // conv1 work items per slot: 32 output columns x all rows
    static constexpr int TILES2 = kG * OW2 / 16;    // conv2 M tiles per slot
    static constexpr int FEAT = PH2 * PW2 * C2;     // 400 | 256
    static constexpr int KS1 = FEAT / 16;
    static constexpr int P1_ROW = PW1 * C1;                         // halves
    // image stride of the pooled conv1 map, padded so that a conv2 tile that straddles two images keeps hitting
    // distinct banks: stride/2 = 3*OW2 (mod 32) continues the 3-words-per-column progression across the seam
    static constexpr int P1_IMG = PH1 * P1_ROW + (((3 * OW2 - PH1 * P1_ROW / 2) % 32 + 32) % 32) * 2;
    static constexpr int P1_BUF = kG * P1_IMG + 16;                 // + slack for the K-padding over-read
#ifndef SIB_NB1
#define SIB_NB1 3
#endif
#ifndef SIB_NP2
#define SIB_NP2 2
#endif
    // p1 ring slots / p2 buffers. -DSIB_NB1=4 -DSIB_NP2=1 fits (one p2 buffer is enough for the tcgen05 head, h2 overlays
    // h1) and was measured: 238.0 M img/s against 238.0 with 3 / 2 -- the conv1 warps' wait for a p1 slot is imbalance
    // among themselves (see SIB_DYN_SPANS), not ring depth.
    static constexpr int NB1 = SIB_NB1;
    static constexpr int NP2 = SIB_NP2;                             // p2 buffers (pairs)
    static constexpr int P2_STRIDE = FEAT + 8;                      // halves; +8 keeps A-fragment loads conflict-free
    // tcgen05 dense head: p2 / h1 / h2 are B operands (K-major, no swizzle: [img/8][k/8][img%8][k%8]), the weights A operands
    // ([m/8][k/8][m%8][k%8]); *_SBO = bytes between 8-row groups, 128 bytes between K chunks
    static constexpr int F1_SBO = FEAT / 8 * 128, F2_SBO = 128 / 8 * 128, F3_SBO = 96 / 8 * 128;
    // The B operands (p2, h1, h2) are WRITTEN by warps whose lanes differ in the K chunk: with the chunks 128 bytes
    // apart every lane of a store hit the same four banks (p2's 16-byte stores took 22 wavefronts instead of 4, the
    // head's 2-byte stores 4 instead of 1: 12% of the kernel's shared-memory wavefronts, profiles/ncu_tc_r02_final.txt).
    // 144 bytes between K chunks moves consecutive chunks to consecutive 16-byte bank groups; the descriptor's leading
    // byte offset takes any multiple of 16.
#ifndef SIB_BK
#define SIB_BK 144
#endif
    static constexpr int BK = SIB_BK;                                     // bytes between the K chunks of a B operand
    static constexpr int P2_SBO = FEAT / 8 * BK, H1_SBO = 128 / 8 * BK, H2_SBO = 96 / 8 * BK;  // bytes between its 8-image groups
    static constexpr int F1_GROUPS = 15, F2_GROUPS = 11, F3_GROUPS = 2;   // row groups that are stored (120 | 84 -> 88 | 10 -> 16 rows)
    static constexpr int P2_BUF = SIB_CONV2_UMMA ? 2 * P2_SBO / 2 : kPair * P2_STRIDE;  // halves
    static constexpr int DBG_P2 = kPair * P2_STRIDE;               // the stage dump keeps the padded [img][feature] layout
    static constexpr int H1_STRIDE = 136, H2_STRIDE = 104;
    // Input staging. H % 8 == 0: two register batches of 4 rows (2 scoreboard slots, no shared-memory cost).
    // Otherwise: a cp.async ring of PD rows per warp (in-order groups, costs ~10 shared-memory wavefronts a row).
    static constexpr bool REG_STAGE = (H % (2 * SIB_BATCH_ROWS) == 0) && SIB_REG_STAGE;
    static constexpr int PD = REG_STAGE ? 2 * SIB_BATCH_ROWS : 7;   // rows in flight per conv1 warp
    static constexpr int NCHUNK = (40 * C + 3) / 4;                 // 16-byte chunks per staged span row
    static constexpr int RING_ROW = ((40 * C + 16 + 7) / 8) * 8;    // halves per staged row incl. zero slack
    static_assert(kG * OW1 % 32 == 0 && kG * OW2 % 16 == 0 && FEAT % 16 == 0 && (SIB_CONV1_UMMA || H % PD == 0) && ((!REG_STAGE && !SIB_CONV1_UMMA) || H % (2 * kBR) == 0), "tiling");
    static_assert(K1ROW <= 15 && NCHUNK <= 32 && OW1 % 2 == 0 && OW2 % 2 == 0 && W % 4 == 0 && ROWF % 4 == 0, "layout");
    // packed weight image (32-bit words)
    static constexpr int W_C1 = 0;                               // [2 variants][5 kh][32 lanes][2]
    static constexpr int W_C2 = W_C1 + 2 * 5 * 64;               // [5 kh][2 ks][2 nt][32][2]
    static constexpr int NT1 = 15;                               // fc1 N tiles: 120 outputs, no padding tile
    static constexpr int W_F1 = W_C2 + 5 * 2 * 2 * 64;           // [KS1][NT1][32][2]
    static constexpr int W_F2 = W_F1 + (SIB_CONV2_UMMA ? F1_GROUPS * F1_SBO / 4 : KS1 * NT1 * 64);  // [8 ks][12 nt][32][2]
    static constexpr int W_F3 = W_F2 + (SIB_CONV2_UMMA ? F2_GROUPS * F2_SBO / 4 : 8 * 12 * 64);     // [6 ks][2 nt][32][2]
    static constexpr int W_BIAS = W_F3 + (SIB_CONV2_UMMA ? F3_GROUPS * F3_SBO / 4 : 6 * 2 * 64);    // fp32: c1[8] c2[16] f1[128] f2[96] f3[16]
    static constexpr int W_C1U = W_BIAS + 8 + 16 + 128 + 96 + 16;  // conv1 as tcgen05 B operands: [even|odd][5 kh][16 n][16 k] halves
    // conv2 as tcgen05 B operands: per K step the kernel rows 4,3,2,1,0 and a block of zeros stacked along N (96 x 16 halves)
    static constexpr int B2U_KS_BYTES = 6 * 512;
    static constexpr int W_C2U = W_C1U + 2 * 5 * 128;
    static constexpr int W_TOTAL = W_C2U + 2 * B2U_KS_BYTES / 4;
    // shared memory map (bytes)
    static constexpr int S_FCW = 0;                                        // fc1|fc2|fc3 fragments
    static constexpr int S_FCW_BYTES = (W_BIAS - W_F1) * 4;
    static constexpr int S_BIAS = S_FCW + S_FCW_BYTES;                     // 264 floats
    static constexpr int S_B1U = ((S_BIAS + 264 * 4 + 127) / 128) * 128;   // conv1 B operands for tcgen05 (5 KB)
    static constexpr int S_B2U = S_B1U;                                    // conv2 B operands for tcgen05 (5 KB); one variant at a time
    static constexpr int S_P1 = SIB_CONV2_UMMA ? S_B2U + 2 * B2U_KS_BYTES : SIB_CONV1_UMMA ? S_B1U + 2 * 5 * 512 : S_BIAS + 264 * 4;
    static constexpr int S_P2 = S_P1 + NB1 * P1_BUF * 2;
    static constexpr int S_H1 = S_P2 + NP2 * P2_BUF * 2;
    static constexpr int NHB = SIB_FC_IN_CONV2 ? 2 : 1;                    // h1/h2 buffers: dense heads of two pairs may overlap
    static constexpr int S_H2 = SIB_CONV2_UMMA ? S_H1 : S_H1 + NHB * kPair * H1_STRIDE * 2;          // tcgen05 head: h2 overlays h1
    static constexpr int S_FCS = SIB_CONV2_UMMA ? S_H1 + 2 * H1_SBO : S_H2 + NHB * kPair * H2_STRIDE * 2;  // tcgen05 head: logits [16][10] fp32, lane <-> image transpose
    static constexpr int S_RING = S_FCS + (SIB_CONV2_UMMA ? kPair * 10 * 4 : 0);
    static constexpr int STAGE_ROW = NCHUNK * 16;                          // bytes of one fp32 span row
    static constexpr int URING_ROW = 2 * W * C + 16;                       // halves: Umma<N>::UROW
    static constexpr int S_STAGE = S_RING + kConv1Warps * 2 * (SIB_CONV1_UMMA ? URING_ROW : RING_ROW) * 2; // cp.async landing zone
    static constexpr int S_BAR = ((S_STAGE + ((REG_STAGE || SIB_CONV1_UMMA) ? 0 : kConv1Warps * PD * STAGE_ROW) + 7) / 8) * 8;
    static constexpr int B_P1F = 0, B_P1E = NB1, B_P2F = 2 * NB1, B_P2E = 2 * NB1 + 2;
    static constexpr int B_UA = 2 * NB1 + 4, B_UDF = B_UA + 16, B_UDE = B_UDF + 8;  // tcgen05 conv1: a_full[8] a_free[8], d_full[4] (+4 unused), d_empty[4] (+4 unused)
    static constexpr int B_U2 = 2 * NB1 + 4;                               // tcgen05 conv2: a_full[16] a_free[16] d_full[8] d_empty[8]
    static constexpr int B_FC = B_U2 + 48;                                 // tcgen05 dense head: f_full[3] h1_full h2_full
    static constexpr int B_FCW = SIB_CONV1_UMMA ? B_UDE + 8 : SIB_CONV2_UMMA ? B_FC + 5 : 2 * NB1 + 4;  // fc weight image has landed (bulk copy)
    static constexpr int N_BAR = B_FCW + 1;
    static constexpr int S_TMEM = S_BAR + N_BAR * 8;                       // TMEM base address handed out by tcgen05.alloc
    static constexpr int S_TOTAL = S_TMEM + 8;
    static_assert(S_P1 % 16 == 0 && S_P2 % 16 == 0 && S_H1 % 16 == 0 && S_RING % 16 == 0 && S_STAGE % 16 == 0 && S_BAR % 8 == 0, "align");
    static_assert(S_TOTAL <= 227 * 1024, "shared memory");
};

// ---- small PTX helpers --------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Waits for the phase with the given parity to complete. A bounded spin: a protocol bug traps instead of hanging
// the GPU (the trap surfaces as a launch failure on the host).
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0, spin = 0;
    long long t0 = 0;
    for (;;) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (ok) return;
        if ((++spin & 1023u) == 0) {  // watchdog: ~2 s at 2 GHz
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 4000000000ll) __trap();
        }
    }
}

template <typename T> struct Ops;
template <> struct Ops<__half> {
    __device__ static __forceinline__ uint32_t pack(float lo, float hi) {
        __half2 h = __floats2half2_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&h);
    }
    __device__ static __forceinline__ void mma(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    static uint16_t host_bits(float x) { return __half_as_ushort(__float2half_rn(x)); }
    static float host_float(uint16_t b) { return __half2float(__ushort_as_half(b)); }
};
template <> struct Ops<__nv_bfloat16> {
    __device__ static __forceinline__ uint32_t pack(float lo, float hi) {
        __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&h);
    }
    __device__ static __forceinline__ void mma(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    static uint16_t host_bits(float x) { return __bfloat16_as_ushort(__float2bfloat16_rn(x)); }
    static float host_float(uint16_t b) { return __bfloat162float(__ushort_as_bfloat16(b)); }
};

__device__ __forceinline__ uint32_t lds32(const uint16_t* p) { return *reinterpret_cast<const uint32_t*>(p); }
// cp.async (LDGSTS): global -> shared without a register stop-over. Its commit/wait groups complete in order, so a
// deep prefetch queue costs no scoreboard slots (8 individually tracked LDGs do not fit the 6 slots a warp has:
// profiles/ncu_tc_r01_notes.md shows the register-prefetch version paying one full memory latency per row).
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory"); }
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ float4 ldg_stream(const float* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

struct TcParams {
    const uint32_t* wpack;
    const float* in;
    int n;
    float* probs;
    float* logits;   // may be null
    int32_t* argmax; // may be null
    float* margin;   // may be null
    int32_t* list;   // guard compaction list (may be null)
    int32_t* count;
    float tau;
    float* dbg;      // may be null: CTA 0 dumps p1/p2/h1/h2 of its first 16 images (tests)
    int rounds;      // SIB_SPREAD_TAIL: rounds in which every CTA runs a whole pair (set by launch_inst from n and the grid)
    int per;         //                  images per CTA in the round after them (0: the full rounds took everything)
};

// Which images are the Qi-th pair of this CTA: [base, lim), at most kPair of them; false = the CTA has no Qi-th pair.
// Whole pairs are dealt round-robin (pair = blockIdx.x + Qi * gridDim.x). SIB_SPREAD_TAIL cuts what is left after the
// last round that fills every CTA into equal runs of images, one run per CTA: the last round then costs every CTA the
// same fraction of a pair instead of a whole pair on some CTAs and nothing on the rest (65 536 images: 100 of 148 CTAs
// ran a 28th pair), and a launch of fewer images than 16 x the SM count is spread over all the CTAs it can use (16 images:
// 16 CTAs with one image each, not one CTA walking 16 images through its pipeline).
#ifndef SIB_SPREAD_TAIL
#define SIB_SPREAD_TAIL SIB_CONV2_UMMA
#endif
__device__ __forceinline__ bool pair_images(const TcParams& P, int Qi, int& base, int& lim) {
#if SIB_SPREAD_TAIL
    if (Qi < P.rounds) {
        base = (blockIdx.x + Qi * gridDim.x) * kPair;
        lim = base + kPair;
        return true;
    }
    base = P.rounds * (int)gridDim.x * kPair + blockIdx.x * P.per;
    lim = min(base + P.per, P.n);
    return Qi == P.rounds && base < lim;
#else
    base = (blockIdx.x + Qi * gridDim.x) * kPair;
    lim = min(base + kPair, P.n);
    return base < P.n;
#endif
}

// max that hands a NaN on (fmaxf drops it): an overflowed half-precision activation turns into Inf - Inf = NaN one layer
// later, and that NaN has to reach the logits so that the margin guard sends the image to the fp32 path
__device__ __forceinline__ float fmax_nan(float a, float b) {
    float d;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}

__device__ __forceinline__ float fmax_nan3(float a, float b, float c) {
    float d;
    asm("max.NaN.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

#if SIB_CONV1_UMMA || SIB_CONV2_UMMA
#include "sib_tc_umma.cuh"
#endif
#if SIB_CONV1_UMMA
#include "sib_tc_conv1_umma.cuh"
#endif
#if SIB_CONV2_UMMA
#include "sib_tc_conv2_umma.cuh"
#include "sib_tc_fc_umma.cuh"
#endif

// ---- conv1 role (mma.sync form) ------------------------------------------------------------------------------

template <typename N, typename T, bool DBG>
__device__ __forceinline__ void conv1_role(const TcParams& P, unsigned char* smem, int cw, int lane) {
    const int g = lane >> 2, t = lane & 3;
    uint16_t* ring = reinterpret_cast<uint16_t*>(smem + N::S_RING) + cw * 2 * N::RING_ROW;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + N::S_BAR);
    uint64_t* p1_full = bars + N::B_P1F;   // [NB1]
    uint64_t* p1_empty = bars + N::B_P1E;  // [NB1]
    // B fragments of conv1: even-column tiles and odd-column tiles (K shifted by one element so that odd columns
    // read 4-byte aligned pairs from the staged row)
    uint32_t bE[5][2], bO[5][2];
#pragma unroll
    for (int kh = 0; kh < 5; kh++) {
        const uint2 e = *reinterpret_cast<const uint2*>(P.wpack + N::W_C1 + (0 * 5 + kh) * 64 + lane * 2);
        const uint2 o = *reinterpret_cast<const uint2*>(P.wpack + N::W_C1 + (1 * 5 + kh) * 64 + lane * 2);
        bE[kh][0] = e.x; bE[kh][1] = e.y; bO[kh][0] = o.x; bO[kh][1] = o.y;
    }
    const float* biasp = reinterpret_cast<const float*>(P.wpack + N::W_BIAS);
    const float bias0 = biasp[2 * t], bias1 = biasp[2 * t + 1];

    // spans in CTA order: local slot lq = 0,1,2.. ; slot lq -> the CTA's pair Qi = lq/2 (pair_images); lim = one past the
    // pair's last image
    auto span_src = [&](int gs, int& img0, int& lim, bool& valid) {
        const int lq = gs / N::SPANS;
        int base;
        valid = pair_images(P, lq >> 1, base, lim);
        img0 = base + (lq & 1) * kG;
    };
    // Every lane copies one 16-byte chunk of the span row; lanes past the last chunk repeat it (same bytes to the
    // same address), which keeps the whole staging path free of divergent branches.
    const int chunk = lane < N::NCHUNK ? lane : N::NCHUNK - 1;
    auto chunk_ptr = [&](int img0, int s, int lim) -> const float* {
        const int p0 = 32 * s;
        const int u_start = (p0 / N::OW1) * N::W + p0 % N::OW1;
        const int F0 = u_start * N::C + 4 * chunk;
        int img = img0 + F0 / N::ROWF;
        img = min(img, lim - 1);  // tail slots re-read the pair's last image; their outputs are never written
        return P.in + (size_t)img * N::IMG + F0 % N::ROWF;
    };

    const uint32_t stage = smem_u32(smem + N::S_STAGE) + cw * N::PD * N::STAGE_ROW + chunk * 16;
    // SIB_DYN_SPANS: spans are claimed from a CTA-wide counter instead of being dealt round-robin, so the conv1 warps of a
    // sub-partition whose tensor pipe is less loaded (no conv2 warp there) end up taking more of them
    uint32_t* span_ctr = reinterpret_cast<uint32_t*>(smem + N::S_TMEM + 4);  // (the first word holds the TMEM base address)
    auto claim = [&]() -> int {
        int v = 0;
        if (lane == 0) v = (int)atomicAdd(span_ctr, 1u);
        return __shfl_sync(0xffffffffu, v, 0);
    };
    int gs = SIB_DYN_SPANS ? claim() : cw;
    int img0, lim;
    bool valid;
    span_src(gs, img0, lim, valid);
    if (!valid) return;
    const float* src = chunk_ptr(img0, gs % N::SPANS, lim);
    float4 preA[kBR], preB[kBR];  // REG_STAGE: two batches of kBR consecutive rows, one 16-byte chunk per lane
    if constexpr (N::REG_STAGE) {
#pragma unroll
        for (int r = 0; r < kBR; r++) preA[r] = ldg_stream(src + r * N::ROWF);
#pragma unroll
        for (int r = 0; r < kBR; r++) preB[r] = ldg_stream(src + (kBR + r) * N::ROWF);
    } else {
#pragma unroll
        for (int r = 0; r < N::PD; r++) {
            cp_async16(stage + r * N::STAGE_ROW, src + r * N::ROWF);
            cp_async_commit();
        }
    }

    while (valid) {
        const int s = gs % N::SPANS, lq = gs / N::SPANS;
        const int buf = lq % N::NB1, use = lq / N::NB1;
        uint16_t* p1 = reinterpret_cast<uint16_t*>(smem + N::S_P1) + buf * N::P1_BUF;
        // Next span (for the tail prefetch). The claim's atomic is issued here and its answer is taken a few rows into
        // the span (resolve_next): taken at once, its latency sat on every span's critical path -- 5.6% of the conv1
        // warps' stall samples (profiles/ncu_tc_r02_final.txt, the atomicAdd line).
        int n_img0, n_lim, n_gs = 0, n_raw = 0;
        bool n_valid = false;
        const float* n_src = src;
        if (SIB_DYN_SPANS && lane == 0) n_raw = (int)atomicAdd(span_ctr, 1u);
        auto resolve_next = [&]() {
            n_gs = SIB_DYN_SPANS ? __shfl_sync(0xffffffffu, n_raw, 0) : gs + kConv1Warps;
            span_src(n_gs, n_img0, n_lim, n_valid);
            n_src = n_valid ? chunk_ptr(n_img0, n_gs % N::SPANS, n_lim) : src;
#if SIB_L2_BULK_PREFETCH
            // One span (~7 us) ahead of its row loads, the one or two images the next span reads are pulled into L2 by a
            // single bulk prefetch: the row loads then wait an L2 hit, not DRAM (the fp32 -> half conversion of a row
            // was 4.5% of the conv1 warps' stall samples, waiting for its load).
            if (n_valid && lane == 0) {
                const int s2 = n_gs % N::SPANS;
                const int first = min(n_img0 + (32 * s2) / N::OW1, n_lim - 1), last = min(n_img0 + (32 * s2 + 31) / N::OW1, n_lim - 1);
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(P.in + (size_t)first * N::IMG),
                             "r"((uint32_t)((last - first + 1) * N::IMG * 4)) : "memory");
            }
#endif
        };
#ifndef SIB_RESOLVE_AT
#define SIB_RESOLVE_AT 1   // measured: 1 -> 260.0 us per 65 536 images, 3 -> 262.0, 8 -> 269.2
#endif
        constexpr int kResolveAt = N::REG_STAGE ? SIB_RESOLVE_AT : -1;  // the cp.async ring asks for n_valid from its first row on
        if (kResolveAt < 0) resolve_next();
        // A span whose first column already lies past the last image has nothing to compute (the tail of a batch:
        // with 1 image in a launch, 13 of the pair's 14 spans). It only keeps the slot protocol going. Image indices
        // grow with gs, so every later span of this warp is empty as well and the prefetch state no longer matters.
        if ((SIB_DIAG_SKIP & 1) || (SIB_SKIP_EMPTY && img0 + (32 * s) / N::OW1 >= lim)) {
            if (kResolveAt >= 0) resolve_next();
            mbar_wait(&p1_empty[buf], (use & 1) ^ 1);
            mbar_arrive(&p1_full[buf]);
            gs = n_gs; valid = n_valid; src = n_src; img0 = n_img0; lim = n_lim;
            continue;
        }

        // fragment offsets into a staged row (halves)
        const int p0 = 32 * s;
        const int u_start = (p0 / N::OW1) * N::W + p0 % N::OW1;
        auto uoff = [&](int p) { return ((p / N::OW1) * N::W + p % N::OW1 - u_start) * N::C; };
        const int offE0 = uoff(p0 + 2 * g) + 2 * t, offE1 = uoff(p0 + 2 * g + 16) + 2 * t;
        // odd columns sit C elements further in the same image row and read one element earlier: a constant C-1
        constexpr int kOdd = N::C - 1;
        // pooled output columns of this lane
        const int pp0 = 16 * s + g, pp1 = pp0 + 8;
        const int o0 = (pp0 / N::PW1) * N::P1_IMG + (pp0 % N::PW1) * N::C1 + 2 * t;
        const int o1 = (pp1 / N::PW1) * N::P1_IMG + (pp1 % N::PW1) * N::C1 + 2 * t;
        const bool dump = DBG && P.dbg && blockIdx.x == 0 && lq < 2;

        uint32_t aE[5][4], aO[5][4];
        float prev[4];
        // stage(r): row r's cp.async group has landed -> fp32 to half into ring[r&1]; refill the landing slot
        auto stage_row = [&](int r) {
            float4 v;
            if constexpr (N::REG_STAGE) {
                v = ((r / kBR) & 1) ? preB[r % kBR] : preA[r % kBR];
            } else {
                cp_async_wait<N::PD - 1>();
                v = lds128(stage + (r % N::PD) * N::STAGE_ROW);  // bytes this lane copied itself
            }
            uint2 pk;
            pk.x = Ops<T>::pack(v.x, v.y);
            pk.y = Ops<T>::pack(v.z, v.w);
            *reinterpret_cast<uint2*>(ring + (r & 1) * N::RING_ROW + 4 * chunk) = pk;
            // refill: same span while rows remain, else the first rows of this warp's next span (n_src is a valid
            // address even when there is no next span, so the loads need no branch)
            auto row_ptr = [&](int row) { return row < N::H ? src + row * N::ROWF : n_src + (row - N::H) * N::ROWF; };
            if constexpr (N::REG_STAGE) {
                if (r % kBR == kBR - 1) {  // batch drained: fetch the batch two ahead in one burst
#pragma unroll
                    for (int j = 0; j < kBR; j++) {
                        const float4 nv = ldg_stream(row_ptr(r + kBR + 1 + j));
                        if ((r / kBR) & 1) preB[j] = nv; else preA[j] = nv;
                    }
                }
            } else {
                if (r + N::PD < N::H || n_valid) cp_async16(stage + (r % N::PD) * N::STAGE_ROW, row_ptr(r + N::PD));
                cp_async_commit();  // one group per row, empty or not, keeps the wait depth constant
            }
            // and pull the row two windows ahead into L2 (one 128-byte line per 8 chunks)
            if (SIB_L2_PREFETCH && (lane & 7) == 0 && (r + 2 * N::PD < N::H || n_valid)) prefetch_l2(row_ptr(r + 2 * N::PD));
        };
        // frags(r): im2col A fragments of staged row r for the even-column and the odd-column tile
        auto load_frags = [&](int r) {
            __syncwarp();
            const uint16_t* rr = ring + (r & 1) * N::RING_ROW;
            uint32_t(&e)[4] = aE[r % 5];
            uint32_t(&o)[4] = aO[r % 5];
            e[0] = lds32(rr + offE0); e[1] = lds32(rr + offE1);
            if constexpr (N::PACK2) {  // k 8..15 of a fragment come from the next row's k 0..7; odd tile = even tile (kOdd = 0)
                e[2] = e[3] = 0u; o[0] = e[0]; o[1] = e[1]; o[2] = o[3] = 0u;
            } else {
                e[2] = lds32(rr + offE0 + 8); e[3] = lds32(rr + offE1 + 8);
                o[0] = lds32(rr + offE0 + kOdd); o[1] = lds32(rr + offE1 + kOdd);
                o[2] = lds32(rr + offE0 + kOdd + 8); o[3] = lds32(rr + offE1 + kOdd + 8);
            }
        };

        stage_row(0);
        mbar_wait(&p1_empty[buf], (use & 1) ^ 1);
        load_frags(0);
        // Software pipeline: while the mma of output row h-4 run, row h+1 is staged and its fragments are fetched,
        // so neither the shared-memory round trip nor the mma dependency chains sit on the critical path.
#pragma unroll
        for (int h = 0; h < N::H; h++) {
            if (h == kResolveAt) resolve_next();
            if (h + 1 < N::H) stage_row(h + 1);
            // the accumulators start at the bias (same for the 4 positions of a pool window, so max(.) + b == max(. + b))
            float cE[4], cO[4];
#pragma unroll
            for (int i = 0; i < 4; i++) cE[i] = cO[i] = SIB_BIAS_IN_ACC1 ? ((i & 1) ? bias1 : bias0) : 0.f;
            float dE[4] = {0.f, 0.f, 0.f, 0.f}, dO[4] = {0.f, 0.f, 0.f, 0.f};
            if (h >= 4) {
                const int oh = h - 4;
                if constexpr (N::PACK2) {
                    // fragment of kernel rows (2s, 2s+1) = k 0..7 of input row oh+2s | k 0..7 of input row oh+2s+1
#pragma unroll
                    for (int sp = 0; sp < 3; sp++) {
                        const uint32_t(&r0)[4] = aE[(oh + 2 * sp) % 5];
                        const uint32_t(&r1)[4] = aE[(oh + 2 * sp + (sp < 2 ? 1 : 0)) % 5];  // row 5 does not exist: zero weights
                        const uint32_t a2[4] = {r0[0], r0[1], r1[0], r1[1]};
                        Ops<T>::mma(cE, a2, bE[sp]);
                        Ops<T>::mma(cO, a2, bO[sp]);
                    }
                } else {
                // kernel row 0 first: it reads the window slot that row h+1 is about to overwrite
                Ops<T>::mma(cE, aE[(oh + 0) % 5], bE[0]); Ops<T>::mma(cO, aO[(oh + 0) % 5], bO[0]);
                Ops<T>::mma(SIB_SPLIT_ACC ? dE : cE, aE[(oh + 3) % 5], bE[3]); Ops<T>::mma(SIB_SPLIT_ACC ? dO : cO, aO[(oh + 3) % 5], bO[3]);
                Ops<T>::mma(cE, aE[(oh + 1) % 5], bE[1]); Ops<T>::mma(cO, aO[(oh + 1) % 5], bO[1]);
                Ops<T>::mma(SIB_SPLIT_ACC ? dE : cE, aE[(oh + 4) % 5], bE[4]); Ops<T>::mma(SIB_SPLIT_ACC ? dO : cO, aO[(oh + 4) % 5], bO[4]);
                Ops<T>::mma(cE, aE[(oh + 2) % 5], bE[2]); Ops<T>::mma(cO, aO[(oh + 2) % 5], bO[2]);
                }
            }
            if (h + 1 < N::H) load_frags(h + 1);
            if (h >= 4) {
                const int oh = h - 4;
                float m[4];
#pragma unroll
                for (int i = 0; i < 4; i++)  // pool over the column pair (x + 0.f is not a no-op to the compiler: no dead adds)
                    m[i] = SIB_SPLIT_ACC ? fmaxf(cE[i] + dE[i], cO[i] + dO[i]) : SIB_CONV1_NANMAX ? ((oh & 1) ? fmax_nan3(cE[i], cO[i], prev[i]) : fmax_nan(cE[i], cO[i])) : fmaxf(cE[i], cO[i]);
                if ((oh & 1) == 0) {
#pragma unroll
                    for (int i = 0; i < 4; i++) prev[i] = m[i];
                } else {
                    const int ph = oh >> 1;
                    // relu(max(.)+b) == max over the window of relu(.+b)
                    const float pb0 = SIB_BIAS_IN_ACC1 ? 0.f : bias0, pb1 = SIB_BIAS_IN_ACC1 ? 0.f : bias1;
                    const float v0 = SIB_CONV1_NANMAX ? fmax_nan(m[0] + pb0, 0.f) : SIB_BIAS_IN_ACC1 ? fmaxf(fmaxf(m[0], prev[0]), 0.f) : fmaxf(fmaxf(m[0], prev[0]) + pb0, 0.f);
                    const float v1 = SIB_CONV1_NANMAX ? fmax_nan(m[1] + pb1, 0.f) : SIB_BIAS_IN_ACC1 ? fmaxf(fmaxf(m[1], prev[1]), 0.f) : fmaxf(fmaxf(m[1], prev[1]) + pb1, 0.f);
                    const float v2 = SIB_CONV1_NANMAX ? fmax_nan(m[2] + pb0, 0.f) : SIB_BIAS_IN_ACC1 ? fmaxf(fmaxf(m[2], prev[2]), 0.f) : fmaxf(fmaxf(m[2], prev[2]) + pb0, 0.f);
                    const float v3 = SIB_CONV1_NANMAX ? fmax_nan(m[3] + pb1, 0.f) : SIB_BIAS_IN_ACC1 ? fmaxf(fmaxf(m[3], prev[3]), 0.f) : fmaxf(fmaxf(m[3], prev[3]) + pb1, 0.f);
                    if (t < 3) {
                        *reinterpret_cast<uint32_t*>(p1 + o0 + ph * N::P1_ROW) = Ops<T>::pack(v0, v1);
                        *reinterpret_cast<uint32_t*>(p1 + o1 + ph * N::P1_ROW) = Ops<T>::pack(v2, v3);
                        if (dump) {  // dense [img][ph][pw][c] for the test, not the padded ring layout
                            float* d = P.dbg + lq * kG * N::PH1 * N::P1_ROW + ph * N::P1_ROW + 2 * t;
                            const int d0 = (pp0 / N::PW1) * N::PH1 * N::P1_ROW + (pp0 % N::PW1) * N::C1;
                            const int d1 = (pp1 / N::PW1) * N::PH1 * N::P1_ROW + (pp1 % N::PW1) * N::C1;
                            d[d0] = v0; d[d0 + 1] = v1; d[d1] = v2; d[d1 + 1] = v3;
                        }
                    }
                }
            }
        }
        mbar_arrive(&p1_full[buf]);
        gs = n_gs;
        valid = n_valid;
        src = n_src;
        img0 = n_img0;
        lim = n_lim;
    }
}

// ---- conv2 role ---------------------------------------------------------------------------------------------

template <typename N, typename T, bool DBG>
__device__ __forceinline__ void fc_pair(const TcParams& P, unsigned char* smem, int lane, int Qi);  // below

template <typename N, typename T, bool DBG>
__device__ __forceinline__ void conv2_role(const TcParams& P, unsigned char* smem, int w2, int lane) {
    const int g = lane >> 2, t = lane & 3;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + N::S_BAR);
    uint64_t* p1_full = bars + N::B_P1F;
    uint64_t* p1_empty = bars + N::B_P1E;
    uint64_t* p2_full = bars + N::B_P2F;
    uint64_t* p2_empty = bars + N::B_P2E;
    uint32_t bW[5][2][2][2];  // [kh][k step][n tile][reg]: the whole 150x16 filter, register-resident
    auto load_filter = [&]() {
#pragma unroll
        for (int kh = 0; kh < 5; kh++)
#pragma unroll
            for (int ks = 0; ks < 2; ks++)
#pragma unroll
                for (int nt = 0; nt < 2; nt++) {
                    const uint2 v = *reinterpret_cast<const uint2*>(P.wpack + N::W_C2 + ((kh * 2 + ks) * 2 + nt) * 64 + lane * 2);
                    bW[kh][ks][nt][0] = v.x; bW[kh][ks][nt][1] = v.y;
                }
    };
    load_filter();
    const float* biasp = reinterpret_cast<const float*>(P.wpack + N::W_BIAS) + 8;
    const float bias00 = biasp[2 * t], bias01 = biasp[2 * t + 1], bias10 = biasp[8 + 2 * t], bias11 = biasp[8 + 2 * t + 1];
    const int nPairs = (P.n + kPair - 1) / kPair;

    for (int gi = w2;; gi += kConv2Warps) {  // M tiles in CTA order, dealt round-robin to the conv2 warps
        const int lq = gi / N::TILES2, tile = gi % N::TILES2;
        const int Qi = lq >> 1;
        if ((int)(blockIdx.x + Qi * gridDim.x) >= nPairs) break;
        const int buf1 = lq % N::NB1, use1 = lq / N::NB1;
        const int buf2 = Qi & 1, use2 = Qi >> 1;
        const uint16_t* p1 = reinterpret_cast<const uint16_t*>(smem + N::S_P1) + buf1 * N::P1_BUF;
        uint16_t* p2 = reinterpret_cast<uint16_t*>(smem + N::S_P2) + buf2 * N::P2_BUF;
        // tile row r <-> output column P = 16*tile + 2*(r%8) + r/8: fragment rows g and g+8 are a pooling pair
        const int Pe = 16 * tile + 2 * g;
        const int imgl = Pe / N::OW2, ow = Pe % N::OW2;
        const int offA0 = imgl * N::P1_IMG + ow * N::C1 + 2 * t;  // even column (rows g)
        const int offA1 = offA0 + N::C1;                           // odd column  (rows g+8)
        const int o2 = ((lq & 1) * kG + imgl) * N::P2_STRIDE + (ow >> 1) * N::C2 + 2 * t;
        const bool dump = DBG && P.dbg && blockIdx.x == 0 && Qi == 0;

        mbar_wait(&p1_full[buf1], use1 & 1);
        mbar_wait(&p2_empty[buf2], (use2 & 1) ^ 1);
        // a tile whose first column lies past the last image: nothing to compute, keep the slot protocol going
        if (SIB_SKIP_EMPTY && !SIB_FC_IN_CONV2 && (int)(blockIdx.x + Qi * gridDim.x) * kPair + (lq & 1) * kG + (16 * tile) / N::OW2 >= P.n) {  // warp-uniform
            mbar_arrive(&p1_empty[buf1]);
            mbar_arrive(&p2_full[buf2]);
            continue;
        }

        uint32_t a[5][2][4];
        float prev[4] = {0.f, 0.f, 0.f, 0.f};
        auto load_frags = [&](int r) {
            const uint16_t* row = p1 + r * N::P1_ROW;
#pragma unroll
            for (int ks = 0; ks < 2; ks++) {
                uint32_t(&f)[4] = a[r % 5][ks];
                f[0] = lds32(row + offA0 + ks * 16);
                f[1] = lds32(row + offA1 + ks * 16);
                f[2] = lds32(row + offA0 + ks * 16 + 8);
                f[3] = lds32(row + offA1 + ks * 16 + 8);
            }
        };
#pragma unroll
        for (int r = 0; r < 4; r++) load_frags(r);
#pragma unroll
        for (int r = 4; r < N::PH1; r++) {
            load_frags(r);  // consumed by the kh=4 mma at the end of this iteration's chains
            const int oh = r - 4;
            // 4 independent accumulator chains: (n tile) x (k step)
            float c0[4], c1[4], d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 4; i++) {  // SIB_BIAS_IN_ACC2: the bias rides in the accumulator
                c0[i] = SIB_BIAS_IN_ACC2 ? ((i & 1) ? bias01 : bias00) : 0.f;
                c1[i] = SIB_BIAS_IN_ACC2 ? ((i & 1) ? bias11 : bias10) : 0.f;
            }
#if SIB_CONV2_SPLIT
            // 8 chains: kernel rows 1 and 3 accumulate apart from rows 0, 2, 4 (dependent mma twice as far apart)
            float e0[4] = {0.f, 0.f, 0.f, 0.f}, f0[4] = {0.f, 0.f, 0.f, 0.f}, e1[4] = {0.f, 0.f, 0.f, 0.f}, f1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int kh = 0; kh < 5; kh++) {
                Ops<T>::mma((kh & 1) ? e0 : c0, a[(oh + kh) % 5][0], bW[kh][0][0]);
                Ops<T>::mma((kh & 1) ? e1 : c1, a[(oh + kh) % 5][0], bW[kh][0][1]);
                Ops<T>::mma((kh & 1) ? f0 : d0, a[(oh + kh) % 5][1], bW[kh][1][0]);
                Ops<T>::mma((kh & 1) ? f1 : d1, a[(oh + kh) % 5][1], bW[kh][1][1]);
            }
#pragma unroll
            for (int i = 0; i < 4; i++) { c0[i] += e0[i]; c1[i] += e1[i]; d0[i] += f0[i]; d1[i] += f1[i]; }
#else
#pragma unroll
            for (int kh = 0; kh < 5; kh++) {
                Ops<T>::mma(c0, a[(oh + kh) % 5][0], bW[kh][0][0]);
                Ops<T>::mma(c1, a[(oh + kh) % 5][0], bW[kh][0][1]);
                Ops<T>::mma(d0, a[(oh + kh) % 5][1], bW[kh][1][0]);
                Ops<T>::mma(d1, a[(oh + kh) % 5][1], bW[kh][1][1]);
            }
#endif
            // pool over the column pair (rows g / g+8 of the tile)
            const float h[4] = {fmax_nan(c0[0] + d0[0], c0[2] + d0[2]), fmax_nan(c0[1] + d0[1], c0[3] + d0[3]),
                                fmax_nan(c1[0] + d1[0], c1[2] + d1[2]), fmax_nan(c1[1] + d1[1], c1[3] + d1[3])};
            if ((oh & 1) == 0) {
#pragma unroll
                for (int i = 0; i < 4; i++) prev[i] = h[i];
            } else {
                const int ph = oh >> 1;
                const float v0 = SIB_BIAS_IN_ACC2 ? fmax_nan(fmax_nan(h[0], prev[0]), 0.f) : fmax_nan(fmax_nan(h[0], prev[0]) + bias00, 0.f);
                const float v1 = SIB_BIAS_IN_ACC2 ? fmax_nan(fmax_nan(h[1], prev[1]), 0.f) : fmax_nan(fmax_nan(h[1], prev[1]) + bias01, 0.f);
                const float v2 = SIB_BIAS_IN_ACC2 ? fmax_nan(fmax_nan(h[2], prev[2]), 0.f) : fmax_nan(fmax_nan(h[2], prev[2]) + bias10, 0.f);
                const float v3 = SIB_BIAS_IN_ACC2 ? fmax_nan(fmax_nan(h[3], prev[3]), 0.f) : fmax_nan(fmax_nan(h[3], prev[3]) + bias11, 0.f);
                uint16_t* dst = p2 + o2 + ph * N::PW2 * N::C2;
                *reinterpret_cast<uint32_t*>(dst) = Ops<T>::pack(v0, v1);      // channels 2t, 2t+1
                *reinterpret_cast<uint32_t*>(dst + 8) = Ops<T>::pack(v2, v3);  // channels 8+2t, 9+2t
                if (dump) {
                    float* dd = P.dbg + 2 * kG * N::PH1 * N::P1_ROW + o2 + ph * N::PW2 * N::C2;
                    dd[0] = v0; dd[1] = v1; dd[8] = v2; dd[9] = v3;
                }
            }
        }
        mbar_arrive(&p1_empty[buf1]);
        mbar_arrive(&p2_full[buf2]);
#if SIB_FC_IN_CONV2
        {   // was that this warp's last tile of pair Qi? then warp (Qi % kConv2Warps) runs the pair's dense head
            static_assert(2 * N::TILES2 >= 2 * kConv2Warps, "every conv2 warp has tiles in every pair");
            const int nQi = ((gi + kConv2Warps) / N::TILES2) >> 1;
            if ((nQi != Qi || (int)(blockIdx.x + nQi * gridDim.x) >= nPairs) && Qi % kConv2Warps == w2) {
                fc_pair<N, T, DBG>(P, smem, lane, Qi);
                load_filter();  // the filter registers were free for the dense head's accumulators
            }
        }
#endif
    }
}

// ---- fc role (one warp) -------------------------------------------------------------------------------------

// the dense head of pair Qi (16 images): p2[Qi & 1] -> probabilities, argmax, margin
template <typename N, typename T, bool DBG>
__device__ __forceinline__ void fc_pair(const TcParams& P, unsigned char* smem, int lane, int Qi) {
    const int g = lane >> 2, t = lane & 3;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + N::S_BAR);
    uint64_t* p2_full = bars + N::B_P2F;
    uint64_t* p2_empty = bars + N::B_P2E;
    const uint32_t* w1 = reinterpret_cast<const uint32_t*>(smem + N::S_FCW);
    const uint32_t* w2 = w1 + (N::W_F2 - N::W_F1);
    const uint32_t* w3 = w1 + (N::W_F3 - N::W_F1);
    const float* bias = reinterpret_cast<const float*>(smem + N::S_BIAS);  // c1[8] c2[16] f1[128] f2[96] f3[16]
    uint16_t* h1 = reinterpret_cast<uint16_t*>(smem + N::S_H1) + (Qi % N::NHB) * kPair * N::H1_STRIDE;
    uint16_t* h2 = reinterpret_cast<uint16_t*>(smem + N::S_H2) + (Qi % N::NHB) * kPair * N::H2_STRIDE;
    {
        const int pair = blockIdx.x + Qi * gridDim.x;
        const int buf = Qi & 1, use = Qi >> 1;
        const uint16_t* p2 = reinterpret_cast<const uint16_t*>(smem + N::S_P2) + buf * N::P2_BUF;
        const bool dump = DBG && P.dbg && blockIdx.x == 0 && Qi == 0;
        mbar_wait(&p2_full[buf], use & 1);
        mbar_wait(bars + N::B_FCW, 0);  // the weight image is in (completes once; later waits fall through)
        if (SIB_DIAG_SKIP & 2) { mbar_arrive(&p2_empty[buf]); return; }
        // ---- fc1: [16 x FEAT] . [FEAT x 120]: 15 N tiles = 15 independent accumulator chains
        // (h1 columns 120..127 stay zero from the set-up; fc2's K padding multiplies them by zero weights)
        {
            float acc[N::NT1][4];
#pragma unroll
            for (int j = 0; j < N::NT1; j++) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
            const uint16_t* a0p = p2 + g * N::P2_STRIDE + 2 * t;
            const uint16_t* a1p = a0p + 8 * N::P2_STRIDE;
#pragma unroll 1
            for (int ks = 0; ks < N::KS1; ks++) {
                uint32_t a[4];
                a[0] = lds32(a0p + ks * 16); a[1] = lds32(a1p + ks * 16); a[2] = lds32(a0p + ks * 16 + 8); a[3] = lds32(a1p + ks * 16 + 8);
#pragma unroll
                for (int j = 0; j < N::NT1; j++) {
                    const uint2 bv = *reinterpret_cast<const uint2*>(w1 + (ks * N::NT1 + j) * 64 + lane * 2);
                    const uint32_t b[2] = {bv.x, bv.y};
                    Ops<T>::mma(acc[j], a, b);
                }
            }
            mbar_arrive(&p2_empty[buf]);  // p2 fully consumed by this lane
#pragma unroll
            for (int j = 0; j < N::NT1; j++) {
                const int n0 = j * 8 + 2 * t;
                const float b0 = bias[24 + n0], b1 = bias[24 + n0 + 1];
                const float v0 = fmax_nan(acc[j][0] + b0, 0.f), v1 = fmax_nan(acc[j][1] + b1, 0.f);
                const float v2 = fmax_nan(acc[j][2] + b0, 0.f), v3 = fmax_nan(acc[j][3] + b1, 0.f);
                *reinterpret_cast<uint32_t*>(h1 + g * N::H1_STRIDE + n0) = Ops<T>::pack(v0, v1);
                *reinterpret_cast<uint32_t*>(h1 + (g + 8) * N::H1_STRIDE + n0) = Ops<T>::pack(v2, v3);
                if (dump) {
                    float* d = P.dbg + 2 * kG * N::PH1 * N::P1_ROW + N::DBG_P2;
                    d[g * 128 + n0] = v0; d[g * 128 + n0 + 1] = v1; d[(g + 8) * 128 + n0] = v2; d[(g + 8) * 128 + n0 + 1] = v3;
                }
            }
        }
        __syncwarp();
        // ---- fc2: [16 x 128] . [128 x 96]
        {
            float acc[12][4];
#pragma unroll
            for (int j = 0; j < 12; j++) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
            const uint16_t* a0p = h1 + g * N::H1_STRIDE + 2 * t;
            const uint16_t* a1p = a0p + 8 * N::H1_STRIDE;
#pragma unroll 1
            for (int ks = 0; ks < 8; ks++) {
                uint32_t a[4];
                a[0] = lds32(a0p + ks * 16); a[1] = lds32(a1p + ks * 16); a[2] = lds32(a0p + ks * 16 + 8); a[3] = lds32(a1p + ks * 16 + 8);
#pragma unroll
                for (int j = 0; j < 12; j++) {
                    const uint2 bv = *reinterpret_cast<const uint2*>(w2 + (ks * 12 + j) * 64 + lane * 2);
                    const uint32_t b[2] = {bv.x, bv.y};
                    Ops<T>::mma(acc[j], a, b);
                }
            }
#pragma unroll
            for (int j = 0; j < 12; j++) {
                const int n0 = j * 8 + 2 * t;
                const float b0 = bias[152 + n0], b1 = bias[152 + n0 + 1];
                const float v0 = fmax_nan(acc[j][0] + b0, 0.f), v1 = fmax_nan(acc[j][1] + b1, 0.f);
                const float v2 = fmax_nan(acc[j][2] + b0, 0.f), v3 = fmax_nan(acc[j][3] + b1, 0.f);
                *reinterpret_cast<uint32_t*>(h2 + g * N::H2_STRIDE + n0) = Ops<T>::pack(v0, v1);
                *reinterpret_cast<uint32_t*>(h2 + (g + 8) * N::H2_STRIDE + n0) = Ops<T>::pack(v2, v3);
            }
        }
        __syncwarp();
        // ---- fc3 + softmax / argmax / top1-top2 margin
        {
            float acc[2][4];
#pragma unroll
            for (int j = 0; j < 2; j++) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
            const uint16_t* a0p = h2 + g * N::H2_STRIDE + 2 * t;
            const uint16_t* a1p = a0p + 8 * N::H2_STRIDE;
#pragma unroll
            for (int ks = 0; ks < 6; ks++) {
                uint32_t a[4];
                a[0] = lds32(a0p + ks * 16); a[1] = lds32(a1p + ks * 16); a[2] = lds32(a0p + ks * 16 + 8); a[3] = lds32(a1p + ks * 16 + 8);
#pragma unroll
                for (int j = 0; j < 2; j++) {
                    const uint2 bv = *reinterpret_cast<const uint2*>(w3 + (ks * 2 + j) * 64 + lane * 2);
                    const uint32_t b[2] = {bv.x, bv.y};
                    Ops<T>::mma(acc[j], a, b);
                }
            }
            // lane (g,t) holds, for image g (half 0) and g+8 (half 1): classes 2t,2t+1 and, when t==0, 8,9
#pragma unroll
            for (int hrow = 0; hrow < 2; hrow++) {
                const int img = pair * kPair + g + hrow * 8;
                float x[4];
                x[0] = acc[0][hrow * 2 + 0] + bias[248 + 2 * t];
                x[1] = acc[0][hrow * 2 + 1] + bias[248 + 2 * t + 1];
                x[2] = t == 0 ? acc[1][hrow * 2 + 0] + bias[248 + 8] : -INFINITY;
                x[3] = t == 0 ? acc[1][hrow * 2 + 1] + bias[248 + 9] : -INFINITY;
                const int cls[4] = {2 * t, 2 * t + 1, 8, 9};
                float mx = x[0];
                int am = cls[0];
#pragma unroll
                for (int i = 1; i < 4; i++)
                    if (x[i] > mx) { mx = x[i]; am = cls[i]; }
                // quad reduction with first-maximum tie-break on the class index
#pragma unroll
                for (int o = 1; o <= 2; o <<= 1) {
                    const float ox = __shfl_xor_sync(0xffffffffu, mx, o);
                    const int oa = __shfl_xor_sync(0xffffffffu, am, o);
                    if (ox > mx || (ox == mx && oa < am)) { mx = ox; am = oa; }
                }
                float second = -INFINITY;
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if (cls[i] != am) second = fmaxf(second, x[i]);  // unused slots hold -inf
                second = fmaxf(second, __shfl_xor_sync(0xffffffffu, second, 1));
                second = fmaxf(second, __shfl_xor_sync(0xffffffffu, second, 2));
                float e[4], s = 0.f;
#pragma unroll
                for (int i = 0; i < 4; i++) { e[i] = expf(x[i] - mx); s += e[i]; }  // exp(-inf) = 0 for the unused slots
                s += __shfl_xor_sync(0xffffffffu, s, 1);
                s += __shfl_xor_sync(0xffffffffu, s, 2);
                if (img < P.n) {
                    float* pr = P.probs + (size_t)img * 10;
                    *reinterpret_cast<float2*>(pr + 2 * t) = make_float2(e[0] / s, e[1] / s);
                    if (t == 0) *reinterpret_cast<float2*>(pr + 8) = make_float2(e[2] / s, e[3] / s);
                    if (P.logits) {
                        float* lg = P.logits + (size_t)img * 10;
                        *reinterpret_cast<float2*>(lg + 2 * t) = make_float2(x[0], x[1]);
                        if (t == 0) *reinterpret_cast<float2*>(lg + 8) = make_float2(x[2], x[3]);
                    }
                    if (t == 0) {
                        if (P.argmax) P.argmax[img] = am;
                        const float mg = mx - second;
                        if (P.margin) P.margin[img] = mg;
                        if (P.list && (!(mg >= P.tau * fmaxf(1.f, fabsf(mx) * (1.f / 12.f))) || !(fabsf(mx) <= 3.0e38f) || !(s <= 3.0e38f))) {  // guard: recompute in fp32 (threshold grows with the logits; a NaN margin is listed too, and so is a NaN anywhere among the logits: it makes the softmax sum NaN)
                            const int pos = atomicAdd(P.count, 1);
                            if (pos < P.n) P.list[pos] = img;
                        }
                    }
                }
            }
        }
        __syncwarp();  // h1/h2 are rewritten by the next pair
    }
}

// ---- fc role (one warp): every pair's dense head ----------------------------------------------------------------------
template <typename N, typename T, bool DBG>
__device__ __forceinline__ void fc_role(const TcParams& P, unsigned char* smem, int lane) {
    const int nPairs = (P.n + kPair - 1) / kPair;
    for (int Qi = 0; (int)(blockIdx.x + Qi * gridDim.x) < nPairs; Qi++) fc_pair<N, T, DBG>(P, smem, lane, Qi);
}

// arrivals per p1 slot on p1_empty / p2_full
template <typename N>
__host__ __device__ constexpr int kConv2Arrivals() {
#if SIB_CONV2_UMMA
    return U2<N>::NWORK * 32;  // one worker set
#else
    return N::TILES2 * 32;
#endif
}

template <typename N, typename T, bool DBG>
__global__ void __launch_bounds__(kThreadsTc, 1) forward_tc_kernel(TcParams P) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // ---- one-time set-up: fc weight fragments + biases into shared memory, zero the rings, init barriers
    {
#if !SIB_TMA_WEIGHTS
        const uint4* src = reinterpret_cast<const uint4*>(P.wpack + N::W_F1);
        uint4* dst = reinterpret_cast<uint4*>(smem + N::S_FCW);
        for (int i = threadIdx.x; i < N::S_FCW_BYTES / 16; i += blockDim.x) dst[i] = src[i];
#endif
        float* b = reinterpret_cast<float*>(smem + N::S_BIAS);
        for (int i = threadIdx.x; i < 264; i += blockDim.x) b[i] = reinterpret_cast<const float*>(P.wpack + N::W_BIAS)[i];
        uint32_t* z = reinterpret_cast<uint32_t*>(smem + N::S_P1);
        for (int i = threadIdx.x; i < (N::S_BAR - N::S_P1) / 4; i += blockDim.x) z[i] = 0u;  // p1,p2,h1,h2,rings
#if SIB_CONV1_UMMA
        {
            const uint4* bs = reinterpret_cast<const uint4*>(P.wpack + N::W_C1U);
            uint4* bd = reinterpret_cast<uint4*>(smem + N::S_B1U);
            for (int i = threadIdx.x; i < 2 * 5 * 512 / 16; i += blockDim.x) bd[i] = bs[i];
        }
        __syncwarp();
        if (warp == 0) tmem_alloc_all(reinterpret_cast<uint32_t*>(smem + N::S_TMEM));
#endif
#if SIB_CONV2_UMMA
        {
            const uint4* bs = reinterpret_cast<const uint4*>(P.wpack + N::W_C2U);
            uint4* bd = reinterpret_cast<uint4*>(smem + N::S_B2U);
            for (int i = threadIdx.x; i < 2 * N::B2U_KS_BYTES / 16; i += blockDim.x) bd[i] = bs[i];
        }
        __syncwarp();
        if (warp == 0) tmem_alloc_all(reinterpret_cast<uint32_t*>(smem + N::S_TMEM));
#endif
        if (threadIdx.x == 0) {
            uint64_t* bars = reinterpret_cast<uint64_t*>(smem + N::S_BAR);
            *reinterpret_cast<uint32_t*>(smem + N::S_TMEM + 4) = 0u;  // the conv1 span counter (SIB_DYN_SPANS)
#if SIB_CONV2_UMMA
            {
                using U = U2<N>;
                for (int i = 0; i < U::NA; i++) mbar_init(&bars[N::B_U2 + i], U::NWORK * 32);                   // a_full : every lane of every worker
                for (int i = 0; i < U::NA; i++) mbar_init(&bars[N::B_U2 + U::NA + i], 1);                       // a_free : one tcgen05.commit
                for (int i = 0; i < U::NDP; i++) mbar_init(&bars[N::B_U2 + 2 * U::NA + i], 1);                  // d_full : one tcgen05.commit
                for (int i = 0; i < U::NDP; i++) mbar_init(&bars[N::B_U2 + 2 * U::NA + U::NDP + i], U::NWORK * 32);  // d_empty: every lane of every worker
                for (int i = 0; i < 3; i++) mbar_init(&bars[N::B_FC + i], 1);  // f_full: fc1 | fc2 | fc3 mma retired (one tcgen05.commit)
                mbar_init(&bars[N::B_FC + 3], 4 * 32);                         // h1_full: the four lane quadrants' warps
                mbar_init(&bars[N::B_FC + 4], 3 * 32);                         // h2_full: quadrants 0..2 (84 -> 96 features)
            }
#endif
#if SIB_CONV1_UMMA
            for (int i = 0; i < 8; i++) mbar_init(&bars[N::B_UA + i], 128);     // a_full : every thread of the team that owns the pair
            for (int i = 8; i < 16; i++) mbar_init(&bars[N::B_UA + i], 2);      // a_free : one tcgen05.commit per issuer
            for (int i = 0; i < 8; i++) mbar_init(&bars[N::B_UDF + i], 1);      // d_full : one tcgen05.commit
            for (int i = 0; i < 8; i++) mbar_init(&bars[N::B_UDE + i], 128);    // d_empty: every thread of the team that owns the pair
#endif
            for (int i = 0; i < N::NB1; i++) {
                mbar_init(&bars[N::B_P1F + i], SIB_CONV1_UMMA ? 256 : N::SPANS * 32);  // p1_full : every conv1 lane working on the slot
                mbar_init(&bars[N::B_P1E + i], (SIB_CONV2_UMMA ? 2 : 1) * kConv2Arrivals<N>());  // p1_empty: every lane of every conv2 tile | of both worker sets
            }
            for (int i = 0; i < 2; i++) {
                mbar_init(&bars[N::B_P2F + i], (SIB_CONV2_UMMA ? 4 : 2) * kConv2Arrivals<N>());  // p2_full : two slots of conv2 tiles | both worker sets
                mbar_init(&bars[N::B_P2E + i], SIB_CONV2_UMMA ? 1 : 32);  // p2_empty: every lane of the warp that runs the pair's dense head | fc1's tcgen05.commit
            }
            mbar_init(&bars[N::B_FCW], 1);
#if SIB_TMA_WEIGHTS
            // The 121 KB of dense-layer fragments come in by bulk async copies (TMA unit, no register stop-over) that
            // complete on an mbarrier: only the dense head waits for them, the conv warps start at once.
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[N::B_FCW])), "r"((uint32_t)N::S_FCW_BYTES) : "memory");
            constexpr int kChunk = 16384;
            for (int off = 0; off < N::S_FCW_BYTES; off += kChunk) {
                const int bytes = N::S_FCW_BYTES - off < kChunk ? N::S_FCW_BYTES - off : kChunk;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem + N::S_FCW + off)),
                             "l"(reinterpret_cast<const char*>(P.wpack + N::W_F1) + off), "r"((uint32_t)bytes), "r"(smem_u32(&bars[N::B_FCW])) : "memory");
            }
#else
            mbar_arrive(&bars[N::B_FCW]);
#endif
        }
    }
#if SIB_CONV1_UMMA
    tc5_fence_before();
    __syncthreads();
    tc5_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem + N::S_TMEM);
    static_assert(kConv1Warps == 8, "tcgen05 conv1: two teams of 4 warps (one warp per TMEM lane quadrant)");
#ifndef SIB_ROLE_MASK
#define SIB_ROLE_MASK 15   // diagnostic: compile only some roles (1 worker, 2 conv2, 4 fc, 8 issuer) to read their register needs
#endif
    if (warp < 8) { if (SIB_ROLE_MASK & 1) conv1u_worker<N, T, DBG>(P, smem, warp >> 2, warp & 3, lane, tmem); }
    else if (warp < kConv1Warps + kConv2Warps) { if (SIB_ROLE_MASK & 2) conv2_role<N, T, DBG>(P, smem, warp - kConv1Warps, lane); }
    else if (warp < kConv1Warps + kConv2Warps + kFcWarps) { if (SIB_ROLE_MASK & 4) fc_role<N, T, DBG>(P, smem, lane); }
    else if (SIB_ROLE_MASK & 8) conv1u_issuer<N, T>(P, smem, warp - kConv1Warps - kConv2Warps - kFcWarps, tmem);
    tc5_fence_before();
    __syncthreads();
    if (warp == 0) tmem_free_all(tmem);
#elif SIB_CONV2_UMMA
    tc5_fence_before();
    __syncthreads();
    tc5_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem + N::S_TMEM);
    // 20 warps = 5 warpgroups: conv1 x 12 | workers set 0 (q 0..2), conv2 issuer | workers set 1 (q 0..2), dense head (q 3).
    // A worker's / the head's TMEM lane quadrant is its hardware warp id mod 4. The launch gives every thread 96
    // registers; the last two warpgroups hand theirs down to 64 and the conv1 warpgroups take 128 (setmaxnreg).
    static_assert(kConv1Warps == 12 && kConv2Warps == 6 && kIssueWarps == 1 && kFcWarps == 1 && !SIB_TAIL_FIRST && !SIB_FC_IN_CONV2,
                  "tcgen05 conv2: warp layout");
    if (warp < kConv1Warps) {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 128;");
        conv1_role<N, T, DBG>(P, smem, warp, lane);
    } else {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 48;");
        const int w = warp - kConv1Warps;  // 0..7
        if ((w & 3) < 3) conv2u_worker<N, T, DBG>(P, smem, w & 3, w >> 2, lane, tmem);
        else if (w == 3) conv2u_issuer<N, T>(P, smem, tmem);
        else fcu_head<N, T, DBG>(P, smem, lane, tmem);
    }
    tc5_fence_before();
    __syncthreads();
    if (warp == 0) tmem_free_all(tmem);
#else
    __syncthreads();
    // role of hardware warp `warp`: with SIB_TAIL_FIRST the conv2 / fc warps take the lowest warp ids (sub-partition
    // placement is unchanged: ids differ by a multiple of 4)
    const int rw = SIB_TAIL_FIRST ? (warp + kConv1Warps) % kWarps : warp;
    if (rw < kConv1Warps) conv1_role<N, T, DBG>(P, smem, rw, lane);
    else if (rw < kConv1Warps + kConv2Warps) conv2_role<N, T, DBG>(P, smem, rw - kConv1Warps, lane);
    else if (kFcWarps) fc_role<N, T, DBG>(P, smem, lane);
#endif
}

// ---- host side: weight packing ----------------------------------------------------------------------------------

// B fragment of mma.m16n8k16 (col layout): lane (g = lane/4, t = lane%4) holds {B[2t][g], B[2t+1][g]} and
// {B[2t+8][g], B[2t+9][g]}; `at(k, n)` returns the fp32 weight or 0 outside the real matrix.
template <typename T, typename F>
void pack_b(uint32_t* dst, F at) {
    for (int lane = 0; lane < 32; lane++) {
        const int g = lane >> 2, t = lane & 3;
        for (int r = 0; r < 2; r++) {
            const int k = 2 * t + 8 * r;
            const uint32_t lo = Ops<T>::host_bits(at(k, g)), hi = Ops<T>::host_bits(at(k + 1, g));
            dst[lane * 2 + r] = lo | (hi << 16);
        }
    }
}

template <typename N, typename T>
std::vector<uint32_t> pack_weights(const ModelDesc& m) {
    std::vector<uint32_t> w(N::W_TOTAL, 0u);
    const float* wt = m.weights.data();
    const LayerDesc* L = m.layer;
    const float* f1 = wt + L[0].w_off;  // [5][5][C][6]
    for (int v = 0; v < 2; v++)
        for (int kh = 0; kh < 5; kh++)
            pack_b<T>(&w[N::W_C1 + (v * 5 + kh) * 64], [&](int k, int n) -> float {
                if (N::PACK2) {  // entry `kh` < 3 is the pair of kernel rows (2kh, 2kh+1): k 0..7 | k 8..15
                    const int row = 2 * kh + (k >> 3), j = (k & 7) - v;
                    if (kh > 2 || row > 4 || j < 0 || j >= 5 || n >= N::C1) return 0.f;
                    return f1[(row * 5 + j) * N::C1 + n];
                }
                const int j = k - v;  // odd-column tiles read one element earlier, so their K is shifted by one
                if (j < 0 || j >= N::K1ROW || n >= N::C1) return 0.f;
                return f1[((kh * 5 + j / N::C) * N::C + j % N::C) * N::C1 + n];
            });
    const float* f2 = wt + L[1].w_off;  // [5][5][6][16]
    for (int kh = 0; kh < 5; kh++)
        for (int ks = 0; ks < 2; ks++)
            for (int nt = 0; nt < 2; nt++)
                pack_b<T>(&w[N::W_C2 + ((kh * 2 + ks) * 2 + nt) * 64], [&](int k, int n) -> float {
                    const int j = ks * 16 + k;
                    if (j >= 5 * N::C1) return 0.f;
                    return f2[((kh * 5 + j / N::C1) * N::C1 + j % N::C1) * N::C2 + nt * 8 + n];
                });
    auto dense = [&](int off, int ksteps, int ntiles, const LayerDesc& D) {
        const float* k = wt + D.w_off;  // [in][out]
        for (int ks = 0; ks < ksteps; ks++)
            for (int nt = 0; nt < ntiles; nt++)
                pack_b<T>(&w[off + (ks * ntiles + nt) * 64], [&](int kk, int n) -> float {
                    const int i = ks * 16 + kk, j = nt * 8 + n;
                    return (i < (int)D.a && j < (int)D.b) ? k[i * D.b + j] : 0.f;
                });
    };
#if SIB_CONV2_UMMA
    // tcgen05 A operands: the TRANSPOSED dense kernels (row m = output feature), K-major, no swizzle
    auto dense_a = [&](int off, int groups, int sbo, const LayerDesc& D) {
        const float* k = wt + D.w_off;  // [in][out]
        uint16_t* dst = reinterpret_cast<uint16_t*>(&w[off]);
        for (int m = 0; m < groups * 8; m++)
            for (int i = 0; i < sbo / 16; i++) {
                const float x = (m < (int)D.b && i < (int)D.a) ? k[i * D.b + m] : 0.f;
                dst[(m / 8) * (sbo / 2) + (i / 8) * 64 + (m % 8) * 8 + (i % 8)] = Ops<T>::host_bits(x);
            }
    };
    dense_a(N::W_F1, N::F1_GROUPS, N::F1_SBO, L[2]);
    dense_a(N::W_F2, N::F2_GROUPS, N::F2_SBO, L[3]);
    dense_a(N::W_F3, N::F3_GROUPS, N::F3_SBO, L[4]);
#else
    dense(N::W_F1, N::KS1, N::NT1, L[2]);
    dense(N::W_F2, 8, 12, L[3]);
    dense(N::W_F3, 6, 2, L[4]);
#endif
    // conv1 as tcgen05 B operands: canonical K-major no-swizzle tiles [n/8][k/8][n%8][k%8] of halves
    for (int v = 0; v < 2; v++)
        for (int kh = 0; kh < 5; kh++) {
            uint16_t* dst = reinterpret_cast<uint16_t*>(&w[N::W_C1U + (v * 5 + kh) * 128]);
            for (int n = 0; n < 16; n++)
                for (int k = 0; k < 16; k++) {
                    const int j = k - v;
                    const float x = (j < 0 || j >= N::K1ROW || n >= N::C1) ? 0.f : f1[((kh * 5 + j / N::C) * N::C + j % N::C) * N::C1 + n];
                    dst[(n / 8) * 128 + (k / 8) * 64 + (n % 8) * 8 + (k % 8)] = Ops<T>::host_bits(x);
                }
        }
    // conv2 as tcgen05 B operands: per K step, blocks of 16 rows (= output channels) for the kernel rows 4,3,2,1,0, then
    // 16 rows of zeros; k = kw * 6 + c of that kernel row (sib_tc_conv2_umma.cuh: the issuer's N = 32 windows)
    for (int ks = 0; ks < 2; ks++) {
        uint16_t* dst = reinterpret_cast<uint16_t*>(&w[N::W_C2U + ks * N::B2U_KS_BYTES / 4]);
        for (int blk = 0; blk < 5; blk++)
            for (int nn = 0; nn < 16; nn++)
                for (int k = 0; k < 16; k++) {
                    const int kh = 4 - blk, n = blk * 16 + nn, j = ks * 16 + k;
                    float x = j >= 5 * N::C1 ? 0.f : f2[((kh * 5 + j / N::C1) * N::C1 + j % N::C1) * N::C2 + nn];
                    if (kh == 0 && j >= 30) {  // K slots 30, 31 meet the constant 1.0 of every A row: the bias, head and tail
                        const float bv = wt[L[1].b_off + nn];
                        const uint16_t hb = Ops<T>::host_bits(bv);
                        const float hi = Ops<T>::host_float(hb);
                        x = j == 30 ? hi : bv - hi;
                    }
                    dst[(n / 8) * 128 + (k / 8) * 64 + (n % 8) * 8 + (k % 8)] = Ops<T>::host_bits(x);
                }
    }
    float* b = reinterpret_cast<float*>(&w[N::W_BIAS]);
    std::memcpy(b, wt + L[0].b_off, 6 * 4);
    std::memcpy(b + 8, wt + L[1].b_off, 16 * 4);
    std::memcpy(b + 24, wt + L[2].b_off, 120 * 4);
    std::memcpy(b + 152, wt + L[3].b_off, 84 * 4);
    std::memcpy(b + 248, wt + L[4].b_off, 10 * 4);
    return w;
}

template <typename N, typename T, bool DBG>
int launch_inst(const TcParams& p, cudaStream_t st) {
    static SmemLimit limit;  // function attributes are per device
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    if (int rc = limit.ensure(dev, N::S_TOTAL, [](int b) {
            return (int)cudaFuncSetAttribute(forward_tc_kernel<N, T, DBG>, cudaFuncAttributeMaxDynamicSharedMemorySize, b);
        }))
        return rc;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    TcParams q = p;
#if SIB_SPREAD_TAIL
    // one CTA per SM, or per image when there are fewer images than SMs; the stage dump is CTA 0's first pair, so a
    // launch that dumps runs on one CTA (grid 1: whole pairs in order, then the remainder -- the plain order)
    const int grid = DBG ? 1 : (p.n < sms ? p.n : sms);
    q.rounds = p.n / kPair / grid;
    const int rem = p.n - q.rounds * grid * kPair;  // < grid * kPair
    q.per = (rem + grid - 1) / grid;
#else
    const int pairs = (p.n + kPair - 1) / kPair;
    const int grid = pairs < sms ? pairs : sms;
#endif
    forward_tc_kernel<N, T, DBG><<<grid, kThreadsTc, N::S_TOTAL, st>>>(q);
    return (int)cudaGetLastError();
}

template <typename N, typename T>
int launch(const TcWeights&, const TcParams& p, cudaStream_t st) {
    return p.dbg ? launch_inst<N, T, true>(p, st) : launch_inst<N, T, false>(p, st);
}

using NetMnist = Net<28, 28, 1>;
using NetCifar = Net<32, 32, 3>;

}  // namespace

// The two LeNet-5 presets take the fused persistent kernel of this file, every other model of the family the any-shape
// kernel (sib_kernels_tcgen.cu), if its activations fit shared memory.
static bool fused_shape(const ModelDesc& m) { return m.preset == 0 || m.preset == 1; }

bool tc_supported(const ModelDesc& m) {
    if (fused_shape(m)) return true;
    TcgPlan pl;
    return tcg_plan(m, &pl, nullptr);
}

int tc_prepare(const ModelDesc& m, int precision, TcWeights* out) {
    std::vector<uint32_t> w;
    const bool bf = precision == SIB_PREC_BF16;
    if (!fused_shape(m)) {
        std::string why;
        if (!tcg_plan(m, &out->plan, &why)) { set_error("tensor-core path: " + why + "; open this model with SIB_PREC_FP32"); return SIB_EINVAL; }
        if (!bf)
            for (float x : m.weights)
                if (!(std::fabs(x) < 65504.f)) { set_error("a weight exceeds the half-precision range: open the engine with SIB_PREC_BF16 or SIB_PREC_FP32"); return SIB_EINVAL; }
        w = tcg_pack(m, precision, &out->plan);
        out->generic = true;
        out->bytes = w.size() * 4;
        out->preset = m.preset;
        out->precision = precision;
        if (cudaMalloc(&out->d_packed, out->bytes) != cudaSuccess ||
            cudaMemcpy(out->d_packed, w.data(), out->bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
            set_error("tensor-core weight upload failed");
            return SIB_ECUDA;
        }
        return SIB_OK;
    }
    if (!bf)  // half precision tops out at 65504: a larger weight would become Inf and every image NaN
        for (float x : m.weights)
            if (!(std::fabs(x) < 65504.f)) { set_error("a weight exceeds the half-precision range: open the engine with SIB_PREC_BF16 or SIB_PREC_FP32"); return SIB_EINVAL; }
    if (m.preset == 0) w = bf ? pack_weights<NetMnist, __nv_bfloat16>(m) : pack_weights<NetMnist, __half>(m);
    else w = bf ? pack_weights<NetCifar, __nv_bfloat16>(m) : pack_weights<NetCifar, __half>(m);
    out->bytes = w.size() * 4;
    out->preset = m.preset;
    out->precision = precision;
    if (cudaMalloc(&out->d_packed, out->bytes) != cudaSuccess ||
        cudaMemcpy(out->d_packed, w.data(), out->bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
        set_error("tensor-core weight upload failed");
        return SIB_ECUDA;
    }
    return SIB_OK;
}

void tc_release(TcWeights* w) {
    if (w && w->d_packed) cudaFree(w->d_packed);
    if (w) *w = TcWeights{};
}

size_t tc_debug_floats(int preset) {
    return preset == 0 ? 2 * kG * NetMnist::PH1 * NetMnist::P1_ROW + NetMnist::DBG_P2 + 16 * 128
                       : 2 * kG * NetCifar::PH1 * NetCifar::P1_ROW + NetCifar::DBG_P2 + 16 * 128;
}

int launch_forward_tc(const TcWeights& w, const float* d_in, int n, float* d_probs, float* d_logits, int32_t* d_argmax,
                      GuardScratch* guard, float tau, float* d_dbg, void* stream) {
    if (n <= 0) return 0;
    if (w.generic) return launch_forward_tcgen(w, d_in, n, d_probs, d_logits, d_argmax, guard, tau, stream);
    cudaStream_t st = (cudaStream_t)stream;
    TcParams p{};
    p.wpack = static_cast<const uint32_t*>(w.d_packed);
    p.in = d_in; p.n = n; p.probs = d_probs; p.logits = d_logits; p.argmax = d_argmax;
    p.tau = tau; p.dbg = d_dbg;
    if (guard) {
        p.margin = guard->d_margin; p.list = guard->d_list; p.count = guard->d_count;
        if (!guard->self_reset) {  // shared scratch: the list starts empty because this call says so
            cudaError_t e = cudaMemsetAsync(guard->d_count, 0, sizeof(int32_t), st);
            if (e != cudaSuccess) return (int)e;
        }  // else: the previous launch's guard kernel left the counter at zero
    }
    const bool bf = w.precision == SIB_PREC_BF16;
    if (w.preset == 0) return bf ? launch<NetMnist, __nv_bfloat16>(w, p, st) : launch<NetMnist, __half>(w, p, st);
    return bf ? launch<NetCifar, __nv_bfloat16>(w, p, st) : launch<NetCifar, __half>(w, p, st);
}

}  // namespace sib
