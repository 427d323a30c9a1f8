// sib_tc_conv2_umma.cuh -- conv2 + bias + ReLU + 2x2 pool on the 5th-gen tensor cores (tcgen05.mma, A and D in TMEM).
// Included by sib_kernels_tc.cu (namespace sib::{anonymous}) when SIB_CONV2_UMMA is on (the default).
//
// Why conv2: its N is exactly 16 and a kernel row's K is 5 x 6 = 30 -> 32, the native M128 x N16 x K16 tile, and on the
// mma.sync path it was the stage that set the pace (3 warps ~90% busy, 125 of the 402 HMMA per image, 8 shared-memory
// fragment loads per 16 output positions and row). With the A operand in TMEM one mma retires in 9 clk
// (profiles/tcgen05_probe2_r01.jsonl), and a pooled conv1 row written to TMEM once serves all five kernel rows:
//     D(oh) += A(r) . B(kh)   for oh = r - kh,  kh = 0..4
// so the im2col expansion (kw only) exists once per input row, in TMEM, and never in shared memory.
//
// Shape. One 8-image p1 slot = one M tile. TMEM lane = (image, output column): a worker warp owns the lanes of its
// lane quadrant = the columns of 3 (cifar) or 4 (mnist) images; 30 or 32 live lanes. Work goes by PAIRS of p1 rows.
//   worker warps (two sets of q = 0..NWORK-1; set s takes the row pairs rp with rp % 2 == s), per tile:
//        first every row pair of the set: 2 x 15 LDS.32 (halves [6 ow, 6 ow + 30) of both rows; K slots 30, 31 are 1.0
//        and meet the bias in kernel row 0's B tile) -> one tcgen05.st x32 -> a_full[rp]; the p1 slot goes back to conv1
//        after the set's last row pair (both sets arrive on p1_empty);
//        then every output-row pair of the set: tcgen05.ld x32 of both rows -> max (NaN-propagating) -> lane-pair
//        exchange (the pool's other column) -> ReLU -> one 16-byte store of 8 channels into p2, which is the dense head's
//        B operand (144 bytes between K chunks: the lanes of that store differ in the K chunk); between the two phases
//        and after them the set's share of the dense head's epilogues (sib_tc_fc_umma.cuh)
//   issuer warp: per p1 row pair 12 mma (the even row through the stacked filter blocks 4|2|0 with N = 32, so one mma adds
//        it into BOTH rows of an output pair; the odd row through 3|1 and, alone, 0) + tcgen05.commit -> a_free[rp],
//        d_full[output pair rp - 2]
// A lone warp retires about one instruction per 3.5 clk on this dependent code, so the roles are split so that no warp
// carries more than ~1000 instructions per tile (first version: the third worker also issued; it ran 1500 clk per row
// and the whole kernel at half speed, profiles/ncu_tc_r02_notes.md), barriers are addressed by their shared-memory
// offset (no generic-to-shared conversion per wait), and the row-pair loops are ROLLED: unrolled, the kernel's roles
// together overflowed the instruction cache (72% hit rate).
// TMEM: 14 A rows x 16 columns (slot = row; a tile's rows go in while the previous tile's last mma still run) +
// a ring of 8 output-row pairs x 2 x 16 columns = 480 of 512 columns; the dense head's accumulator takes the last 16.
#pragma once

template <typename N>
struct U2 {
    static constexpr int IPQ = 32 / N::OW2;                    // images per TMEM lane quadrant (3 | 4)
    static constexpr int NWORK = (kG + IPQ - 1) / IPQ;         // warps that hold images (3 | 2)
    static constexpr int ROWS = N::PH1, OROWS = N::OH2, OP = N::PH2, RP = ROWS / 2;  // p1 rows, output rows, output pairs, p1 row pairs
    static constexpr int NA = 16, NDP = 8;                     // A row slots (ROWS used) / D row pairs resident in TMEM
    static constexpr int TM_A = 0, TM_D = ROWS * 16;
    static_assert(2 * NWORK <= kConv2Warps && ROWS <= NA && TM_D + NDP * 32 <= 512 && OROWS == 2 * OP && ROWS == OROWS + 4 && ROWS % 2 == 0 && RP == OP + 2, "conv2 umma tiling");
    static_assert(5 * N::C1 <= 32 && N::C2 == 16 && (N::OW2 % 2) == 0 && (N::P2_STRIDE * 2) % 16 == 0, "conv2 umma layout");
};

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
                 "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
                 "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7]), "=f"(r[8]),
                   "=f"(r[9]), "=f"(r[10]), "=f"(r[11]), "=f"(r[12]), "=f"(r[13]), "=f"(r[14]), "=f"(r[15])
                 : "r"(taddr) : "memory");
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
                 "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
                 "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
                 "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
                 "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
                 "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7]), "=f"(r[8]),
                   "=f"(r[9]), "=f"(r[10]), "=f"(r[11]), "=f"(r[12]), "=f"(r[13]), "=f"(r[14]), "=f"(r[15]), "=f"(r[16]),
                   "=f"(r[17]), "=f"(r[18]), "=f"(r[19]), "=f"(r[20]), "=f"(r[21]), "=f"(r[22]), "=f"(r[23]), "=f"(r[24]),
                   "=f"(r[25]), "=f"(r[26]), "=f"(r[27]), "=f"(r[28]), "=f"(r[29]), "=f"(r[30]), "=f"(r[31])
                 : "r"(taddr) : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem]^T with the B descriptor given as two words (the high word is the same for every tile)
__device__ __forceinline__ void umma_ts2(uint32_t d, uint32_t a, uint32_t bdesc_lo, uint32_t bdesc_hi, uint32_t idesc, uint32_t accumulate) {
    if (SIB_DIAG_SKIP & 8) return;  // diagnostic: everything but the tensor work
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 bd;\n\tsetp.ne.b32 p, %5, 0;\n\tmov.b64 bd, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], bd, %4, {%6, %6, %6, %6}, p;\n\t}\n" ::"r"(d), "r"(a), "r"(bdesc_lo),
        "r"(bdesc_hi), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
template <typename T> __device__ __forceinline__ constexpr uint32_t umma_idesc_m128(uint32_t n) {
    constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((n >> 3) << 17) | ((128u >> 4) << 24);
}

// barriers by shared-memory address
__device__ __forceinline__ void mbar_arrive_s(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_s(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __noinline__ void mbar_spin_s(uint32_t bar, uint32_t parity) {  // slow path: bounded spin, traps on a protocol bug
    uint32_t spin = 0;
    long long t0 = 0;
    while (!mbar_try_s(bar, parity)) {
        if ((++spin & 1023u) == 0) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 4000000000ll) __trap();
        }
    }
}
__device__ __forceinline__ void mbar_wait_s(uint32_t bar, uint32_t parity) {
    if (!mbar_try_s(bar, parity)) mbar_spin_s(bar, parity);
}
__device__ __forceinline__ void umma_commit_s(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async();  // sib_tc_fc_umma.cuh

template <typename N>
struct U2Bars {  // shared-memory addresses of the stage's barriers
    uint32_t p1_full, p1_empty, p2_full, p2_empty, a_full, a_free, d_full, d_empty;
    __device__ __forceinline__ explicit U2Bars(unsigned char* smem) {
        const uint32_t b = smem_u32(smem + N::S_BAR);
        p1_full = b + N::B_P1F * 8; p1_empty = b + N::B_P1E * 8; p2_full = b + N::B_P2F * 8; p2_empty = b + N::B_P2E * 8;
        a_full = b + N::B_U2 * 8; a_free = a_full + U2<N>::NA * 8; d_full = a_free + U2<N>::NA * 8; d_empty = d_full + U2<N>::NDP * 8;
    }
};

// does slot lq of this CTA hold work? (the launch's last pair may be half empty)
template <typename N>
__device__ __forceinline__ int conv2u_slot(const TcParams& P, int lq) {  // -1: past the end, 0: empty slot, 1: work
    int base, lim;
    if (!pair_images(P, lq >> 1, base, lim)) return -1;
    return (SIB_SKIP_EMPTY && base + (lq & 1) * kG >= lim) ? 0 : 1;
}

template <typename N, typename T, bool DBG, int STAGE>
__device__ __forceinline__ void fcu_epilogue(const TcParams& P, unsigned char* smem, int q, int lane, uint32_t tmem, int Qi);  // sib_tc_fc_umma.cuh

// ---- worker warps: q = this warp's TMEM lane quadrant (hardware warp 4k + q) ------------------------------------------
// Besides conv2 they run the dense head's epilogues for their quadrant (sib_tc_fc_umma.cuh): the head of pair Qi-1 is
// computed while the first slot of pair Qi passes through here, so its two steps are slotted in after rows 4 and 9.
// Two sets of workers share the row-pair steps of every tile: set s takes the steps rp with rp % 2 == s.
template <typename N, typename T, bool DBG>
__device__ __forceinline__ void conv2u_worker(const TcParams& P, unsigned char* smem, int q, int set, int lane, uint32_t tmem) {
    using U = U2<N>;
    const U2Bars<N> B(smem);
    const bool fc1_mine = set == 1, fc2_mine = set == 0;  // which set runs the dense head's epilogues for this quadrant
    const int nimg = max(0, min(U::IPQ, kG - q * U::IPQ));
    const int nlive = nimg * N::OW2;
    if (SIB_DIAG_SKIP & 4) {  // diagnostic: conv1 alone (the p1 slots are recycled unread)
        for (int lq = 0; conv2u_slot<N>(P, lq) >= 0; lq++) {
            if (nimg == 0) continue;
            mbar_wait_s(B.p1_full + (lq % N::NB1) * 8, (lq / N::NB1) & 1);
            mbar_arrive_s(B.p1_empty + (lq % N::NB1) * 8);
        }
        return;
    }
    if (nimg == 0) {  // no conv2 columns in this quadrant (mnist, q = 2): only the dense head's epilogues
        for (int Qi = 0; conv2u_slot<N>(P, 2 * Qi) >= 0; Qi++) {
            if (fc1_mine) fcu_epilogue<N, T, DBG, 1>(P, smem, q, lane, tmem, Qi);
            if (fc2_mine) fcu_epilogue<N, T, DBG, 2>(P, smem, q, lane, tmem, Qi);
        }
        return;
    }
    const bool live = lane < nlive;                  // lanes past the last column shadow it: same loads, no stores
    const int l = min(lane, nlive - 1);
    const int imgl = q * U::IPQ + l / N::OW2, ow = l % N::OW2;
    const int odd = lane & 1;                        // the pool pair is (lane, lane ^ 1); this lane stores channels 8*odd..+7
    const uint32_t offA = smem_u32(smem + N::S_P1) + (imgl * N::P1_IMG + ow * N::C1) * 2;            // + buf, + row
    const int o2 = imgl * N::P2_STRIDE + (ow >> 1) * N::C2 + odd * 8;  // stage dump: padded [img][feature] layout
    // p2 is the dense head's B operand: [img/8][k/8][img%8][k%8], BK bytes per k/8; this lane's 8 channels are one 16-byte chunk of it
    const uint32_t offP2 = smem_u32(smem + N::S_P2) + ((ow >> 1) * 2 + odd) * N::BK + imgl * 16;
    const uint32_t tm_lane = tmem + ((uint32_t)(q * 32) << 16);
    // The bias rides in the mma: K slots 30, 31 of every A row hold 1.0 and kernel row 0's B tile holds the bias there,
    // split into a half-precision head and tail (b = hi + lo to ~22 bits); every output row meets kernel row 0 once.
    const uint32_t ones = Ops<T>::pack(1.f, 1.f);

    int seq = 0;  // tiles that went through the tensor core so far
    int last = -1;
    for (int lq = 0;; lq++) {
        const int work = conv2u_slot<N>(P, lq);
        if (work < 0) break;
        const int Qi = lq >> 1;
        const int buf1 = lq % N::NB1, use1 = lq / N::NB1;
        const int buf2 = Qi % N::NP2, use2 = Qi / N::NP2;
        mbar_wait_s(B.p1_full + buf1 * 8, use1 & 1);
        if (work == 0) {  // nothing to compute, keep the slot protocol going
            mbar_arrive_s(B.p1_empty + buf1 * 8);
            mbar_arrive_s(B.p2_full + buf2 * 8);
            continue;
        }
        const uint32_t src0 = offA + buf1 * N::P1_BUF * 2;
        const uint32_t p2 = offP2 + buf2 * N::P2_BUF * 2 + (lq & 1) * N::P2_SBO;
        const bool dump = DBG && P.dbg && blockIdx.x == 0 && Qi == 0;
        const uint32_t par = seq & 1;

        // Rolled loops (the unrolled form was 36 KB of code; the kernel's roles together overflowed the instruction
        // cache: 72% hit rate, profiles/ncu_tc_r02_notes.md). First this set's row pairs go into TMEM back to back, so
        // the p1 slot returns to the conv1 warps as early as possible (they were waiting for it 20% of their time when
        // rows and epilogues alternated); then the previous pair's fc1 epilogue, this set's output pairs, fc2's.
#pragma unroll 1
        for (int rp = set; rp < U::RP; rp += 2) {
            mbar_wait_s(B.a_free + rp * 8, par ^ 1);  // the mma that read these row slots' previous tenants are done
            tc5_fence_after();
            const uint32_t src = src0 + rp * (2 * N::P1_ROW * 2);
            uint32_t a[32];
#pragma unroll
            for (int k = 0; k < 15; k++) asm volatile("ld.shared.b32 %0, [%1];" : "=r"(a[k]) : "r"(src + k * 4));
#pragma unroll
            for (int k = 0; k < 15; k++) asm volatile("ld.shared.b32 %0, [%1];" : "=r"(a[16 + k]) : "r"(src + N::P1_ROW * 2 + k * 4));
            a[15] = a[31] = ones;
            tmem_st32(tm_lane + U::TM_A + rp * 32, a);
            tmem_wait_st();
            tc5_fence_before();
            mbar_arrive_s(B.a_full + rp * 8);
        }
        // Both sets hand the slot back, each after its own last load from it. (Only the set that stored the slot's last
        // rows used to: the other set is at most a tile behind and conv1 needs ~1 us to reach the first row it might still
        // want, so no run ever showed a difference -- but nothing in the protocol ruled it out.)
        mbar_arrive_s(B.p1_empty + buf1 * 8);
        const bool head_due = (lq & 1) == 0 && Qi > 0;  // the previous pair's dense head
        if (head_due && fc1_mine) fcu_epilogue<N, T, DBG, 1>(P, smem, q, lane, tmem, Qi - 1);
        // first write into this pair's p2 buffer: fc1 of the pair that used it before must have read it (its mma ran
        // under the row stores above)
        if ((lq & 1) == 0) mbar_wait_s(B.p2_empty + buf2 * 8, (use2 & 1) ^ 1);
#pragma unroll 1
        for (int j = set; j < U::OP; j += 2) {
            const int gq = seq * U::OP + j, dp = gq & (U::NDP - 1);
            mbar_wait_s(B.d_full + dp * 8, (gq / U::NDP) & 1);
            tc5_fence_after();
            float dd[32];
            tmem_ld32(tm_lane + U::TM_D + dp * 32, dd);
            tmem_wait_ld();
            tc5_fence_before();
            mbar_arrive_s(B.d_empty + dp * 8);
            float v[8];
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const float lo = fmax_nan(dd[c], dd[16 + c]), hi = fmax_nan(dd[c + 8], dd[24 + c]);
                const float other = __shfl_xor_sync(0xffffffffu, odd ? lo : hi, 1);  // what the pair lane stores, it gets
                v[c] = fmax_nan(fmax_nan(odd ? hi : lo, other), 0.f);                 // the bias is in already; relu(max(.)) == max relu(.)
            }
            if (live) {
                uint4 pk;
                pk.x = Ops<T>::pack(v[0], v[1]); pk.y = Ops<T>::pack(v[2], v[3]);
                pk.z = Ops<T>::pack(v[4], v[5]); pk.w = Ops<T>::pack(v[6], v[7]);
                asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(p2 + j * N::PW2 * 2 * N::BK), "r"(pk.x), "r"(pk.y), "r"(pk.z), "r"(pk.w) : "memory");
                if (dump) {
                    float* dg = P.dbg + 2 * kG * N::PH1 * N::P1_ROW + (lq & 1) * kG * N::P2_STRIDE + o2 + j * N::PW2 * N::C2;
#pragma unroll
                    for (int c = 0; c < 8; c++) dg[c] = v[c];
                }
            }
        }
        if (head_due && fc2_mine) fcu_epilogue<N, T, DBG, 2>(P, smem, q, lane, tmem, Qi - 1);
        fence_proxy_async();  // p2 was written by ordinary stores and is read by the tensor core
        mbar_arrive_s(B.p2_full + buf2 * 8);
        seq++;
        last = Qi;
    }
    if (last >= 0) {  // the last pair's dense head
        if (fc1_mine) fcu_epilogue<N, T, DBG, 1>(P, smem, q, lane, tmem, last);
        if (fc2_mine) fcu_epilogue<N, T, DBG, 2>(P, smem, q, lane, tmem, last);
    }
}

// ---- issuer warp ---------------------------------------------------------------------------------------------------------
template <typename N, typename T>
__device__ __forceinline__ void conv2u_issuer(const TcParams& P, unsigned char* smem, uint32_t tmem) {
    using U = U2<N>;
    const U2Bars<N> B(smem);
    // B operands: the five kernel rows' 16 x 16 tiles stacked along N in the order 4,3,2,1,0 and a block of zeros,
    // per K step ([n/8][k/8][n%8][k%8], 256 bytes per 8 rows). An N = 32 mma that starts at block i multiplies by
    // [B(4-i) | B(3-i)] and so adds one p1 row into BOTH output rows of a D pair (16 + 16 accumulator columns) at once:
    // 12 mma per row pair instead of 20. Block 4 with N = 32 is [B0 | 0]: the pair's first mma, which clears it.
    const uint32_t b_lo = (smem_u32(smem + N::S_B2U) >> 4) & 0x3FFF;
    const uint32_t b_hi = (256u >> 4) | (1u << 14);  // stride 256 bytes between 8-row groups, descriptor version 1 (bit 46)
    constexpr uint32_t LBO = (128u >> 4) << 16;       // next K chunk: 128 bytes
    constexpr uint32_t KS = N::B2U_KS_BYTES >> 4, BLK = 512u >> 4;
    const uint32_t i16 = umma_idesc_m128<T>(16), i32 = umma_idesc_m128<T>(32);
    if (SIB_DIAG_SKIP & 4) return;
    int seq = 0;
    for (int lq = 0;; lq++) {
        const int work = conv2u_slot<N>(P, lq);
        if (work < 0) break;
        if (work == 0) continue;
        const uint32_t par = seq & 1;
#pragma unroll 1
        for (int p = 0; p < U::RP; p++) {  // p1 rows 2p, 2p+1 -> output pairs p, p-1, p-2
            mbar_wait_s(B.a_full + p * 8, par);
            if (p < U::OP) {  // output pair p starts here: its accumulator columns must have been read
                const int gq = seq * U::OP + p;
                mbar_wait_s(B.d_empty + (gq & (U::NDP - 1)) * 8, ((gq / U::NDP) & 1) ^ 1);
            }
            tc5_fence_after();
            if (elect_one()) {
                const uint32_t a_e = tmem + U::TM_A + p * 32, a_o = a_e + 16;
#pragma unroll
                for (int i = 0; i < 3; i++) {  // pair p - i: even row through blocks 4 | 2 | 0, odd row through 3 | 1 | (B4 alone)
                    const int j = p - i;
                    if (j < 0 || j >= U::OP) continue;
                    const uint32_t d = tmem + U::TM_D + ((seq * U::OP + j) & (U::NDP - 1)) * 32;
                    const uint32_t be = (b_lo + (4 - 2 * i) * BLK) | LBO, bo = (b_lo + (i < 2 ? 3 - 2 * i : 0) * BLK) | LBO;
                    umma_ts2(d, a_e, be, b_hi, i32, i ? 1u : 0u);
                    umma_ts2(d, a_e + 8, be + KS, b_hi, i32, 1u);
                    umma_ts2(i < 2 ? d : d + 16, a_o, bo, b_hi, i < 2 ? i32 : i16, 1u);
                    umma_ts2(i < 2 ? d : d + 16, a_o + 8, bo + KS, b_hi, i < 2 ? i32 : i16, 1u);
                }
                umma_commit_s(B.a_free + p * 8);
                if (p >= 2) {  // output pair p-2 is complete once everything up to here has retired
                    const int gq = seq * U::OP + p - 2;
                    umma_commit_s(B.d_full + (gq & (U::NDP - 1)) * 8);
                }
            }
            __syncwarp();
        }
        seq++;
    }
}
