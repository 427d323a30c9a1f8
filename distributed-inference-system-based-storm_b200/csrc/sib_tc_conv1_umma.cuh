// sib_tc_conv1_umma.cuh -- conv1 + bias + ReLU + 2x2 pool on the 5th-gen tensor cores (tcgen05.mma, operands in TMEM).
// Included by sib_kernels_tc.cu (namespace sib::{anonymous}) when SIB_CONV1_UMMA is on.
//
// Why: on the mma.sync path conv1 is 245 of the network's 402 HMMA per image and the legacy tensor path tops out
// at one m16n8k16 per 8.5 clk per sub-partition. tcgen05.mma M=128 N=16 K=16 with the A operand in TMEM retires in
// 9 clk (tools/tcgen05_probe.cu), 3.8x the MAC rate, and its im2col rows can be written from registers, so
// the operand still never exists in shared memory.
//
// Shape. One 8-image slot is in flight. Thread q of a 4-warp team (= the 4 TMEM lane quadrants) owns the output
// columns 2q and 2q+1 of the slot's column space, i.e. TMEM lane q of an "even" and an "odd" M=128 tile (odd columns
// read their 16 K values one element early against filters shifted by one: every shared-memory read stays 4-byte
// aligned, and the 2x2 pool never leaves the thread). A warp holds the column pairs of exactly two images
// (28 | 24 live lanes): with one image seam per warp the 3-words-per-lane shared-memory reads never share a bank.
// Everything moves in ROW PAIRS so that one mbarrier round trip, one tcgen05.wait and one fence serve two rows,
// and two teams of warps alternate pairs so that their latency chains overlap (profiles/ncu_tc_r01_notes.md):
//   worker warps (2 teams x 4; team t owns the row pairs of parity t):
//     build    : stage input rows 2i, 2i+1 (fp32 -> half, warp-private) -> LDS.32 -> 4 x tcgen05.st (A_e | A_o of
//                both rows) -> arrive on a_full[pair % 8]
//     epilogue : four pairs later, tcgen05.ld D of both rows and both tiles -> max over the 2x2 window -> +bias,
//                ReLU -> p1
//   issuer warps (2; issuer w owns the output pairs of parity w): once pairs j..j+2 are in TMEM:
//                       D_t(oh) = sum_kh A_t(oh+kh) . B_t,kh for oh = 2j, 2j+1 and t = even | odd (20 mma)
//                       -> tcgen05.commit -> d_full[j % 4], a_free[pairs j, j+1]
// TMEM: 16 input rows x (8 + 8) columns + 4 output-row pairs x 2 x (16 + 16) columns = 512 columns, all of it.
#pragma once

#ifndef SIB_U_EARLY_EPI
#define SIB_U_EARLY_EPI 0   // 1: start the epilogue's TMEM loads before the build, collect them after it (measured: -7%)
#endif
#ifndef SIB_U_PF
#define SIB_U_PF 0          // 1: L2 prefetch of the images two slots ahead, once per slot (measured: -7%); 0: none
#endif

template <typename N>
struct Umma {
    static constexpr int NPOS = kG * N::OW1, NPAIR = NPOS / 2;       // output columns / column pairs per slot row
    static constexpr int PPW = NPAIR / 4;                             // pairs per warp = the pairs of two images
    static constexpr int PXW = 2 * N::W;                              // input pixels a warp stages per row: two image rows
    static constexpr int NCH = PXW * N::C / 4;                        // 16-byte chunks of a staged warp row
    static constexpr int CPL = (NCH + 31) / 32;                       // chunks per lane
    static constexpr int UROW = PXW * N::C + 16;                      // halves per staged row incl. zero slack
    static constexpr int NAP = 8, NDP = 4;                            // input-row pairs / output-row pairs resident in TMEM
    static constexpr int TM_A = 0, TM_D = NAP * 32;                   // TMEM columns: A rows 16 each, D row pairs 64 each
    static constexpr int HP = N::H / 2, OP = N::OH1 / 2;              // input / output row pairs per slot
    static_assert(NPAIR % 4 == 0 && PPW <= 32 && PPW == N::OW1 && UROW % 8 == 0 && UROW == N::URING_ROW, "umma tiling");
    static_assert(N::H % 4 == 0 && N::OH1 % 2 == 0 && kBR == 2 && TM_D + NDP * 64 <= 512, "umma pipeline");
};

// this lane's column pair inside the slot and inside the warp's staged row (lanes >= PPW shadow the last live lane;
// their tile rows are computed and never used)
template <typename N>
struct UmmaLane {
    bool live;
    int img_l, ow, offE;
    __device__ __forceinline__ UmmaLane(int quad, int lane) {
        using U = Umma<N>;
        live = lane < U::PPW;
        const int pe = 2 * (U::PPW * quad + min(lane, U::PPW - 1));
        img_l = pe / N::OW1;
        ow = pe % N::OW1;
        offE = ((img_l - 2 * quad) * N::W + ow) * N::C;  // even column: halves offE .. offE+15 of the staged row
    }
};

// ---- worker warps: two teams of 4 warps; team t owns the input-row pairs and output-row pairs of parity t -----------

template <typename N, typename T, bool DBG>
__device__ __forceinline__ void conv1u_worker(const TcParams& P, unsigned char* smem, int team, int quad, int lane, uint32_t tmem) {
    using U = Umma<N>;
    uint16_t* ring = reinterpret_cast<uint16_t*>(smem + N::S_RING) + (team * 4 + quad) * 2 * U::UROW;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + N::S_BAR);
    uint64_t* p1_full = bars + N::B_P1F;
    uint64_t* p1_empty = bars + N::B_P1E;
    uint64_t* a_full = bars + N::B_UA;
    uint64_t* a_free = bars + N::B_UA + U::NAP;
    uint64_t* d_full = bars + N::B_UDF;
    uint64_t* d_empty = bars + N::B_UDE;
    const float* biasp = reinterpret_cast<const float*>(P.wpack + N::W_BIAS);
    float bias[N::C1];
#pragma unroll
    for (int c = 0; c < N::C1; c++) bias[c] = biasp[c];
    const uint32_t tm_lane = tmem + ((uint32_t)(quad * 32) << 16);  // this warp's TMEM lane quadrant
    const int nPairs = (P.n + kPair - 1) / kPair;
    const UmmaLane<N> L(quad, lane);
    constexpr int kOdd = N::C - 1;  // odd column: one element early (see header)
    const int o_p1 = L.img_l * N::P1_IMG + (L.ow >> 1) * N::C1;
    const int o_dbg = L.img_l * N::PH1 * N::P1_ROW + (L.ow >> 1) * N::C1;

    // chunk c of the warp row -> (image, offset in the image row); the warp stages images 2*quad and 2*quad+1.
    // Lanes past the last chunk repeat it (same load, same store, same value): no predicates, no divergence.
    int c_img[U::CPL], c_rem[U::CPL], c_sts[U::CPL];
#pragma unroll
    for (int i = 0; i < U::CPL; i++) {
        const int c = min(lane + 32 * i, U::NCH - 1);
        const int F0 = 2 * quad * N::ROWF + 4 * c;
        c_img[i] = F0 / N::ROWF;
        c_rem[i] = F0 % N::ROWF;
        c_sts[i] = 4 * c;
    }
    auto row_src = [&](int img0, int i) -> const float* {
        return P.in + (size_t)min(img0 + c_img[i], P.n - 1) * N::IMG + c_rem[i];
    };
    auto slot_img0 = [&](int lq, bool& valid) {
        const int pr = blockIdx.x + (lq >> 1) * gridDim.x;
        valid = pr < nPairs;
        return pr * kPair + (lq & 1) * kG;
    };

    // Output rows 2j, 2j+1 in two halves, so that the TMEM loads travel while the next pair is being built:
    // fetch = wait for the mma, start the loads; finish = max over the 2x2 window -> +bias, ReLU -> p1 row j.
    float e0[8], o0[8], e1[8], o1[8];
    auto epi_fetch = [&](int gq) {  // gq = running number of the output pair (slot * OP + j)
        const int dp = gq & (U::NDP - 1);
        mbar_wait(&d_full[dp], (gq / U::NDP) & 1);
        tc5_fence_after();
        const uint32_t td = tm_lane + U::TM_D + dp * 64;
        tmem_ld8(td, e0);
        tmem_ld8(td + 16, o0);
        tmem_ld8(td + 32, e1);
        tmem_ld8(td + 48, o1);
    };
    auto epi_finish = [&](int gq, int j, uint16_t* p1, int lq) {
        const int dp = gq & (U::NDP - 1);
        const bool dump = DBG && P.dbg && blockIdx.x == 0 && lq < 2;
        tmem_wait_ld();
        tc5_fence_before();
        mbar_arrive(&d_empty[dp]);
        float v[N::C1];
#pragma unroll
        for (int c = 0; c < N::C1; c++) v[c] = fmax_nan(fmax_nan(fmax_nan(e0[c], o0[c]), fmax_nan(e1[c], o1[c])) + bias[c], 0.f);
        uint16_t* dst = p1 + o_p1 + j * N::P1_ROW;  // shadow lanes hold the last live lane's values: same store
#pragma unroll
        for (int c = 0; c < N::C1; c += 2) *reinterpret_cast<uint32_t*>(dst + c) = Ops<T>::pack(v[c], v[c + 1]);
        if (dump) {
            float* d = P.dbg + lq * kG * N::PH1 * N::P1_ROW + o_dbg + j * N::P1_ROW;
#pragma unroll
            for (int c = 0; c < N::C1; c++) d[c] = v[c];
        }
    };
    float4 preA[2][U::CPL], preB[2][U::CPL];  // this team's next two row pairs
    const float* fp[U::CPL];                  // where the team's next fetch (two rows) starts
    int frow = 0;                             // ... and its row inside that slot
    int sd = 0;                               // slots finished
    uint16_t* pend_p1 = nullptr;              // p1 buffer of the slot whose last output pair is still owed
    for (int lq = 0;; lq++, sd++) {
        bool valid, n_valid, nn_valid;
        const int img0 = slot_img0(lq, valid);
        if (!valid) break;
        const int n_img0 = slot_img0(lq + 1, n_valid);
        const float* n_src[U::CPL];
#pragma unroll
        for (int i = 0; i < U::CPL; i++) n_src[i] = row_src(n_valid ? n_img0 : img0, i);
        if (sd == 0) {  // very first slot: fill both prefetch batches (pairs team, team + 2)
#pragma unroll
            for (int i = 0; i < U::CPL; i++) {
                const float* s0 = row_src(img0, i) + 2 * team * N::ROWF;
#pragma unroll
                for (int r = 0; r < 2; r++) {
                    preA[r][i] = ldg_stream(s0 + r * N::ROWF);
                    preB[r][i] = ldg_stream(s0 + (4 + r) * N::ROWF);
                }
                fp[i] = s0 + 8 * N::ROWF;
            }
            frow = 2 * team + 8;
        }
        {   // the two images this warp stages two slots from now -> L2 (24 KB | 6 KB contiguous, once per slot)
            const int nn_img0 = slot_img0(lq + 2, nn_valid);
            if (SIB_U_PF && nn_valid && nn_img0 + 2 * quad + 1 < P.n) {
                const char* base = reinterpret_cast<const char*>(P.in + (size_t)(nn_img0 + 2 * quad) * N::IMG);
                for (int k = lane * 128; k < 2 * N::IMG * 4; k += 32 * 128) prefetch_l2(base + k);
            }
        }
        const int buf = lq % N::NB1, use = lq / N::NB1;
        uint16_t* p1 = reinterpret_cast<uint16_t*>(smem + N::S_P1) + buf * N::P1_BUF;

        // one row pair: registers `pre` -> staged halves -> A_e | A_o of both rows in TMEM; refill `pre` 4 pairs ahead;
        // the epilogue of output pair ip-4 (if any) rides along
        auto build = [&](int ip, float4 (&pre)[2][U::CPL], bool epi) {
            const int gp = sd * U::HP + ip, slot = gp & (U::NAP - 1);
            if (SIB_U_EARLY_EPI && epi) epi_fetch(sd * U::OP + ip - 4);
            mbar_wait(&a_free[slot], ((gp / U::NAP) & 1) ^ 1);  // the mma that read this slot's previous tenant are done
            tc5_fence_after();
            __syncwarp();  // every lane has finished reading the staged rows of the previous pair
#pragma unroll
            for (int r = 0; r < 2; r++)
#pragma unroll
                for (int i = 0; i < U::CPL; i++) {
                    const float4 v = pre[r][i];
                    uint2 pk;
                    pk.x = Ops<T>::pack(v.x, v.y);
                    pk.y = Ops<T>::pack(v.z, v.w);
                    *reinterpret_cast<uint2*>(ring + r * U::UROW + c_sts[i]) = pk;
                }
#pragma unroll
            for (int i = 0; i < U::CPL; i++) {
#pragma unroll
                for (int r = 0; r < 2; r++) pre[r][i] = ldg_stream(fp[i] + r * N::ROWF);
                fp[i] += 4 * N::ROWF;
            }
            frow += 4;
            if (frow >= N::H) {  // the next fetch belongs to the next slot
                frow -= N::H;
#pragma unroll
                for (int i = 0; i < U::CPL; i++) fp[i] = n_src[i] + frow * N::ROWF;
            }
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 2; r++) {
                uint32_t ae[8], ao[8];
                const uint16_t* rr = ring + r * U::UROW + L.offE;
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    ae[k] = lds32(rr + 2 * k);
                    ao[k] = lds32(rr + kOdd + 2 * k);
                }
                tmem_st8(tm_lane + U::TM_A + slot * 32 + r * 16, ae);
                tmem_st8(tm_lane + U::TM_A + slot * 32 + r * 16 + 8, ao);
            }
            if (SIB_U_EARLY_EPI && epi) epi_finish(sd * U::OP + ip - 4, ip - 4, p1, lq);
            tmem_wait_st();
            tc5_fence_before();
            mbar_arrive(&a_full[slot]);
            if (!SIB_U_EARLY_EPI && epi) { epi_fetch(sd * U::OP + ip - 4); epi_finish(sd * U::OP + ip - 4, ip - 4, p1, lq); }
        };
        // p0 holds this slot's first pair of the team, p1r the second; each build refills its batch 4 pairs ahead
        auto run_slot = [&](float4 (&p0)[2][U::CPL], float4 (&p1r)[2][U::CPL]) {
            build(team, p0, false);
            if (sd > 0) {  // the previous slot's last output pair: its mma ran under the build above (no drain per slot)
                epi_fetch(sd * U::OP - 2 + team);
                epi_finish(sd * U::OP - 2 + team, U::OP - 2 + team, pend_p1, lq - 1);
                mbar_arrive(&p1_full[(lq - 1) % N::NB1]);
            }
            build(team + 2, p1r, false);
            mbar_wait(&p1_empty[buf], (use & 1) ^ 1);  // only the epilogues below touch p1
#pragma unroll 1
            for (int ip = team + 4; ip < U::HP; ip += 4) {
                build(ip, p0, true);
                if (ip + 2 < U::HP) build(ip + 2, p1r, true);
            }
        };
        // with HP % 4 == 2 a slot ends after an odd number of builds, so the two batches swap roles every slot
        if (U::HP % 4 == 2 && (sd & 1)) run_slot(preB, preA); else run_slot(preA, preB);
        pend_p1 = p1;  // the team's last output pair (OP-2+team) is finished by the next slot, or after the loop
    }
    if (sd > 0) {
        const int lq = sd - 1;
        epi_fetch(sd * U::OP - 2 + team);
        epi_finish(sd * U::OP - 2 + team, U::OP - 2 + team, pend_p1, lq);
        mbar_arrive(&p1_full[lq % N::NB1]);
    }
}

// ---- issuer warps (2): issuer w feeds the tensor core the output-row pairs of parity w --------------------------------
// (one warp issuing all 20 mma + commits of every pair was busy 80% of the time and set the pace: r01 profile)

template <typename N, typename T>
__device__ __forceinline__ void conv1u_issuer(const TcParams& P, unsigned char* smem, int w, uint32_t tmem) {
    using U = Umma<N>;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + N::S_BAR);
    uint64_t* a_full = bars + N::B_UA;
    uint64_t* a_free = bars + N::B_UA + U::NAP;
    uint64_t* d_full = bars + N::B_UDF;
    uint64_t* d_empty = bars + N::B_UDE;
    const uint32_t b_base = smem_u32(smem + N::S_B1U);  // [even|odd][kh] 16x16 halves each, 512 bytes apart
    const uint32_t idesc = umma_idesc_m128_n16<T>();
    const int nPairs = (P.n + kPair - 1) / kPair;
    auto wait_pair = [&](int gp) { mbar_wait(&a_full[gp & (U::NAP - 1)], (gp / U::NAP) & 1); };
    int sd = 0;
    for (int lq = 0;; lq++, sd++) {
        if ((int)(blockIdx.x + (lq >> 1) * gridDim.x) >= nPairs) break;
        for (int j = w; j < U::OP; j += 2) {
            const int gq = sd * U::OP + j, dp = gq & (U::NDP - 1);
            const int gp = sd * U::HP + j;  // input pair j of this slot; output rows 2j, 2j+1 read pairs j..j+2
            if (j < 2) wait_pair(gp);       // later: pair j was this issuer's j+2 one step ago
            wait_pair(gp + 1);              // the other team's pair: completion order across teams is not fixed
            wait_pair(gp + 2);
            mbar_wait(&d_empty[dp], ((gq / U::NDP) & 1) ^ 1);  // the previous tenant of this D pair has been read
            tc5_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int r = 0; r < 2; r++)
#pragma unroll
                    for (int t = 0; t < 2; t++)
#pragma unroll
                        for (int kh = 0; kh < 5; kh++) {
                            const int row = (2 * gp + r + kh) & (2 * U::NAP - 1);  // A ring position of input row 2j+r+kh
                            umma_ts(tmem + U::TM_D + dp * 64 + r * 32 + t * 16, tmem + U::TM_A + row * 16 + t * 8,
                                    umma_bdesc(b_base + (t * 5 + kh) * 512), idesc, kh > 0 ? 1u : 0u);
                        }
                umma_commit(&d_full[dp]);
                // Input pair p is read by the output pairs p-2, p-1, p. a_free[p] counts two arrivals, one per issuer,
                // each made after that issuer's last output pair that reads p: pair j itself and pair j+1 ...
                umma_commit(&a_free[gp & (U::NAP - 1)]);
                umma_commit(&a_free[(gp + 1) & (U::NAP - 1)]);
                // ... plus the pairs at the slot's ends that this issuer reads early or not at all
                if (j == 1) umma_commit(&a_free[(gp - 1) & (U::NAP - 1)]);
                if (j + 2 >= U::OP)
                    for (int p = j + 2; p < U::HP; p++) umma_commit(&a_free[(gp - j + p) & (U::NAP - 1)]);
            }
            __syncwarp();
        }
    }
}
