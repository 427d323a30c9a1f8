// sib_tc_conv1_umma.cuh -- conv1 + bias + ReLU + 2x2 pool on the 5th-gen tensor cores (tcgen05.mma, operands in TMEM).
// Included by sib_kernels_tc.cu (namespace sib::{anonymous}) when SIB_CONV1_UMMA is on.
//
// Why: on the mma.sync path conv1 is 245 of the network's 402 HMMA per image and the legacy tensor path tops out
// at one m16n8k16 per 8.5 clk per sub-partition. tcgen05.mma M=128 N=16 K=16 with the A operand in TMEM retires in
// 9 clk (tools/tcgen05_probe.cu), 3.8x the MAC rate, and its im2col rows can be written from registers, so the
// operand still never exists in shared memory.
//
// Shape. A warpgroup (4 warps = the 4 TMEM lane quadrants) owns one 8-image slot. Thread q owns the output columns
// 2q and 2q+1 of the slot's column space (224 | 192 columns a row), i.e. TMEM lane q of an "even" and an "odd"
// M=128 tile (odd columns read their 16 K values one element early against filters shifted by one: every
// shared-memory read stays 4-byte aligned, and the 2x2 pool never leaves the thread).
//   per input row h:  stage the row (fp32 -> half, warp-private)  ->  LDS.32 -> 2 x tcgen05.st (A_e(h), A_o(h))
//                     -> arrive on a_full[row % 8]
//   issuer warp    :  once rows oh..oh+4 are in TMEM: D_t(oh) = sum_kh A_t(oh+kh) . B_t,kh  (10 mma, t = even|odd)
//                     -> tcgen05.commit -> d_full[oh % 4]
//   kLag rows later:  tcgen05.ld D_e, D_o -> max (column pair) -> max with the previous row -> +bias, ReLU -> p1
// The epilogue trails the A rows by kLag = 2 rows so that the issuer's wake-up + 10 mma + commit (about 250 clk) runs
// under the next row's staging instead of being waited for; its tcgen05.ld is issued first in a row and collected
// after the row's tcgen05.st, so the two TMEM round trips overlap. A warp holds the column pairs of exactly two images
// (28 | 24 live lanes): with one image seam per warp the 3-words-per-lane shared-memory reads never share a bank.
// TMEM per warpgroup: 8 input rows x (8 + 8) columns + 4 output rows x (16 + 16) columns = 256 of the 512 columns.
#pragma once

template <typename N>
struct Umma {
    static constexpr int NPOS = kG * N::OW1, NPAIR = NPOS / 2;       // output columns / column pairs per slot row
    static constexpr int PPW = NPAIR / 4;                             // pairs per warp = the pairs of two images
    static constexpr int PXW = 2 * N::W;                              // input pixels a warp stages per row: two image rows
    static constexpr int NCH = PXW * N::C / 4;                        // 16-byte chunks of a staged warp row
    static constexpr int CPL = (NCH + 31) / 32;                       // chunks per lane
    static constexpr int UROW = PXW * N::C + 16;                      // halves per staged row incl. zero slack
    static constexpr int NA = 8, ND = 4, kLag = 2;                    // A rows / D rows resident in TMEM, epilogue lag
    static constexpr int TM_WG = 256, TM_A = 0, TM_D = NA * 16;       // TMEM columns per warpgroup
    static_assert(NPAIR % 4 == 0 && PPW <= 32 && PPW == N::OW1 && UROW % 8 == 0 && UROW == N::URING_ROW, "umma tiling");
    static_assert(N::H % 4 == 0 && N::OH1 % 4 == 0 && kBR == 2 && kLag <= NA - 5 && 2 * TM_WG <= 512, "umma pipeline");
};

// ---- worker warps (2 warpgroups) ------------------------------------------------------------------------------------

template <typename N, typename T, bool DBG>
__device__ __forceinline__ void conv1u_worker(const TcParams& P, unsigned char* smem, int wg, int quad, int lane, uint32_t tmem) {
    using U = Umma<N>;
    uint16_t* ring = reinterpret_cast<uint16_t*>(smem + N::S_RING) + (wg * 4 + quad) * 2 * U::UROW;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + N::S_BAR);
    uint64_t* p1_full = bars + N::B_P1F;
    uint64_t* p1_empty = bars + N::B_P1E;
    uint64_t* a_full = bars + N::B_UA + wg * U::NA;
    uint64_t* d_full = bars + N::B_UDF + wg * U::ND;
    uint64_t* d_empty = bars + N::B_UDE + wg * U::ND;
    const float* biasp = reinterpret_cast<const float*>(P.wpack + N::W_BIAS);
    float bias[N::C1];
#pragma unroll
    for (int c = 0; c < N::C1; c++) bias[c] = biasp[c];
    const uint32_t tm_lane = tmem + ((uint32_t)(quad * 32) << 16) + wg * U::TM_WG;  // this warp's lane quadrant, WG's columns
    const int nPairs = (P.n + kPair - 1) / kPair;

    // this lane's column pair (lanes >= PPW shadow the last live lane; their tile rows are never read back)
    const bool live = lane < U::PPW;
    const int q = U::PPW * quad + min(lane, U::PPW - 1);
    const int pe = 2 * q;
    const int img_l = pe / N::OW1, ow = pe % N::OW1;
    const int offE = ((img_l - 2 * quad) * N::W + ow) * N::C;  // even column: halves offE .. offE+15 of the staged row
    constexpr int kOdd = N::C - 1;                               // odd column: one element early (see header)
    const int o_p1 = img_l * N::P1_IMG + (ow >> 1) * N::C1;
    const int o_dbg = img_l * N::PH1 * N::P1_ROW + (ow >> 1) * N::C1;

    // chunk c of the warp row -> (image, offset in the image row); the warp stages images 2*quad and 2*quad+1
    int c_img[U::CPL], c_rem[U::CPL];
    bool c_on[U::CPL];
#pragma unroll
    for (int i = 0; i < U::CPL; i++) {
        const int c = lane + 32 * i;
        c_on[i] = c < U::NCH;
        const int F0 = 2 * quad * N::ROWF + 4 * min(c, U::NCH - 1);
        c_img[i] = F0 / N::ROWF;
        c_rem[i] = F0 % N::ROWF;
    }
    auto row_src = [&](int img0, int i) -> const float* {
        return P.in + (size_t)min(img0 + c_img[i], P.n - 1) * N::IMG + c_rem[i];
    };

    float4 preA[kBR][U::CPL], preB[kBR][U::CPL];
    int sd = 0;  // slots this warpgroup has finished
    for (int lq = wg;; lq += 2, sd++) {
        const int pair = blockIdx.x + (lq >> 1) * gridDim.x;
        if (pair >= nPairs) break;
        const int img0 = pair * kPair + (lq & 1) * kG;
        const int n_pair = blockIdx.x + ((lq + 2) >> 1) * gridDim.x;
        const bool n_valid = n_pair < nPairs;
        const int n_img0 = n_pair * kPair + ((lq + 2) & 1) * kG;
        const float* src[U::CPL];
        const float* n_src[U::CPL];
#pragma unroll
        for (int i = 0; i < U::CPL; i++) {
            src[i] = row_src(img0, i);
            n_src[i] = n_valid ? row_src(n_img0, i) : src[i];
        }
        auto row_ptr = [&](int i, int row) { return row < N::H ? src[i] + row * N::ROWF : n_src[i] + (row - N::H) * N::ROWF; };
        if (sd == 0) {  // very first slot: fill both prefetch batches
#pragma unroll
            for (int r = 0; r < kBR; r++)
#pragma unroll
                for (int i = 0; i < U::CPL; i++) {
                    preA[r][i] = ldg_stream(row_ptr(i, r));
                    preB[r][i] = ldg_stream(row_ptr(i, kBR + r));
                }
        }
        const int buf = lq % N::NB1, use = lq / N::NB1;
        uint16_t* p1 = reinterpret_cast<uint16_t*>(smem + N::S_P1) + buf * N::P1_BUF;
        const bool dump = DBG && P.dbg && blockIdx.x == 0 && lq < 2;
        mbar_wait(&p1_empty[buf], (use & 1) ^ 1);
        const int rbase = sd * N::H, gbase = sd * N::OH1;   // running A-row / D-row numbers (both multiples of 4)

        float prev[N::C1];
        // the epilogue of output row oh in two halves: fetch (wait for the mma, start the TMEM loads) and finish
        auto epi_fetch = [&](int oh, int b, float (&de)[8], float (&dod)[8]) {
            mbar_wait(&d_full[b], ((gbase + oh) >> 2) & 1);
            tc5_fence_after();
            tmem_ld8(tm_lane + U::TM_D + b * 32, de);
            tmem_ld8(tm_lane + U::TM_D + b * 32 + 16, dod);
        };
        auto epi_finish = [&](int oh, int b, bool odd, const float (&de)[8], const float (&dod)[8]) {
            tmem_wait_ld();
            tc5_fence_before();
            mbar_arrive(&d_empty[b]);
            if (!odd) {
#pragma unroll
                for (int c = 0; c < N::C1; c++) prev[c] = fmaxf(de[c], dod[c]);
            } else {
                float v[N::C1];
#pragma unroll
                for (int c = 0; c < N::C1; c++) v[c] = fmaxf(fmaxf(fmaxf(de[c], dod[c]), prev[c]) + bias[c], 0.f);
                if (live) {
                    uint16_t* dst = p1 + o_p1 + (oh >> 1) * N::P1_ROW;
#pragma unroll
                    for (int c = 0; c < N::C1; c += 2) *reinterpret_cast<uint32_t*>(dst + c) = Ops<T>::pack(v[c], v[c + 1]);
                    if (dump) {
                        float* d = P.dbg + lq * kG * N::PH1 * N::P1_ROW + o_dbg + (oh >> 1) * N::P1_ROW;
#pragma unroll
                        for (int c = 0; c < N::C1; c++) d[c] = v[c];
                    }
                }
            }
        };

#pragma unroll 1
        for (int h0 = 0; h0 < N::H; h0 += 4) {
#pragma unroll
            for (int j = 0; j < 4; j++) {  // everything that depends on h % 4 is a compile-time constant below
                const int h = h0 + j;
                const int oh = h - 4 - U::kLag;
                const bool epi = oh >= 0;
                constexpr int kL4 = U::kLag & 3;
                const int b = (j + 4 - kL4) & 3;            // = oh % 4 (H, OH1 and 4 are all multiples of 4)
                const bool odd = ((j + U::kLag) & 1) != 0;  // = oh % 2
                float de[8], dod[8];
                if (epi) epi_fetch(oh, b, de, dod);
                // ---- stage input row h (warp-private): fp32 chunks from the prefetch registers -> half
                uint16_t* rr = ring + (j & 1) * U::UROW;
#pragma unroll
                for (int i = 0; i < U::CPL; i++) {
                    const float4 v = ((j / kBR) & 1) ? preB[j % kBR][i] : preA[j % kBR][i];
                    if (c_on[i]) {
                        uint2 pk;
                        pk.x = Ops<T>::pack(v.x, v.y);
                        pk.y = Ops<T>::pack(v.z, v.w);
                        *reinterpret_cast<uint2*>(rr + 4 * (lane + 32 * i)) = pk;
                    }
                }
                if (j % kBR == kBR - 1) {  // batch drained: fetch the batch two ahead (may belong to the next slot)
#pragma unroll
                    for (int jj = 0; jj < kBR; jj++)
#pragma unroll
                        for (int i = 0; i < U::CPL; i++) {
                            const float4 nv = ldg_stream(row_ptr(i, h + kBR + 1 + jj));
                            if ((j / kBR) & 1) preB[jj][i] = nv; else preA[jj][i] = nv;
                        }
                }
                if ((lane & 7) == 0 && c_on[0] && (h + 16 < N::H || n_valid)) prefetch_l2(row_ptr(0, h + 16));
                __syncwarp();
                // ---- im2col rows of this lane's two columns -> TMEM (A_e(h) | A_o(h))
                uint32_t ae[8], ao[8];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    ae[k] = lds32(rr + offE + 2 * k);
                    ao[k] = lds32(rr + offE + kOdd + 2 * k);
                }
                const int slot = (rbase + h) & (U::NA - 1);
                tmem_st8(tm_lane + U::TM_A + slot * 16, ae);
                tmem_st8(tm_lane + U::TM_A + slot * 16 + 8, ao);
                if (epi) epi_finish(oh, b, odd, de, dod);
                tmem_wait_st();
                tc5_fence_before();
                mbar_arrive(&a_full[slot]);
            }
        }
#pragma unroll
        for (int oh = N::OH1 - U::kLag; oh < N::OH1; oh++) {  // the slot's last output rows
            float de[8], dod[8];
            epi_fetch(oh, oh & 3, de, dod);
            epi_finish(oh, oh & 3, (oh & 1) != 0, de, dod);
        }
        mbar_arrive(&p1_full[buf]);
    }
}

// ---- issuer warp (one per warpgroup): the only thread that touches tcgen05.mma -----------------------------------------

template <typename N, typename T>
__device__ __forceinline__ void conv1u_issuer(const TcParams& P, unsigned char* smem, int wg, uint32_t tmem) {
    using U = Umma<N>;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + N::S_BAR);
    uint64_t* a_full = bars + N::B_UA + wg * U::NA;
    uint64_t* d_full = bars + N::B_UDF + wg * U::ND;
    uint64_t* d_empty = bars + N::B_UDE + wg * U::ND;
    const uint32_t b_base = smem_u32(smem + N::S_B1U);  // [even|odd][kh] 16x16 halves each, 512 bytes apart
    const uint32_t idesc = umma_idesc_m128_n16<T>();
    const uint32_t tm = tmem + wg * U::TM_WG;
    const int nPairs = (P.n + kPair - 1) / kPair;
    int sd = 0;
    for (int lq = wg;; lq += 2, sd++) {
        if ((int)(blockIdx.x + (lq >> 1) * gridDim.x) >= nPairs) break;
        for (int h = 4; h < N::H; h++) {
            const int oh = h - 4, gd = sd * N::OH1 + oh, b = gd & (U::ND - 1);
            const int Rg = sd * N::H + h;
            mbar_wait(&a_full[Rg & (U::NA - 1)], (Rg / U::NA) & 1);  // row h is in TMEM (arrivals are in row order)
            mbar_wait(&d_empty[b], ((gd / U::ND) & 1) ^ 1);          // the previous tenant of this D buffer has been read
            tc5_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int t = 0; t < 2; t++)
#pragma unroll
                    for (int kh = 0; kh < 5; kh++) {
                        const int slot = (Rg - 4 + kh) & (U::NA - 1);
                        umma_ts(tm + U::TM_D + b * 32 + t * 16, tm + U::TM_A + slot * 16 + t * 8,
                                umma_bdesc(b_base + (t * 5 + kh) * 512), idesc, kh > 0 ? 1u : 0u);
                    }
                umma_commit(&d_full[b]);
            }
            __syncwarp();
        }
    }
}
