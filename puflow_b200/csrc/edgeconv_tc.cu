// Fused EdgeConv block on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM).
//
// Same contract as pf_edgeconv (edgeconv.cu; replaces FeatureExtractUnit.forward,
// /root/reference/modules/discrete/interpflow.py:223-248) for K in {8, 16, 32} and (G, C_out) in {(8, 32), (16, 64),
// (32, 128)} -- all six blocks of the path; ncu shows the FFMA version issue-bound
// (profiles/r01_ncu_edgeconv_v0_ffma.md): the hidden-channel contractions of the dense block
// (6 G^2 + 4 G C_out MACs per edge) run as tcgen05.mma.kind::f16 with an error-compensated fp16 split so the result
// stays fp32-class:
//     x = hi + lo (two fp16, |x - hi - lo| <~ 2^-22 |x|),   a.w ~= a_hi w_hi + a_lo w_hi + a_hi w_lo   (fp32 accumulate)
// The dropped a_lo w_lo term is 2^-22 relative.  Tolerance stated in tests/test_gpu_kernels.py.
//
// One persistent CTA per SM, three kinds of warps (round-2 structure; round 1 had the epilogue warps issue their own
// gathers, 3,000 instructions per warp per tile of which a third were cp.async / LDG address arithmetic and waits):
//   NWG epilogue warpgroups (128 threads)  one 128-edge tile (128 / K points x K neighbours) each; CUDA-core work only:
//                                           P_i + Q_j, LeakyReLU, fp16 hi/lo split, max over the K neighbours
//   NWG MMA warps                          one elected thread issues the tcgen05.mma of "its" warpgroup's tile
//   NWG producer warps                     fetch every gathered operand of the tile (neighbour terms Q_j, centre terms P_i)
//                                           with cp.async into a 2-slot ring of shared memory, one stage ahead of use
// The three talk through mbarriers only (no block or named barriers): ring slot full (cp.async.mbarrier.arrive) / empty,
// operands ready (epilogue -> MMA), MMA done (tcgen05.commit).  The warpgroups run independent tiles, so one tile's MMAs
// and TMEM round trips hide behind the other's CUDA-core work.  Per tile (9 ring stages for (32, 128)):
//   layer 0      h0 = lrelu(P0_i + Q0_j)                                  (gather only)
//   layers 1..3  every new 32-channel slice h_m (fp16 hi/lo) goes to tensor memory as the A operand of ONE wide MMA
//                group that adds its contribution to ALL later layers at once:
//                    D[128 edges x 32 (3 - m)] += h_m . [W_{m+1,m}; ..; W_{3,m}]^T      (M=128, N=96/64/32, K=32)
//                so layer l's accumulator is complete right after h_{l-1}'s group (K = 32 on the critical path, not
//                32 l) and no dense-layer MMA reads its A operand from shared memory: ncu showed the kernel bound by
//                the L1TEX data pipe -- 2,200 tensor-core + 2,800 LSU shared-memory wavefronts per tile against 7,000
//                cycles -- and the re-read hi/lo activations were half of the tensor-core share.
//                -> TMEM -> +P+Q, lrelu -> fp16 hi/lo -> TMEM slice (dense) and K-slice of H in shared memory (conv_out)
//   conv_out     D^T[128 ch x 128 edges] = Wo[128 x 128] . H^T  (M=128, N=128, K=128): transposed so that a
//                thread owns one output channel and the max over the K neighbours is a register max; the neighbour
//                term is read from the ring (rows = edges, lanes = channels: conflict-free) and the output store is
//                coalesced.  Wo (hi/lo) lives in TMEM as the A operand (tcgen05.mma with A from tensor memory).
// TMA was measured for the gathers and rejected: cp.async.bulk.tensor ... tile::gather4 sustains one 128-byte row per
// ~18 cycles per SM (tools/probes/gather4_probe.cu: 7 B/cycle/SM; 33 B/cycle/SM even with 1 KB rows), a tile needs 640.
// Weights are pre-split and pre-laid-out by puflow_b200/packing.py (pack_edgeconv_tc): the dense-layer weights
// arrive with one TMA bulk copy per CTA, Wo is written to TMEM once per CTA with tcgen05.st.
#include "common.cuh"
#include "puflow_b200.h"
#include "tc.cuh"

// Template parameters: G growth rate, CO output channels, NWG tiles in flight per CTA, KN neighbours per point.
// GP = max(G, 16) is the K-slice width of one dense layer inside the activation operand (tcgen05 needs K % 16 == 0 and
// N % 16 == 0): for G = 8 the upper half of every slice is zero in both operands.
// For CO < 128 the transposed conv_out still runs with M = 128: row r of the A operand is Wo[r % CO], and thread r
// takes channel r % CO for the edges [CO * (r / CO), CO * (r / CO + 1)) of the tile, so no epilogue thread idles.
// Text below uses the (32, 128, K = 16) numbers.

#ifdef PF_PROF
__device__ long long g_prof[32];
#endif

namespace {

using namespace pf;

constexpr int TM = 128;                         // edges per tile

constexpr int ilog2(int v) { return v <= 1 ? 0 : 1 + ilog2(v / 2); }

template <int G, int CO, int NWG, int KN>
struct Cfg {
    static constexpr int GP = G < 16 ? 16 : G;
    static constexpr int PT = TM / KN;                                   // points per tile
    static constexpr int R = 4 * G + CO, R2 = 2 * R;
    static constexpr int KT = 4 * GP;                                    // K width of the activation operand
    static constexpr int NT = NWG * 128 + NWG * 32 + NWG * 32;           // epilogue | MMA warps | producer warps
    static constexpr int CPR = G / 4, LCPR = ilog2(CPR);                 // 16-byte chunks per dense Q row
    static constexpr int NCH = CO / 32;                                  // conv_out steps (32 edges per thread each)
    static constexpr int STAGES = 4 + NCH;                               // ring stages per tile
    // conv_out MMA groups: the four 32-edge column blocks of D^T are issued OW at a time (one commit each), NQ groups per tile
    static constexpr int OW = NCH == 4 ? 1 : 4, NQ = 4 / OW;
    // ---- shared memory map (bytes) ----
    static constexpr uint32_t WD_BYTES = 24u * GP * GP;                  // W1..W3, hi and lo
    static constexpr uint32_t WO_PART = 128u * KT * 2;                   // blob: [Wd | Wo hi [128][KT/2 words] | Wo lo]
    static constexpr uint32_t BLOB = WD_BYTES + 2 * WO_PART;
    static constexpr uint32_t A_OFF = WD_BYTES, A_PART = 128u * KT * 2;  // A[wg][hi|lo]
    // ring slot = neighbour part (dense stage: 128 rows x 4G bytes; conv_out step: 4096 / CO rows x 4 CO bytes = 16 KB)
    //           + centre part (the P columns of the stage for the tile's points: 16 KB / KN)
    static constexpr uint32_t QPART = 16384, PPART = 16384 / KN, SLOT = QPART + PPART;
    static constexpr uint32_t QS_OFF = A_OFF + 2 * NWG * A_PART;
    static constexpr uint32_t BAR_OFF = QS_OFF + 2 * NWG * SLOT;
    static constexpr int NBAR = 1 + 10 * NWG;                            // weights | per wg: mma, ready, full[2], empty[2], conv[4]
    static constexpr uint32_t SMEM_BYTES = BAR_OFF + 8 * NBAR + 16;
    static constexpr uint32_t A_SBO = KT * 16;
    // ---- tensor memory map (columns): per warpgroup 128 (conv_out accumulator; the dense accumulators of layers 1..3
    // alias its first 3 GP columns -- never live together) + GP (the current activation slice, fp16 pairs: hi | lo, the
    // A operand of the dense MMAs), then Wo hi / lo as the A operand of conv_out ----
    static constexpr uint32_t TM_SLICE = 128, TM_WG = 128 + GP, TM_WO_HI = NWG * TM_WG, TM_WO_LO = TM_WO_HI + KT / 2,
                              TM_COLS = 512;
    static_assert(TM_WO_LO + KT / 2 <= 512, "tensor memory budget");
    static_assert(SMEM_BYTES <= 232448, "shared memory budget");
    static_assert(KN == 8 || KN == 16 || KN == 32, "K in {8, 16, 32}");
    // dense weights by K-slice m = 0..2: rows = the GP (3 - m) output channels of layers m+1..3, K = GP; [hi | lo]
    __host__ __device__ static constexpr uint32_t wd_part(int m) { return 2u * GP * GP * (3 - m); }
    __host__ __device__ static constexpr uint32_t wd_off(int m) { return 4u * GP * GP * (3 * m - m * (m - 1) / 2); }
};

template <int G, int CO, int NWG, int KN>
__global__ void __launch_bounds__(Cfg<G, CO, NWG, KN>::NT, 1)
    edgeconv_tc_kernel(const float* __restrict__ pq, const int32_t* __restrict__ idx, const uint8_t* __restrict__ wblob,
                       float* __restrict__ out, uint8_t* __restrict__ out_split, int N, int tiles_per_patch,
                       long long total_tiles, float slope, int prec) {
    using C = Cfg<G, CO, NWG, KN>;
    constexpr int GP = C::GP, R = C::R, R2 = C::R2, KT = C::KT, CPR = C::CPR, LCPR = C::LCPR, NCH = C::NCH, PT = C::PT;
    constexpr int OW = C::OW, NQ = C::NQ;
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + C::BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wbar + C::NBAR);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int role = warp < 4 * NWG ? 0 : (warp < 5 * NWG ? 1 : 2);       // epilogue | MMA | producer
    const int wg = role == 0 ? warp >> 2 : (role == 1 ? warp - 4 * NWG : warp - 5 * NWG);
    const int wq = warp & 3;                                              // TMEM lane quadrant (epilogue warps)
    const int e = tid & 127;                                              // tile row (edge) / conv_out row of this thread
    uint64_t* mbar = wbar + 1 + 10 * wg;                                  // dense MMA group done
    uint64_t* ready = mbar + 1;                                           // operands ready (one arrival per epilogue warp)
    uint64_t* qfull = mbar + 2;                                           // [2] ring slot filled (32 cp.async arrivals)
    uint64_t* qempty = mbar + 4;                                          // [2] ring slot consumed (one arrival per warp)
    uint64_t* cbar = mbar + 6;                                            // [4] conv_out step c done (one barrier per step: all
                                                                          // steps are committed back to back, a waiter must
                                                                          // never fall two phases behind one barrier)

    if (tid == 0) {
        mbar_init(wbar, 1);
        for (int i = 0; i < NWG; i++) {
            uint64_t* b = wbar + 1 + 10 * i;
            mbar_init(b, 1), mbar_init(b + 1, 4);
            mbar_init(b + 2, 32), mbar_init(b + 3, 32), mbar_init(b + 4, 4), mbar_init(b + 5, 4);
            for (int c = 0; c < 4; c++) mbar_init(b + 6 + c, 1);
        }
        mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, C::TM_COLS);
    if (G < GP) {   // the unused upper half of every K slice must read as zero
        for (uint32_t i = tid; i < 2 * NWG * C::A_PART / 16; i += C::NT)
            reinterpret_cast<uint4*>(smem + C::A_OFF)[i] = make_uint4(0, 0, 0, 0);
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t lane_sel = (uint32_t)(wq * 32) << 16;
    if (tid == 0) {
        mbar_expect_tx(wbar, C::WD_BYTES);
        bulk_g2s_chunked(smem, wblob, C::WD_BYTES, wbar);
    }
    if (warp < 4) {   // Wo -> TMEM: thread = lane = conv_out row, two fp16 per column, KT / 2 columns per part
#pragma unroll 1
        for (int part = 0; part < 2; part++)
#pragma unroll 1
            for (int c32 = 0; c32 < KT / 64; c32++) {
                const uint4* src = reinterpret_cast<const uint4*>(wblob + C::WD_BYTES + (size_t)part * C::WO_PART +
                                                                  (size_t)e * (KT * 2) + c32 * 128);
                uint32_t r[32];
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const uint4 v = __ldg(src + i);
                    r[4 * i] = v.x, r[4 * i + 1] = v.y, r[4 * i + 2] = v.z, r[4 * i + 3] = v.w;
                }
                tc::st32(tmem_base + lane_sel + (part ? C::TM_WO_LO : C::TM_WO_HI) + 32 * c32, r);
            }
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();

    const uint32_t a_hi = sbase + C::A_OFF + (uint32_t)wg * 2 * C::A_PART, a_lo = a_hi + C::A_PART;
    const uint32_t ring = sbase + C::QS_OFF + (uint32_t)wg * 2 * C::SLOT;
    const uint8_t* ring_p = smem + C::QS_OFF + (size_t)wg * 2 * C::SLOT;
    const uint32_t d_out = tmem_base + (uint32_t)wg * C::TM_WG;
    const uint32_t d_dense = d_out;                                       // aliases the first GP columns of d_out
    auto tile_coords = [&](long long t, int& patch, int& p0) {
        const uint32_t tt = (uint32_t)t, tpp = (uint32_t)tiles_per_patch;
        patch = (int)(tt / tpp);
        p0 = (int)(tt - (uint32_t)patch * tpp) * PT;
    };
    const long long stride = (long long)gridDim.x * NWG;
    long long t = (long long)blockIdx.x * NWG + wg;

    if (role == 1) {
        // ================= MMA-issuing warp of warpgroup wg: one elected thread (elect_one, not lane == 0: tc.cuh) waits
        // for "operands ready", issues, commits =================
        const uint32_t idesc_o = tc::idesc_f16(128, 32 * OW);
        const bool split3 = prec == PF_PREC_FP32_SPLIT;     // PF_PREC_FP16: the hi . hi products alone
        const uint64_t da_hi = tc::smem_desc(a_hi, C::A_SBO), da_lo = tc::smem_desc(a_lo, C::A_SBO);
        // D[128 edges x GP (3 - m)] (+)= h_m . [W_{m+1,m}; ..; W_{3,m}]^T, h_m from tensor memory (8 columns per K-step).
        // 3 split products, the two small cross terms first: the accumulator adder truncates, so the hi . hi pass -- the
        // additions at full magnitude -- comes last.
        auto issue_group = [&](int m) {
            const uint32_t wl = sbase + C::wd_off(m);
            const uint64_t db_hi = tc::smem_desc(wl, 16u * GP), db_lo = tc::smem_desc(wl + C::wd_part(m), 16u * GP);
            const uint32_t idesc = tc::idesc_f16(128, GP * (3 - m));
            const uint32_t d = d_dense + GP * m, h_hi = d_out + C::TM_SLICE, h_lo = h_hi + GP / 2;
            if (split3) {
#pragma unroll
                for (int s = 0; s < GP / 16; s++) tc::mma_f16_ts(d, h_lo + 8u * s, db_hi + 16u * s, idesc, m > 0 || s > 0);
#pragma unroll
                for (int s = 0; s < GP / 16; s++) tc::mma_f16_ts(d, h_hi + 8u * s, db_lo + 16u * s, idesc, 1);
            }
#pragma unroll
            for (int s = 0; s < GP / 16; s++) tc::mma_f16_ts(d, h_hi + 8u * s, db_hi + 16u * s, idesc, split3 || m > 0 || s > 0);
        };
        // conv_out group q: D^T[128 rows x 32 OW edges] = Wo . H^T for the column blocks [OW q, OW q + OW) (32 edges each;
        // epilogue step c of edge group g reads block g NCH + c); Wo from TMEM (8 columns per K-step), H^T rows 32 cb ..
        // of the shared-memory operand (4 row groups = 4 SBO further per block).  One commit per group.  (32, 128): four
        // N = 32 groups, the epilogue of step c runs while the tensor pipe computes step c + 1 (measured: one N = 128 chain
        // 1,500 cycles of waiting per tile; two N = 64 groups 3.22 ms against 2.91).  C_out < 128: a step reads blocks that
        // are not adjacent, so the whole tile is ONE N = 128 group (a quarter of the instructions: (8, 32) 1.04 -> 0.96 ms).
        auto issue_out = [&](int q) {
            const uint32_t wo_hi = tmem_base + C::TM_WO_HI, wo_lo = tmem_base + C::TM_WO_LO;
            const uint32_t cb = (uint32_t)(q * OW), d = d_out + 32u * cb;
            const uint64_t bh = da_hi + (uint64_t)(4u * cb * (C::A_SBO >> 4)), bl = da_lo + (uint64_t)(4u * cb * (C::A_SBO >> 4));
            if (split3) {
#pragma unroll
                for (int s = 0; s < KT / 16; s++) tc::mma_f16_ts(d, wo_lo + 8u * s, bh + 16u * s, idesc_o, s > 0);
#pragma unroll
                for (int s = 0; s < KT / 16; s++) tc::mma_f16_ts(d, wo_hi + 8u * s, bl + 16u * s, idesc_o, 1);
            }
#pragma unroll
            for (int s = 0; s < KT / 16; s++) tc::mma_f16_ts(d, wo_hi + 8u * s, bh + 16u * s, idesc_o, split3 || s > 0);
        };
        tc::mbar_wait_bounded(wbar, 0);   // dense-layer weights have landed
        if (tc::elect_one()) {
            uint32_t rphase = 0;
            PROF_DECL
            for (; t < total_tiles; t += stride) {
#pragma unroll
                for (int m = 0; m < 3; m++) {
                    PROF_TICK(1)
                    tc::mbar_wait_bounded(ready, rphase);
                    rphase ^= 1;
                    tc::fence_after_sync();
                    PROF_TICK(0)
                    issue_group(m);
                    tc::commit(mbar);
                }
                PROF_TICK(1)
                tc::mbar_wait_bounded(ready, rphase);
                rphase ^= 1;
                tc::fence_after_sync();
                PROF_TICK(0)
#pragma unroll
                for (int q = 0; q < NQ; q++) {
                    issue_out(q);
                    tc::commit(cbar + q);
                }
            }
            PROF_TICK(1)
            if (wg == 0) { PROF_DUMP(16, 2) }
        }
        __syncwarp();
    } else if (role == 2) {
        // ================= producer warp of warpgroup wg: every gathered operand of the tile, one ring stage ahead ======
        // Dense stage l: neighbour rows Q[j][l G .. l G + G) of the 128 edges (row e at e * 4G bytes, 16-byte chunk c at
        // position c ^ swz(e): conflict-free LDS.128 per thread-row) + the tile's PT centre rows P[i][l G ..).
        // conv_out step s: Q[j][4G .. 4G + CO) of the 4096 / CO edges the epilogue threads read in that step (slot row
        // 32 grp + u <- edge CO grp + 32 s + u; chunk c at c ^ ((row & 3) << LCPR) so that the 32 cp.async of one
        // instruction spread over all banks) + P[i][4G ..) of their points.
        // Lane: chunk c0 = lane % CPR of rows rl + RPP it (it < ROWS): the same ROWS neighbours in every stage, so their
        // row pointers are computed once per tile; the indices of the NEXT tile are fetched while this one streams.
        constexpr int ROWS = 4 * CPR, RPP = 32 / CPR;
        const int c0 = lane % CPR, rl = lane / CPR;
        const float* rowp[ROWS];
        int jn[ROWS];
        auto load_j = [&](long long tt) {
            int patch, p0;
            tile_coords(tt, patch, p0);
#pragma unroll
            for (int it = 0; it < ROWS; it++) {
                const int ed = rl + RPP * it;
                const int pi = min(p0 + ed / KN, N - 1);
                jn[it] = __ldg(idx + ((size_t)patch * N + pi) * KN + (ed % KN));
            }
        };
        if (t < total_tiles) load_j(t);
        uint32_t st = 0;                                                  // ring stage counter: slot st & 1, phase (st >> 1) & 1
        for (; t < total_tiles; t += stride) {
            int patch, p0;
            tile_coords(t, patch, p0);
            const float* ppq = pq + (size_t)patch * N * R2;
#pragma unroll
            for (int it = 0; it < ROWS; it++) rowp[it] = ppq + (size_t)min(max(jn[it], 0), N - 1) * R2 + R + c0 * 4;
            if (t + stride < total_tiles) load_j(t + stride);
            // ---- dense stages ----
#pragma unroll
            for (int l = 0; l < 4; l++, st++) {
                const uint32_t slot = ring + (st & 1) * C::SLOT;
                tc::mbar_wait_bounded(qempty + (st & 1), ((st >> 1) & 1) ^ 1);
#pragma unroll
                for (int it = 0; it < ROWS; it++) {
                    const int ed = rl + RPP * it;
                    const int swz = ((ed * CPR) >> 3) & (CPR - 1);
                    tc::cp_async16(slot + ed * (16 * CPR) + ((c0 ^ swz) << 4), rowp[it] + l * G);
                }
#pragma unroll
                for (int k = 0; k < (PT * CPR + 31) / 32; k++) {
                    const int o = lane + 32 * k, pt = o / CPR, c = o % CPR;
                    if ((PT * CPR) % 32 == 0 || o < PT * CPR)
                        tc::cp_async16(slot + C::QPART + pt * (16 * CPR) + ((c ^ (pt & (CPR - 1))) << 4),
                                       ppq + (size_t)min(p0 + pt, N - 1) * R2 + l * G + c * 4);
                }
                tc::cp_async_arrive_noinc(qfull + (st & 1));
            }
            // ---- conv_out steps ----
#pragma unroll
            for (int s = 0; s < NCH; s++, st++) {
                const uint32_t slot = ring + (st & 1) * C::SLOT;
                tc::mbar_wait_bounded(qempty + (st & 1), ((st >> 1) & 1) ^ 1);
#pragma unroll
                for (int it = 0; it < ROWS; it++) {
                    if ((it / CPR) % NCH != s) continue;                  // edges of another step
                    const int grp = (it / CPR) / NCH;
                    const int sr = 32 * grp + rl + RPP * (it % CPR);      // slot row
                    const int sw = (rl & 3) << LCPR;                      // sr & 3 == rl & 3 (RPP % 4 == 0)
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        tc::cp_async16(slot + sr * (4 * CO) + (((c0 + CPR * q) ^ sw) << 4), rowp[it] + 4 * G + CPR * q * 4);
                }
                constexpr int POPS = 1024 / KN, CPO = CO / 4;             // 16-byte chunks of the centre part, per point row
#pragma unroll
                for (int k = 0; k < (POPS + 31) / 32; k++) {
                    const int o = lane + 32 * k, prow = o / CPO, c = o % CPO;
                    const int sr = prow * KN;                             // first slot row (edge) of that point
                    const int pt = (CO * (sr / 32) + 32 * s + (sr % 32)) / KN;
                    if (POPS % 32 == 0 || o < POPS)
                        tc::cp_async16(slot + C::QPART + prow * (4 * CO) + c * 16,
                                       ppq + (size_t)min(p0 + pt, N - 1) * R2 + 4 * G + c * 4);
                }
                tc::cp_async_arrive_noinc(qfull + (st & 1));
            }
        }
        tc::cp_async_wait<0>();
    } else {
        // ================= epilogue warpgroup =================
        uint8_t* a_hi_p = smem + C::A_OFF + (size_t)wg * 2 * C::A_PART;
        uint8_t* a_lo_p = a_hi_p + C::A_PART;
        const uint32_t a_row = (uint32_t)(e >> 3) * C::A_SBO + (uint32_t)(e & 7) * 16;   // byte offset of row e, chunk 0
        const int ptl = e / KN;                                           // this edge's point within the tile
        const uint32_t q_row = (uint32_t)e * (16 * CPR), q_swz = (uint32_t)((e * CPR) >> 3) & (CPR - 1);
        const uint32_t p_row = C::QPART + (uint32_t)ptl * (16 * CPR), p_swz = (uint32_t)ptl & (CPR - 1);
        // conv_out: this thread's channel and edge group; ring offsets of its channel for the four values of (row & 3)
        const int ch = e % CO, grp = e / CO;
        uint32_t o_off[4];
#pragma unroll
        for (int v = 0; v < 4; v++)
            o_off[v] = (uint32_t)(32 * grp) * (4 * CO) + ((((uint32_t)ch >> 2) ^ ((uint32_t)v << LCPR)) << 4) + (ch & 3) * 4;
        const uint32_t op_off = C::QPART + (uint32_t)(32 * grp / KN) * (4 * CO) + ch * 4;
        uint32_t phase = 0, cphase = 0, st = 0;
        // pre-split output through a register transpose when a thread owns 8 points of a tile (the (32, 128), K = 16 blocks)
        constexpr bool SPLIT_T = CO == 128 && KN == 16;
        uint32_t sp[8] = {0, 0, 0, 0, 0, 0, 0, 0};

        // h[G] (fp32) of layer l -> fp16 hi/lo.  l < 3: this thread's lane of the tensor-memory slice (the A operand of dense
        // group l) is written and published FIRST; the copy into K-slice l of H in shared memory (conv_out's B operand)
        // follows off the critical path.  l == 3: the shared-memory copy completes H: generic -> async proxy fence (it
        // covers the earlier slices' stores of this thread as well), then publish.
        auto store_slice = [&](int l, const float (&h)[G]) {
            uint32_t hi[GP / 2], lo[GP / 2];
            const bool split3 = prec == PF_PREC_FP32_SPLIT;        // single-pass mode: no lo halves are computed or stored
            if (split3) {
#pragma unroll
                for (int u = 0; u < G / 2; u++) tc::split2(h[2 * u], h[2 * u + 1], hi[u], lo[u]);
            } else {
#pragma unroll
                for (int u = 0; u < G / 2; u++) hi[u] = tc::cvt_f16x2_sat(h[2 * u], h[2 * u + 1]), lo[u] = 0;
            }
#pragma unroll
            for (int u = G / 2; u < GP / 2; u++) hi[u] = lo[u] = 0;
            auto publish = [&]() {
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(ready);
            };
            if (l < 3) {
                const uint32_t ts = d_out + lane_sel + C::TM_SLICE;
                if constexpr (GP == 32) {
                    tc::st16_nowait(ts, hi);
                    if (split3) tc::st16_nowait(ts + 16, lo);
                } else {
                    tc::st8_nowait(ts, hi);
                    if (split3) tc::st8_nowait(ts + 8, lo);
                }
                tc::tmem_st_wait();
                publish();
            }
#pragma unroll
            for (int c = 0; c < G / 8; c++) {
                const uint32_t off = a_row + (uint32_t)(l * (GP / 8) + c) * 128;
                *reinterpret_cast<uint4*>(a_hi_p + off) = make_uint4(hi[4 * c], hi[4 * c + 1], hi[4 * c + 2], hi[4 * c + 3]);
                if (split3) *reinterpret_cast<uint4*>(a_lo_p + off) = make_uint4(lo[4 * c], lo[4 * c + 1], lo[4 * c + 2], lo[4 * c + 3]);
            }
            if (l == 3) {
                tc::fence_async_smem();
                publish();
            }
        };
        // q[G] = Q_j + P_i of the dense stage in ring slot `slot`
        auto read_pq = [&](const uint8_t* slot, float (&q)[G]) {
#pragma unroll
            for (int c = 0; c < CPR; c++) {
                const float4 v = *reinterpret_cast<const float4*>(slot + q_row + ((c ^ q_swz) << 4));
                const float4 p = *reinterpret_cast<const float4*>(slot + p_row + ((c ^ p_swz) << 4));
                q[4 * c] = v.x + p.x, q[4 * c + 1] = v.y + p.y, q[4 * c + 2] = v.z + p.z, q[4 * c + 3] = v.w + p.w;
            }
        };
        // this warp has read what it needs from ring slot b (the values are in registers): the producer may refill it.
        // Dense stages release right after the read, a whole stage before their results are published -- with a 2-slot ring
        // that is what lets the producer run a stage ahead.
        auto release = [&](uint32_t b) {
            __syncwarp();
            if (lane == 0) mbar_arrive(qempty + b);
        };
        tc::mbar_wait_bounded(wbar, 0);   // (ordering only: nothing below reads the weights)

        PROF_DECL
        for (; t < total_tiles; t += stride) {
            int patch, p0;
            tile_coords(t, patch, p0);
            // ---- layer 0: h0 = lrelu(P0_i + Q0_j) ----
            {
                const uint32_t b = st & 1;
                tc::mbar_wait_bounded(qfull + b, (st >> 1) & 1);
                st++;
                PROF_TICK(0)
                float q[G];
                read_pq(ring_p + b * C::SLOT, q);
                release(b);
#pragma unroll
                for (int c = 0; c < G; c++) q[c] = lrelu(q[c], slope);
                store_slice(0, q);
                PROF_TICK(1)
            }
            // ---- layers 1..3 ----
#pragma unroll 1
            for (int l = 1; l < 4; l++) {
                const uint32_t b = st & 1;
                tc::mbar_wait_bounded(qfull + b, (st >> 1) & 1);
                st++;
                PROF_TICK(2)
                float q[G];
                read_pq(ring_p + b * C::SLOT, q);
                release(b);
                PROF_TICK(3)
                tc::mbar_wait_bounded(mbar, phase);
                phase ^= 1;
                tc::fence_after_sync();
                PROF_TICK(4)
                float acc[G];
                const uint32_t dl = d_dense + lane_sel + (uint32_t)(GP * (l - 1));   // layer l: columns [GP (l-1), GP l)
                if constexpr (G == 32) {
                    tc::ld32(dl, acc);
                } else if constexpr (G == 16) {
                    tc::ld16(dl, acc);
                } else {
                    float a16[16];
                    tc::ld16(dl, a16);
#pragma unroll
                    for (int c = 0; c < G; c++) acc[c] = a16[c];
                }
#pragma unroll
                for (int c = 0; c < G; c++) acc[c] = lrelu(fmaf(acc[c], tc::WINV, q[c]), slope);
                store_slice(l, acc);
                PROF_TICK(5)
            }
            // ---- conv_out (transposed) + max over the KN neighbours ----
#pragma unroll
            for (int c = 0; c < NCH; c++) {                               // 32 edges = 32 / KN points per step
                const uint32_t b = st & 1;
                tc::mbar_wait_bounded(qfull + b, (st >> 1) & 1);
                st++;
                PROF_TICK(6)
                if (c % (NCH / NQ) == 0) {
                    tc::mbar_wait_bounded(cbar + c / (NCH / NQ), cphase);
                    tc::fence_after_sync();
                }
                PROF_TICK(7)
                const uint8_t* slot = ring_p + b * C::SLOT;
                float acc[32];
                tc::ld32(d_out + lane_sel + 32 * (grp * NCH + c), acc);
#pragma unroll
                for (int u = 0; u < 32; u++)
                    acc[u] = fmaf(acc[u], tc::WINV, *reinterpret_cast<const float*>(slot + o_off[u & 3] + u * (4 * CO)));
                float pv[32 / KN];
#pragma unroll
                for (int h = 0; h < 32 / KN; h++) pv[h] = *reinterpret_cast<const float*>(slot + op_off + h * (4 * CO));
                release(b);
                float val[32 / KN];
#pragma unroll
                for (int h = 0; h < 32 / KN; h++) {
                    float m = acc[KN * h];
#pragma unroll
                    for (int u = 1; u < KN; u++) m = fmaxf(m, acc[KN * h + u]);
                    val[h] = m + pv[h];
                    const int pp = p0 + (CO * grp + 32 * c) / KN + h;
                    if (pp < N) {
                        const size_t grow = (size_t)patch * N + pp;
                        if (out) out[grow * CO + ch] = val[h];
                        if (!SPLIT_T && out_split) {   // fp16 hi/lo in the tile-blocked operand layout of linear_split.cu (K = CO)
                            uint32_t hi, lo;
                            tc::split2(val[h], 0.f, hi, lo);
                            uint8_t* tile = out_split + (grow >> 7) * (size_t)(CO * 512) + (size_t)(ch >> 3) * 2048 +
                                            (grow & 127) * 16 + (ch & 7) * 2;
                            *reinterpret_cast<uint16_t*>(tile) = (uint16_t)(hi & 0xffffu);
                            *reinterpret_cast<uint16_t*>(tile + CO * 256) = (uint16_t)(lo & 0xffffu);
                        }
                    }
                }
                if constexpr (SPLIT_T) {   // this step's two points -> (hi | lo << 16) of this thread's channel
                    uint32_t hi, lo;
                    tc::split2(val[0], val[1], hi, lo);
                    sp[2 * c] = __byte_perm(hi, lo, 0x5410), sp[2 * c + 1] = __byte_perm(hi, lo, 0x7632);
                }
                PROF_TICK(8)
            }
            cphase ^= 1;
            if constexpr (SPLIT_T) {
                if (out_split) {
                    // sp[q] = this channel's value for point q of the tile.  8 x 8 transpose inside each group of 8 lanes
                    // (= 8 consecutive channels): lane r ends up with the 8 channels of point r, i.e. one aligned 16-byte
                    // piece of the operand tile per half -- 2 STG.128 per thread and tile instead of 16 two-byte stores
                    // (ncu: 21 of 32 bytes per sector used, +0.2 ms per launch).
#pragma unroll
                    for (int m = 1; m < 8; m <<= 1) {
                        const bool up = (lane & m) != 0;
#pragma unroll
                        for (int i = 0; i < 8; i++)
                            if (!(i & m)) {
                                const uint32_t recv = __shfl_xor_sync(0xffffffffu, up ? sp[i] : sp[i + m], m);
                                if (up) sp[i] = recv; else sp[i + m] = recv;
                            }
                    }
                    const int pp = p0 + (lane & 7);
                    if (pp < N) {
                        const size_t grow = (size_t)patch * N + pp;
                        uint8_t* tile = out_split + (grow >> 7) * (size_t)(CO * 512) + (size_t)(ch >> 3) * 2048 + (grow & 127) * 16;
                        *reinterpret_cast<uint4*>(tile) = make_uint4(__byte_perm(sp[0], sp[1], 0x5410), __byte_perm(sp[2], sp[3], 0x5410),
                                                                     __byte_perm(sp[4], sp[5], 0x5410), __byte_perm(sp[6], sp[7], 0x5410));
                        *reinterpret_cast<uint4*>(tile + CO * 256) =
                            make_uint4(__byte_perm(sp[0], sp[1], 0x7632), __byte_perm(sp[2], sp[3], 0x7632),
                                       __byte_perm(sp[4], sp[5], 0x7632), __byte_perm(sp[6], sp[7], 0x7632));
                    }
                }
            }
            // The next tile's layer 0 overwrites H and (layer 1) the accumulator: both are free -- this thread's conv_out
            // tcgen05.ld have completed, and its next `ready` arrival (after fence::before_thread_sync) orders them
            // before the MMA warp's next issue.
        }
        if (tid == 0) { PROF_DUMP(0, 10) }
    }

    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem_base, C::TM_COLS);
}

template <int G, int CO, int NWG, int KN>
int launch(const float* pq, const int32_t* idx, const void* wtc, float* out, void* out_split, int Bp, int N, float slope,
           int prec, cudaStream_t st) {
    using C = Cfg<G, CO, NWG, KN>;
    constexpr int PT = C::PT;
    auto kern = edgeconv_tc_kernel<G, CO, NWG, KN>;
    if (int rc = pf_smem_optin(kern, C::SMEM_BYTES)) return rc;
    const int tiles_per_patch = (N + PT - 1) / PT;
    const long long total = (long long)Bp * tiles_per_patch;
    int dev = 0, sms = 148;
    PF_CUDA(cudaGetDevice(&dev));
    PF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const long long want = (total + NWG - 1) / NWG;
    const int grid = (int)(want < sms ? want : sms);
    kern<<<grid, C::NT, C::SMEM_BYTES, st>>>(pq, idx, static_cast<const uint8_t*>(wtc), out, static_cast<uint8_t*>(out_split),
                                             N, tiles_per_patch, total, slope, prec);
    return pf_launch_status("edgeconv_tc_kernel");
}

}  // namespace

static int edgeconv_tc_dispatch(const float* pq, const int32_t* idx, const void* wtc, float* out, void* out_split, int Bp,
                                int N, int K, int G_, int C_out, float slope, int precision, void* stream) {
    PF_CHECK_ARG(precision == PF_PREC_FP32_SPLIT || precision == PF_PREC_FP16, "pf_edgeconv_tc: precision %d (PF_PREC_FP32_SPLIT or PF_PREC_FP16)", precision);
    PF_CHECK_ARG(slope >= 0.f && slope <= 1.f, "LeakyReLU slope %g outside [0, 1]", (double)slope);
    PF_CHECK_ARG(Bp >= 0 && N > 0, "pf_edgeconv_tc: bad sizes Bp=%d N=%d", Bp, N);
    if (Bp == 0) return 0;
    PF_CHECK_ARG(pq && idx && wtc && (out || out_split), "pf_edgeconv_tc: null pointer");
    PF_CHECK_ARG((long long)Bp * N < 0x7fffffffLL, "pf_edgeconv_tc: Bp*N overflows int32");
    PF_CHECK_ARG(N <= 65536, "pf_edgeconv_tc: N=%d > 65536 points per patch", N);
    PF_CHECK_ARG(K == 8 || K == 16 || K == 32, "pf_edgeconv_tc: K=%d not in {8, 16, 32}", K);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(pq) | reinterpret_cast<uintptr_t>(wtc) | reinterpret_cast<uintptr_t>(out) |
                   reinterpret_cast<uintptr_t>(out_split)) & 15) == 0,
                 "pf_edgeconv_tc: pointers must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
#define PF_EC_CASE(g, c, nwg)                                                                                     \
    if (G_ == g && C_out == c) {                                                                                  \
        if (K == 16) return launch<g, c, nwg, 16>(pq, idx, wtc, out, out_split, Bp, N, slope, precision, st);                  \
        if (K == 8) return launch<g, c, nwg, 8>(pq, idx, wtc, out, out_split, Bp, N, slope, precision, st);                    \
        return launch<g, c, nwg, 32>(pq, idx, wtc, out, out_split, Bp, N, slope, precision, st);                               \
    }
    PF_EC_CASE(32, 128, 2)
    PF_EC_CASE(16, 64, 3)
    PF_EC_CASE(8, 32, 3)
#undef PF_EC_CASE
    PF_CHECK_ARG(false, "pf_edgeconv_tc: (G, C_out) = (%d, %d) not in {(8,32), (16,64), (32,128)}", G_, C_out);
    return 0;
}

#ifdef PF_PROF
extern "C" __attribute__((visibility("default"))) int pf_debug_edgeconv_prof(long long* host32) {
    return cudaMemcpyFromSymbol(host32, g_prof, sizeof(long long) * 32) == cudaSuccess ? 0 : 1;
}
#endif

extern "C" int pf_edgeconv_tc(const float* pq, const int32_t* idx, const void* wtc, float* out, int Bp, int N, int K,
                              int G_, int C_out, float slope, int precision, void* stream) {
    return edgeconv_tc_dispatch(pq, idx, wtc, out, nullptr, Bp, N, K, G_, C_out, slope, precision, stream);
}

extern "C" int pf_edgeconv_tc_split(const float* pq, const int32_t* idx, const void* wtc, float* out, void* out_split, int Bp,
                                    int N, int K, int G_, int C_out, float slope, int precision, void* stream) {
    return edgeconv_tc_dispatch(pq, idx, wtc, out, out_split, Bp, N, K, G_, C_out, slope, precision, stream);
}
