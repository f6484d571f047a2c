// Latent-space kNN-weighted interpolation on the tensor cores (tcgen05.mma, accumulators in TMEM).
//
// Same result as pf_interp (interp.cu; replaces InterpolationModule.forward,
// /root/reference/modules/discrete/interpflow.py:162-186, with DistanceEncoder :85-120, KnnContextEncoder
// :123-134 and WeightEstimationUnit :137-159 fused).  CUDA-event timing shows the FFMA kernel FP32-pipe-bound
// (743 GFLOP in 35.7 ms at config 2, 104 MB of HBM traffic), so the per-edge MLPs go to tcgen05 with the fp16
// hi/lo split of edgeconv_tc.cu (3 MMAs per product, fp32 accumulation, fp32-class result).
//
// The input of this module is 3-D coordinates, so nothing is gathered from pre-projected features: every conv
// that reads coordinates takes them from a 16-wide "geo" K-slice of the activation operand
//     geo = [x_i | x_j | x_j - x_i | |x_i - x_j| | 1 | 0..]   (bias rides on the constant 1)
// with the reference's unfolded weights (puflow_b200/packing.py pack_interp_tc).  A tile is 128 edges = 16 points
// x 8 neighbours; X[128 edges][K = 208] = geo(16) | dense block h (128) | distance encoder (64), K-major fp16 hi/lo
// in shared memory, is the B operand of the composed 208 -> 128 layer t0.
//
// Round-2 structure (per-stage clock64() tables in profiles/r02_interp.md).  One tile is a chain of 11 dependent
// tensor-core round trips (~900 cycles each through mbarriers, whatever the MMA does), and the 8 dense layers were 60 %
// of it: each a round trip of two epilogue warpgroups (TMEM load, LeakyReLU, split, STS, fence.proxy.async) plus
// 2 (l + 1) small MMAs at ~36 cycles of issue each on the critical path (145 tcgen05.mma per tile from one thread).  A
// second tile in flight does not fit (X is 106 KB of shared memory, and 2 x (t0 + dense accumulators) exceed the 512
// TMEM columns), so the chain itself was shortened and everything else moved off it:
//   warpgroup 0  the dense chain alone.  Layer l's accumulator (16 columns) -> LeakyReLU -> fp16 hi/lo -> a 16-column
//                slice of TENSOR MEMORY (tcgen05.st, no proxy fence), the A operand of ONE wide MMA group that adds
//                h_l's contribution to ALL later layers at once:
//                    D[128 edges x 16 (7 - l)] += h_l . [W_{l+1,l}; ..; W_{7,l}]^T     (3 MMAs, K = 16)
//                (the geo slice feeds all eight layers the same way, N = 128), so the critical path of a layer is
//                three instructions and K = 16, whatever l.  It skips the softmax stage, starts the next tile's chain
//                while warpgroup 1 finishes the current tile, and does its half of that softmax when the chain is done;
//   warpgroup 1  everything off that path: the distance encoder (d1, d2), the shared-memory copies of h_l / d2 that t0
//                reads as its B operand (it re-reads the same accumulator columns; its stores and proxy fences cost the
//                chain nothing), the softmax + output stage.  Each publish has its own mbarrier (one completion per
//                tile), so the chain may run ahead of it safely;
//   MMA warp     one elected thread: dense groups as their slices arrive; behind each group up to three items of the
//                work that does not gate the chain (d2, and the 13 K-steps of t0^T = M12g . X^T), chosen by TESTING
//                (mbarrier.test_wait) which of warpgroup 1's publishes have happened -- never waiting for them.  The
//                head of the NEXT tile (geo part of all dense layers, d1) is issued right behind the logits MMAs: the geo
//                slice has a buffer of its own and is written during the tail.
// Then, both warpgroups (column halves of a stage):
//   9  t0^T[128 ch x 128 edges] (M12g = mlp.0 o (mlp.6 | conv_out) composed on the host, resident in TMEM as the A
//      operand); thread = channel; t0 is written back as an M-major (edge-contiguous) operand
//  10  t1 = lrelu(t0.W3 + b3)  (A = t0 M-major, N=64, K=128)      11  logits = t1.W6 + b6  (N=32, K=64)
//  then softmax over the 8 neighbours (8 consecutive lanes) and the weighted sum of the neighbours' latents.
// Narrow products (d1, d2, t1) use the stacked form: hi and lo of a weight matrix are adjacent row groups of ONE K-major
// operand, a_hi . [b_hi | b_lo] is a single MMA of width 2n and only a_lo . b_hi remains.
// 4.17 -> 3.59 ms at config 2 (warpgroup 0 also copies h6 / h7 and does its softmax half one tile late, in the idle time
// at the end of the next chain); measured and rejected on the way: test_wait spinning on the chain's two waits (no change),
// warpgroup 0 writing the X copies itself from its hi/lo registers (the chain's stage grows by 165 cycles: 3.75 ms),
// d1 of the next tile before the softmax (3.75 ms), four / five instead of three side items per gap (3.62 / 3.97 ms).
#include "common.cuh"
#include "puflow_b200.h"
#include "tc.cuh"

namespace {

using namespace pf;

constexpr int KN = 8, PT = 16, TM = 128;
constexpr int NT = 288;                                   // 2 x 4 epilogue warps + 1 MMA-issuing warp
constexpr float S_CTX = 0.05f, S_DEF = 0.01f;             // interpflow.py:206,217 / nn.LeakyReLU() default :94,97,147,150

// ---- shared memory map (bytes) ----
constexpr uint32_t W_D1 = 0, W_D2 = 4096, W_H = 20480, W_3 = 57344, W_6 = 90112, B_D3 = 98304, B_3 = 98560, B_6 = 98816;
constexpr uint32_t W_BYTES = 98944;
// dense block, by K-slice: Wg = the geo columns of all eight layers [128][16] (hi, lo), then S_m = the h_m columns of
// layers m+1..7 [16 (7 - m)][16] (hi, lo), m = 0..6
constexpr uint32_t WG_PART = 4096;
__host__ __device__ constexpr uint32_t ws_off(int m) { return W_H + 2 * WG_PART + 1024u * (7 * m - m * (m - 1) / 2); }
__host__ __device__ constexpr uint32_t ws_part(int m) { return 512u * (7 - m); }
constexpr uint32_t X_OFF = 99072, KX = 208, X_SBO = KX * 16, X_PART = 128 * KX * 2;    // 53,248 per part
constexpr uint32_t T0_OFF = X_OFF, T0_PART = 32768;       // t0 M-major [128 ch][128 edges], aliases X (dead by then)
constexpr uint32_t T1_OFF = X_OFF + 65536, T1_PART = 16384, T1_SBO = 64 * 16;   // t1 K-major [128 edges][64], behind t0
// the geo slice has a buffer of its own ([128][16] K-major, hi | lo): the NEXT tile's slice is written, and its first MMAs
// are issued, under the tail of the current tile (whose t0 / t1 operands alias X)
constexpr uint32_t GEO_OFF = X_OFF + 2 * X_PART, GEO_PART = 4096, GEO_SBO = 256;   // 205,568
constexpr uint32_t BAR_OFF = GEO_OFF + 2 * GEO_PART;      // 213,760
// mbarriers: weights | ready8 | readyA | mbarT | geoR | acc[8] | xbar[8] | dbar[2]
constexpr int NBAR = 5 + 8 + 8 + 2;
constexpr uint32_t SMEM_BYTES = BAR_OFF + 8 * NBAR + 16;
constexpr uint32_t COL_H = 16, COL_D = 144;               // K offsets inside X
// ---- tensor memory map (columns): M12g hi | lo (A operand of t0), T0 = the distance encoder's accumulators (stacked
// form, 128) until d2 has been read, then t0's; DENSE = the eight dense layers' accumulators (16 columns each), later
// t1's; SLICE = the current h_l as fp16 pairs (hi 8 columns | lo 8 columns); LG = the logits (DENSE and T0 already
// belong to the next tile by then) ----
constexpr uint32_t TM_M12_HI = 0, TM_M12_LO = 104, TM_T0 = 208, TM_DENSE = 336, TM_SLICE = 464, TM_LG = 480, TM_COLS = 512;
constexpr uint32_t M12_GLOBAL = W_BYTES, M12_PART = 128 * 104 * 4;

#ifdef PF_PROF   // dev builds only: per-stage clock64() totals of CTA 0 (epilogue thread 0 and the MMA thread), tc.cuh
__device__ long long g_prof[32];
#endif

template <int RP>   // logits evaluated per edge (R rounded up to 4 / 8 / 16 / 32)
__global__ void __launch_bounds__(NT, 1)
    interp_tc_kernel(const float* __restrict__ xyz, const float* __restrict__ z, const int32_t* __restrict__ idx,
                     int ldidx, const uint8_t* __restrict__ wblob, float* __restrict__ out, int N, int R,
                     int tiles_per_patch, long long total_tiles, int prec) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + BAR_OFF);
    uint64_t* ready8 = wbar + 1;      // both warpgroups published (t0, t1): 8 warp arrivals
    uint64_t* readyA = wbar + 2;      // warpgroup 0 wrote the slice h_l to tensor memory: 4 warp arrivals
    uint64_t* mbarT = wbar + 3;       // tcgen05.commit: t0 complete / t1 / logits
    uint64_t* geoR = wbar + 4;        // both warpgroups wrote the geo slice of the NEXT tile: 8 warp arrivals
    uint64_t* accbar = wbar + 5;      // [8] tcgen05.commit: dense layer l's accumulator complete (once per tile each)
    uint64_t* xbar = wbar + 13;       // [8] published in X: d1 | d2 (+ copies of h0, h1) | copy of h2 | .. | h7 (h6, h7: warpgroup 0)
    uint64_t* dbar = wbar + 21;       // [2] tcgen05.commit: d1 / d2 accumulator complete
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wbar + NBAR);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool is_mma_warp = warp == 8;
    const int e = tid & 127;          // row of the tile (TMEM lane)
    const int half = (warp >> 2) & 1;  // epilogue warpgroup: 0 = dense chain, 1 = side work; column halves elsewhere

    if (tid == 0) {
        mbar_init(wbar, 1), mbar_init(ready8, 8), mbar_init(readyA, 4), mbar_init(mbarT, 1), mbar_init(geoR, 8);
        for (int i = 0; i < 8; i++) mbar_init(accbar + i, 1), mbar_init(xbar + i, 4);
        mbar_init(dbar, 1), mbar_init(dbar + 1, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, TM_COLS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    if (tid == 0) {
        mbar_expect_tx(wbar, W_BYTES);
        bulk_g2s_chunked(smem, wblob, W_BYTES, wbar);
    }
    if (!is_mma_warp) {   // M12g -> TMEM: thread = lane = output channel; 104 columns per part (32+32+32+8)
        {
            const int part = half;
            const uint32_t* src = reinterpret_cast<const uint32_t*>(wblob + M12_GLOBAL + (size_t)part * M12_PART) + e * 104;
            const uint32_t dst = tmem_base + lane_sel + (part ? TM_M12_LO : TM_M12_HI);
#pragma unroll 1
            for (int c = 0; c < 3; c++) {
                uint32_t r[32];
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + 32 * c) + i);
                    r[4 * i] = v.x, r[4 * i + 1] = v.y, r[4 * i + 2] = v.z, r[4 * i + 3] = v.w;
                }
                tc::st32(dst + 32 * c, r);
            }
            {
                const uint4 v0 = __ldg(reinterpret_cast<const uint4*>(src + 96)), v1 = __ldg(reinterpret_cast<const uint4*>(src + 96) + 1);
                tc::st8(dst + 96, v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w);
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();

    const uint32_t x_hi = sbase + X_OFF, x_lo = x_hi + X_PART;
    const long long stride = gridDim.x;
    long long t = blockIdx.x;
    const bool split = prec == PF_PREC_FP32_SPLIT;           // PF_PREC_FP16: the activations' lo products are dropped

    if (is_mma_warp) {
        tc::mbar_wait_bounded(wbar, 0);
        if (tc::elect_one()) {
            // Stacked form for the narrow products: a_hi . [b_hi | b_lo] is one MMA of width 2n (big term in columns [0, n),
            // a_hi . b_lo in [n, 2n)), a_lo . b_hi accumulates into [n, 2n); the epilogue adds the two halves.
            auto issue2 = [&](uint32_t d, uint64_t ah, uint64_t al, uint32_t a_step, uint64_t bh, uint32_t b_step, int ks,
                              uint32_t idesc_2n, uint32_t idesc_n, uint32_t n) {
#pragma unroll
                for (int s = 0; s < ks; s++) tc::mma_f16(d, ah + a_step * s, bh + b_step * s, idesc_2n, s > 0);
                if (split) {
#pragma unroll
                    for (int s = 0; s < ks; s++) tc::mma_f16(d + n, al + a_step * s, bh + b_step * s, idesc_n, 1);
                }
            };
            auto wdesc = [&](uint32_t off, uint32_t kt) { return tc::smem_desc(sbase + off, kt * 16); };
            const uint64_t xh = tc::smem_desc(x_hi, X_SBO), xl = tc::smem_desc(x_lo, X_SBO);
            const uint64_t gh = tc::smem_desc(sbase + GEO_OFF, GEO_SBO), gl = tc::smem_desc(sbase + GEO_OFF + GEO_PART, GEO_SBO);
            const uint32_t d_t0 = tmem_base + TM_T0, d_dense = tmem_base + TM_DENSE, slice = tmem_base + TM_SLICE;
            const uint32_t d_lg = tmem_base + TM_LG;
            constexpr uint32_t i64 = tc::idesc_f16(128, 64), i128 = tc::idesc_f16(128, 128), i32 = tc::idesc_f16(128, 32),
                               AT = 1u << 15;
            uint32_t p8 = 0, pA = 0, pG = 0, tp = 0;         // parities: ready8, readyA, geoR (running), per-tile barriers
            PROF_DECL
            auto wait_bar = [&](uint64_t* bar, uint32_t parity) {
                PROF_TICK(1)
                tc::mbar_wait_bounded(bar, parity);
                tc::fence_after_sync();
                PROF_TICK(0)
            };
            // One K-step (16 columns of X) of t0^T = M12g . X^T, A from tensor memory (8 columns per K-step), 3 split
            // products, the small cross terms first.  K-step 0 = geo, 1 + j = h_j, 9..12 = d2.
            const uint32_t mh = tmem_base + TM_M12_HI, ml = tmem_base + TM_M12_LO;
            auto t0_step = [&](int s, bool first) {
                const uint64_t bh = s == 0 ? gh : xh + 16u * s, bl = s == 0 ? gl : xl + 16u * s;
                tc::mma_f16_ts(d_t0, ml + 8u * s, bh, i128, first ? 0u : 1u);
                if (split) tc::mma_f16_ts(d_t0, mh + 8u * s, bl, i128, 1);
                tc::mma_f16_ts(d_t0, mh + 8u * s, bh, i128, 1);
            };
            // The head of a tile, issued under the tail of the tile before it: the geo part of all eight dense layers (N = 128)
            // and d1.  DENSE (t1's accumulator, read before t1 was published) and T0 are free by then.
            auto issue_head = [&]() {
                wait_bar(geoR, pG), pG ^= 1;
                const uint64_t bh = wdesc(W_H, 16), bl = wdesc(W_H + WG_PART, 16);
                tc::mma_f16(d_dense, gh, bl, i128, 0);
                if (split) tc::mma_f16(d_dense, gl, bh, i128, 1);
                tc::mma_f16(d_dense, gh, bh, i128, 1);
                tc::commit(accbar + 0);
                issue2(d_t0, gh, gl, 16, wdesc(W_D1, 16), 16, 1, i128, i64, 64);
                tc::commit(dbar + 0);
            };
            // Work that does not gate the dense chain, in the order warpgroup 1 makes it possible (xbar[i] complete = d1 | d2,
            // h0, h1 | h2 | .. | h7 are in X): item 0 = d2 = d1 . Wd3 (A = X columns [144, 208)), items 1..13 = t0's K-steps
            // geo, h0, h1, d2 x 4, h2 .. h7.  Behind every chain group up to `budget` items whose inputs are there are issued;
            // the barriers are only TESTED (mbarrier.test_wait), so a late warpgroup 1 never stalls the chain.
            constexpr int NSIDE = 14;
            int side = 0, xdone = 0;
            auto side_work = [&](int budget, bool drain) {
                while (side < NSIDE && budget > 0) {
                    const int need = side == 0 ? 0 : (side <= 7 ? 1 : side - 6);
                    while (xdone <= need) {
                        if (drain) tc::mbar_wait_bounded(xbar + xdone, tp);
                        else if (!mbar_test(xbar + xdone, tp)) break;
                        xdone++;
                        tc::fence_after_sync();
                    }
                    if (xdone <= need) break;
                    if (side == 0) {
                        issue2(d_t0, xh + (COL_D / 8) * 8, xl + (COL_D / 8) * 8, 16, wdesc(W_D2, 64), 16, 4, i128, i64, 64);
                        tc::commit(dbar + 1);
                        budget -= 3;
                    } else {
                        t0_step(side <= 3 ? side - 1 : (side <= 7 ? side + 5 : side - 5), side == 1);
                        budget -= 1;
                    }
                    side++;
                }
            };
            if (t < total_tiles) issue_head();
            for (; t < total_tiles; t += stride, tp ^= 1) {
                side = 0, xdone = 0;
                // dense groups m = 0..6: h_m (tensor-memory slice) x the h_m columns of layers m+1..7; behind each, work
                // that does not gate the chain (its inputs were published at least a stage earlier)
#pragma unroll
                for (int m = 0; m < 7; m++) {
                    wait_bar(readyA, pA), pA ^= 1;
                    {
                        const uint64_t bh = wdesc(ws_off(m), 16), bl = wdesc(ws_off(m) + ws_part(m), 16);
                        const uint32_t d = d_dense + 16u * (m + 1), idesc = tc::idesc_f16(128, 16 * (7 - m));
                        tc::mma_f16_ts(d, slice, bl, idesc, 1);
                        if (split) tc::mma_f16_ts(d, slice + 8, bh, idesc, 1);
                        tc::mma_f16_ts(d, slice, bh, idesc, 1);
                        tc::commit(accbar + m + 1);
                    }
                    side_work(3, false);
                }
                side_work(1 << 20, true);                   // whatever is left: at least h7's K-step
                tc::commit(mbarT);
                // 10: t1 = t0 . W3^T, A = t0 M-major (8-edge groups 128 B apart, 8-channel groups 2048 B apart)
                wait_bar(ready8, p8), p8 ^= 1;
                issue2(d_dense, tc::smem_desc(sbase + T0_OFF, 128, 2048), tc::smem_desc(sbase + T0_OFF + T0_PART, 128, 2048), 256,
                       wdesc(W_3, 128), 16, 8, i128 | AT, i64 | AT, 64);
                tc::commit(mbarT);
                // 11: logits = t1 . W6^T: three split products into one 32-column accumulator of their own
                wait_bar(ready8, p8), p8 ^= 1;
                {
                    const uint64_t ah = tc::smem_desc(sbase + T1_OFF, T1_SBO), al = tc::smem_desc(sbase + T1_OFF + T1_PART, T1_SBO);
                    const uint64_t bh = wdesc(W_6, 64), bl = bh + (uint64_t)((32u * 64u * 2u) >> 4);
#pragma unroll
                    for (int s = 0; s < 4; s++) tc::mma_f16(d_lg, ah + 16u * s, bl + 16u * s, i32, s > 0);
                    if (split) {
#pragma unroll
                        for (int s = 0; s < 4; s++) tc::mma_f16(d_lg, al + 16u * s, bh + 16u * s, i32, 1);
                    }
#pragma unroll
                    for (int s = 0; s < 4; s++) tc::mma_f16(d_lg, ah + 16u * s, bh + 16u * s, i32, 1);
                }
                tc::commit(mbarT);
                if (t + stride < total_tiles) issue_head();
            }
            PROF_TICK(1)
            PROF_DUMP(16, 2)
        }
        __syncwarp();
    } else {
        // ---- epilogue warps: thread e = edge row of the tile (point e / 8, neighbour e % 8) ----
        uint8_t* xh_p = smem + X_OFF;
        uint8_t* xl_p = xh_p + X_PART;
        const uint32_t x_row = (uint32_t)(e >> 3) * X_SBO + (uint32_t)(e & 7) * 16;
        const float* bd3 = reinterpret_cast<const float*>(smem + B_D3);
        const float* b3 = reinterpret_cast<const float*>(smem + B_3);
        const float* b6 = reinterpret_cast<const float*>(smem + B_6);
        const uint32_t d_t0 = tmem_base + TM_T0 + lane_sel, d_a = tmem_base + TM_DENSE + lane_sel,
                       slice = tmem_base + TM_SLICE + lane_sel, d_lg = tmem_base + TM_LG + lane_sel;
        uint32_t phase = 0, tp = 0;                          // mbarT (running), per-tile barriers
        // v[8] -> fp16 hi/lo -> 16-byte chunk `chunk` of this thread's row of X
        auto store8 = [&](uint32_t chunk, const float* v) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int u = 0; u < 4; u++) tc::split2(v[2 * u], v[2 * u + 1], hi[u], lo[u]);
            *reinterpret_cast<uint4*>(xh_p + x_row + chunk * 128) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(xl_p + x_row + chunk * 128) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        };
        // This warp's operand rows are written and its accumulator reads are done.  No block barrier: epilogue threads never
        // read each other's shared-memory writes (only the MMA does), and the MMA warp waits for the warps' arrivals.
        auto publish = [&](uint64_t* bar) {
            tc::fence_async_smem();
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar);
        };
        auto wait_bar = [&](uint64_t* bar, uint32_t parity) {
            tc::mbar_wait_bounded(bar, parity);
            tc::fence_after_sync();
        };
        auto wait_tail = [&]() {
            wait_bar(mbarT, phase);
            phase ^= 1;
        };
        // dense layer l of this thread's edge: accumulator -> LeakyReLU (the bias rode on geo's constant 1)
        auto dense_act = [&](int l, float (&a)[16]) {
            tc::ld16(d_a + 16u * l, a);
#pragma unroll
            for (int u = 0; u < 16; u++) a[u] = lrelu(a[u] * tc::WINV, S_CTX);
        };
        tc::mbar_wait_bounded(wbar, 0);   // biases live in the weight blob

        // neighbour index and coordinates of a tile are fetched one tile ahead (two dependent global round trips)
        auto tile_rows = [&](long long tt, int& p0_, size_t& row_i_) {
            const int patch_ = (int)((uint32_t)tt / (uint32_t)tiles_per_patch);
            p0_ = (int)((uint32_t)tt - (uint32_t)patch_ * (uint32_t)tiles_per_patch) * PT;
            row_i_ = (size_t)patch_ * N + min(p0_ + (e >> 3), N - 1);
            return (size_t)patch_ * N;
        };
        int p0_n = 0, j_n = 0;
        size_t row_i_n = 0, base_n = 0;
        float xi_n[3] = {0.f, 0.f, 0.f}, xj_n[3] = {0.f, 0.f, 0.f};
        // geo slice of the tile whose coordinates are in xi_n / xj_n -> its own buffer; warpgroup 0 writes columns 0..7,
        // warpgroup 1 columns 8..15
        auto geo_publish = [&]() {
            const float dx = xi_n[0] - xj_n[0], dy = xi_n[1] - xj_n[1], dz = xi_n[2] - xj_n[2];
            const float dist = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
            const float g0[8] = {xi_n[0], xi_n[1], xi_n[2], xj_n[0], xj_n[1], xj_n[2], -dx, -dy};
            const float g1[8] = {-dz, dist, 1.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            const float* g = half == 0 ? g0 : g1;
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int u = 0; u < 4; u++) tc::split2(g[2 * u], g[2 * u + 1], hi[u], lo[u]);
            uint8_t* dst = smem + GEO_OFF + (uint32_t)(e >> 3) * GEO_SBO + (uint32_t)half * 128 + (uint32_t)(e & 7) * 16;
            *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(dst + GEO_PART) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            publish(geoR);
        };
        if (t < total_tiles) {
            base_n = tile_rows(t, p0_n, row_i_n);
            j_n = min(max(__ldg(idx + row_i_n * ldidx + (e & 7)), 0), N - 1);
#pragma unroll
            for (int a = 0; a < 3; a++) xi_n[a] = __ldg(xyz + row_i_n * 3 + a), xj_n[a] = __ldg(xyz + (base_n + j_n) * 3 + a);
            geo_publish();
        }
        // this warpgroup's half of the R logits of a tile whose logits are complete: r in [half RP/2, half RP/2 + RP/2)
        auto softmax_half = [&](bool writer, float* dst, float z0, float z1, float z2) {
            constexpr int RH = RP / 2;
            const int r0 = half * RH;
            float lg[RH];
            if constexpr (RH == 16) {
                tc::ld16(d_lg + r0, lg);
            } else {
                float a8[8];
                tc::ld8(d_lg + r0, a8);
#pragma unroll
                for (int r = 0; r < RH; r++) lg[r] = a8[r];
            }
#pragma unroll
            for (int rr = 0; rr < RH; rr++) {
                const int r = r0 + rr;
                float v = fmaf(lg[rr], tc::WINV, b6[r]);
                float mx = v;
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                const float ex = expf(v - mx);
                float sum = ex;
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                const float w = ex * (1.f / sum);
                float o0 = w * z0, o1 = w * z1, o2 = w * z2;
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) {
                    o0 += __shfl_xor_sync(0xffffffffu, o0, o);
                    o1 += __shfl_xor_sync(0xffffffffu, o1, o);
                    o2 += __shfl_xor_sync(0xffffffffu, o2, o);
                }
                if (writer && r < R) dst[r * 3] = o0, dst[r * 3 + 1] = o1, dst[r * 3 + 2] = o2;
            }
        };
        bool pend = false, pend_writer = false;                            // warpgroup 0: softmax half of the previous tile
        float* pend_dst = out;
        float pz0 = 0.f, pz1 = 0.f, pz2 = 0.f;
        PROF_DECL
        for (; t < total_tiles; t += stride, tp ^= 1) {
            const int p0 = p0_n;
            const size_t row_i = row_i_n, row_j = base_n + j_n;
            const bool has_next = t + stride < total_tiles;
            int j_raw = 0;
            if (has_next) {
                base_n = tile_rows(t + stride, p0_n, row_i_n);
                j_raw = __ldg(idx + row_i_n * ldidx + (e & 7));             // consumed after the first chain stage below
#pragma unroll
                for (int a = 0; a < 3; a++) xi_n[a] = __ldg(xyz + row_i_n * 3 + a);
            }
            PROF_TICK(0)
            if (half == 0) {
                // ---- warpgroup 0: the dense chain.  h_l -> tensor-memory slice (hi 8 columns | lo 8 columns) ----
#pragma unroll 1
                for (int l = 0; l < 7; l++) {
                    wait_bar(accbar + l, tp);
                    PROF_TICK(1)
                    float a[16];
                    dense_act(l, a);
                    uint32_t hi[8], lo[8];
#pragma unroll
                    for (int u = 0; u < 8; u++) tc::split2(a[2 * u], a[2 * u + 1], hi[u], lo[u]);
                    tc::st8_nowait(slice, hi);
                    tc::st8_nowait(slice + 8, lo);
                    tc::tmem_st_wait();
                    tc::fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(readyA);
                    if (l == 6) {                                            // h6's copy into X from the halves at hand
                        uint8_t* ph = xh_p + x_row + (COL_H / 8 + 12) * 128;
                        uint8_t* pl = xl_p + x_row + (COL_H / 8 + 12) * 128;
                        *reinterpret_cast<uint4*>(ph) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                        *reinterpret_cast<uint4*>(ph + 128) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
                        *reinterpret_cast<uint4*>(pl) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                        *reinterpret_cast<uint4*>(pl + 128) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
                        publish(xbar + 6);
                    }
                    PROF_TICK(2)
                    if (l == 0 && has_next) {                                // second hop of the prefetch: x_j of the next tile
                        j_n = min(max(j_raw, 0), N - 1);
#pragma unroll
                        for (int a2 = 0; a2 < 3; a2++) xj_n[a2] = __ldg(xyz + (base_n + j_n) * 3 + a2);
                    }
                }
                // the chain is done; what is left costs nobody anything: h7's copy into X (warpgroup 1 is the busier one),
                // then this warpgroup's softmax half of the PREVIOUS tile
                {
                    wait_bar(accbar + 7, tp);
                    float a[16];
                    dense_act(7, a);
                    store8(COL_H / 8 + 14, a), store8(COL_H / 8 + 15, a + 8);
                    publish(xbar + 7);
                }
                if (pend) {
                    softmax_half(pend_writer, pend_dst, pz0, pz1, pz2);
                    pend = false;
                }
            } else {
                // ---- warpgroup 1: distance encoder and the shared-memory copies of the slices (B operand of t0) ----
                auto d_stage = [&](bool with_bias) {                         // d1 / d2 -> X[:, 144:208)
#pragma unroll 1
                    for (int c = 0; c < 2; c++) {
                        float a[32], b[32];                                  // big half, small half (stacked form)
                        tc::ld32(d_t0 + 32 * c, a), tc::ld32(d_t0 + 64 + 32 * c, b);
#pragma unroll
                        for (int u = 0; u < 32; u++)
                            a[u] = lrelu(fmaf(a[u] + b[u], tc::WINV, with_bias ? bd3[32 * c + u] : 0.f), S_DEF);
#pragma unroll
                        for (int u = 0; u < 4; u++) store8(COL_D / 8 + 4 * c + u, a + 8 * u);
                    }
                };
                auto copy_slice = [&](int l) {
                    wait_bar(accbar + l, tp);
                    float a[16];
                    dense_act(l, a);
                    store8(COL_H / 8 + 2 * l, a), store8(COL_H / 8 + 2 * l + 1, a + 8);
                };
                wait_bar(dbar + 0, tp);
                d_stage(false);
                publish(xbar + 0);
                if (has_next) {
                    j_n = min(max(j_raw, 0), N - 1);
#pragma unroll
                    for (int a2 = 0; a2 < 3; a2++) xj_n[a2] = __ldg(xyz + (base_n + j_n) * 3 + a2);
                }
                copy_slice(0), copy_slice(1);
                wait_bar(dbar + 1, tp);
                d_stage(true);
                publish(xbar + 1);
#pragma unroll 1
                for (int l = 2; l < 6; l++) {                               // (h6 and h7: warpgroup 0)
                    copy_slice(l);
                    publish(xbar + l);
                }
            }
            PROF_TICK(3)
            // ---- 9: t0^T: thread = channel e; write t0 back as an M-major operand (aliases X, dead now) ----
            wait_tail();
            PROF_TICK(4)
            {
                uint8_t* th = smem + T0_OFF + (uint32_t)(e >> 3) * 2048 + (uint32_t)(e & 7) * 16;
#pragma unroll 1
                for (int c = 2 * half; c < 2 * half + 2; c++) {
                    float a[32];
                    tc::ld32(d_t0 + 32 * c, a);
#pragma unroll
                    for (int u = 0; u < 32; u++) a[u] = lrelu(a[u] * tc::WINV, S_DEF);
#pragma unroll
                    for (int g = 0; g < 4; g++) {
                        uint32_t hi[4], lo[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) tc::split2(a[8 * g + 2 * u], a[8 * g + 2 * u + 1], hi[u], lo[u]);
                        *reinterpret_cast<uint4*>(th + (4 * c + g) * 128) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                        *reinterpret_cast<uint4*>(th + T0_PART + (4 * c + g) * 128) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    }
                }
            }
            publish(ready8);
            PROF_TICK(5)
            // ---- 10: t1 (thread = edge again) -> K-major operand ----
            wait_tail();
            PROF_TICK(6)
            {
                uint8_t* t1h = smem + T1_OFF + (uint32_t)(e >> 3) * T1_SBO + (uint32_t)(e & 7) * 16;
                {
                    const int c = half;
                    float a[32], b[32];
                    tc::ld32(d_a + 32 * c, a), tc::ld32(d_a + 64 + 32 * c, b);
#pragma unroll
                    for (int u = 0; u < 32; u++) a[u] = lrelu(fmaf(a[u] + b[u], tc::WINV, b3[32 * c + u]), S_DEF);
#pragma unroll
                    for (int g = 0; g < 4; g++) {
                        uint32_t hi[4], lo[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) tc::split2(a[8 * g + 2 * u], a[8 * g + 2 * u + 1], hi[u], lo[u]);
                        *reinterpret_cast<uint4*>(t1h + (4 * c + g) * 128) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                        *reinterpret_cast<uint4*>(t1h + T1_PART + (4 * c + g) * 128) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    }
                }
            }
            publish(ready8);
            if (has_next) geo_publish();                                   // the next tile's head runs under this tile's last stage
            PROF_TICK(7)
            // ---- 11: logits -> softmax over the 8 neighbours -> weighted latent sum.  Warpgroup 1 does its half of the logits
            // now; warpgroup 0 goes straight to the next tile's dense chain (whose head is already in flight) and does its half
            // in the idle time at the end of that chain -- the logits' accumulator columns are not reused before ----
            if (half == 1) {
                const float z0 = __ldg(z + row_j * 3), z1 = __ldg(z + row_j * 3 + 1), z2 = __ldg(z + row_j * 3 + 2);
                wait_tail();
                PROF_TICK(8)
                softmax_half((e & 7) == 0 && (p0 + (e >> 3)) < N, out + row_i * R * 3, z0, z1, z2);
            } else if (has_next) {
                phase ^= 1;                                                // the logits phase of mbarT, not waited for
                pend = true, pend_writer = (e & 7) == 0 && (p0 + (e >> 3)) < N, pend_dst = out + row_i * R * 3;
                pz0 = __ldg(z + row_j * 3), pz1 = __ldg(z + row_j * 3 + 1), pz2 = __ldg(z + row_j * 3 + 2);
            } else {
                const float z0 = __ldg(z + row_j * 3), z1 = __ldg(z + row_j * 3 + 1), z2 = __ldg(z + row_j * 3 + 2);
                wait_tail();
                softmax_half((e & 7) == 0 && (p0 + (e >> 3)) < N, out + row_i * R * 3, z0, z1, z2);
            }
            // no barrier at the tile end either: every MMA that reads X / t0 / t1 has completed (the logits commit was waited
            // for); the next tile's head, already issued, touches only the geo buffer, DENSE and T0
            PROF_TICK(9)
        }
        if (tid == 0) { PROF_DUMP(0, 10) }
        if (tid == 128) { PROF_DUMP(20, 10) }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem_base, TM_COLS);
}

template <int RP>
int launch(const float* xyz, const float* z, const int32_t* idx, int ldidx, const uint8_t* w, float* out, int Bp, int N,
           int R, int prec, cudaStream_t st) {
    auto kern = interp_tc_kernel<RP>;
    if (int rc = pf_smem_optin(kern, SMEM_BYTES)) return rc;
    const int tiles_per_patch = (N + PT - 1) / PT;
    const long long total = (long long)Bp * tiles_per_patch;
    int dev = 0, sms = 148;
    PF_CUDA(cudaGetDevice(&dev));
    PF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = (int)(total < sms ? total : sms);
    kern<<<grid, NT, SMEM_BYTES, st>>>(xyz, z, idx, ldidx, w, out, N, R, tiles_per_patch, total, prec);
    return pf_launch_status("interp_tc_kernel");
}

}  // namespace

#ifdef PF_PROF
extern "C" __attribute__((visibility("default"))) int pf_debug_interp_prof(long long* host32) {
    return cudaMemcpyFromSymbol(host32, g_prof, sizeof(long long) * 32) == cudaSuccess ? 0 : 1;
}
#endif

extern "C" int pf_interp_tc(const float* xyz, const float* z, const int32_t* idx, int ldidx, const void* wtc, float* out,
                            int Bp, int N, int R, int precision, void* stream) {
    PF_CHECK_ARG(precision == PF_PREC_FP32_SPLIT || precision == PF_PREC_FP16, "pf_interp_tc: precision %d (PF_PREC_FP32_SPLIT or PF_PREC_FP16)", precision);
    PF_CHECK_ARG(xyz && z && idx && wtc && out, "pf_interp_tc: null pointer");
    PF_CHECK_ARG(Bp >= 0 && N >= KN && ldidx >= KN, "pf_interp_tc: bad sizes Bp=%d N=%d ldidx=%d", Bp, N, ldidx);
    PF_CHECK_ARG(R >= 1 && R <= 32, "pf_interp_tc: R=%d outside [1, 32]", R);
    PF_CHECK_ARG((long long)Bp * N * R < 0x7fffffffLL, "pf_interp_tc: Bp*N*R overflows int32");
    PF_CHECK_ARG((reinterpret_cast<uintptr_t>(wtc) & 15) == 0, "pf_interp_tc: wtc must be 16-byte aligned");
    if (Bp == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const uint8_t* w = static_cast<const uint8_t*>(wtc);
    if (R <= 4) return launch<4>(xyz, z, idx, ldidx, w, out, Bp, N, R, precision, st);
    if (R <= 8) return launch<8>(xyz, z, idx, ldidx, w, out, Bp, N, R, precision, st);
    if (R <= 16) return launch<16>(xyz, z, idx, ldidx, w, out, Bp, N, R, precision, st);
    return launch<32>(xyz, z, idx, ldidx, w, out, Bp, N, R, precision, st);
}
