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
// x 8 neighbours; activation operand X[128 edges][K = 208] = geo(16) | dense block h (128) | distance encoder (64),
// K-major fp16 hi/lo in shared memory.  Per tile (all MMAs M = 128, issued by a dedicated warp):
//   1  d1 = lrelu(geo.Wd1g)  (N=64)        and  h0 = lrelu(geo.Wh0)         (N=16)
//   2  d2 = lrelu(d1.Wd3+bd3) (N=64, K=64) and  h1 = lrelu([geo|h0].Wh1)    (N=16, K=32)
//   3..8  h_l = lrelu([geo|h_<l].Wh_l)  (N=16, K=16+16l)
//   9  t0^T[128 ch x 128 edges] = M12g . X^T  (K=208; M12g = mlp.0 o (mlp.6 | conv_out) composed on the host, resident
//      in TMEM as the A operand); thread = channel; t0 is written back as an M-major (edge-contiguous) operand
//      -- its 13 K-steps are issued as their X columns become final, behind the MMAs of stages 3..8, so only h6 / h7 remain
//  10  t1 = lrelu(t0.W3 + b3)  (A = t0 M-major, N=64, K=128)      11  logits = t1.W6 + b6  (N=32, K=64)
//  then softmax over the 8 neighbours (8 consecutive lanes) and the weighted sum of the neighbours' latents.
// Per-stage clock64() totals (tools/dev_interp_prof.py, -DPF_INTERP_PROF) showed the MMA-issuing thread, not the tensor
// pipe, bounding the N = 16 layers (69 cycles per tcgen05.mma issued from an `if (lane == 0)` region: an ELECT/branch
// loop around every instruction), hence elect_one() and the 2-MMA split form below.
#include "common.cuh"
#include "puflow_b200.h"
#include "tc.cuh"

namespace {

using namespace pf;

constexpr int KN = 8, PT = 16, TM = 128;
constexpr int NT = 288;                                   // 2 x 4 epilogue warps (column halves of one tile) + 1 MMA-issuing warp
constexpr float S_CTX = 0.05f, S_DEF = 0.01f;             // interpflow.py:206,217 / nn.LeakyReLU() default :94,97,147,150

// ---- shared memory map (bytes) ----
constexpr uint32_t W_D1 = 0, W_D2 = 4096, W_H = 20480, W_3 = 57344, W_6 = 90112, B_D3 = 98304, B_3 = 98560, B_6 = 98816;
constexpr uint32_t W_BYTES = 98944;
__host__ __device__ constexpr uint32_t wh_off(int l) { return W_H + 1024u * (l * (l + 1) / 2); }
__host__ __device__ constexpr uint32_t wh_part(int l) { return 512u * (l + 1); }
constexpr uint32_t X_OFF = 99072, KX = 208, X_SBO = KX * 16, X_PART = 128 * KX * 2;    // 53,248 per part
constexpr uint32_t T0_OFF = X_OFF, T0_PART = 32768;       // t0 M-major [128 ch][128 edges], aliases X (dead by then)
constexpr uint32_t T1_OFF = X_OFF + 65536, T1_PART = 16384, T1_SBO = 64 * 16;
constexpr uint32_t BAR_OFF = X_OFF + 2 * X_PART;          // 205,568
constexpr uint32_t SMEM_BYTES = BAR_OFF + 64;
constexpr uint32_t COL_H = 16, COL_D = 144;               // K offsets inside X
// ---- tensor memory map (columns) ----
constexpr uint32_t TM_M12_HI = 0, TM_M12_LO = 104, TM_DA = 208, TM_DB = 336, TM_DC = 464, TM_COLS = 512;   // DA 128, DB 128, DC 32
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
    uint64_t* mbar = wbar + 1;
    uint64_t* ready = wbar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wbar + 3);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool is_mma_warp = warp == 8;
    const int e = tid & 127;          // row of the tile (TMEM lane)
    const int half = (warp >> 2) & 1;  // which half of every stage's columns this epilogue warpgroup handles

    if (tid == 0) {
        mbar_init(wbar, 1), mbar_init(mbar, 1), mbar_init(ready, 8);   // ready: one arrive per epilogue warp
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

    if (is_mma_warp) {
        tc::mbar_wait_bounded(wbar, 0);
        if (tc::elect_one()) {
            // Split products.  3-pass form (t0): al.bh, ah.bl, then ah.bh (full magnitude last, see tc.cuh).  2-MMA form for
            // the narrow layers: hi and lo of a weight matrix are adjacent 8-row groups of ONE K-major operand, so
            // ah . [bh | bl] is a single MMA of width 2n (big term in columns [0, n), ah.bl in [n, 2n)), and al . bh
            // accumulates into [n, 2n); the epilogue adds the two halves.  A third fewer instructions to issue (the
            // N = 16 layers are issue-bound, not tensor-bound) and the small terms never meet the big accumulator.
            auto issue2 = [&](uint32_t d, uint64_t ah, uint64_t al, uint32_t a_step, uint64_t bh, uint32_t b_step, int ks,
                              uint32_t idesc_2n, uint32_t idesc_n, uint32_t n) {
#pragma unroll
                for (int s = 0; s < ks; s++) tc::mma_f16(d, ah + a_step * s, bh + b_step * s, idesc_2n, s > 0);
                if (prec == PF_PREC_FP32_SPLIT) {     // reduced precision: the activations' lo products are dropped
#pragma unroll
                    for (int s = 0; s < ks; s++) tc::mma_f16(d + n, al + a_step * s, bh + b_step * s, idesc_n, 1);
                }
            };
            auto wdesc = [&](uint32_t off, uint32_t kt) { return tc::smem_desc(sbase + off, kt * 16); };
            const uint64_t xh = tc::smem_desc(x_hi, X_SBO), xl = tc::smem_desc(x_lo, X_SBO);
            const uint32_t d_a = tmem_base + TM_DA, d_b = tmem_base + TM_DB, d_c = tmem_base + TM_DC;
            constexpr uint32_t i64 = tc::idesc_f16(128, 64), i16 = tc::idesc_f16(128, 16), i128 = tc::idesc_f16(128, 128),
                               i32 = tc::idesc_f16(128, 32), AT = 1u << 15;
            uint32_t rphase = 0;
            PROF_DECL
            auto wait_ready = [&]() {
                PROF_TICK(1)
                tc::mbar_wait_bounded(ready, rphase);
                rphase ^= 1;
                tc::fence_after_sync();
                PROF_TICK(0)
            };
            // One K-step (16 columns of X) of t0^T = M12g . X^T, A from tensor memory (8 columns per K-step): the cross terms go
            // to the d-branch's accumulator (dead after stage 2), the hi . hi term to d_a; the epilogue adds the two.  The 13
            // K-steps are issued as their X columns become final, BEHIND the dense chain's own MMAs of each stage, so that
            // the tensor pipe works on them while the epilogue warps turn the chain around; only h6 / h7 are left for stage 9.
            const uint32_t mh = tmem_base + TM_M12_HI, ml = tmem_base + TM_M12_LO;
            auto t0_step = [&](int s, bool first) {
                tc::mma_f16_ts(d_b, ml + 8u * s, xh + 16u * s, i128, first ? 0u : 1u);
                if (prec == PF_PREC_FP32_SPLIT) tc::mma_f16_ts(d_b, mh + 8u * s, xl + 16u * s, i128, 1);
                tc::mma_f16_ts(d_a, mh + 8u * s, xh + 16u * s, i128, first ? 0u : 1u);
            };
            for (; t < total_tiles; t += stride) {
                // 1: d1, h0 from the geo slice
                wait_ready();
                issue2(d_b, xh, xl, 16, wdesc(W_D1, 16), 16, 1, i128, i64, 64);
                issue2(d_c, xh, xl, 16, wdesc(wh_off(0), 16), 16, 1, i32, i16, 16);
                tc::commit(mbar);
                // 2: d2 (A = X columns [144, 208)), h1
                wait_ready();
                issue2(d_b, xh + (COL_D / 8) * 8, xl + (COL_D / 8) * 8, 16, wdesc(W_D2, 64), 16, 4, i128, i64, 64);
                issue2(d_c, xh, xl, 16, wdesc(wh_off(1), 32), 16, 2, i32, i16, 16);
                tc::commit(mbar);
                // 3..8: h_l
#pragma unroll
                for (int l = 2; l < 8; l++) {
                    wait_ready();
                    issue2(d_c, xh, xl, 16, wdesc(wh_off(l), 16 + 16 * l), 16, l + 1, i32, i16, 16);
                    tc::commit(mbar);
                    // t0 K-steps whose columns are final: 0 = geo, 1 + j = h_j (published before stage j + 2), 9..12 = d2
                    // (published before this loop; d_b is free from here on)
                    if (l == 2) t0_step(0, true), t0_step(1, false);
                    else if (l < 7) t0_step(l - 1, false), t0_step(l + 6, false);
                    else t0_step(6, false);
                }
                // 9: the rest of t0^T: h6, h7
                wait_ready();
                t0_step(7, false), t0_step(8, false);
                tc::commit(mbar);
                // 10: t1 = t0 . W3^T, A = t0 M-major (8-edge groups 128 B apart, 8-channel groups 2048 B apart)
                wait_ready();
                issue2(d_a, tc::smem_desc(sbase + T0_OFF, 128, 2048), tc::smem_desc(sbase + T0_OFF + T0_PART, 128, 2048), 256,
                       wdesc(W_3, 128), 16, 8, i128 | AT, i64 | AT, 64);
                tc::commit(mbar);
                // 11: logits = t1 . W6^T
                wait_ready();
                issue2(d_a, tc::smem_desc(sbase + T1_OFF, T1_SBO), tc::smem_desc(sbase + T1_OFF + T1_PART, T1_SBO), 16,
                       wdesc(W_6, 64), 16, 4, i64, i32, 32);
                tc::commit(mbar);
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
        const uint32_t d_a = tmem_base + TM_DA + lane_sel, d_b = tmem_base + TM_DB + lane_sel, d_c = tmem_base + TM_DC + lane_sel;
        uint32_t phase = 0;
        // v[8] -> fp16 hi/lo -> 16-byte chunk `chunk` of this thread's row of X
        auto store8 = [&](uint32_t chunk, const float* v) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int u = 0; u < 4; u++) tc::split2(v[2 * u], v[2 * u + 1], hi[u], lo[u]);
            *reinterpret_cast<uint4*>(xh_p + x_row + chunk * 128) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(xl_p + x_row + chunk * 128) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        };
        // This warp's operand rows are written and its accumulator reads are done.  No block barrier: epilogue threads never
        // read each other's shared-memory writes (only the MMA does), and the MMA warp waits for all 8 warps' arrivals.
        auto publish = [&]() {
            tc::fence_async_smem();
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(ready);
        };
        auto wait_mma = [&]() {
            tc::mbar_wait_bounded(mbar, phase);
            phase ^= 1;
            tc::fence_after_sync();
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
        if (t < total_tiles) {
            base_n = tile_rows(t, p0_n, row_i_n);
            j_n = min(max(__ldg(idx + row_i_n * ldidx + (e & 7)), 0), N - 1);
#pragma unroll
            for (int a = 0; a < 3; a++) xi_n[a] = __ldg(xyz + row_i_n * 3 + a), xj_n[a] = __ldg(xyz + (base_n + j_n) * 3 + a);
        }
        PROF_DECL
        for (; t < total_tiles; t += stride) {
            const int p0 = p0_n;
            const size_t row_i = row_i_n, row_j = base_n + j_n;
            const float xi0 = xi_n[0], xi1 = xi_n[1], xi2 = xi_n[2], xj0 = xj_n[0], xj1 = xj_n[1], xj2 = xj_n[2];
            const bool has_next = t + stride < total_tiles;
            int j_raw = 0;
            if (has_next) {
                base_n = tile_rows(t + stride, p0_n, row_i_n);
                j_raw = __ldg(idx + row_i_n * ldidx + (e & 7));             // consumed after stage 2 below
#pragma unroll
                for (int a = 0; a < 3; a++) xi_n[a] = __ldg(xyz + row_i_n * 3 + a);
            }
            // ---- geo slice ----
            {
                const float dx = xi0 - xj0, dy = xi1 - xj1, dz = xi2 - xj2;
                const float dist = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
                const float g0[8] = {xi0, xi1, xi2, xj0, xj1, xj2, -dx, -dy};
                const float g1[8] = {-dz, dist, 1.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (half == 0) store8(0, g0); else store8(1, g1);
            }
            publish();
            PROF_TICK(0)
            // ---- 1: d1 -> X[:, 144:208), h0 -> X[:, 16:32) ----
            wait_mma();
            PROF_TICK(1)
            {
                const int c = half;
                float a[32], b[32];                                      // big half, small half (see issue2)
                tc::ld32(d_b + 32 * c, a), tc::ld32(d_b + 64 + 32 * c, b);
#pragma unroll
                for (int u = 0; u < 32; u++) a[u] = lrelu((a[u] + b[u]) * tc::WINV, S_DEF);
#pragma unroll
                for (int u = 0; u < 4; u++) store8(COL_D / 8 + 4 * c + u, a + 8 * u);
            }
            {
                float a[8];
                tc::ld8_sum(d_c + 8 * half, d_c + 16 + 8 * half, a);
#pragma unroll
                for (int u = 0; u < 8; u++) a[u] = lrelu(a[u] * tc::WINV, S_CTX);
                store8(COL_H / 8 + half, a);
            }
            publish();
            PROF_TICK(2)
            // ---- 2: d2 (overwrites d1), h1 ----
            wait_mma();
            PROF_TICK(3)
            {
                const int c = half;
                float a[32], b[32];
                tc::ld32(d_b + 32 * c, a), tc::ld32(d_b + 64 + 32 * c, b);
#pragma unroll
                for (int u = 0; u < 32; u++) a[u] = lrelu(fmaf(a[u] + b[u], tc::WINV, bd3[32 * c + u]), S_DEF);
#pragma unroll
                for (int u = 0; u < 4; u++) store8(COL_D / 8 + 4 * c + u, a + 8 * u);
            }
            if (has_next) {                                              // second hop of the prefetch: x_j of the next tile
                j_n = min(max(j_raw, 0), N - 1);
#pragma unroll
                for (int a = 0; a < 3; a++) xj_n[a] = __ldg(xyz + (base_n + j_n) * 3 + a);
            }
#pragma unroll 1
            for (int l = 1; l < 8; l++) {
                if (l > 1) {
                    wait_mma();
                    PROF_TICK(5)
                }
                float a[8];
                tc::ld8_sum(d_c + 8 * half, d_c + 16 + 8 * half, a);
#pragma unroll
                for (int u = 0; u < 8; u++) a[u] = lrelu(a[u] * tc::WINV, S_CTX);
                store8(COL_H / 8 + 2 * l + half, a);
                publish();
                if (l > 1) { PROF_TICK(6) } else { PROF_TICK(4) }
            }
            // ---- 9: t0^T: thread = channel e; write t0 back as an M-major operand (aliases X, dead now) ----
            wait_mma();
            PROF_TICK(7)
            {
                uint8_t* th = smem + T0_OFF + (uint32_t)(e >> 3) * 2048 + (uint32_t)(e & 7) * 16;
#pragma unroll 1
                for (int c = 2 * half; c < 2 * half + 2; c++) {
                    float a[32], b[32];                                  // hi . hi part, cross terms (t0_step)
                    tc::ld32(d_a + 32 * c, a), tc::ld32(d_b + 32 * c, b);
#pragma unroll
                    for (int u = 0; u < 32; u++) a[u] = lrelu((a[u] + b[u]) * tc::WINV, S_DEF);
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
            publish();
            PROF_TICK(8)
            // ---- 10: t1 (thread = edge again) -> K-major operand ----
            wait_mma();
            PROF_TICK(9)
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
            publish();
            PROF_TICK(10)
            // ---- 11: logits -> softmax over the 8 neighbours (8 consecutive lanes) -> weighted latent sum ----
            const float z0 = __ldg(z + row_j * 3), z1 = __ldg(z + row_j * 3 + 1), z2 = __ldg(z + row_j * 3 + 2);
            wait_mma();
            PROF_TICK(11)
            {
                constexpr int RH = RP / 2;                                 // logits per warpgroup: r in [r0, r0 + RH)
                const int r0 = half * RH;
                float lg[RH];
                if constexpr (RH == 16) {
                    float a16[16], s16[16];
                    tc::ld16(d_a + r0, a16), tc::ld16(d_a + 32 + r0, s16);
#pragma unroll
                    for (int r = 0; r < RH; r++) lg[r] = a16[r] + s16[r];
                } else {
                    float a8[8], s8[8];
                    tc::ld8(d_a + r0, a8), tc::ld8(d_a + 32 + r0, s8);
#pragma unroll
                    for (int r = 0; r < RH; r++) lg[r] = a8[r] + s8[r];
                }
                const bool writer = (e & 7) == 0 && (p0 + (e >> 3)) < N;
                float* dst = out + row_i * R * 3;
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
            }
            // no barrier at the tile end either: every MMA that reads X / t0 / t1 has completed (the logits commit was waited
            // for), and the next tile's first MMA is issued only after all 8 warps have arrived again
            PROF_TICK(12)
        }
        if (tid == 0) { PROF_DUMP(0, 13) }
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
