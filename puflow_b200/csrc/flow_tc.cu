// Glow-style flow stack on the 5th-gen tensor cores: same contract as flow.cu (pf_flow_fwd / pf_flow_inv /
// pf_flow_cond; replaces FlowBlock.forward / inverse and PointInterpFlow.f / g,
// /root/reference/modules/discrete/interpflow.py:66-82,302-321; LinearA1D :22-43; coupling.py:55-58,82-85,108-151).
//
// ncu on the FFMA kernels (profiles/r01_ncu_v3_tc.md): FP32 FMA pipe 40-42 % busy, DRAM 3-11 %: they are bound by the
// 64x64 hidden layer of LinearA1D (4,096 of the ~4,500 FMAs per point per block).  Here that layer runs as
// tcgen05.mma.kind::f16 on fp16 hi/lo split operands (3 MMAs per product, fp32 accumulation in TMEM: fp32-class
// result, tolerance stated in tests/test_gpu_kernels.py), everything else stays per-thread fp32 with the same
// operation order as flow.cu.
//
// A warpgroup (128 threads) owns a 128-point tile: thread = point = TMEM lane, the point's 3 coordinates stay in
// registers through all blocks.  Per block: layer 0 (conditioning part from `raw`, one 256-bit load per 8 channels,
// + W1h . h1) -> LeakyReLU -> fp16 hi/lo -> this thread's row of the A operand in shared memory; one thread issues
// 12 MMAs (M=128, N=64, K=64, x3 split terms); tcgen05.ld of the thread's 64 accumulators -> +b2, LeakyReLU, the
// 64 -> (3 - tdim) output layer as per-thread FMAs.  NWG warpgroups per CTA work on independent tiles, so one
// group's MMA / TMEM latency is covered by the others' CUDA-core work.  W2 (hi/lo) of all blocks + the small fp32
// parameters arrive with one TMA bulk copy per (persistent) CTA.
#include "common.cuh"
#include "puflow_b200.h"
#include "tc.cuh"

namespace {

using namespace pf;

constexpr int H = 64, TP = 128;
constexpr float SLOPE = 0.01f;                 // nn.LeakyReLU() default (interpflow.py:34,36)
constexpr int MAXB = 8;
constexpr uint32_t W_PART = H * H * 2;         // one 64x64 fp16 operand
constexpr uint32_t A_PART = TP * H * 2, A_WG = 2 * A_PART;     // A[hi|lo] of one warpgroup: 128 rows x 64 halfs each
constexpr uint32_t A_SBO = H * 16;             // 8-row groups are K*16 bytes apart (tc.cuh layout)

__device__ __forceinline__ void ldg256(const float* p, float (&v)[8]) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(p));
}

// Per-warpgroup state of the "64 -> 64 hidden layer on a 128-row tile" building block.
struct Wg {
    uint8_t* a_row_hi;      // this thread's row of the hi operand (chunk 0); lo is A_PART further; chunk c at + c * 128
    uint32_t a_hi;          // shared-memory address of the hi operand
    uint32_t d_tmem;        // accumulator columns of this warpgroup
    uint32_t a_tmem;        // flow kernels: the A operand lives in tensor memory (hi: 32 columns, lo: the next 32)
    uint32_t lane_sel;      // TMEM lane quadrant of this warp
    uint64_t* mbar;         // MMA-done barrier of this warpgroup
    uint32_t phase;
    int bar_id, e;
    int prec;               // PF_PREC_FP32_SPLIT: three split products; PF_PREC_FP16: the hi . hi product alone
};

// 64 activations (already through LeakyReLU) -> this thread's operand row.  fp16 saturates at 65,504, and exp(s) of the
// affine injector can make intermediate coordinates (hence activations) arbitrarily large, so a row whose largest
// magnitude exceeds 2^14 is scaled by a power of two (exact) before the split; the caller multiplies the row's
// accumulators by the returned factor (the layer is linear), which also takes out the weights' 2^10 (tc::WINV).
__device__ __forceinline__ float store_row(const Wg& g, float (&h)[64]) {
    float amax = 0.f;
#pragma unroll
    for (int k = 0; k < 64; k++) amax = fmaxf(amax, fabsf(h[k]));
    const int ex = (__float_as_int(amax) >> 23) & 0xff;       // biased exponent (NaN/Inf rows stay NaN/Inf)
    float inv = tc::WINV;                                     // the weights carry a factor 1 / WINV (tc.cuh)
    if (ex > 127 + 13) {
        const int k = min(ex - 140, 120);
        const float sc = __int_as_float((127 - k) << 23);
        inv = __int_as_float((127 + k) << 23) * tc::WINV;
#pragma unroll
        for (int i = 0; i < 64; i++) h[i] *= sc;
    }
#pragma unroll
    for (int c = 0; c < 8; c++) {
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int u = 0; u < 4; u++) tc::split2(h[8 * c + 2 * u], h[8 * c + 2 * u + 1], hi[u], lo[u]);
        *reinterpret_cast<uint4*>(g.a_row_hi + c * 128) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(g.a_row_hi + A_PART + c * 128) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    return inv;
}

// acc[64] = A[row of this thread] . W^T, W (hi at w_addr, lo W_PART further) in the tc.cuh operand layout.
// All 128 threads of the warpgroup call this after store_row().
__device__ __forceinline__ void hidden_mma(Wg& g, uint32_t w_addr, float (&acc)[64]) {
    tc::fence_async_smem();
    tc::fence_before_sync();
    tc::named_bar(g.bar_id, 128);        // operand rows written; previous accumulators read by everyone
    if (g.e < 32 && tc::elect_one()) {       // first warp of the warpgroup, one elected lane (tc.cuh)
        tc::fence_after_sync();
        const uint32_t idesc = tc::idesc_f16(128, H);
        const uint64_t da_hi = tc::smem_desc(g.a_hi, A_SBO), da_lo = tc::smem_desc(g.a_hi + A_PART, A_SBO);
        const uint64_t db_hi = tc::smem_desc(w_addr, H * 16), db_lo = tc::smem_desc(w_addr + W_PART, H * 16);
        if (g.prec == PF_PREC_FP32_SPLIT) {
#pragma unroll
            for (int s = 0; s < H / 16; s++) tc::mma_f16(g.d_tmem, da_lo + 16u * s, db_hi + 16u * s, idesc, s > 0);
#pragma unroll
            for (int s = 0; s < H / 16; s++) tc::mma_f16(g.d_tmem, da_hi + 16u * s, db_lo + 16u * s, idesc, 1);
        }
#pragma unroll
        for (int s = 0; s < H / 16; s++)       // full magnitude last
            tc::mma_f16(g.d_tmem, da_hi + 16u * s, db_hi + 16u * s, idesc, (g.prec == PF_PREC_FP32_SPLIT || s > 0) ? 1u : 0u);
        tc::commit(g.mbar);
    }
    tc::mbar_wait_bounded(g.mbar, g.phase);
    g.phase ^= 1;
    tc::fence_after_sync();
    float lo32[32], hi32[32];
    tc::ld32(g.d_tmem + g.lane_sel, lo32);
    tc::ld32(g.d_tmem + g.lane_sel + 32, hi32);
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = lo32[i], acc[32 + i] = hi32[i];
}

// The same two steps with the A operand in TENSOR memory (flow forward / inverse; round 2).  ncu on the shared-memory
// form (profiles/r02_ncu_step_full.md): L1TEX data pipe 66-70 % busy, tensor pipe 11-19 %: every 64-wide activation row was
// written to shared memory (16 STS.128 per thread), made visible to the async proxy, and read back by the 12 MMAs' A
// fetches.  tcgen05.st puts the row into this thread's own TMEM lane (2 fp16 per 32-bit column, K order) instead: no
// shared-memory traffic for A at all, no proxy fence.
__device__ __forceinline__ float store_row_t(const Wg& g, float (&h)[64]) {
    float amax = 0.f;
#pragma unroll
    for (int k = 0; k < 64; k++) amax = fmaxf(amax, fabsf(h[k]));
    const int ex = (__float_as_int(amax) >> 23) & 0xff;       // biased exponent (NaN/Inf rows stay NaN/Inf)
    float inv = tc::WINV;                                     // the weights carry a factor 1 / WINV (tc.cuh)
    if (ex > 127 + 13) {
        const int k = min(ex - 140, 120);
        const float sc = __int_as_float((127 - k) << 23);
        inv = __int_as_float((127 + k) << 23) * tc::WINV;
#pragma unroll
        for (int i = 0; i < 64; i++) h[i] *= sc;
    }
#pragma unroll
    for (int q = 0; q < 2; q++) {
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int u = 0; u < 16; u++) tc::split2(h[32 * q + 2 * u], h[32 * q + 2 * u + 1], hi[u], lo[u]);
        tc::st16_nowait(g.a_tmem + g.lane_sel + 16 * q, hi);
        tc::st16_nowait(g.a_tmem + g.lane_sel + 32 + 16 * q, lo);
    }
    return inv;
}
__device__ __forceinline__ void hidden_mma_t(Wg& g, uint32_t w_addr, float (&acc)[64]) {
    tc::tmem_st_wait();
    tc::fence_before_sync();
    tc::named_bar(g.bar_id, 128);        // operand rows written; previous accumulators read by everyone
    if (g.e < 32 && tc::elect_one()) {
        tc::fence_after_sync();
        const uint32_t idesc = tc::idesc_f16(128, H);
        const uint32_t a_hi = g.a_tmem, a_lo = g.a_tmem + 32;
        const uint64_t db_hi = tc::smem_desc(w_addr, H * 16), db_lo = tc::smem_desc(w_addr + W_PART, H * 16);
        if (g.prec == PF_PREC_FP32_SPLIT) {
#pragma unroll
            for (int s = 0; s < H / 16; s++) tc::mma_f16_ts(g.d_tmem, a_lo + 8u * s, db_hi + 16u * s, idesc, s > 0);
#pragma unroll
            for (int s = 0; s < H / 16; s++) tc::mma_f16_ts(g.d_tmem, a_hi + 8u * s, db_lo + 16u * s, idesc, 1);
        }
#pragma unroll
        for (int s = 0; s < H / 16; s++)       // full magnitude last
            tc::mma_f16_ts(g.d_tmem, a_hi + 8u * s, db_hi + 16u * s, idesc, (g.prec == PF_PREC_FP32_SPLIT || s > 0) ? 1u : 0u);
        tc::commit(g.mbar);
    }
    tc::mbar_wait_bounded(g.mbar, g.phase);
    g.phase ^= 1;
    tc::fence_after_sync();
    float lo32[32], hi32[32];
    tc::ld32(g.d_tmem + g.lane_sel, lo32);
    tc::ld32(g.d_tmem + g.lane_sel + 32, hi32);
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = lo32[i], acc[32 + i] = hi32[i];
}

__device__ __forceinline__ void wg_init(Wg& g, uint8_t* smem, uint32_t a_off, uint64_t* mbars, uint32_t tmem_base,
                                        int tid, int prec) {
    g.prec = prec;
    const int warp = tid >> 5, wg = warp >> 2;
    g.e = tid & 127;
    uint8_t* a = smem + a_off + (size_t)wg * A_WG;
    g.a_hi = smem_u32(a);
    g.a_row_hi = a + (uint32_t)(g.e >> 3) * A_SBO + (uint32_t)(g.e & 7) * 16;
    g.d_tmem = tmem_base + (uint32_t)wg * H;
    g.lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    g.mbar = mbars + wg;
    g.phase = 0;
    g.bar_id = 1 + wg;
}

// ---- flow forward / inverse --------------------------------------------------------------------------------
// tc blob per block (PF_FLOW_TC_BLOCK bytes): W2 hi [64 out][64 in] | W2 lo | the first 384 floats of the fp32 flow
// blob (flow.cu header: actnorm, W, Winv, tdim, reverse, w1h, b2, w4, b4)
constexpr int NWG = 3, NT = NWG * 128;
constexpr uint32_t FBLK = PF_FLOW_TC_BLOCK;

template <bool INVERSE>
__global__ void __launch_bounds__(NT, 1)
    flow_tc_kernel(const float* __restrict__ xin, const float* __restrict__ raw, const float* __restrict__ sb,
                   const uint8_t* __restrict__ wtc, float* __restrict__ xout, float* __restrict__ ssum, int64_t M,
                   int R, int nblocks, int ldraw, int prec) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t w_bytes = (uint32_t)nblocks * FBLK;
    const uint32_t a_off = (w_bytes + 127u) & ~127u;
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + a_off);      // no operand tile: A lives in tensor memory
    uint64_t* mbars = wbar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbars + NWG);

    const int tid = threadIdx.x, warp = tid >> 5, wg = warp >> 2;
    if (tid == 0) {
        mbar_init(wbar, 1);
        for (int i = 0; i < NWG; i++) mbar_init(mbars + i, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    if (tid == 0) {
        mbar_expect_tx(wbar, w_bytes);
        bulk_g2s_chunked(smem, wtc, w_bytes, wbar);
    }
    Wg g;
    wg_init(g, smem, a_off, mbars, *tmem_slot, tid, prec);
    g.d_tmem = *tmem_slot + (uint32_t)wg * 128;      // per warpgroup: 64 accumulator columns, then the A operand (hi 32 | lo 32)
    g.a_tmem = g.d_tmem + 64;
    const uint32_t sbase = smem_u32(smem);
    const int64_t rows = INVERSE ? M * R : M;
    const int64_t ntiles = (rows + TP - 1) / TP;
    tc::mbar_wait_bounded(wbar, 0);

    for (int64_t t = (int64_t)blockIdx.x * NWG + wg; t < ntiles; t += (int64_t)gridDim.x * NWG) {
        const int64_t r = t * TP + g.e;
        const int64_t row = min(r, rows - 1);
        const int64_t m = INVERSE ? row / R : row;             // input point that conditions this row
        float x0, x1, x2, ss = 0.f;
        {
            const float* s = xin + (size_t)row * 3;
            x0 = s[0], x1 = s[1], x2 = s[2];
        }
#pragma unroll 1
        for (int it = 0; it < nblocks; it++) {
            const int blk = INVERSE ? nblocks - 1 - it : it;
            const float* fw = reinterpret_cast<const float*>(smem + (size_t)blk * FBLK + 2 * W_PART);
            const int tdim = (int)fw[27];
            const bool rev = fw[28] != 0.f;
            const float* rawrow = raw + ((size_t)blk * M + m) * ldraw;
            const float4 v0 = __ldg(reinterpret_cast<const float4*>(sb + ((size_t)blk * M + m) * 8));
            const float4 v1 = __ldg(reinterpret_cast<const float4*>(sb + ((size_t)blk * M + m) * 8) + 1);
            const float s0 = v0.x, s1 = v0.y, s2 = v0.z, b0 = v0.w, b1 = v1.x, b2 = v1.y;
            if (!INVERSE) {
                // ActNorm.forward: x * exp(logs) + bias ; inv1x1.forward: z_i = sum_j W[i][j] x_j
                const float a0 = fmaf(x0, fw[0], fw[3]), a1 = fmaf(x1, fw[1], fw[4]), a2 = fmaf(x2, fw[2], fw[5]);
                x0 = fw[9] * a0 + fw[10] * a1 + fw[11] * a2;
                x1 = fw[12] * a0 + fw[13] * a1 + fw[14] * a2;
                x2 = fw[15] * a0 + fw[16] * a1 + fw[17] * a2;
            } else {
                // coupling2.inverse: z * exp(s) + b ; permutate2.inverse: reverse
                const float a0 = fmaf(x0, expf(s0), b0), a1 = fmaf(x1, expf(s1), b1), a2 = fmaf(x2, expf(s2), b2);
                x0 = rev ? a2 : a0, x1 = a1, x2 = rev ? a0 : a2;
            }
            // layer 0: a1 = lrelu(u1 + W1h . h1)   (W1h rows beyond tdim are zero)
            const float h1b = (tdim == 2) ? x1 : 0.f;
            float inv;
            {
                float h[64];
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    float u[8];
                    ldg256(rawrow + 8 * c, u);
#pragma unroll
                    for (int k = 0; k < 8; k++) h[8 * c + k] = u[k];
                }
#pragma unroll
                for (int c = 0; c < 16; c++) {
                    const float4 wa = *reinterpret_cast<const float4*>(fw + 32 + 4 * c);
                    const float4 wb = *reinterpret_cast<const float4*>(fw + 96 + 4 * c);
                    h[4 * c + 0] = lrelu(fmaf(wb.x, h1b, fmaf(wa.x, x0, h[4 * c + 0])), SLOPE);
                    h[4 * c + 1] = lrelu(fmaf(wb.y, h1b, fmaf(wa.y, x0, h[4 * c + 1])), SLOPE);
                    h[4 * c + 2] = lrelu(fmaf(wb.z, h1b, fmaf(wa.z, x0, h[4 * c + 2])), SLOPE);
                    h[4 * c + 3] = lrelu(fmaf(wb.w, h1b, fmaf(wa.w, x0, h[4 * c + 3])), SLOPE);
                }
                inv = store_row_t(g, h);
            }
            // layer 2 on the tensor cores, layer 4 per thread
            float acc[64];
            hidden_mma_t(g, sbase + (uint32_t)blk * FBLK, acc);
            float o0 = fw[352], o1 = fw[353];
#pragma unroll
            for (int k = 0; k < H; k++) {
                const float a = lrelu(fmaf(acc[k], inv, fw[160 + k]), SLOPE);
                const float2 w = *reinterpret_cast<const float2*>(fw + 224 + 2 * k);
                o0 = fmaf(a, w.x, o0), o1 = fmaf(a, w.y, o1);
            }
            if (!INVERSE) {
                // coupling1.forward (additive): h2 -= bias_net ; reverse ; coupling2.forward
                if (tdim == 1) x1 -= o0, x2 -= o1;
                else x2 -= o0;
                const float a0 = rev ? x2 : x0, a1 = x1, a2 = rev ? x0 : x2;
                x0 = (a0 - b0) * expf(-s0);
                x1 = (a1 - b1) * expf(-s1);
                x2 = (a2 - b2) * expf(-s2);
                ss += (s0 + s1) + s2;
            } else {
                // coupling1.inverse: h2 += bias_net ; inv1x1.inverse ; ActNorm.inverse
                if (tdim == 1) x1 += o0, x2 += o1;
                else x2 += o0;
                const float a0 = fw[18] * x0 + fw[19] * x1 + fw[20] * x2;
                const float a1 = fw[21] * x0 + fw[22] * x1 + fw[23] * x2;
                const float a2 = fw[24] * x0 + fw[25] * x1 + fw[26] * x2;
                x0 = (a0 - fw[3]) * fw[6];
                x1 = (a1 - fw[4]) * fw[7];
                x2 = (a2 - fw[5]) * fw[8];
            }
        }
        if (r < rows) {
            float* d = xout + (size_t)r * 3;
            d[0] = x0, d[1] = x1, d[2] = x2;
            if (!INVERSE && ssum) ssum[r] = ss;
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(*tmem_slot, 512);
}

// ---- conditioning: coupling2.scale_net / bias_net hidden + output layers per input point ------------------------
// tc blob per block (PF_COND_TC_BLOCK bytes): Ws2 hi | Ws2 lo | Wb2 hi | Wb2 lo | floats 8192.. of the fp32 cond blob
// (bs2[64] | bb2[64] | ws4[64][4] | wb4[64][4] | bs4[4] | bb4[4])
constexpr int NWG_C = 4, NT_C = NWG_C * 128;
constexpr uint32_t CBLK = PF_COND_TC_BLOCK;

__global__ void __launch_bounds__(NT_C, 1)
    flow_cond_tc_kernel(const float* __restrict__ raw, const uint8_t* __restrict__ wtc, float* __restrict__ sb,
                        int64_t M, int prec) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr uint32_t a_off = (CBLK + 127u) & ~127u;
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + a_off + NWG_C * A_WG);
    uint64_t* mbars = wbar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbars + NWG_C);

    const int tid = threadIdx.x, warp = tid >> 5, wg = warp >> 2;
    const int blk = blockIdx.y;
    if (tid == 0) {
        mbar_init(wbar, 1);
        for (int i = 0; i < NWG_C; i++) mbar_init(mbars + i, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 256);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    if (tid == 0) {
        mbar_expect_tx(wbar, CBLK);
        bulk_g2s_chunked(smem, wtc + (size_t)blk * CBLK, CBLK, wbar);
    }
    Wg g;
    wg_init(g, smem, a_off, mbars, *tmem_slot, tid, prec);
    const uint32_t sbase = smem_u32(smem);
    const float* cw = reinterpret_cast<const float*>(smem + 4 * W_PART);
    const int64_t ntiles = (M + TP - 1) / TP;
    tc::mbar_wait_bounded(wbar, 0);

    for (int64_t t = (int64_t)blockIdx.x * NWG_C + wg; t < ntiles; t += (int64_t)gridDim.x * NWG_C) {
        const int64_t r = t * TP + g.e;
        const int64_t m = min(r, M - 1);
        const float* rawrow = raw + ((size_t)blk * M + m) * 192 + 64;
        float o[2][3];
#pragma unroll
        for (int net = 0; net < 2; net++) {
            float inv;
            {
                float h[64];
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    float u[8];
                    ldg256(rawrow + net * 64 + 8 * c, u);
#pragma unroll
                    for (int k = 0; k < 8; k++) h[8 * c + k] = lrelu(u[k], SLOPE);
                }
                inv = store_row(g, h);
            }
            float acc[64];
            hidden_mma(g, sbase + (uint32_t)net * 2 * W_PART, acc);
            const float* b2 = cw + net * 64;
            const float* w4 = cw + 128 + net * 256;
            const float* b4 = cw + 640 + net * 4;
            float o0 = b4[0], o1 = b4[1], o2 = b4[2];
#pragma unroll
            for (int k = 0; k < H; k++) {
                const float a = lrelu(fmaf(acc[k], inv, b2[k]), SLOPE);
                const float4 w = *reinterpret_cast<const float4*>(w4 + 4 * k);
                o0 = fmaf(a, w.x, o0), o1 = fmaf(a, w.y, o1), o2 = fmaf(a, w.z, o2);
            }
            o[net][0] = o0, o[net][1] = o1, o[net][2] = o2;
        }
        if (r < M) {
            float4* dst = reinterpret_cast<float4*>(sb + ((size_t)blk * M + r) * 8);
            dst[0] = make_float4(o[0][0], o[0][1], o[0][2], o[1][0]);
            dst[1] = make_float4(o[1][1], o[1][2], 0.f, 0.f);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(*tmem_slot, 256);
}

// ---- fused conditioning chain per block: FeatMergeUnit + the conditioning first layers + the coupling2 nets ------
//   c[M][C]  ->  h = relu(c.W1 + b1) [C/2]  ->  raw = h.(W2.[W1c|Ws0|Wb0]) [192]   (conv2 has no bias and no
//   activation, interpflow.py:251-258, so it is composed with the first layers of the three LinearA1D nets on the host)
//   -> raw1[M][64] = raw[:, 0:64] (what coupling1 needs) and sb[M][8] from raw[:, 64:192] as in flow_cond_tc_kernel.
// One kernel per block instead of pf_linear x3 + pf_flow_cond: c is read once, cs / raw never reach HBM.
// chain blob (bytes): W1 hi|lo [C/2 out][C in] | W2c hi|lo [192 out][C/2 in] | Ws2 hi|lo | Wb2 hi|lo |
//                     fp32: b1[64] | cond tail (bs2[64] bb2[64] ws4[64][4] wb4[64][4] bs4[4] bb4[4])
__device__ __forceinline__ void stg256(float* p, const float (&v)[8]) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
                 "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
                 : "memory");
}
// operands published -> one thread issues the 3 split products D[:, 0:n) = A[:, 0:16 ks] . W^T -> commit; everybody waits.
// A in tensor memory at a_hi / a_lo (2 fp16 per column, 8 columns per K-step), see store_row_t above.
__device__ __forceinline__ void issue3_t(Wg& g, uint32_t d_tmem, uint32_t a_hi, uint32_t a_lo, uint32_t w_addr, uint32_t w_part,
                                         uint32_t w_sbo, int ks, int n) {
    tc::tmem_st_wait();
    tc::fence_before_sync();
    tc::named_bar(g.bar_id, 128);
    if (g.e < 32 && tc::elect_one()) {
        tc::fence_after_sync();
        const uint32_t idesc = tc::idesc_f16(128, n);
        const uint64_t db_hi = tc::smem_desc(w_addr, w_sbo), db_lo = tc::smem_desc(w_addr + w_part, w_sbo);
        const bool split = g.prec == PF_PREC_FP32_SPLIT;
        if (split) {
#pragma unroll 1
            for (int s = 0; s < ks; s++) tc::mma_f16_ts(d_tmem, a_lo + 8u * s, db_hi + 16u * s, idesc, s > 0 ? 1u : 0u);
#pragma unroll 1
            for (int s = 0; s < ks; s++) tc::mma_f16_ts(d_tmem, a_hi + 8u * s, db_lo + 16u * s, idesc, 1);
        }
#pragma unroll 1
        for (int s = 0; s < ks; s++) tc::mma_f16_ts(d_tmem, a_hi + 8u * s, db_hi + 16u * s, idesc, (split || s > 0) ? 1u : 0u);
        tc::commit(g.mbar);
    }
    tc::mbar_wait_bounded(g.mbar, g.phase);
    g.phase ^= 1;
    tc::fence_after_sync();
}
// 16 values (K-steps' worth) of this thread's operand row -> columns [8 * cp, 8 * cp + 8) of a_hi / a_lo, plain saturating split
__device__ __forceinline__ void store_pair_t(const Wg& g, uint32_t a_hi, uint32_t a_lo, int cp, const float (&u)[8], const float (&v)[8]) {
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 4; i++) tc::split2(u[2 * i], u[2 * i + 1], hi[i], lo[i]), tc::split2(v[2 * i], v[2 * i + 1], hi[4 + i], lo[4 + i]);
    tc::st8_nowait(a_hi + g.lane_sel + 8 * cp, hi);
    tc::st8_nowait(a_lo + g.lane_sel + 8 * cp, lo);
}

constexpr int NWG_CH = 2, NT_CH = NWG_CH * 128;
__host__ __device__ constexpr uint32_t chain_w1_part(int C) { return (uint32_t)C * C; }          // (C/2) x C halfs
__host__ __device__ constexpr uint32_t chain_w2_part(int C) { return 192u * C; }                 // 192 x (C/2) halfs
__host__ __device__ constexpr uint32_t chain_bytes(int C) {
    return 2 * chain_w1_part(C) + 2 * chain_w2_part(C) + 4 * W_PART + (64 + 648) * 4;
}

template <int C>
__global__ void __launch_bounds__(NT_CH, 1)
    cond_chain_tc_kernel(const float* __restrict__ cin, const uint8_t* __restrict__ wtc, float* __restrict__ raw1,
                         float* __restrict__ sb, int64_t M, int prec) {
    constexpr int H1 = C / 2;
    constexpr uint32_t W1P = chain_w1_part(C), W2P = chain_w2_part(C), WBYTES = chain_bytes(C);
    constexpr uint32_t OFF_W2 = 2 * W1P, OFF_NET = OFF_W2 + 2 * W2P, OFF_F = OFF_NET + 4 * W_PART;
    constexpr uint32_t a_off = (WBYTES + 127u) & ~127u;
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + a_off);      // no operand tiles: every A operand lives in tensor memory
    uint64_t* mbars = wbar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbars + NWG_CH);

    const int tid = threadIdx.x, warp = tid >> 5, wg = warp >> 2;
    if (tid == 0) {
        mbar_init(wbar, 1);
        for (int i = 0; i < NWG_CH; i++) mbar_init(mbars + i, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    if (tid == 0) {
        mbar_expect_tx(wbar, WBYTES);
        bulk_g2s_chunked(smem, wtc, WBYTES, wbar);
    }
    Wg g;
    wg_init(g, smem, a_off, mbars, *tmem_slot, tid, prec);
    g.d_tmem = *tmem_slot + (uint32_t)wg * 256;          // per warpgroup: raw at +0 (192 columns), scratch at +192 (64)
    const uint32_t d_raw = g.d_tmem, d_s = g.d_tmem + 192;
    // A operands alias accumulator columns that are dead at that point: stage 1 (K = C) in raw[0:128) (raw is produced by
    // stage 2), stage 2 in the scratch columns (stage 1's accumulators are in registers by then), stage 3 in raw[0:64) (already
    // stored to HBM).  A warp only ever touches its own 32 lanes, so warps of a group need no ordering among themselves.
    const uint32_t a1_hi = d_raw, a1_lo = d_raw + 64;
    const uint32_t sbase = smem_u32(smem);
    const float* fb1 = reinterpret_cast<const float*>(smem + OFF_F);
    const float* cw = fb1 + 64;
    const int64_t ntiles = (M + TP - 1) / TP;
    tc::mbar_wait_bounded(wbar, 0);

    for (int64_t t = (int64_t)blockIdx.x * NWG_CH + wg; t < ntiles; t += (int64_t)gridDim.x * NWG_CH) {
        const int64_t r = t * TP + g.e;
        const float* crow = cin + (size_t)min(r, M - 1) * C;
        // ---- stage 1: h = relu(c . W1 + b1), K = C in slices of at most 64 through the 64-wide operand buffer ----
#pragma unroll
        for (int cp = 0; cp < C / 16; cp++) {
            float u[8], v[8];
            ldg256(crow + 16 * cp, u), ldg256(crow + 16 * cp + 8, v);
            store_pair_t(g, a1_hi, a1_lo, cp, u, v);
        }
        issue3_t(g, d_s, a1_hi, a1_lo, sbase, W1P, C * 16, C / 16, H1);
        float inv;
        {
            float h[64];
            if (H1 == 64) {
                float lo32[32], hi32[32];
                tc::ld32(d_s + g.lane_sel, lo32), tc::ld32(d_s + g.lane_sel + 32, hi32);
#pragma unroll
                for (int i = 0; i < 32; i++) h[i] = lo32[i], h[32 + i] = hi32[i];
            } else if (H1 == 32) {
                float lo32[32];
                tc::ld32(d_s + g.lane_sel, lo32);
#pragma unroll
                for (int i = 0; i < 32; i++) h[i] = lo32[i], h[32 + i] = 0.f;
            } else {
                float lo16[16];
                tc::ld16(d_s + g.lane_sel, lo16);
#pragma unroll
                for (int i = 0; i < 64; i++) h[i] = i < 16 ? lo16[i < 16 ? i : 0] : 0.f;
            }
#pragma unroll
            for (int i = 0; i < H1; i++) h[i] = fmaxf(fmaf(h[i], tc::WINV, fb1[i]), 0.f);
            g.a_tmem = d_s;
            inv = store_row_t(g, h);                    // columns beyond H1 are zero and never read by the MMAs
        }
        // ---- stage 2: raw = h . W2c  (N = 192) ----
        issue3_t(g, d_raw, d_s, d_s + 32, sbase + OFF_W2, W2P, H1 * 16, H1 / 16, 192);
        {
            float* dst = raw1 + (size_t)r * 64;
#pragma unroll
            for (int q = 0; q < 2; q++) {
                float v[32];
                tc::ld32(d_raw + g.lane_sel + 32 * q, v);
                if (r < M) {
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const float o[8] = {v[8 * c] * inv, v[8 * c + 1] * inv, v[8 * c + 2] * inv, v[8 * c + 3] * inv,
                                            v[8 * c + 4] * inv, v[8 * c + 5] * inv, v[8 * c + 6] * inv, v[8 * c + 7] * inv};
                        stg256(dst + 32 * q + 8 * c, o);
                    }
                }
            }
        }
        // ---- stage 3: the two coupling2 nets on raw[:, 64:128] / raw[:, 128:192] ----
        float o[2][3];
#pragma unroll
        for (int net = 0; net < 2; net++) {
            float inv2;
            {
                float h[64], lo32[32], hi32[32];
                tc::ld32(d_raw + g.lane_sel + 64 + 64 * net, lo32), tc::ld32(d_raw + g.lane_sel + 96 + 64 * net, hi32);
#pragma unroll
                for (int i = 0; i < 32; i++) h[i] = lrelu(lo32[i] * inv, SLOPE), h[32 + i] = lrelu(hi32[i] * inv, SLOPE);
                g.a_tmem = d_raw;
                inv2 = store_row_t(g, h);
            }
            issue3_t(g, d_s, d_raw, d_raw + 32, sbase + OFF_NET + (uint32_t)net * 2 * W_PART, W_PART, H * 16, H / 16, H);
            float acc[64];
            {
                float lo32[32], hi32[32];
                tc::ld32(d_s + g.lane_sel, lo32), tc::ld32(d_s + g.lane_sel + 32, hi32);
#pragma unroll
                for (int i = 0; i < 32; i++) acc[i] = lo32[i], acc[32 + i] = hi32[i];
            }
            const float* b2 = cw + net * 64;
            const float* w4 = cw + 128 + net * 256;
            const float* b4 = cw + 640 + net * 4;
            float o0 = b4[0], o1 = b4[1], o2 = b4[2];
#pragma unroll
            for (int k = 0; k < H; k++) {
                const float a = lrelu(fmaf(acc[k], inv2, b2[k]), SLOPE);
                const float4 w = *reinterpret_cast<const float4*>(w4 + 4 * k);
                o0 = fmaf(a, w.x, o0), o1 = fmaf(a, w.y, o1), o2 = fmaf(a, w.z, o2);
            }
            o[net][0] = o0, o[net][1] = o1, o[net][2] = o2;
        }
        if (r < M) {
            float4* dst = reinterpret_cast<float4*>(sb + (size_t)r * 8);
            dst[0] = make_float4(o[0][0], o[0][1], o[0][2], o[1][0]);
            dst[1] = make_float4(o[1][1], o[1][2], 0.f, 0.f);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(*tmem_slot, 512);
}

template <int C>
int chain_launch(const float* c, const uint8_t* w, float* raw1, float* sb, int64_t M, int sms, int prec, cudaStream_t st) {
    constexpr size_t smem = ((chain_bytes(C) + 127u) & ~127u) + 64;      // weights + barriers
    if (int rc = pf_smem_optin(cond_chain_tc_kernel<C>, smem)) return rc;
    const int64_t ntiles = (M + TP - 1) / TP;
    int grid = sms;
    if ((int64_t)grid * NWG_CH > ntiles) grid = (int)((ntiles + NWG_CH - 1) / NWG_CH);
    cond_chain_tc_kernel<C><<<grid, NT_CH, smem, st>>>(c, w, raw1, sb, M, prec);
    return pf_launch_status("cond_chain_tc_kernel");
}

int sm_count() {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
}

}  // namespace

#define PF_CHECK_PREC(p, who) \
    PF_CHECK_ARG((p) == PF_PREC_FP32_SPLIT || (p) == PF_PREC_FP16, who ": precision %d (PF_PREC_FP32_SPLIT or PF_PREC_FP16)", (p))

extern "C" int pf_flow_cond_tc(const float* raw, const void* cond_tc, float* sb, int64_t M, int nblocks, int precision,
                               void* stream) {
    PF_CHECK_PREC(precision, "pf_flow_cond_tc");
    if (M == 0) return 0;
    PF_CHECK_ARG(raw && cond_tc && sb, "pf_flow_cond_tc: null pointer");
    PF_CHECK_ARG(M >= 0 && nblocks >= 1 && nblocks <= MAXB, "pf_flow_cond_tc: bad sizes M=%lld nblocks=%d", (long long)M,
                 nblocks);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(raw) | reinterpret_cast<uintptr_t>(cond_tc) |
                   reinterpret_cast<uintptr_t>(sb)) & 31) == 0, "pf_flow_cond_tc: pointers must be 32-byte aligned");
    if (M == 0) return 0;
    constexpr size_t smem = ((CBLK + 127u) & ~127u) + NWG_C * A_WG + 64;
    if (int rc = pf_smem_optin(flow_cond_tc_kernel, smem)) return rc;
    const int64_t ntiles = (M + TP - 1) / TP;
    int workers = sm_count() / nblocks;
    if (workers < 1) workers = 1;
    if ((int64_t)workers * NWG_C > ntiles) workers = (int)((ntiles + NWG_C - 1) / NWG_C);
    dim3 grid((unsigned)workers, (unsigned)nblocks);
    flow_cond_tc_kernel<<<grid, NT_C, smem, (cudaStream_t)stream>>>(raw, static_cast<const uint8_t*>(cond_tc), sb, M, precision);
    return pf_launch_status("flow_cond_tc_kernel");
}

static int flow_tc_launch(bool inverse, const float* xin, const float* raw, int ldraw, const float* sb,
                          const void* flow_tc, float* xout, float* ssum, int64_t M, int R, int nblocks, int precision,
                          void* stream) {
    PF_CHECK_PREC(precision, "pf_flow_tc");
    PF_CHECK_ARG(ldraw >= 64 && ldraw % 8 == 0, "pf_flow_tc: ldraw=%d must be a multiple of 8, >= 64", ldraw);
    if (M == 0) return 0;
    PF_CHECK_ARG(xin && raw && sb && flow_tc && xout, "pf_flow_tc: null pointer");
    PF_CHECK_ARG(M >= 0 && R >= 1 && nblocks >= 1 && nblocks <= 6, "pf_flow_tc: bad sizes M=%lld R=%d nblocks=%d (<= 6)",
                 (long long)M, R, nblocks);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(raw) | reinterpret_cast<uintptr_t>(flow_tc) |
                   reinterpret_cast<uintptr_t>(sb)) & 31) == 0, "pf_flow_tc: pointers must be 32-byte aligned");
    if (M == 0) return 0;
    const size_t smem = (((size_t)nblocks * FBLK + 127u) & ~(size_t)127u) + 64;       // weights + barriers: the A operand is in tensor memory
    const int64_t rows = inverse ? M * R : M;
    const int64_t ntiles = (rows + TP - 1) / TP;
    int grid = sm_count();
    if ((int64_t)grid * NWG > ntiles) grid = (int)((ntiles + NWG - 1) / NWG);
    const uint8_t* w = static_cast<const uint8_t*>(flow_tc);
    if (inverse) {
        if (int rc = pf_smem_optin(flow_tc_kernel<true>, smem)) return rc;
        flow_tc_kernel<true><<<grid, NT, smem, (cudaStream_t)stream>>>(xin, raw, sb, w, xout, nullptr, M, R, nblocks, ldraw, precision);
    } else {
        if (int rc = pf_smem_optin(flow_tc_kernel<false>, smem)) return rc;
        flow_tc_kernel<false><<<grid, NT, smem, (cudaStream_t)stream>>>(xin, raw, sb, w, xout, ssum, M, 1, nblocks, ldraw, precision);
    }
    return pf_launch_status("flow_tc_kernel");
}

extern "C" int pf_flow_fwd_tc(const float* x, const float* raw, int ldraw, const float* sb, const void* flow_tc, float* z,
                              float* ssum, int64_t M, int nblocks, int precision, void* stream) {
    return flow_tc_launch(false, x, raw, ldraw, sb, flow_tc, z, ssum, M, 1, nblocks, precision, stream);
}

extern "C" int pf_flow_inv_tc(const float* zin, const float* raw, int ldraw, const float* sb, const void* flow_tc,
                              float* x, int64_t M, int R, int nblocks, int precision, void* stream) {
    return flow_tc_launch(true, zin, raw, ldraw, sb, flow_tc, x, nullptr, M, R, nblocks, precision, stream);
}

extern "C" int pf_cond_chain_tc(const float* c, int C, const void* chain_tc, float* raw1, float* sb, int64_t M,
                                int precision, void* stream) {
    PF_CHECK_PREC(precision, "pf_cond_chain_tc");
    if (M == 0) return 0;
    PF_CHECK_ARG(c && chain_tc && raw1 && sb, "pf_cond_chain_tc: null pointer");
    PF_CHECK_ARG(M >= 0, "pf_cond_chain_tc: bad size M=%lld", (long long)M);
    PF_CHECK_ARG(C == 32 || C == 64 || C == 128, "pf_cond_chain_tc: C=%d not in {32, 64, 128}", C);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(c) | reinterpret_cast<uintptr_t>(chain_tc) |
                   reinterpret_cast<uintptr_t>(raw1) | reinterpret_cast<uintptr_t>(sb)) & 31) == 0,
                 "pf_cond_chain_tc: pointers must be 32-byte aligned");
    if (M == 0) return 0;
    const uint8_t* w = static_cast<const uint8_t*>(chain_tc);
    cudaStream_t st = (cudaStream_t)stream;
    const int sms = sm_count();
    switch (C) {
        case 32: return chain_launch<32>(c, w, raw1, sb, M, sms, precision, st);
        case 64: return chain_launch<64>(c, w, raw1, sb, M, sms, precision, st);
        default: return chain_launch<128>(c, w, raw1, sb, M, sms, precision, st);
    }
}
