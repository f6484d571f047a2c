// Fused EdgeConv block on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM).
//
// Same contract as pf_edgeconv (edgeconv.cu; replaces FeatureExtractUnit.forward,
// /root/reference/modules/discrete/interpflow.py:223-248) for the (G, C_out, K) = (32, 128, 16) blocks that
// ncu shows FFMA-issue-bound (profiles/r01_ncu_edgeconv_v0_ffma.md): the hidden-channel contractions of the
// dense block (6 G^2 + 4 G C_out MACs per edge) run as tcgen05.mma.kind::f16 with an error-compensated fp16
// split so the result stays fp32-class:
//     x = hi + lo (two fp16, |x - hi - lo| <~ 2^-22 |x|),   a.w ~= a_hi w_hi + a_lo w_hi + a_hi w_lo   (fp32 accumulate)
// The dropped a_lo w_lo term is 2^-22 relative.  Tolerance stated in tests/test_gpu_kernels.py.
//
// One CTA per SM (persistent), 2 warpgroups (WG) of 128 threads; each WG owns one 128-edge tile
// (8 points x 16 neighbours) at a time, and the two WGs ping-pong: while one runs its epilogue on the CUDA
// cores the other's MMAs occupy the tensor pipe.  Per tile:
//   layer 0      h0 = lrelu(P0_i + Q0_j)                                  (gather only)
//   layers 1..3  D[128 edges x 32]  = H[128 x 32l] . W_l^T   (M=128, N=32, K=32l)   -> TMEM -> +P+Q, lrelu
//                -> fp16 hi/lo -> next K-slice of H in shared memory (UMMA K-major layout, tc.cuh)
//   conv_out     D^T[128 ch x 128 edges] = Wo[128 x 128] . H^T  (M=128, N=128, K=128): transposed so that a
//                thread owns one output channel and the max over the 16 neighbours is a register max; the Q_j
//                gather and the output store are then coalesced across the warp.
//                Wo (hi/lo) lives in TMEM as the A operand (tcgen05.mma with A from tensor memory), which frees
//                64 KB of shared memory for the gather staging below.
// Gathers never stall the math: the Q_j rows of layer l+1 (and P_i / Q_j / neighbour indices of the NEXT tile) are
// fetched with cp.async into XOR-swizzled staging buffers while layer l computes.
// Weights are pre-split and pre-laid-out by puflow_b200/packing.py (pack_edgeconv_tc): the dense-layer weights
// arrive with one TMA bulk copy per CTA, Wo is written to TMEM once per CTA with tcgen05.st.
#include "common.cuh"
#include "puflow_b200.h"
#include "tc.cuh"

namespace {

using namespace pf;

constexpr int G = 32, CO = 128, KN = 16;
constexpr int R = 4 * G + CO, R2 = 2 * R;      // 256, 512
constexpr int TM = 128, PT = TM / KN;          // edges / points per tile
constexpr int NT = 320;                        // 2 epilogue warpgroups + one MMA-issuing warp per warpgroup

// ---- shared memory map (bytes) ----
// W_l (l = 1..3): 32 rows x 32l halfs per part; hi part at wd_off(l), lo part at wd_off(l) + wd_part(l)
__host__ __device__ constexpr uint32_t wd_part(int l) { return 2048u * l; }
__host__ __device__ constexpr uint32_t wd_off(int l) { return 4096u * (l * (l - 1) / 2); }
constexpr uint32_t WD_BYTES = 24576;
constexpr uint32_t WO_GLOBAL_OFF = WD_BYTES, WO_PART = 32768;            // blob: [Wd | Wo hi [128][64 words] | Wo lo]
constexpr uint32_t A_OFF = WD_BYTES, A_PART = 32768;                     // A[wg][hi|lo]
constexpr uint32_t QS_OFF = A_OFF + 4 * A_PART, QS_BUF = 16384;          // qs[wg][2]: 128 rows x 128 B, XOR-swizzled
constexpr uint32_t PS_OFF = QS_OFF + 4 * QS_BUF, PS_BUF = 4096;          // ps[wg]: 8 points x 128 floats (dense P)
constexpr uint32_t JB_OFF = PS_OFF + 2 * PS_BUF, JB_BUF = 256;           // jbuf[wg][2][128] uint16 (patch-local j)
constexpr uint32_t BAR_OFF = JB_OFF + 4 * JB_BUF;
constexpr uint32_t SMEM_BYTES = BAR_OFF + 64;                            // 230,464 <= 232,448
constexpr uint32_t A_SBO = 128 * 16;                                     // Kt = 128 halfs

// ---- tensor memory map (columns) ----
constexpr uint32_t TM_COLS = 512;
constexpr uint32_t TM_WG = 160, TM_OUT = 32;                             // per WG: dense D at +0 (32 cols), out D at +32 (128)
constexpr uint32_t TM_WO_HI = 320, TM_WO_LO = 384;                       // Wo as A operand: 128 lanes x 64 cols each

#ifdef PF_TC_TIMING
__device__ long long g_tc_timing[64];
#define TSTAMP(k)                                                          \
    do {                                                                   \
        if (blockIdx.x == 0 && tid == 0) {                                 \
            long long now_ = clock64();                                    \
            g_tc_timing[k] += now_ - tprev_;                               \
            tprev_ = now_;                                                 \
        }                                                                  \
    } while (0)
#else
#define TSTAMP(k)
#endif

__global__ void __launch_bounds__(NT, 1)
    edgeconv_tc_kernel(const float* __restrict__ pq, const int32_t* __restrict__ idx, const uint8_t* __restrict__ wblob,
                       float* __restrict__ out, int N, int tiles_per_patch, long long total_tiles, float slope) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + BAR_OFF);
    uint64_t* mbar = wbar + 1;                                            // [2] MMA done, one per WG
    uint64_t* ready = wbar + 3;                                           // [2] operands ready, one per WG
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + BAR_OFF + 40);

    const int tid = threadIdx.x, warp = tid >> 5;
    const bool is_mma_warp = warp >= 8;
    const int wg = is_mma_warp ? warp - 8 : warp >> 2, wq = warp & 3;    // warpgroup, TMEM lane quadrant
    const int e = tid & 127;                                              // tile row (edge) / output channel of this thread

    if (tid == 0) {
        mbar_init(wbar, 1);
        mbar_init(mbar, 1);
        mbar_init(mbar + 1, 1);
        mbar_init(ready, 1);
        mbar_init(ready + 1, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, TM_COLS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t lane_sel = (uint32_t)(wq * 32) << 16;
    if (tid == 0) {
        mbar_expect_tx(wbar, WD_BYTES);
        bulk_g2s_chunked(smem, wblob, WD_BYTES, wbar);
    }
    if (warp < 4) {   // Wo -> TMEM: thread = lane = output channel, two fp16 per column
#pragma unroll 1
        for (int part = 0; part < 2; part++)
#pragma unroll 1
            for (int half = 0; half < 2; half++) {
                const uint4* src = reinterpret_cast<const uint4*>(wblob + WO_GLOBAL_OFF + (size_t)part * WO_PART +
                                                                  (size_t)e * 256 + half * 128);
                uint32_t r[32];
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const uint4 v = __ldg(src + i);
                    r[4 * i] = v.x, r[4 * i + 1] = v.y, r[4 * i + 2] = v.z, r[4 * i + 3] = v.w;
                }
                tc::st32(tmem_base + lane_sel + (part ? TM_WO_LO : TM_WO_HI) + 32 * half, r);
            }
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();

    const uint32_t a_hi = sbase + A_OFF + (uint32_t)wg * 2 * A_PART, a_lo = a_hi + A_PART;
    uint8_t* a_hi_p = smem + A_OFF + (size_t)wg * 2 * A_PART;
    uint8_t* a_lo_p = a_hi_p + A_PART;
    const uint32_t a_row = (uint32_t)(e >> 3) * A_SBO + (uint32_t)(e & 7) * 16;   // byte offset of row e, chunk 0
    const uint32_t qs_addr = sbase + QS_OFF + (uint32_t)wg * 2 * QS_BUF;
    const uint8_t* qs_p = smem + QS_OFF + (size_t)wg * 2 * QS_BUF;
    const uint32_t ps_addr = sbase + PS_OFF + (uint32_t)wg * PS_BUF;
    const uint8_t* ps_p = smem + PS_OFF + (size_t)wg * PS_BUF;
    uint16_t* jb = reinterpret_cast<uint16_t*>(smem + JB_OFF) + wg * 2 * TM;
    const uint32_t d_dense = tmem_base + (uint32_t)wg * TM_WG;
    const uint32_t d_out = d_dense + TM_OUT;
    const uint32_t idesc_d = tc::idesc_f16(128, G), idesc_o = tc::idesc_f16(128, 128);
    const int bar_id = 1 + wg;
    uint32_t phase = 0;

    // ---- staging helpers (cp.async; every thread of the WG issues its share) ----
    // Q columns [col, col+32) of the 128 neighbour rows -> qs[buf]; row r chunk c at r*128 + ((c ^ (r & 7)) * 16)
    auto stage_q = [&](const float* patch_pq, const uint16_t* jl, int col, int buf) {
        const int c = e & 7;
#pragma unroll
        for (int it = 0; it < 8; it++) {
            const int r = it * 16 + (e >> 3);
            tc::cp_async16(qs_addr + buf * QS_BUF + r * 128 + ((c ^ (r & 7)) << 4),
                           patch_pq + (size_t)jl[r] * R2 + R + col + c * 4);
        }
    };
    // dense P columns [0, 128) of the tile's 8 points -> ps (8 x 512 B)
    auto stage_p = [&](const float* patch_pq, int p0) {
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int i = it * 128 + e, pt = i >> 5, c16 = i & 31;
            tc::cp_async16(ps_addr + pt * 512 + c16 * 16, patch_pq + (size_t)min(p0 + pt, N - 1) * R2 + c16 * 4);
        }
    };
    auto read_q = [&](int buf, float (&q)[32]) {
        const uint8_t* row = qs_p + buf * QS_BUF + e * 128;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const float4 v = *reinterpret_cast<const float4*>(row + ((c ^ (e & 7)) << 4));
            q[4 * c] = v.x, q[4 * c + 1] = v.y, q[4 * c + 2] = v.z, q[4 * c + 3] = v.w;
        }
    };
    auto add_p = [&](int l, float (&q)[32]) {
        const uint8_t* row = ps_p + (e >> 4) * 512 + l * 128;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            const float4 v = *reinterpret_cast<const float4*>(row + c * 16);
            q[4 * c] += v.x, q[4 * c + 1] += v.y, q[4 * c + 2] += v.z, q[4 * c + 3] += v.w;
        }
    };
    // h[32] (fp32) -> fp16 hi/lo into K-slice l of this thread's row of the activation operand
    auto store_slice = [&](int l, const float (&h)[32]) {
#pragma unroll
        for (int c = 0; c < 4; c++) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int u = 0; u < 4; u++) tc::split2(h[8 * c + 2 * u], h[8 * c + 2 * u + 1], hi[u], lo[u]);
            const uint32_t off = a_row + (uint32_t)(4 * l + c) * 128;
            *reinterpret_cast<uint4*>(a_hi_p + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(a_lo_p + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
    };
    auto tile_coords = [&](long long t, int& patch, int& p0) {
        const uint32_t tt = (uint32_t)t, tpp = (uint32_t)tiles_per_patch;
        patch = (int)(tt / tpp);
        p0 = (int)(tt - (uint32_t)patch * tpp) * PT;
    };
    auto load_j = [&](int patch, int p0) {
        const int pi = min(p0 + (e >> 4), N - 1);
        const int j = __ldg(idx + ((size_t)patch * N + pi) * KN + (e & 15));
        return (uint16_t)min(max(j, 0), N - 1);
    };
    const uint64_t da_hi = tc::smem_desc(a_hi, A_SBO), da_lo = tc::smem_desc(a_lo, A_SBO);
    // D[128 edges x 32] = H[:, 0:32l] . W_l^T   (3 split products; K-step s is 256 B = 16 descriptor units further)
    auto issue_dense = [&](int l) {
        const uint32_t wl = sbase + wd_off(l);
        const uint64_t db_hi = tc::smem_desc(wl, 512u * l), db_lo = tc::smem_desc(wl + wd_part(l), 512u * l);
        const int ks = 2 * l;
        uint32_t acc = 0;
#pragma unroll 1
        for (int s = 0; s < ks; s++, acc = 1) tc::mma_f16(d_dense, da_hi + 16u * s, db_hi + 16u * s, idesc_d, acc);
#pragma unroll 1
        for (int s = 0; s < ks; s++) tc::mma_f16(d_dense, da_lo + 16u * s, db_hi + 16u * s, idesc_d, 1);
#pragma unroll 1
        for (int s = 0; s < ks; s++) tc::mma_f16(d_dense, da_hi + 16u * s, db_lo + 16u * s, idesc_d, 1);
    };
    // D^T[128 ch x 128 edges] = Wo . H^T, Wo from TMEM (8 columns per K-step)
    auto issue_out = [&]() {
        const uint32_t wo_hi = tmem_base + TM_WO_HI, wo_lo = tmem_base + TM_WO_LO;
#pragma unroll
        for (int s = 0; s < 8; s++) tc::mma_f16_ts(d_out, wo_hi + 8u * s, da_hi + 16u * s, idesc_o, s > 0);
#pragma unroll
        for (int s = 0; s < 8; s++) tc::mma_f16_ts(d_out, wo_lo + 8u * s, da_hi + 16u * s, idesc_o, 1);
#pragma unroll
        for (int s = 0; s < 8; s++) tc::mma_f16_ts(d_out, wo_hi + 8u * s, da_lo + 16u * s, idesc_o, 1);
    };

    const long long stride = (long long)gridDim.x * 2;
    long long t = (long long)blockIdx.x * 2 + wg;
    if (is_mma_warp) {
        // ---- MMA-issuing warp of warpgroup wg: one thread; waits for "operands ready", issues, commits ----
        if ((tid & 31) == 0) {
            tc::mbar_wait_bounded(wbar, 0);   // dense-layer weights have landed
            uint32_t rphase = 0;
            for (; t < total_tiles; t += stride) {
#pragma unroll 1
                for (int l = 1; l < 4; l++) {
                    tc::mbar_wait_bounded(ready + wg, rphase);
                    rphase ^= 1;
                    tc::fence_after_sync();
                    issue_dense(l);
                    tc::commit(mbar + wg);
                }
                tc::mbar_wait_bounded(ready + wg, rphase);
                rphase ^= 1;
                tc::fence_after_sync();
                issue_out();
                tc::commit(mbar + wg);
            }
        }
    } else {
    int tb = 0;
    if (t < total_tiles) {   // prologue: indices, P and Q0 of the first tile
        int patch, p0;
        tile_coords(t, patch, p0);
        jb[e] = load_j(patch, p0);
        tc::named_bar(bar_id, 128);
        const float* ppq = pq + (size_t)patch * N * R2;
        stage_p(ppq, p0);
        stage_q(ppq, jb, 0, 0);
        tc::cp_async_commit();
        tc::cp_async_wait<0>();
        tc::named_bar(bar_id, 128);
    }
    tc::mbar_wait_bounded(wbar, 0);   // dense-layer weights have landed

#ifdef PF_TC_TIMING
    long long tprev_ = clock64();
#endif
    for (; t < total_tiles; t += stride, tb ^= 1) {
        int patch, p0;
        tile_coords(t, patch, p0);
        TSTAMP(0);
        const float* ppq = pq + (size_t)patch * N * R2;
        const uint16_t* jl = jb + tb * TM;
        const bool has_next = t + stride < total_tiles;
        int patch_n = 0, p0_n = 0;
        uint16_t j_next = 0;
        if (has_next) {
            tile_coords(t + stride, patch_n, p0_n);
            j_next = load_j(patch_n, p0_n);
        }
        const float* ppq_n = pq + (size_t)patch_n * N * R2;

        // ---- layer 0: h0 = lrelu(P0_i + Q0_j) ----
        stage_q(ppq, jl, G, 1);
        tc::cp_async_commit();
        {
            float q[32];
            read_q(0, q);
            add_p(0, q);
#pragma unroll
            for (int c = 0; c < 32; c++) q[c] = lrelu(q[c], slope);
            store_slice(0, q);
        }
        TSTAMP(1);
        tc::cp_async_wait<0>();
        tc::fence_async_smem();
        tc::fence_before_sync();
        tc::named_bar(bar_id, 128);
        if (e == 0) mbar_arrive(ready + wg);
        TSTAMP(2);

        // ---- layers 1..3 ----
#pragma unroll 1
        for (int l = 1; l < 4; l++) {
            TSTAMP(3);
            if (l < 3)
                stage_q(ppq, jl, (l + 1) * G, (l + 1) & 1);
            else if (has_next)
                stage_q(ppq_n, jb + (tb ^ 1) * TM, 0, 0);                 // next tile's Q0
            tc::cp_async_commit();
            if (l == 1 && has_next) jb[(tb ^ 1) * TM + e] = j_next;       // published by this layer's closing barrier
            float q[32];
            read_q(l & 1, q);
            add_p(l, q);
            TSTAMP(4);
            tc::mbar_wait_bounded(mbar + wg, phase);
            phase ^= 1;
            tc::fence_after_sync();
            TSTAMP(5);
            float acc[32];
            tc::ld32(d_dense + lane_sel, acc);
            TSTAMP(6);
#pragma unroll
            for (int c = 0; c < 32; c++) acc[c] = lrelu(acc[c] + q[c], slope);
            store_slice(l, acc);
            TSTAMP(7);
            tc::cp_async_wait<0>();
            TSTAMP(8);
            tc::fence_async_smem();
            tc::fence_before_sync();
            tc::named_bar(bar_id, 128);
            if (e == 0) mbar_arrive(ready + wg);
            TSTAMP(9);
        }

        // ---- conv_out (transposed) + max over the 16 neighbours ----
        TSTAMP(10);
        if (has_next) stage_p(ppq_n, p0_n);
        tc::cp_async_commit();
        {
            const int ch = e;                                             // output channel of this thread
            const float* qcol = ppq + R + 4 * G + ch;
            float qva[32], qvb[32];
            auto load_qchunk = [&](int c, float (&qv)[32]) {
#pragma unroll
                for (int v8 = 0; v8 < 4; v8++) {
                    const uint4 jj = *reinterpret_cast<const uint4*>(jl + 32 * c + 8 * v8);
                    const uint32_t w[4] = {jj.x, jj.y, jj.z, jj.w};
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        qv[8 * v8 + 2 * u] = __ldg(qcol + (size_t)(w[u] & 0xffffu) * R2);
                        qv[8 * v8 + 2 * u + 1] = __ldg(qcol + (size_t)(w[u] >> 16) * R2);
                    }
                }
            };
            load_qchunk(0, qva);
            load_qchunk(1, qvb);
            TSTAMP(11);
            tc::mbar_wait_bounded(mbar + wg, phase);
            phase ^= 1;
            tc::fence_after_sync();
            TSTAMP(12);
#pragma unroll
            for (int c = 0; c < 4; c++) {                                 // 32 edges = 2 points per chunk
                float acc[32];
                tc::ld32(d_out + lane_sel + 32 * c, acc);
#pragma unroll
                for (int u = 0; u < 32; u++) acc[u] += (c & 1) ? qvb[u] : qva[u];
                if (c < 2) {
                    if (c & 1) load_qchunk(c + 2, qvb); else load_qchunk(c + 2, qva);
                }
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    float m = acc[16 * h];
#pragma unroll
                    for (int u = 1; u < 16; u++) m = fmaxf(m, acc[16 * h + u]);
                    const int pp = p0 + 2 * c + h;
                    if (pp < N) out[((size_t)patch * N + pp) * CO + ch] = m + __ldg(ppq + (size_t)pp * R2 + 4 * G + ch);
                }
            }
        }
        TSTAMP(13);
        tc::cp_async_wait<0>();
        tc::fence_before_sync();
        tc::named_bar(bar_id, 128);          // D, H and jbuf[tb] are free; ps / qs[0] / jbuf[tb^1] hold the next tile
    }
    }   // epilogue warpgroups

    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem_base, TM_COLS);
}

}  // namespace

#ifdef PF_TC_TIMING
extern "C" __attribute__((visibility("default"))) int pf_tc_timing(long long* host64) {
    cudaDeviceSynchronize();
    cudaError_t e = cudaMemcpyFromSymbol(host64, g_tc_timing, sizeof(long long) * 64);
    long long z[64] = {0};
    cudaMemcpyToSymbol(g_tc_timing, z, sizeof(z));
    return (int)e;
}
#endif

extern "C" int pf_edgeconv_tc(const float* pq, const int32_t* idx, const void* wtc, float* out, int Bp, int N, int K,
                              int G_, int C_out, float slope, void* stream) {
    PF_CHECK_ARG(pq && idx && wtc && out, "pf_edgeconv_tc: null pointer");
    PF_CHECK_ARG(Bp >= 0 && N > 0, "pf_edgeconv_tc: bad sizes Bp=%d N=%d", Bp, N);
    PF_CHECK_ARG((long long)Bp * N < 0x7fffffffLL, "pf_edgeconv_tc: Bp*N overflows int32");
    PF_CHECK_ARG(N <= 65536, "pf_edgeconv_tc: N=%d > 65536 points per patch", N);
    PF_CHECK_ARG(K == KN && G_ == G && C_out == CO, "pf_edgeconv_tc: only (G, C_out, K) = (32, 128, 16), got (%d, %d, %d)",
                 G_, C_out, K);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(pq) | reinterpret_cast<uintptr_t>(wtc) | reinterpret_cast<uintptr_t>(out)) &
                  15) == 0,
                 "pf_edgeconv_tc: pointers must be 16-byte aligned");
    if (Bp == 0) return 0;
    if (int rc = pf_smem_optin(edgeconv_tc_kernel, SMEM_BYTES)) return rc;
    const int tiles_per_patch = (N + PT - 1) / PT;
    const long long total = (long long)Bp * tiles_per_patch;
    int dev = 0, sms = 148;
    PF_CUDA(cudaGetDevice(&dev));
    PF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = (int)((total + 1) / 2 < sms ? (total + 1) / 2 : sms);
    edgeconv_tc_kernel<<<grid, NT, SMEM_BYTES, (cudaStream_t)stream>>>(pq, idx, static_cast<const uint8_t*>(wtc), out, N,
                                                                       tiles_per_patch, total, slope);
    return pf_launch_status("edgeconv_tc_kernel");
}
