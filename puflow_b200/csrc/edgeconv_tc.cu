// Fused EdgeConv block on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM).
//
// Same contract as pf_edgeconv (edgeconv.cu; replaces FeatureExtractUnit.forward,
// /root/reference/modules/discrete/interpflow.py:223-248) for K = 16 and (G, C_out) in {(8, 32), (16, 64), (32, 128)} --
// all six blocks of the path; ncu shows the FFMA version issue-bound (profiles/r01_ncu_edgeconv_v0_ffma.md): the
// hidden-channel contractions of the
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

// Template parameters: G growth rate, CO output channels, NWG epilogue warpgroups per CTA (each with its own
// MMA-issuing warp).  GP = max(G, 16) is the K-slice width of one dense layer inside the activation operand
// (tcgen05 needs K % 16 == 0 and N % 16 == 0): for G = 8 the upper half of every slice is zero in both operands.
// For CO < 128 the transposed conv_out still runs with M = 128: row r of the A operand is Wo[r % CO], and thread r
// takes channel r % CO for the edges [CO * (r / CO), CO * (r / CO + 1)) of the tile, so no epilogue thread idles.
// Text below uses the (32, 128) numbers.

#ifdef PF_PROF
__device__ long long g_prof[32];
#endif

namespace {

using namespace pf;

constexpr int KN = 16;
constexpr int TM = 128, PT = TM / KN;          // edges / points per tile

template <int G, int CO, int NWG>
struct Cfg {
    static constexpr int GP = G < 16 ? 16 : G;
    static constexpr int R = 4 * G + CO, R2 = 2 * R;
    static constexpr int KT = 4 * GP;                                    // K width of the activation operand
    static constexpr int NT = NWG * 128 + NWG * 32;
    static constexpr int QCH = G / 4;                                    // 16-byte chunks per staged Q row
    // ---- shared memory map (bytes) ----
    static constexpr uint32_t WD_BYTES = 24u * GP * GP;                  // W1..W3, hi and lo
    static constexpr uint32_t WO_PART = 128u * KT * 2;                   // blob: [Wd | Wo hi [128][KT/2 words] | Wo lo]
    static constexpr uint32_t BLOB = WD_BYTES + 2 * WO_PART;
    static constexpr uint32_t A_OFF = WD_BYTES, A_PART = 128u * KT * 2;  // A[wg][hi|lo]
    static constexpr uint32_t QS_OFF = A_OFF + 2 * NWG * A_PART, QS_BUF = 128u * G * 4;   // qs[wg][2], XOR-swizzled rows
    static constexpr uint32_t PS_OFF = QS_OFF + 2 * NWG * QS_BUF, PS_BUF = 8u * 4 * G * 4;  // ps[wg]: 8 points x 4G floats
    static constexpr uint32_t JB_OFF = PS_OFF + NWG * PS_BUF, JB_BUF = 256;               // jbuf[wg][2][128] uint16
    static constexpr uint32_t BAR_OFF = JB_OFF + 2 * NWG * JB_BUF;
    static constexpr uint32_t SMEM_BYTES = BAR_OFF + 8 * (1 + 2 * NWG) + 16;
    static constexpr uint32_t A_SBO = KT * 16;
    // ---- tensor memory map (columns): per warpgroup 128 (conv_out accumulator; the dense accumulator aliases its
    // first GP columns -- the two are never live together), then Wo hi / lo as the A operand of conv_out ----
    static constexpr uint32_t TM_WG = 128, TM_WO_HI = NWG * 128, TM_WO_LO = TM_WO_HI + KT / 2, TM_COLS = 512;
    static_assert(TM_WO_LO + KT / 2 <= 512, "tensor memory budget");
    static_assert(SMEM_BYTES <= 232448, "shared memory budget");
    __host__ __device__ static constexpr uint32_t wd_part(int l) { return 2u * GP * GP * l; }
    __host__ __device__ static constexpr uint32_t wd_off(int l) { return 4u * GP * GP * (l * (l - 1) / 2); }
};

template <int G, int CO, int NWG>
__global__ void __launch_bounds__(Cfg<G, CO, NWG>::NT, 1)
    edgeconv_tc_kernel(const float* __restrict__ pq, const int32_t* __restrict__ idx, const uint8_t* __restrict__ wblob,
                       float* __restrict__ out, uint8_t* __restrict__ out_split, int N, int tiles_per_patch,
                       long long total_tiles, float slope) {
    using C = Cfg<G, CO, NWG>;
    constexpr int GP = C::GP, R = C::R, R2 = C::R2, KT = C::KT, QCH = C::QCH;
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + C::BAR_OFF);
    uint64_t* mbar = wbar + 1;                                            // [NWG] MMA done
    uint64_t* ready = mbar + NWG;                                         // [NWG] operands ready
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(ready + NWG);

    const int tid = threadIdx.x, warp = tid >> 5;
    const bool is_mma_warp = warp >= 4 * NWG;
    const int wg = is_mma_warp ? warp - 4 * NWG : warp >> 2, wq = warp & 3;   // warpgroup, TMEM lane quadrant
    const int e = tid & 127;                                              // tile row (edge) / conv_out row of this thread

    if (tid == 0) {
        mbar_init(wbar, 1);
        for (int i = 0; i < NWG; i++) mbar_init(mbar + i, 1), mbar_init(ready + i, 1);
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
    uint8_t* a_hi_p = smem + C::A_OFF + (size_t)wg * 2 * C::A_PART;
    uint8_t* a_lo_p = a_hi_p + C::A_PART;
    const uint32_t a_row = (uint32_t)(e >> 3) * C::A_SBO + (uint32_t)(e & 7) * 16;   // byte offset of row e, chunk 0
    const uint32_t qs_addr = sbase + C::QS_OFF + (uint32_t)wg * 2 * C::QS_BUF;
    const uint8_t* qs_p = smem + C::QS_OFF + (size_t)wg * 2 * C::QS_BUF;
    const uint32_t ps_addr = sbase + C::PS_OFF + (uint32_t)wg * C::PS_BUF;
    const uint8_t* ps_p = smem + C::PS_OFF + (size_t)wg * C::PS_BUF;
    uint16_t* jb = reinterpret_cast<uint16_t*>(smem + C::JB_OFF) + wg * 2 * TM;
    const uint32_t d_out = tmem_base + (uint32_t)wg * C::TM_WG;
    const uint32_t d_dense = d_out;                                       // aliases the first GP columns of d_out
    const uint32_t idesc_d = tc::idesc_f16(128, GP), idesc_o = tc::idesc_f16(128, 128);
    const int bar_id = 1 + wg;
    uint32_t phase = 0;

    // ---- staging helpers (cp.async; every thread of the WG issues its share) ----
    // Q columns [col, col+G) of the 128 neighbour rows -> qs[buf]; row r chunk c at r*(16 QCH) + ((c ^ (r % QCH)) * 16)
    auto stage_q = [&](const float* patch_pq, const uint16_t* jl, int col, int buf) {
        const int c = e % QCH;
#pragma unroll
        for (int it = 0; it < QCH; it++) {
            const int r = it * (128 / QCH) + e / QCH;
            tc::cp_async16(qs_addr + buf * C::QS_BUF + r * (16 * QCH) + ((c ^ (r & (QCH - 1))) << 4),
                           patch_pq + (size_t)jl[r] * R2 + R + col + c * 4);
        }
    };
    // dense P columns [0, 4G) of the tile's 8 points -> ps (8 x 16G bytes)
    auto stage_p = [&](const float* patch_pq, int p0) {
        constexpr int CHUNKS = 8 * G;                                     // 16-byte chunks in all
#pragma unroll
        for (int it = 0; it < (CHUNKS + 127) / 128; it++) {
            const int i = it * 128 + e, pt = i / G, c16 = i % G;
            if (CHUNKS % 128 == 0 || i < CHUNKS)
                tc::cp_async16(ps_addr + pt * (16 * G) + c16 * 16, patch_pq + (size_t)min(p0 + pt, N - 1) * R2 + c16 * 4);
        }
    };
    auto read_q = [&](int buf, float (&q)[G]) {
        const uint8_t* row = qs_p + buf * C::QS_BUF + e * (16 * QCH);
#pragma unroll
        for (int c = 0; c < QCH; c++) {
            const float4 v = *reinterpret_cast<const float4*>(row + ((c ^ (e & (QCH - 1))) << 4));
            q[4 * c] = v.x, q[4 * c + 1] = v.y, q[4 * c + 2] = v.z, q[4 * c + 3] = v.w;
        }
    };
    auto add_p = [&](int l, float (&q)[G]) {
        const uint8_t* row = ps_p + (e >> 4) * (16 * G) + l * (4 * G);
#pragma unroll
        for (int c = 0; c < QCH; c++) {
            const float4 v = *reinterpret_cast<const float4*>(row + c * 16);
            q[4 * c] += v.x, q[4 * c + 1] += v.y, q[4 * c + 2] += v.z, q[4 * c + 3] += v.w;
        }
    };
    // h[G] (fp32) -> fp16 hi/lo into K-slice l (GP wide) of this thread's row of the activation operand
    auto store_slice = [&](int l, const float (&h)[G]) {
#pragma unroll
        for (int c = 0; c < G / 8; c++) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int u = 0; u < 4; u++) tc::split2(h[8 * c + 2 * u], h[8 * c + 2 * u + 1], hi[u], lo[u]);
            const uint32_t off = a_row + (uint32_t)(l * (GP / 8) + c) * 128;
            *reinterpret_cast<uint4*>(a_hi_p + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(a_lo_p + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
    };
    auto tile_coords = [&](long long t, int& patch, int& p0) {
        const uint32_t tt = (uint32_t)t, tpp = (uint32_t)tiles_per_patch;
        patch = (int)(tt / tpp);
        p0 = (int)(tt - (uint32_t)patch * tpp) * PT;
    };
    auto load_j = [&](int patch, int p0) {                                 // raw index: clamp at the use (clamp_j), not here,
        const int pi = min(p0 + (e >> 4), N - 1);                         // so that the load's latency is not waited for
        return __ldg(idx + ((size_t)patch * N + pi) * KN + (e & 15));
    };
    auto clamp_j = [&](int j) { return (uint16_t)min(max(j, 0), N - 1); };
    const uint64_t da_hi = tc::smem_desc(a_hi, C::A_SBO), da_lo = tc::smem_desc(a_lo, C::A_SBO);
    // D[128 edges x GP] = H[:, 0:GP l] . W_l^T   (3 split products, the two small cross terms first: the accumulator
    // adder truncates, so the additions at full magnitude -- the hi . hi pass -- come last and are as few as possible; K-step s is 256 B = 16 descriptor units further)
    auto issue_dense = [&](int l) {
        const uint32_t wl = sbase + C::wd_off(l);
        const uint64_t db_hi = tc::smem_desc(wl, 16u * GP * l), db_lo = tc::smem_desc(wl + C::wd_part(l), 16u * GP * l);
        const int ks = GP * l / 16;
#pragma unroll
        for (int s = 0; s < ks; s++) tc::mma_f16(d_dense, da_lo + 16u * s, db_hi + 16u * s, idesc_d, s > 0);
#pragma unroll
        for (int s = 0; s < ks; s++) tc::mma_f16(d_dense, da_hi + 16u * s, db_lo + 16u * s, idesc_d, 1);
#pragma unroll
        for (int s = 0; s < ks; s++) tc::mma_f16(d_dense, da_hi + 16u * s, db_hi + 16u * s, idesc_d, 1);
    };
    // D^T[128 rows x 128 edges] = Wo . H^T, Wo from TMEM (8 columns per K-step)
    auto issue_out = [&]() {
        const uint32_t wo_hi = tmem_base + C::TM_WO_HI, wo_lo = tmem_base + C::TM_WO_LO;
#pragma unroll
        for (int s = 0; s < KT / 16; s++) tc::mma_f16_ts(d_out, wo_lo + 8u * s, da_hi + 16u * s, idesc_o, s > 0);
#pragma unroll
        for (int s = 0; s < KT / 16; s++) tc::mma_f16_ts(d_out, wo_hi + 8u * s, da_lo + 16u * s, idesc_o, 1);
#pragma unroll
        for (int s = 0; s < KT / 16; s++) tc::mma_f16_ts(d_out, wo_hi + 8u * s, da_hi + 16u * s, idesc_o, 1);
    };

    const long long stride = (long long)gridDim.x * NWG;
    long long t = (long long)blockIdx.x * NWG + wg;
    if (is_mma_warp) {
        // ---- MMA-issuing warp of warpgroup wg: one elected thread (elect_one, not lane == 0: tc.cuh) waits for
        // "operands ready", issues, commits ----
        tc::mbar_wait_bounded(wbar, 0);   // dense-layer weights have landed
        if (tc::elect_one()) {
            uint32_t rphase = 0;
            PROF_DECL
            for (; t < total_tiles; t += stride) {
#pragma unroll
                for (int l = 1; l < 4; l++) {
                    PROF_TICK(1)
                    tc::mbar_wait_bounded(ready + wg, rphase);
                    rphase ^= 1;
                    tc::fence_after_sync();
                    PROF_TICK(0)
                    issue_dense(l);
                    tc::commit(mbar + wg);
                }
                PROF_TICK(1)
                tc::mbar_wait_bounded(ready + wg, rphase);
                rphase ^= 1;
                tc::fence_after_sync();
                PROF_TICK(0)
                issue_out();
                tc::commit(mbar + wg);
            }
            PROF_TICK(1)
            if (wg == 0) { PROF_DUMP(16, 2) }
        }
        __syncwarp();
    } else {
    int tb = 0;
    if (t < total_tiles) {   // prologue: indices, P and Q0 of the first tile
        int patch, p0;
        tile_coords(t, patch, p0);
        jb[e] = clamp_j(load_j(patch, p0));
        tc::named_bar(bar_id, 128);
        const float* ppq = pq + (size_t)patch * N * R2;
        stage_p(ppq, p0);
        stage_q(ppq, jb, 0, 0);
        tc::cp_async_commit();
        tc::cp_async_wait<0>();
        tc::named_bar(bar_id, 128);
    }
    tc::mbar_wait_bounded(wbar, 0);   // dense-layer weights have landed

    PROF_DECL
    for (; t < total_tiles; t += stride, tb ^= 1) {
        int patch, p0;
        tile_coords(t, patch, p0);
        const float* ppq = pq + (size_t)patch * N * R2;
        const uint16_t* jl = jb + tb * TM;
        const bool has_next = t + stride < total_tiles;
        int patch_n = 0, p0_n = 0;
        int j_next = 0;
        if (has_next) {
            tile_coords(t + stride, patch_n, p0_n);
            j_next = load_j(patch_n, p0_n);
        }
        const float* ppq_n = pq + (size_t)patch_n * N * R2;

        // ---- layer 0: h0 = lrelu(P0_i + Q0_j) ----
        stage_q(ppq, jl, G, 1);
        tc::cp_async_commit();
        {
            float q[G];
            read_q(0, q);
            add_p(0, q);
#pragma unroll
            for (int c = 0; c < G; c++) q[c] = lrelu(q[c], slope);
            store_slice(0, q);
        }
        PROF_TICK(0)
        tc::cp_async_wait<0>();
        tc::fence_async_smem();
        tc::fence_before_sync();
        tc::named_bar(bar_id, 128);
        if (e == 0) mbar_arrive(ready + wg);
        PROF_TICK(1)

        // ---- layers 1..3 ----
#pragma unroll 1
        for (int l = 1; l < 4; l++) {
            if (l < 3)
                stage_q(ppq, jl, (l + 1) * G, (l + 1) & 1);
            else if (has_next)
                stage_q(ppq_n, jb + (tb ^ 1) * TM, 0, 0);                 // next tile's Q0
            tc::cp_async_commit();
            if (l == 1 && has_next) jb[(tb ^ 1) * TM + e] = clamp_j(j_next);       // published by this layer's closing barrier
            float q[G];
            read_q(l & 1, q);
            add_p(l, q);
            PROF_TICK(2)
            tc::mbar_wait_bounded(mbar + wg, phase);
            phase ^= 1;
            tc::fence_after_sync();
            PROF_TICK(3)
            float acc[G];
            if constexpr (G == 32) {
                tc::ld32(d_dense + lane_sel, acc);
            } else if constexpr (G == 16) {
                tc::ld16(d_dense + lane_sel, acc);
            } else {
                float a16[16];
                tc::ld16(d_dense + lane_sel, a16);
#pragma unroll
                for (int c = 0; c < G; c++) acc[c] = a16[c];
            }
#pragma unroll
            for (int c = 0; c < G; c++) acc[c] = lrelu(fmaf(acc[c], tc::WINV, q[c]), slope);
            store_slice(l, acc);
            PROF_TICK(4)
            tc::cp_async_wait<0>();
            tc::fence_async_smem();
            tc::fence_before_sync();
            tc::named_bar(bar_id, 128);
            if (e == 0) mbar_arrive(ready + wg);
            PROF_TICK(5)
        }

        // ---- conv_out (transposed) + max over the 16 neighbours ----
        if (has_next) stage_p(ppq_n, p0_n);
        tc::cp_async_commit();
        {
            constexpr int NCH = CO / 32;                                  // 32-edge (2-point) chunks per thread
            const int ch = e % CO, grp = e / CO;                          // output channel, edge range of this thread
            const float* qcol = ppq + R + 4 * G + ch;
            float qva[32], qvb[32];
            auto load_qchunk = [&](int c, float (&qv)[32]) {
#pragma unroll
                for (int v8 = 0; v8 < 4; v8++) {
                    const uint4 jj = *reinterpret_cast<const uint4*>(jl + 32 * (grp * NCH + c) + 8 * v8);
                    const uint32_t w[4] = {jj.x, jj.y, jj.z, jj.w};
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        qv[8 * v8 + 2 * u] = __ldg(qcol + (size_t)(w[u] & 0xffffu) * R2);
                        qv[8 * v8 + 2 * u + 1] = __ldg(qcol + (size_t)(w[u] >> 16) * R2);
                    }
                }
            };
            load_qchunk(0, qva);
            if (NCH > 1) load_qchunk(1, qvb);
            float pv[2 * NCH];                                            // conv_out P term of this thread's points
#pragma unroll
            for (int i = 0; i < 2 * NCH; i++)
                pv[i] = __ldg(ppq + (size_t)min(p0 + 2 * grp * NCH + i, N - 1) * R2 + 4 * G + ch);
            PROF_TICK(6)
            tc::mbar_wait_bounded(mbar + wg, phase);
            phase ^= 1;
            tc::fence_after_sync();
            PROF_TICK(7)
#pragma unroll
            for (int c = 0; c < NCH; c++) {                               // 32 edges = 2 points per chunk
                float acc[32];
                tc::ld32(d_out + lane_sel + 32 * (grp * NCH + c), acc);
#pragma unroll
                for (int u = 0; u < 32; u++) acc[u] = fmaf(acc[u], tc::WINV, (c & 1) ? qvb[u] : qva[u]);
                if (c + 2 < NCH) {
                    if (c & 1) load_qchunk(c + 2, qvb); else load_qchunk(c + 2, qva);
                }
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    float m = acc[16 * h];
#pragma unroll
                    for (int u = 1; u < 16; u++) m = fmaxf(m, acc[16 * h + u]);
                    const int pp = p0 + 2 * (grp * NCH + c) + h;
                    if (pp < N) {
                        const float val = m + pv[2 * c + h];
                        const size_t grow = (size_t)patch * N + pp;
                        if (out) out[grow * CO + ch] = val;
                        if (out_split) {   // fp16 hi/lo in the tile-blocked operand layout of linear_split.cu (K = CO)
                            uint32_t hi, lo;
                            tc::split2(val, 0.f, hi, lo);
                            uint8_t* tile = out_split + (grow >> 7) * (size_t)(CO * 512) + (size_t)(ch >> 3) * 2048 +
                                            (grow & 127) * 16 + (ch & 7) * 2;
                            *reinterpret_cast<uint16_t*>(tile) = (uint16_t)(hi & 0xffffu);
                            *reinterpret_cast<uint16_t*>(tile + CO * 256) = (uint16_t)(lo & 0xffffu);
                        }
                    }
                }
            }
        }
        PROF_TICK(8)
        tc::cp_async_wait<0>();
        tc::fence_before_sync();
        tc::named_bar(bar_id, 128);          // D, H and jbuf[tb] are free; ps / qs[0] / jbuf[tb^1] hold the next tile
        PROF_TICK(9)
    }
    if (tid == 0) { PROF_DUMP(0, 10) }
    }   // epilogue warpgroups

    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem_base, C::TM_COLS);
}

template <int G, int CO, int NWG>
int launch(const float* pq, const int32_t* idx, const void* wtc, float* out, void* out_split, int Bp, int N, float slope,
           cudaStream_t st) {
    using C = Cfg<G, CO, NWG>;
    auto kern = edgeconv_tc_kernel<G, CO, NWG>;
    if (int rc = pf_smem_optin(kern, C::SMEM_BYTES)) return rc;
    const int tiles_per_patch = (N + PT - 1) / PT;
    const long long total = (long long)Bp * tiles_per_patch;
    int dev = 0, sms = 148;
    PF_CUDA(cudaGetDevice(&dev));
    PF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const long long want = (total + NWG - 1) / NWG;
    const int grid = (int)(want < sms ? want : sms);
    kern<<<grid, C::NT, C::SMEM_BYTES, st>>>(pq, idx, static_cast<const uint8_t*>(wtc), out, static_cast<uint8_t*>(out_split),
                                             N, tiles_per_patch, total, slope);
    return pf_launch_status("edgeconv_tc_kernel");
}

}  // namespace

static int edgeconv_tc_dispatch(const float* pq, const int32_t* idx, const void* wtc, float* out, void* out_split, int Bp,
                                int N, int K, int G_, int C_out, float slope, void* stream) {
    PF_CHECK_ARG(slope >= 0.f && slope <= 1.f, "LeakyReLU slope %g outside [0, 1]", (double)slope);
    PF_CHECK_ARG(Bp >= 0 && N > 0, "pf_edgeconv_tc: bad sizes Bp=%d N=%d", Bp, N);
    if (Bp == 0) return 0;
    PF_CHECK_ARG(pq && idx && wtc && (out || out_split), "pf_edgeconv_tc: null pointer");
    PF_CHECK_ARG((long long)Bp * N < 0x7fffffffLL, "pf_edgeconv_tc: Bp*N overflows int32");
    PF_CHECK_ARG(N <= 65536, "pf_edgeconv_tc: N=%d > 65536 points per patch", N);
    PF_CHECK_ARG(K == KN, "pf_edgeconv_tc: K=%d, only K = 16", K);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(pq) | reinterpret_cast<uintptr_t>(wtc) | reinterpret_cast<uintptr_t>(out) |
                   reinterpret_cast<uintptr_t>(out_split)) & 15) == 0,
                 "pf_edgeconv_tc: pointers must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    if (G_ == 32 && C_out == 128) return launch<32, 128, 2>(pq, idx, wtc, out, out_split, Bp, N, slope, st);
    if (G_ == 16 && C_out == 64) return launch<16, 64, 3>(pq, idx, wtc, out, out_split, Bp, N, slope, st);
    if (G_ == 8 && C_out == 32) return launch<8, 32, 3>(pq, idx, wtc, out, out_split, Bp, N, slope, st);
    PF_CHECK_ARG(false, "pf_edgeconv_tc: (G, C_out) = (%d, %d) not in {(8,32), (16,64), (32,128)}", G_, C_out);
    return 0;
}

#ifdef PF_PROF
extern "C" __attribute__((visibility("default"))) int pf_debug_edgeconv_prof(long long* host32) {
    return cudaMemcpyFromSymbol(host32, g_prof, sizeof(long long) * 32) == cudaSuccess ? 0 : 1;
}
#endif

extern "C" int pf_edgeconv_tc(const float* pq, const int32_t* idx, const void* wtc, float* out, int Bp, int N, int K,
                              int G_, int C_out, float slope, void* stream) {
    return edgeconv_tc_dispatch(pq, idx, wtc, out, nullptr, Bp, N, K, G_, C_out, slope, stream);
}

extern "C" int pf_edgeconv_tc_split(const float* pq, const int32_t* idx, const void* wtc, float* out, void* out_split, int Bp,
                                    int N, int K, int G_, int C_out, float slope, void* stream) {
    return edgeconv_tc_dispatch(pq, idx, wtc, out, out_split, Bp, N, K, G_, C_out, slope, stream);
}
