// Per-point linear layer whose input arrives PRE-SPLIT in the tensor-core operand layout:  out = act(A . W + bias).
//
// Same arithmetic as pf_linear_tc (linear_tc.cu; the P|Q projections the fused EdgeConv kernels gather from,
// /root/reference/modules/discrete/interpflow.py:223-232): fp16 hi/lo split operands, 3 tcgen05.mma per product, fp32
// accumulation in TMEM.  ncu shows pf_linear_tc at 23-30 % of its HBM floor (profiles/r01_ncu_v6_all_tc.md): the
// fp32 -> hi/lo conversion of every A tile sits between the global load and the MMA, with two tiles in flight per SM.
// Here the producer of the activations (pf_edgeconv_tc_split's epilogue) has already written them as fp16 hi/lo in the
// tile-blocked operand layout, so a tile is ONE contiguous TMA bulk copy straight into the MMA's A operand: a
// multi-stage mbarrier pipeline (producer thread -> MMA thread -> 2 epilogue warpgroups) keeps several tiles of loads
// in flight and the CUDA cores only run the epilogue (tcgen05.ld -> bias / activation -> 256-bit row stores).
//
// Split layout (PF_SPLIT_TILE_BYTES(K) = 512 K per 128-row tile, tile t at t * 512 K):
//     hi(r, k) at (k / 8) * 2048 + r * 16 + (k % 8) * 2,   lo at K * 256 + the same        (r = row % 128)
// i.e. 8-row x 16-byte core matrices, adjacent along the rows (SBO = 128), K-chunks 2,048 B apart (LBO = 2048).
#include "common.cuh"
#include "puflow_b200.h"
#include "tc.cuh"

namespace {

using namespace pf;

constexpr int TMR = 128;
constexpr int NT = 320;                        // 2 epilogue warpgroups + producer warp + MMA warp
constexpr uint32_t W_MAX = 65536;              // 128 columns x 128 K x 2 B x (hi, lo)
constexpr uint32_t LBO = 2048;

__device__ __forceinline__ void stg256(float* p, const float (&v)[8]) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
                 "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
                 : "memory");
}

template <int NCH, int S>   // NCH = K / 8, S = A stages
__global__ void __launch_bounds__(NT, 1)
    linear_split_kernel(const uint8_t* __restrict__ asplit, const uint8_t* __restrict__ wblob, const float* __restrict__ bias,
                        float* __restrict__ out, int64_t ldo, int64_t M, int N, int ny, int nworkers, int act, float slope,
                        int prec) {
    constexpr int K = NCH * 8;
    constexpr uint32_t A_PART = NCH * LBO, A_TILE = 2 * A_PART;
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + W_MAX + S * A_TILE);
    uint64_t* full = wbar + 1;          // [S] A tile landed
    uint64_t* empty = full + S;         // [S] MMAs that read the stage have completed
    uint64_t* acc_full = empty + S;     // [2] accumulator of warpgroup g complete
    uint64_t* acc_empty = acc_full + 2; // [2] warpgroup g has read its accumulator
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
    float* sbias = reinterpret_cast<float*>(smem + W_MAX + S * A_TILE + 128);     // [128] bias of this column chunk (zeros if none)

    const int tid = threadIdx.x, warp = tid >> 5;
    const int ychunk = blockIdx.x % ny, worker = blockIdx.x / ny;
    const int n0 = ychunk * 128;
    const int nc = min(128, N - n0);                         // columns of this chunk (multiple of 16)
    const uint32_t w_part = (uint32_t)nc * K * 2;

    if (tid == 0) {
        mbar_init(wbar, 1);
        for (int i = 0; i < S; i++) mbar_init(full + i, 1), mbar_init(empty + i, 1);
        for (int i = 0; i < 2; i++) mbar_init(acc_full + i, 1), mbar_init(acc_empty + i, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 256);
    if (tid < 128) sbias[tid] = (bias && tid < nc) ? __ldg(bias + n0 + tid) : 0.f;
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const int64_t ntiles = (M + TMR - 1) / TMR;
    // this CTA's tiles: worker, worker + nworkers, ...;  tile i uses A stage i % S and accumulator / warpgroup i % 2

    if (warp == 8) {
        // ---- producer: weights once, then one bulk copy per tile ----
        if ((tid & 31) == 0) {
            mbar_expect_tx(wbar, 2 * w_part);
            bulk_g2s_chunked(smem, wblob + (size_t)ychunk * (128 * K * 4), 2 * w_part, wbar);
            int64_t i = 0;
            for (int64_t t = worker; t < ntiles; t += nworkers, i++) {
                const int st = (int)(i % S);
                if (i >= S) tc::mbar_wait_bounded(empty + st, (uint32_t)((i / S - 1) & 1));
                mbar_expect_tx(full + st, A_TILE);
                bulk_g2s_chunked(smem + W_MAX + (size_t)st * A_TILE, asplit + (size_t)t * A_TILE, A_TILE, full + st);
            }
        }
    } else if (warp == 9) {
        // ---- MMA issuer ----
        tc::mbar_wait_bounded(wbar, 0);
        if (tc::elect_one()) {             // not `lane == 0`: see tc.cuh
            const uint32_t idesc = tc::idesc_f16(128, nc);
            const uint64_t db_hi = tc::smem_desc(sbase, K * 16), db_lo = tc::smem_desc(sbase + w_part, K * 16);
            constexpr uint32_t a_step = (2 * LBO) >> 4, b_step = 256 >> 4;
            int64_t i = 0;
            for (int64_t t = worker; t < ntiles; t += nworkers, i++) {
                const int st = (int)(i % S), g = (int)(i & 1);
                tc::mbar_wait_bounded(full + st, (uint32_t)((i / S) & 1));
                if (i >= 2) tc::mbar_wait_bounded(acc_empty + g, (uint32_t)((i / 2 - 1) & 1));
                tc::fence_after_sync();
                const uint32_t a_hi = sbase + W_MAX + (uint32_t)st * A_TILE, a_lo = a_hi + A_PART;
                const uint64_t da_hi = tc::smem_desc(a_hi, 128, LBO), da_lo = tc::smem_desc(a_lo, 128, LBO);
                const uint32_t d_tmem = tmem_base + (uint32_t)g * 128;
                if (prec == PF_PREC_FP32_SPLIT) {
#pragma unroll
                    for (int s = 0; s < NCH / 2; s++) tc::mma_f16(d_tmem, da_lo + a_step * s, db_hi + b_step * s, idesc, s > 0);
#pragma unroll
                    for (int s = 0; s < NCH / 2; s++) tc::mma_f16(d_tmem, da_hi + a_step * s, db_lo + b_step * s, idesc, 1);
                }
#pragma unroll
                for (int s = 0; s < NCH / 2; s++)      // full magnitude last; alone in the single-pass mode
                    tc::mma_f16(d_tmem, da_hi + a_step * s, db_hi + b_step * s, idesc, (prec == PF_PREC_FP32_SPLIT || s > 0) ? 1u : 0u);
                tc::commit(empty + st);
                tc::commit(acc_full + g);
            }
        }
        __syncwarp();
    } else {
        // ---- epilogue warpgroup g: thread = row ----
        const int g = warp >> 2, wq = warp & 3, e = tid & 127;
        const uint32_t lane_sel = (uint32_t)(wq * 32) << 16;
        const uint32_t d_tmem = tmem_base + (uint32_t)g * 128;
        int64_t i = 0;
        for (int64_t t = worker; t < ntiles; t += nworkers, i++) {
            if ((i & 1) != g) continue;
            tc::mbar_wait_bounded(acc_full + g, (uint32_t)((i / 2) & 1));
            tc::fence_after_sync();
            const int64_t row = t * TMR + e;
            float* orow = out + row * ldo + n0;
            for (int c0 = 0; c0 < nc; c0 += 32) {
                float acc[32];
                if (nc - c0 >= 32) {
                    tc::ld32(d_tmem + lane_sel + c0, acc);
                } else {
                    float a16[16];
                    tc::ld16(d_tmem + lane_sel + c0, a16);
#pragma unroll
                    for (int u = 0; u < 16; u++) acc[u] = a16[u], acc[16 + u] = 0.f;
                }
                const int w = min(32, nc - c0);
#pragma unroll
                for (int u = 0; u < 32; u += 8) {
                    if (u < w) {
                        float o[8];
                        const float4 b0 = *reinterpret_cast<const float4*>(sbias + c0 + u);
                        const float4 b1 = *reinterpret_cast<const float4*>(sbias + c0 + u + 4);
                        const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                        for (int v = 0; v < 8; v++) {
                            float x = fmaf(acc[u + v], tc::WINV, bv[v]);
                            if (act == PF_ACT_RELU) x = fmaxf(x, 0.f);
                            else if (act == PF_ACT_LRELU) x = lrelu(x, slope);
                            o[v] = x;
                        }
                        if (row < M) stg256(orow + c0 + u, o);
                    }
                }
            }
            tc::fence_before_sync();
            tc::named_bar(1 + g, 128);
            if (e == 0) mbar_arrive(acc_empty + g);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem_base, 256);
}

template <int NCH, int S>
int launch(const uint8_t* a, const uint8_t* w, const float* bias, float* out, int64_t ldo, int64_t M, int N, int act,
           float slope, int prec, cudaStream_t st) {
    const size_t smem = W_MAX + (size_t)S * 2 * NCH * LBO + 128 + 512;
    auto kern = linear_split_kernel<NCH, S>;
    if (int rc = pf_smem_optin(kern, smem)) return rc;
    int dev = 0, sms = 148;
    PF_CUDA(cudaGetDevice(&dev));
    PF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int ny = (N + 127) / 128;
    const int64_t ntiles = (M + TMR - 1) / TMR;
    int nworkers = sms / ny;
    if (nworkers < 1) nworkers = 1;
    if ((int64_t)nworkers > ntiles) nworkers = (int)ntiles;
    kern<<<nworkers * ny, NT, smem, st>>>(a, w, bias, out, ldo, M, N, ny, nworkers, act, slope, prec);
    return pf_launch_status("linear_split_kernel");
}

}  // namespace

extern "C" int pf_linear_tc_split(const void* a_split, const void* wtc, const float* bias, float* out, int64_t ldo, int64_t M,
                                  int K, int N, int act, float slope, int precision, void* stream) {
    PF_CHECK_ARG(precision == PF_PREC_FP32_SPLIT || precision == PF_PREC_FP16, "pf_linear_tc_split: precision %d (PF_PREC_FP32_SPLIT or PF_PREC_FP16)", precision);
    PF_CHECK_ARG(slope >= 0.f && slope <= 1.f, "LeakyReLU slope %g outside [0, 1]", (double)slope);
    PF_CHECK_ARG(M >= 0 && N > 0, "pf_linear_tc_split: bad sizes M=%lld N=%d", (long long)M, N);
    if (M == 0) return 0;
    PF_CHECK_ARG(a_split && wtc && out, "pf_linear_tc_split: null pointer");
    PF_CHECK_ARG(K == 32 || K == 64 || K == 128, "pf_linear_tc_split: K=%d not in {32,64,128}", K);
    PF_CHECK_ARG(N % 16 == 0, "pf_linear_tc_split: N=%d must be a multiple of 16", N);
    PF_CHECK_ARG(ldo % 8 == 0 && ldo >= N, "pf_linear_tc_split: ldo must be a multiple of 8 and >= N");
    PF_CHECK_ARG((reinterpret_cast<uintptr_t>(a_split) & 127) == 0 &&
                     ((reinterpret_cast<uintptr_t>(wtc) | reinterpret_cast<uintptr_t>(bias)) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(out) & 31) == 0,
                 "pf_linear_tc_split: alignment (a_split 128 B, out 32 B, weights / bias 16 B)");
    PF_CHECK_ARG(act >= 0 && act <= 2, "pf_linear_tc_split: bad activation code %d", act);
    cudaStream_t st = (cudaStream_t)stream;
    const uint8_t* a = static_cast<const uint8_t*>(a_split);
    const uint8_t* w = static_cast<const uint8_t*>(wtc);
    switch (K) {
        case 32: return launch<4, 4>(a, w, bias, out, ldo, M, N, act, slope, precision, st);
        case 64: return launch<8, 4>(a, w, bias, out, ldo, M, N, act, slope, precision, st);
        default: return launch<16, 2>(a, w, bias, out, ldo, M, N, act, slope, precision, st);
    }
}
