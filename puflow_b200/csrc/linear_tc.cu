// Per-point linear layer on the tensor cores:  out = act(A . W + bias), fp32 in / fp32 out.
//
// Same contract as pf_linear (linear.cu) for the K >= 16 layers of the path -- the P|Q projections the fused
// EdgeConv kernels gather from, FeatMergeUnit (/root/reference/modules/discrete/interpflow.py:251-258) and the
// conditioning first layers of LinearA1D (:22-43).  ncu/CUDA events show the FFMA version FP32-pipe-bound
// (K=128, N=512: 134 GFLOP in 3.76 ms = 36 TFLOP/s while its HBM traffic, 2.7 GB, would take 0.4 ms).
// Here the contraction is tcgen05.mma.kind::f16 on fp16 hi/lo split operands (3 MMAs per product, fp32
// accumulation in TMEM; fp32-class result, see edgeconv_tc.cu), which makes the layer HBM-bound.
//
// Grid: persistent CTAs, CTA c works on column chunk c % ny (<= 128 output columns, weights resident in shared
// memory) and on the 128-row tiles  c / ny, c / ny + nworkers, ...  -- the ny CTAs that share an A tile run
// side by side, so the tile is read from HBM once and from L2 afterwards.
// CTA: 2 warpgroups ping-pong over tiles (load A tile fp32 -> split -> shared memory | MMA | epilogue) + one
// MMA-issuing warp per warpgroup.
#include "common.cuh"
#include "puflow_b200.h"
#include "tc.cuh"

namespace {

using namespace pf;

constexpr int TMR = 128;                 // rows per tile
constexpr int NT = 320;
constexpr uint32_t W_MAX = 65536;        // 128 columns x 128 K x 2 B x (hi, lo)

// A tile: chunk c (8 halfs of K) of row r at  c * lbo + (r / 8) * 128 + (r % 8) * 16;  lbo = 2048 + 16 * pad
// (pad chosen so that the 8 lanes of a quarter warp hit distinct banks when they store).
__host__ __device__ constexpr uint32_t a_lbo(int nch) { return 2048u + 16u * (nch >= 8 ? 1 : (nch == 4 ? 2 : 4)); }

template <int NCH>   // NCH = K / 8 in {2, 4, 8, 16}
__global__ void __launch_bounds__(NT, 1)
    linear_tc_kernel(const float* __restrict__ A, int64_t lda, const uint8_t* __restrict__ wblob,
                     const float* __restrict__ bias, float* __restrict__ out, int64_t ldo, int64_t M, int N, int ny,
                     int nworkers, int act, float slope, int prec) {
    constexpr int K = NCH * 8;
    constexpr uint32_t LBO = a_lbo(NCH);
    constexpr uint32_t A_PART = NCH * LBO;
    constexpr uint32_t A_WG = 2 * A_PART > 18432u ? 2 * A_PART : 18432u;   // also the epilogue's staging area
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    uint8_t* a_base = smem + W_MAX;                          // A[wg][hi|lo]
    uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + W_MAX + 2 * A_WG);
    uint64_t* mbar = wbar + 1;
    uint64_t* ready = wbar + 3;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wbar + 5);

    const int tid = threadIdx.x, warp = tid >> 5;
    const bool is_mma_warp = warp >= 8;
    const int wg = is_mma_warp ? warp - 8 : warp >> 2, wq = warp & 3;
    const int e = tid & 127;

    const int ychunk = blockIdx.x % ny, worker = blockIdx.x / ny;
    const int n0 = ychunk * 128;
    const int nc = min(128, N - n0);                         // columns of this chunk (multiple of 16)
    const uint32_t w_part = (uint32_t)nc * K * 2;

    if (tid == 0) {
        mbar_init(wbar, 1);
        mbar_init(mbar, 1), mbar_init(mbar + 1, 1);
        mbar_init(ready, 1), mbar_init(ready + 1, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 256);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    if (tid == 0) {
        mbar_expect_tx(wbar, 2 * w_part);
        bulk_g2s_chunked(smem, wblob + (size_t)ychunk * (128 * K * 4), 2 * w_part, wbar);
    }
    const int64_t ntiles = (M + TMR - 1) / TMR;
    const uint32_t d_tmem = tmem_base + (uint32_t)wg * 128;
    const uint32_t a_hi = sbase + W_MAX + (uint32_t)wg * A_WG, a_lo = a_hi + A_PART;

    if (is_mma_warp) {
        tc::mbar_wait_bounded(wbar, 0);
        if (tc::elect_one()) {             // not `lane == 0`: see tc.cuh
            const uint32_t idesc = tc::idesc_f16(128, nc);
            const uint64_t da_hi = tc::smem_desc(a_hi, 128, LBO), da_lo = tc::smem_desc(a_lo, 128, LBO);
            const uint64_t db_hi = tc::smem_desc(sbase, K * 16), db_lo = tc::smem_desc(sbase + w_part, K * 16);
            constexpr uint32_t a_step = (2 * LBO) >> 4, b_step = 256 >> 4;
            uint32_t rphase = 0;
            for (int64_t t = worker * 2 + wg; t < ntiles; t += (int64_t)nworkers * 2) {
                tc::mbar_wait_bounded(ready + wg, rphase);
                rphase ^= 1;
                tc::fence_after_sync();
                // the small cross terms first: the accumulator adder truncates, so the fewer additions at full magnitude
                // (the hi . hi pass) the better; the single-pass mode issues the hi . hi pass alone
                if (prec == PF_PREC_FP32_SPLIT) {
#pragma unroll
                    for (int s = 0; s < NCH / 2; s++) tc::mma_f16(d_tmem, da_lo + a_step * s, db_hi + b_step * s, idesc, s > 0);
#pragma unroll
                    for (int s = 0; s < NCH / 2; s++) tc::mma_f16(d_tmem, da_hi + a_step * s, db_lo + b_step * s, idesc, 1);
                }
#pragma unroll
                for (int s = 0; s < NCH / 2; s++)
                    tc::mma_f16(d_tmem, da_hi + a_step * s, db_hi + b_step * s, idesc, (prec == PF_PREC_FP32_SPLIT || s > 0) ? 1u : 0u);
                tc::commit(mbar + wg);
            }
        }
        __syncwarp();
    } else {
        uint8_t* a_hi_p = a_base + (size_t)wg * A_WG;
        uint8_t* a_lo_p = a_hi_p + A_PART;
        const uint32_t lane_sel = (uint32_t)(wq * 32) << 16;
        const int bar_id = 1 + wg;
        uint32_t phase = 0;
        for (int64_t t = worker * 2 + wg; t < ntiles; t += (int64_t)nworkers * 2) {
            const int64_t m0 = t * TMR;
            // ---- A tile: fp32 global -> fp16 hi/lo in the tensor-core layout ----
            constexpr int BATCH = NCH < 8 ? NCH : 8;
#pragma unroll
            for (int it0 = 0; it0 < NCH; it0 += BATCH) {
                float4 v[BATCH][2];
#pragma unroll
                for (int u = 0; u < BATCH; u++) {
                    const int i = (it0 + u) * 128 + e, r = i / NCH, c = i % NCH;
                    const int64_t row = m0 + r;
                    if (row < M) {
                        const float4* src = reinterpret_cast<const float4*>(A + row * lda + 8 * c);
                        v[u][0] = __ldg(src), v[u][1] = __ldg(src + 1);
                    } else {
                        v[u][0] = v[u][1] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
#pragma unroll
                for (int u = 0; u < BATCH; u++) {
                    const int i = (it0 + u) * 128 + e, r = i / NCH, c = i % NCH;
                    uint32_t hi[4], lo[4];
                    tc::split2(v[u][0].x, v[u][0].y, hi[0], lo[0]);
                    tc::split2(v[u][0].z, v[u][0].w, hi[1], lo[1]);
                    tc::split2(v[u][1].x, v[u][1].y, hi[2], lo[2]);
                    tc::split2(v[u][1].z, v[u][1].w, hi[3], lo[3]);
                    const uint32_t off = (uint32_t)c * LBO + (uint32_t)(r >> 3) * 128 + (uint32_t)(r & 7) * 16;
                    *reinterpret_cast<uint4*>(a_hi_p + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                    *reinterpret_cast<uint4*>(a_lo_p + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                }
            }
            tc::fence_async_smem();
            tc::fence_before_sync();
            tc::named_bar(bar_id, 128);
            if (e == 0) mbar_arrive(ready + wg);
            // ---- epilogue: thread = row ----
            tc::mbar_wait_bounded(mbar + wg, phase);
            phase ^= 1;
            tc::fence_after_sync();
            // thread = row out of TMEM; bias + activation; transposed through a warp-private staging area (the dead
            // A tile) so that global stores are full 128-byte row segments
            float* stg = reinterpret_cast<float*>(a_hi_p) + wq * (32 * 36);       // 32 rows x 36 floats per warp
            const int lane = tid & 31;
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
                for (int u = 0; u < 32; u += 4) {
                    float4 o = make_float4(acc[u] * tc::WINV, acc[u + 1] * tc::WINV, acc[u + 2] * tc::WINV, acc[u + 3] * tc::WINV);
                    if (u < w) {
                        if (bias) {
                            const float4 b = __ldg(reinterpret_cast<const float4*>(bias + n0 + c0 + u));
                            o.x += b.x, o.y += b.y, o.z += b.z, o.w += b.w;
                        }
                        if (act == PF_ACT_RELU) {
                            o.x = fmaxf(o.x, 0.f), o.y = fmaxf(o.y, 0.f), o.z = fmaxf(o.z, 0.f), o.w = fmaxf(o.w, 0.f);
                        } else if (act == PF_ACT_LRELU) {
                            o.x = lrelu(o.x, slope), o.y = lrelu(o.y, slope), o.z = lrelu(o.z, slope), o.w = lrelu(o.w, slope);
                        }
                    }
                    *reinterpret_cast<float4*>(stg + lane * 36 + u) = o;
                }
                __syncwarp();
                const int c4 = (lane & 7) * 4;
#pragma unroll
                for (int it = 0; it < 8; it++) {
                    const int r = it * 4 + (lane >> 3);
                    const int64_t row = m0 + wq * 32 + r;
                    if (row < M && c4 < w)
                        *reinterpret_cast<float4*>(out + row * ldo + n0 + c0 + c4) =
                            *reinterpret_cast<const float4*>(stg + r * 36 + c4);
                }
                __syncwarp();
            }
            tc::fence_before_sync();
            tc::named_bar(bar_id, 128);      // the staging area is the next tile's A buffer
        }
    }
    if (tid == 0) tc::mbar_wait_bounded(wbar, 0);   // never exit with the bulk copy in flight
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem_base, 256);
}

template <int NCH>
int launch(const float* A, int64_t lda, const uint8_t* w, const float* bias, float* out, int64_t ldo, int64_t M, int N,
           int act, float slope, int prec, cudaStream_t st) {
    const size_t a_wg = 2 * (size_t)NCH * a_lbo(NCH) > 18432 ? 2 * (size_t)NCH * a_lbo(NCH) : 18432;
    const size_t smem = W_MAX + 2 * a_wg + 64;
    auto kern = linear_tc_kernel<NCH>;
    if (int rc = pf_smem_optin(kern, smem)) return rc;
    int dev = 0, sms = 148;
    PF_CUDA(cudaGetDevice(&dev));
    PF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int ny = (N + 127) / 128;
    const int64_t ntiles = (M + TMR - 1) / TMR;
    int nworkers = sms / ny;
    if (nworkers < 1) nworkers = 1;
    if ((int64_t)nworkers * 2 > ntiles) nworkers = (int)((ntiles + 1) / 2);
    kern<<<nworkers * ny, NT, smem, st>>>(A, lda, w, bias, out, ldo, M, N, ny, nworkers, act, slope, prec);
    return pf_launch_status("linear_tc_kernel");
}

}  // namespace

extern "C" int pf_linear_tc(const float* A, int64_t lda, const void* wtc, const float* bias, float* out, int64_t ldo,
                            int64_t M, int K, int N, int act, float slope, int precision, void* stream) {
    PF_CHECK_ARG(precision == PF_PREC_FP32_SPLIT || precision == PF_PREC_FP16, "pf_linear_tc: precision %d (PF_PREC_FP32_SPLIT or PF_PREC_FP16)", precision);
    PF_CHECK_ARG(slope >= 0.f && slope <= 1.f, "LeakyReLU slope %g outside [0, 1]", (double)slope);
    PF_CHECK_ARG(A && wtc && out, "pf_linear_tc: null pointer");
    PF_CHECK_ARG(M >= 0 && N > 0, "pf_linear_tc: bad sizes M=%lld N=%d", (long long)M, N);
    PF_CHECK_ARG(K == 16 || K == 32 || K == 64 || K == 128, "pf_linear_tc: K=%d not in {16,32,64,128}", K);
    PF_CHECK_ARG(N % 16 == 0, "pf_linear_tc: N=%d must be a multiple of 16", N);
    PF_CHECK_ARG(lda % 4 == 0 && ldo % 4 == 0 && lda >= K && ldo >= N, "pf_linear_tc: bad row strides");
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(wtc) | reinterpret_cast<uintptr_t>(out) |
                   reinterpret_cast<uintptr_t>(bias)) & 15) == 0,
                 "pf_linear_tc: pointers must be 16-byte aligned");
    PF_CHECK_ARG(act >= 0 && act <= 2, "pf_linear_tc: bad activation code %d", act);
    if (M == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const uint8_t* w = static_cast<const uint8_t*>(wtc);
    switch (K) {
        case 16: return launch<2>(A, lda, w, bias, out, ldo, M, N, act, slope, precision, st);
        case 32: return launch<4>(A, lda, w, bias, out, ldo, M, N, act, slope, precision, st);
        case 64: return launch<8>(A, lda, w, bias, out, ldo, M, N, act, slope, precision, st);
        default: return launch<16>(A, lda, w, bias, out, ldo, M, N, act, slope, precision, st);
    }
}
