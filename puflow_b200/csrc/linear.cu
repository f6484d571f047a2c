// Per-point linear layer:  out = act(A . W + bias)   (FP32 FFMA, 128 x BN output tile per CTA).
//
// Serves every per-POINT contraction of the path: the P/Q projections the fused EdgeConv and
// interpolation kernels gather from, FeatMergeUnit (/root/reference/modules/discrete/
// interpflow.py:251-258) and the conditioning first layers of LinearA1D (:22-43).
// A tile rows are transposed into shared memory (k-major) so the micro-kernel reads them with
// conflict-free LDS.128; W rows are already k-major.  K is consumed in chunks of 64.
#include "common.cuh"
#include "puflow_b200.h"

namespace {

constexpr int BM = 128, BK = 64, NT = 256;

template <int TO>
__global__ void __launch_bounds__(NT) linear_kernel(const float* __restrict__ A, int64_t lda,
                                                    const float* __restrict__ W, const float* __restrict__ bias,
                                                    float* __restrict__ out, int64_t ldo, int64_t M, int K, int N,
                                                    int act, float slope) {
    constexpr int BN = 16 * TO;
    extern __shared__ __align__(16) float smem[];
    float* hs = smem;                 // [BK][BM]
    float* ws = smem + BK * BM;       // [BK][BN]

    const int tid = threadIdx.x;
    const int64_t m0 = (int64_t)blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;
    const int te = tid & 15, to = tid >> 4;
    const bool vecA = ((lda & 3) == 0) && ((K & 3) == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);

    float acc[8][TO];
    pf::zero_acc<8, TO>(acc);

    for (int k0 = 0; k0 < K; k0 += BK) {
        const int kc = min(BK, K - k0);
        if (k0) __syncthreads();
        if (vecA) {
            const int nq = kc >> 2;
            for (int it = tid; it < BM * nq; it += NT) {
                int m = it & (BM - 1), kq = it >> 7;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (m0 + m < M) v = __ldg(reinterpret_cast<const float4*>(A + (m0 + m) * lda + k0 + 4 * kq));
                hs[(4 * kq + 0) * BM + m] = v.x;
                hs[(4 * kq + 1) * BM + m] = v.y;
                hs[(4 * kq + 2) * BM + m] = v.z;
                hs[(4 * kq + 3) * BM + m] = v.w;
            }
        } else {
            for (int it = tid; it < BM * kc; it += NT) {
                int m = it & (BM - 1), k = it >> 7;
                hs[k * BM + m] = (m0 + m < M) ? __ldg(A + (m0 + m) * lda + k0 + k) : 0.f;
            }
        }
        for (int it = tid; it < kc * (BN / 4); it += NT) {
            int nq = it % (BN / 4), k = it / (BN / 4);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (n0 + 4 * nq < N) v = __ldg(reinterpret_cast<const float4*>(W + (size_t)(k0 + k) * N + n0 + 4 * nq));
            *reinterpret_cast<float4*>(ws + k * BN + 4 * nq) = v;
        }
        __syncthreads();
        pf::tile_gemm<8, TO, 64>(hs + te * 4, BM, kc, ws + to * TO, BN, acc);
    }

    // epilogue: bias + activation, rows {te*4+v, 64+te*4+v}, columns n0 + to*TO ...
    float bv[TO];
#pragma unroll
    for (int o = 0; o < TO; o++) {
        int n = n0 + to * TO + o;
        bv[o] = (bias != nullptr && n < N) ? __ldg(bias + n) : 0.f;
    }
    if (n0 + to * TO < N) {
#pragma unroll
        for (int e = 0; e < 8; e++) {
            int64_t m = m0 + (e < 4 ? te * 4 + e : 64 + te * 4 + (e - 4));
            if (m >= M) continue;
            float v[TO];
#pragma unroll
            for (int o = 0; o < TO; o++) {
                float t = acc[e][o] + bv[o];
                if (act == PF_ACT_RELU) t = fmaxf(t, 0.f);
                else if (act == PF_ACT_LRELU) t = pf::lrelu(t, slope);
                v[o] = t;
            }
            float* dst = out + m * ldo + n0 + to * TO;
            if constexpr (TO % 4 == 0) {
#pragma unroll
                for (int o = 0; o < TO; o += 4)
                    *reinterpret_cast<float4*>(dst + o) = make_float4(v[o], v[o + 1], v[o + 2], v[o + 3]);
            } else {
                *reinterpret_cast<float2*>(dst) = make_float2(v[0], v[1]);
            }
        }
    }
}

}  // namespace

extern "C" int pf_linear(const float* A, int64_t lda, const float* W, const float* bias, float* out, int64_t ldo,
                         int64_t M, int K, int N, int act, float slope, void* stream) {
    PF_CHECK_ARG(slope >= 0.f && slope <= 1.f, "LeakyReLU slope %g outside [0, 1]", (double)slope);
    PF_CHECK_ARG(A && W && out, "pf_linear: null pointer");
    PF_CHECK_ARG(M >= 0 && K > 0 && N > 0, "pf_linear: bad sizes M=%lld K=%d N=%d", (long long)M, K, N);
    PF_CHECK_ARG((N & 3) == 0 && (ldo & 3) == 0, "pf_linear: N=%d and ldo=%lld must be multiples of 4", N,
                 (long long)ldo);
    PF_CHECK_ARG(N <= 32 || (N & 7) == 0, "pf_linear: N=%d > 32 must be a multiple of 8", N);
    PF_CHECK_ARG(lda >= K && ldo >= N, "pf_linear: lda/ldo smaller than the row");
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
                 "pf_linear: W / out must be 16-byte aligned");
    PF_CHECK_ARG(act >= 0 && act <= 2, "pf_linear: act=%d", act);
    if (M == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int64_t gm = (M + BM - 1) / BM;
    PF_CHECK_ARG(gm <= 0x7fffffffLL, "pf_linear: M too large");
    if (N > 32) {
        constexpr size_t smem = (size_t)(BK * BM + BK * 128) * sizeof(float);
        if (int rc = pf_smem_optin(linear_kernel<8>, smem)) return rc;
        dim3 grid((unsigned)gm, (N + 127) / 128);
        linear_kernel<8><<<grid, NT, smem, st>>>(A, lda, W, bias, out, ldo, M, K, N, act, slope);
    } else {
        constexpr size_t smem = (size_t)(BK * BM + BK * 32) * sizeof(float);
        dim3 grid((unsigned)gm, 1);
        linear_kernel<2><<<grid, NT, smem, st>>>(A, lda, W, bias, out, ldo, M, K, N, act, slope);
    }
    return pf_launch_status("linear_kernel");
}
