// Fused EdgeConv block: neighbour gather + dense block + conv_out + max over K, one launch.
//
// Replaces FeatureExtractUnit.forward, /root/reference/modules/discrete/interpflow.py:223-248
// (derive_edge_feat :223-232, dense block :238-240, conv_out :242, max-pool :245), which runs as
// ~17 ATen launches and materialises [B, 3C+4g, N, K] in HBM.
//
// Factoring: the edge feature [x_i, x_j, x_j - x_i] is linear in (x_i, x_j), and the
// dense block re-reads it in every conv through the concat, so W.[x_i; x_j; x_j-x_i] =
// (W1-W3).x_i + (W2+W3).x_j for ALL five convs is two per-POINT projections P, Q (pf_linear),
// gathered here per edge; what is left per edge are the hidden-to-hidden contractions
// (6 g^2 + 4 g C_out MACs), done as register-tiled FP32 FFMA GEMMs on a 128-edge tile whose
// activations live k-major in shared memory.  Weights (<= 88 KB) are staged once per CTA by TMA
// bulk copies.  No edge tensor ever reaches HBM.
//
// Tile geometry: 128 edge slots = PT points x K neighbours.  Slot s -> (point, neighbour):
//   te = (s % 64) / 4, v = s % 4, hi = s / 64;  point = te / TPP, nbr = (te % TPP) * 8 + hi * 4 + v
// with TPP = K / 8 threads per point in the conv_out stage, so that the 8 slots a conv_out thread
// owns (two quads half a tile apart, see tile_gemm) belong to one point and the max over K is a
// register max plus log2(TPP) shuffles.
#include "common.cuh"
#include "puflow_b200.h"

namespace {

constexpr int TM = 128;   // edge slots per tile
constexpr int NT = 256;   // threads per CTA

template <int G, int CO, int KN>
__global__ void __launch_bounds__(NT, 1)
    edgeconv_kernel(const float* __restrict__ pq, const int32_t* __restrict__ idx, const float* __restrict__ wd,
                    const float* __restrict__ wo, float* __restrict__ out, int N, int tiles_per_patch,
                    int tiles_per_cta, float slope) {
    constexpr int R = 4 * G + CO, R2 = 2 * R;
    constexpr int PT = TM / KN;          // points per tile
    constexpr int TPP = 16 / PT;         // conv_out threads per point (= KN / 8)
    constexpr int WD = 6 * G * G, WO = 4 * G * CO;
    constexpr int TOD = G / 8;           // dense-layer outputs per thread
    constexpr int TOO = CO / 16;         // conv_out outputs per thread

    extern __shared__ __align__(16) float smem[];
    float* wds = smem;                   // [WD]
    float* wos = wds + WD;               // [4G][CO]
    float* hs = wos + WO;                // [4G][TM]
    int* jrow = reinterpret_cast<int*>(hs + 4 * G * TM);   // [TM] global row of the neighbour
    int* irow = jrow + TM;                                  // [TM] global row of the own point
    uint64_t* bar = reinterpret_cast<uint64_t*>(irow + TM);

    const int tid = threadIdx.x;
    const int patch = blockIdx.x;
    const int t_begin = blockIdx.y * tiles_per_cta;
    const int t_end = min(t_begin + tiles_per_cta, tiles_per_patch);
    if (t_begin >= t_end) return;

    if (tid == 0) {
        pf::mbar_init(bar, 1);
        pf::mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {
        pf::mbar_expect_tx(bar, (WD + WO) * 4);
        pf::bulk_g2s_chunked(wds, wd, WD * 4, bar);
        pf::bulk_g2s_chunked(wos, wo, WO * 4, bar);
    }
    bool weights_ready = false;

    for (int t = t_begin; t < t_end; t++) {
        const int p0 = t * PT;
        if (tid < TM) {
            int s = tid, te = (s & 63) >> 2, v = s & 3, hi = s >> 6;
            int pi = min(p0 + te / TPP, N - 1);
            int nb = (te % TPP) * 8 + hi * 4 + v;
            int row = patch * N + pi;
            int j = __ldg(idx + (size_t)row * KN + nb);
            j = min(max(j, 0), N - 1);
            irow[s] = row;
            jrow[s] = patch * N + j;
        }
        __syncthreads();

        // ---- layer 0: h0 = lrelu(P0_i + Q0_j), pure gather ----
        {
            constexpr int CH = G / 2;                      // channels per thread (4, 8, 16)
            int s = tid & (TM - 1), c0 = (tid >> 7) * CH;
            const float* pp = pq + (size_t)irow[s] * R2 + c0;
            const float* qq = pq + (size_t)jrow[s] * R2 + R + c0;
            float pv[CH], qv[CH];
            pf::ldg_vec<CH>(pp, pv);
            pf::ldg_vec<CH>(qq, qv);
#pragma unroll
            for (int c = 0; c < CH; c++) hs[(c0 + c) * TM + s] = pf::lrelu(pv[c] + qv[c], slope);
        }
        if (!weights_ready) {
            pf::mbar_wait(bar, 0);
            weights_ready = true;
        }
        __syncthreads();

        // ---- layers 1..3: h_l = lrelu(P_l,i + Q_l,j + Wh_l . h_<l) ----
        {
            const int te = tid & 31, to = tid >> 5;
            const int s0 = te * 4;
#pragma unroll
            for (int l = 1; l < 4; l++) {
                float acc[4][TOD];
                pf::zero_acc<4, TOD>(acc);
                const float* w = wds + G * G * ((l - 1) * l / 2) + to * TOD;
                pf::tile_gemm<4, TOD, 0>(hs + s0, TM, l * G, w, G, acc);
                const int col = l * G + to * TOD;
                float pv[TOD];
                pf::ldg_vec<TOD>(pq + (size_t)irow[s0] * R2 + col, pv);
                float hv[TOD][4];
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    float qv[TOD];
                    pf::ldg_vec<TOD>(pq + (size_t)jrow[s0 + e] * R2 + R + col, qv);
#pragma unroll
                    for (int o = 0; o < TOD; o++) hv[o][e] = pf::lrelu(acc[e][o] + (pv[o] + qv[o]), slope);
                }
#pragma unroll
                for (int o = 0; o < TOD; o++)
                    *reinterpret_cast<float4*>(hs + (col + o) * TM + s0) =
                        make_float4(hv[o][0], hv[o][1], hv[o][2], hv[o][3]);
                __syncthreads();
            }
        }

        // ---- conv_out + max over the K neighbours ----
        {
            const int te = tid & 15, to = tid >> 4;
            float acc[8][TOO];
            pf::zero_acc<8, TOO>(acc);
            pf::tile_gemm<8, TOO, 64>(hs + te * 4, TM, 4 * G, wos + to * TOO, CO, acc);
            const int col = 4 * G + to * TOO;
            float mx[TOO];
#pragma unroll
            for (int e = 0; e < 8; e++) {
                int s = (e < 4) ? te * 4 + e : 64 + te * 4 + (e - 4);
                float qv[TOO];
                pf::ldg_vec<TOO>(pq + (size_t)jrow[s] * R2 + R + col, qv);
#pragma unroll
                for (int o = 0; o < TOO; o++) {
                    float v = acc[e][o] + qv[o];
                    mx[o] = (e == 0) ? v : fmaxf(mx[o], v);
                }
            }
#pragma unroll
            for (int d = 1; d < TPP; d <<= 1)
#pragma unroll
                for (int o = 0; o < TOO; o++) mx[o] = fmaxf(mx[o], __shfl_xor_sync(0xffffffffu, mx[o], d));
            const int pi = p0 + te / TPP;
            if ((te % TPP) == 0 && pi < N) {
                const int row = patch * N + pi;
                float pv[TOO];
                pf::ldg_vec<TOO>(pq + (size_t)row * R2 + col, pv);
                float* dst = out + (size_t)row * CO + to * TOO;
                if constexpr (TOO % 4 == 0) {
#pragma unroll
                    for (int o = 0; o < TOO; o += 4)
                        *reinterpret_cast<float4*>(dst + o) =
                            make_float4(mx[o] + pv[o], mx[o + 1] + pv[o + 1], mx[o + 2] + pv[o + 2], mx[o + 3] + pv[o + 3]);
                } else {
                    *reinterpret_cast<float2*>(dst) = make_float2(mx[0] + pv[0], mx[1] + pv[1]);
                }
            }
        }
        __syncthreads();   // hs / jrow are rewritten by the next tile
    }
}

template <int G, int CO, int KN>
int launch(const float* pq, const int32_t* idx, const float* wd, const float* wo, float* out, int Bp, int N,
           float slope, cudaStream_t st) {
    constexpr int PT = TM / KN;
    constexpr size_t smem = (size_t)(6 * G * G + 4 * G * CO + 4 * G * TM) * sizeof(float) + 2 * TM * sizeof(int) + 16;
    auto kern = edgeconv_kernel<G, CO, KN>;
    if (int rc = pf_smem_optin(kern, smem)) return rc;
    const int tiles_per_patch = (N + PT - 1) / PT;
    // split a patch over several CTAs when there are few patches (one 2,048-pt cloud = 32 patches)
    int split = 1;
    while ((long long)Bp * split < 2 * 148 && split * 2 <= tiles_per_patch) split *= 2;
    const int tiles_per_cta = (tiles_per_patch + split - 1) / split;
    dim3 grid(Bp, (tiles_per_patch + tiles_per_cta - 1) / tiles_per_cta);
    kern<<<grid, NT, smem, st>>>(pq, idx, wd, wo, out, N, tiles_per_patch, tiles_per_cta, slope);
    return pf_launch_status("edgeconv_kernel");
}

template <int G, int CO>
int launch_k(const float* pq, const int32_t* idx, const float* wd, const float* wo, float* out, int Bp, int N, int K,
             float slope, cudaStream_t st) {
    if (K == 8) return launch<G, CO, 8>(pq, idx, wd, wo, out, Bp, N, slope, st);
    if (K == 16) return launch<G, CO, 16>(pq, idx, wd, wo, out, Bp, N, slope, st);
    return launch<G, CO, 32>(pq, idx, wd, wo, out, Bp, N, slope, st);
}

}  // namespace

extern "C" int pf_edgeconv(const float* pq, const int32_t* idx, const float* wd, const float* wo, float* out, int Bp,
                           int N, int K, int G, int C_out, float slope, void* stream) {
    PF_CHECK_ARG(slope >= 0.f && slope <= 1.f, "LeakyReLU slope %g outside [0, 1]", (double)slope);
    PF_CHECK_ARG(pq && idx && wd && wo && out, "pf_edgeconv: null pointer");
    PF_CHECK_ARG(Bp >= 0 && N > 0, "pf_edgeconv: bad sizes Bp=%d N=%d", Bp, N);
    PF_CHECK_ARG((long long)Bp * N < 0x7fffffffLL, "pf_edgeconv: Bp*N overflows int32");
    PF_CHECK_ARG(K == 8 || K == 16 || K == 32, "pf_edgeconv: K=%d not in {8,16,32}", K);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(pq) | reinterpret_cast<uintptr_t>(wd) | reinterpret_cast<uintptr_t>(wo) |
                   reinterpret_cast<uintptr_t>(out)) & 15) == 0,
                 "pf_edgeconv: pointers must be 16-byte aligned");
    if (Bp == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (G == 8 && C_out == 32) return launch_k<8, 32>(pq, idx, wd, wo, out, Bp, N, K, slope, st);
    if (G == 16 && C_out == 64) return launch_k<16, 64>(pq, idx, wd, wo, out, Bp, N, K, slope, st);
    if (G == 32 && C_out == 128) return launch_k<32, 128>(pq, idx, wd, wo, out, Bp, N, K, slope, st);
    pf_set_error("pf_edgeconv: (G, C_out) = (%d, %d) not in {(8,32), (16,64), (32,128)}", G, C_out);
    return PF_EINVAL;
}
