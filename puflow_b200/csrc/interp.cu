// Latent-space kNN-weighted interpolation, fused: context encoders -> weight MLP -> softmax over
// the 8 neighbours -> weighted sum of neighbour latents, written in the row order g() consumes.
//
// Replaces InterpolationModule.forward, /root/reference/modules/discrete/interpflow.py:162-186:
//   DistanceEncoder        :85-120  (10-channel edge feature [x_i, x_j, x_i-x_j, |x_i-x_j|], 3 convs)
//   KnnContextEncoder      :123-134 (FeatureExtractUnit(3->128, k=8, growth 16, no pooling) + concat)
//   WeightEstimationUnit   :137-159 (256->128->64->32, rows [:R] used, :180)
//   softmax / gather / einsum('bnck,bnrk->bncr')  :180-185
//
// Algebra: (i) every conv that reads raw coordinates is linear in (x_i, x_j) and comes in
// as per-point projections P, Q (pf_linear with K = 3); (ii) DistanceEncoder.mlp.6, feat_conv.conv_out
// and WeightEstimationUnit.mlp.0 are composed on the host (no activation between them):
//   t0 = lrelu(M1 . d2 + M2 . h + Pm_i + Qm_j),  M1 = s0 W0[:, :128] Wd6,  M2 = s0 W0[:, 128:] Wout_h.
//
// Tile = 64 edge slots = 8 points x 8 neighbours (slot s = 8 q + nbr); activations k-major in shared
// memory: rows [0,64) d2, rows [64,192) the dense block's h (d1 borrows rows [64,128) first); t0
// overwrites rows [0,128), t1 rows [128,192).  Small weights are staged in shared memory by TMA bulk
// copies; the 192x128 composed matrix is streamed through the read-only path (L1/L2 resident).
//
// Blob layout (floats, PF_INTERP_BLOB = 46304):  0 wdist[64] | 64 wd3[64][64] | 4160 bd3[64] |
//   4224 wdense (layers 1..7, [16 l][16]) | 11392 w3[128][64] | 19584 b3[64] | 19648 w6[64][32] |
//   21696 b6[32] | 21728 m1[64][128] | 29920 m2[128][128]
#include "common.cuh"
#include "puflow_b200.h"

namespace {

constexpr int TM = 64, NT = 256, KN = 8, PT = 8;
constexpr int PQW = 640, PQH = 320;
constexpr int SMEM_W = 21728;                       // floats of the blob staged in shared memory
constexpr int OFF_WD3 = 64, OFF_BD3 = 4160, OFF_DENSE = 4224, OFF_W3 = 11392, OFF_B3 = 19584, OFF_W6 = 19648,
              OFF_B6 = 21696, OFF_M12 = 21728;
constexpr float S_CTX = 0.05f;                      // LeakyReLU(0.05) in FeatureExtractUnit (interpflow.py:206,217)
constexpr float S_DEF = 0.01f;                      // nn.LeakyReLU() default elsewhere (:94,97,147,150)

__global__ void __launch_bounds__(NT, 1)
    interp_kernel(const float* __restrict__ xyz, const float* __restrict__ z, const int32_t* __restrict__ idx,
                  int ldidx, const float* __restrict__ pq, const float* __restrict__ wblob, float* __restrict__ out,
                  int N, int R, int tiles_per_patch, int tiles_per_cta) {
    extern __shared__ __align__(16) float smem[];
    float* ws = smem;                       // [SMEM_W]
    float* hs = ws + SMEM_W;                // [192][TM]
    float* lg = hs + 192 * TM;              // [32][TM] logits
    float* dist = lg + 32 * TM;             // [TM]
    int* jrow = reinterpret_cast<int*>(dist + TM);
    int* irow = jrow + TM;
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
        pf::mbar_expect_tx(bar, SMEM_W * 4);
        pf::bulk_g2s_chunked(ws, wblob, SMEM_W * 4, bar);
    }
    const float* m12 = wblob + OFF_M12;     // [192][128], global
    bool weights_ready = false;

    for (int t = t_begin; t < t_end; t++) {
        const int p0 = t * PT;
        if (tid < TM) {
            const int s = tid, q = s >> 3, nb = s & 7;
            const int pi = min(p0 + q, N - 1);
            const int row = patch * N + pi;
            int j = __ldg(idx + (size_t)row * ldidx + nb);
            j = min(max(j, 0), N - 1);
            const int jr = patch * N + j;
            irow[s] = row;
            jrow[s] = jr;
            const float* a = xyz + (size_t)row * 3;
            const float* b = xyz + (size_t)jr * 3;
            float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
            dist[s] = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
        }
        __syncthreads();
        if (!weights_ready) {
            pf::mbar_wait(bar, 0);
            weights_ready = true;
        }

        // ---- distance encoder layer 0: d1 = lrelu(Pd_i + Qd_j + wdist * |x_i - x_j|) -> rows [64,128)
        {
            const int s = tid & (TM - 1), c0 = (tid >> 6) * 16;
            const float* pp = pq + (size_t)irow[s] * PQW + c0;
            const float* qq = pq + (size_t)jrow[s] * PQW + PQH + c0;
            const float d = dist[s];
            float pv[16], qv[16];
            pf::ldg_vec<16>(pp, pv);
            pf::ldg_vec<16>(qq, qv);
#pragma unroll
            for (int c = 0; c < 16; c++)
                hs[(64 + c0 + c) * TM + s] = pf::lrelu(fmaf(ws[c0 + c], d, pv[c] + qv[c]), S_DEF);
        }
        __syncthreads();
        // ---- distance encoder layer 1: d2 = lrelu(Wd3 . d1 + bd3) -> rows [0,64)
        {
            const int te = tid & 15, to = tid >> 4;
            float acc[4][4];
            pf::zero_acc<4, 4>(acc);
            pf::tile_gemm<4, 4, 0>(hs + 64 * TM + te * 4, TM, 64, ws + OFF_WD3 + to * 4, 64, acc);
#pragma unroll
            for (int o = 0; o < 4; o++) {
                float b = ws[OFF_BD3 + to * 4 + o];
                *reinterpret_cast<float4*>(hs + (to * 4 + o) * TM + te * 4) =
                    make_float4(pf::lrelu(acc[0][o] + b, S_DEF), pf::lrelu(acc[1][o] + b, S_DEF),
                                pf::lrelu(acc[2][o] + b, S_DEF), pf::lrelu(acc[3][o] + b, S_DEF));
            }
        }
        __syncthreads();
        // ---- dense block layer 0: h0 = lrelu(Pe_i + Qe_j) -> rows [64,80)
        {
            const int s = tid & (TM - 1), c0 = (tid >> 6) * 4;
            float pv[4], qv[4];
            pf::ldg_vec<4>(pq + (size_t)irow[s] * PQW + 64 + c0, pv);
            pf::ldg_vec<4>(pq + (size_t)jrow[s] * PQW + PQH + 64 + c0, qv);
#pragma unroll
            for (int c = 0; c < 4; c++) hs[(64 + c0 + c) * TM + s] = pf::lrelu(pv[c] + qv[c], S_CTX);
        }
        __syncthreads();
        // ---- dense block layers 1..7 (growth 16): warps 0-3, 4 slots x 2 outputs per thread
        for (int l = 1; l < 8; l++) {
            if (tid < 128) {
                const int te = tid & 15, to = tid >> 4;         // to in [0,8): outputs 2*to, 2*to+1
                float acc[4][2];
                pf::zero_acc<4, 2>(acc);
                pf::tile_gemm<4, 2, 0>(hs + 64 * TM + te * 4, TM, 16 * l, ws + OFF_DENSE + 128 * (l - 1) * l + to * 2,
                                       16, acc);
                const int col = 64 + 16 * l + to * 2;
                const int s0 = te * 4;
                float pv[2];
                pf::ldg_vec<2>(pq + (size_t)irow[s0] * PQW + col, pv);
                float hv[2][4];
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    float qv[2];
                    pf::ldg_vec<2>(pq + (size_t)jrow[s0 + e] * PQW + PQH + col, qv);
                    hv[0][e] = pf::lrelu(acc[e][0] + (pv[0] + qv[0]), S_CTX);
                    hv[1][e] = pf::lrelu(acc[e][1] + (pv[1] + qv[1]), S_CTX);
                }
                *reinterpret_cast<float4*>(hs + (col + 0) * TM + s0) = make_float4(hv[0][0], hv[0][1], hv[0][2], hv[0][3]);
                *reinterpret_cast<float4*>(hs + (col + 1) * TM + s0) = make_float4(hv[1][0], hv[1][1], hv[1][2], hv[1][3]);
            }
            __syncthreads();
        }
        // ---- composed context layer: t0 = lrelu(M12 . [d2; h] + Pm_i + Qm_j) -> rows [0,128)
        {
            const int te = tid & 7, to = tid >> 3;              // 8 slots x 4 outputs per thread
            float acc[8][4];
            pf::zero_acc<8, 4>(acc);
            pf::tile_gemm<8, 4, 32, true>(hs + te * 4, TM, 192, m12 + to * 4, 128, acc);
            const int col = 192 + to * 4;
            float hv[4][8];
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const int s = (e < 4) ? te * 4 + e : 32 + te * 4 + (e - 4);
                float pv[4], qv[4];
                pf::ldg_vec<4>(pq + (size_t)irow[s] * PQW + col, pv);
                pf::ldg_vec<4>(pq + (size_t)jrow[s] * PQW + PQH + col, qv);
#pragma unroll
                for (int o = 0; o < 4; o++) hv[o][e] = pf::lrelu(acc[e][o] + (pv[o] + qv[o]), S_DEF);
            }
            __syncthreads();                                    // every thread is done reading rows [0,192)
#pragma unroll
            for (int o = 0; o < 4; o++) {
                *reinterpret_cast<float4*>(hs + (to * 4 + o) * TM + te * 4) =
                    make_float4(hv[o][0], hv[o][1], hv[o][2], hv[o][3]);
                *reinterpret_cast<float4*>(hs + (to * 4 + o) * TM + 32 + te * 4) =
                    make_float4(hv[o][4], hv[o][5], hv[o][6], hv[o][7]);
            }
        }
        __syncthreads();
        // ---- t1 = lrelu(W3 . t0 + b3) -> rows [128,192)
        {
            const int te = tid & 15, to = tid >> 4;
            float acc[4][4];
            pf::zero_acc<4, 4>(acc);
            pf::tile_gemm<4, 4, 0>(hs + te * 4, TM, 128, ws + OFF_W3 + to * 4, 64, acc);
#pragma unroll
            for (int o = 0; o < 4; o++) {
                float b = ws[OFF_B3 + to * 4 + o];
                *reinterpret_cast<float4*>(hs + (128 + to * 4 + o) * TM + te * 4) =
                    make_float4(pf::lrelu(acc[0][o] + b, S_DEF), pf::lrelu(acc[1][o] + b, S_DEF),
                                pf::lrelu(acc[2][o] + b, S_DEF), pf::lrelu(acc[3][o] + b, S_DEF));
            }
        }
        __syncthreads();
        // ---- logits (rows [:R] of WeightEstimationUnit.mlp.6)
        for (int it = tid; it < R * TM; it += NT) {
            const int s = it & (TM - 1), r = it >> 6;
            float a = ws[OFF_B6 + r];
#pragma unroll 8
            for (int k = 0; k < 64; k++) a = fmaf(hs[(128 + k) * TM + s], ws[OFF_W6 + k * 32 + r], a);
            lg[r * TM + s] = a;
        }
        __syncthreads();
        // ---- softmax over the 8 neighbours + weighted sum of their latents
        for (int it = tid; it < R * PT; it += NT) {
            const int q = it % PT, r = it / PT;
            if (p0 + q >= N) continue;
            float l[KN], mx = -INFINITY;
#pragma unroll
            for (int n = 0; n < KN; n++) l[n] = lg[r * TM + q * KN + n], mx = fmaxf(mx, l[n]);
            float sum = 0.f;
#pragma unroll
            for (int n = 0; n < KN; n++) l[n] = expf(l[n] - mx), sum += l[n];
            const float inv = 1.f / sum;
            float o0 = 0.f, o1 = 0.f, o2 = 0.f;
#pragma unroll
            for (int n = 0; n < KN; n++) {
                const float w = l[n] * inv;
                const float* zz = z + (size_t)jrow[q * KN + n] * 3;
                o0 = fmaf(w, __ldg(zz), o0), o1 = fmaf(w, __ldg(zz + 1), o1), o2 = fmaf(w, __ldg(zz + 2), o2);
            }
            float* d = out + ((size_t)irow[q * KN] * R + r) * 3;
            d[0] = o0, d[1] = o1, d[2] = o2;
        }
        __syncthreads();    // hs / lg / jrow are rewritten by the next tile
    }
}

}  // namespace

extern "C" int pf_interp(const float* xyz, const float* z, const int32_t* idx, int ldidx, const float* pq,
                         const float* w, float* out, int Bp, int N, int R, void* stream) {
    PF_CHECK_ARG(xyz && z && idx && pq && w && out, "pf_interp: null pointer");
    PF_CHECK_ARG(Bp >= 0 && N >= KN && ldidx >= KN, "pf_interp: bad sizes Bp=%d N=%d ldidx=%d", Bp, N, ldidx);
    PF_CHECK_ARG(R >= 1 && R <= 32, "pf_interp: R=%d outside [1, 32]", R);
    PF_CHECK_ARG((long long)Bp * N * R < 0x7fffffffLL, "pf_interp: Bp*N*R overflows int32");
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(pq) | reinterpret_cast<uintptr_t>(w)) & 15) == 0,
                 "pf_interp: pq / w must be 16-byte aligned");
    if (Bp == 0) return 0;
    constexpr size_t smem = (size_t)(SMEM_W + 192 * TM + 32 * TM + TM) * sizeof(float) + 2 * TM * sizeof(int) + 16;
    if (int rc = pf_smem_optin(interp_kernel, smem)) return rc;
    const int tiles_per_patch = (N + PT - 1) / PT;
    int split = 1;
    while ((long long)Bp * split < 2 * 148 && split * 2 <= tiles_per_patch) split *= 2;
    const int tiles_per_cta = (tiles_per_patch + split - 1) / split;
    dim3 grid(Bp, (tiles_per_patch + tiles_per_cta - 1) / tiles_per_cta);
    interp_kernel<<<grid, NT, smem, (cudaStream_t)stream>>>(xyz, z, idx, ldidx, pq, w, out, N, R, tiles_per_patch,
                                                            tiles_per_cta);
    return pf_launch_status("interp_kernel");
}
