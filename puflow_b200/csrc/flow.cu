// Glow-style flow stack, forward and inverse, all blocks fused per direction.
//
// Replaces FlowBlock.forward / inverse and PointInterpFlow.f / g,
// /root/reference/modules/discrete/interpflow.py:66-82,302-321, built from
//   ActNorm                 /root/reference/modules/flows/normalize.py:30-43
//   InvertibleConv1x1_1D    /root/reference/modules/flows/permutate.py:117-126
//   AffineSpatialCoupling   /root/reference/modules/flows/coupling.py:55-58,82-85,108-118 (additive)
//   reverse permutation     /root/reference/modules/flows/permutate.py:75-80
//   AffineInjectorLayer     /root/reference/modules/flows/coupling.py:127-151 (affine)
// (~25 ATen launches per block per direction in the reference, torch.inverse(W) on every call).
//
// Given the conditioning features the flow is strictly per point.  Everything that depends only on
// the INPUT point is computed once (pf_linear -> raw, pf_flow_cond -> sb) and shared by the forward
// pass and the R inverse copies; the kernels below keep a point's 3 coordinates in registers
// through all blocks and run the 64x64 hidden layer of coupling1's LinearA1D
// (interpflow.py:22-43) as a register-tiled FFMA GEMM over a 128-point tile.
//
// Packed per-block weights (puflow_b200/packing.py), floats:
//   flow blob (PF_FLOW_BLOB = 4480):  0 exp(logs)[3] | 3 bias[3] | 6 exp(-logs)[3] | 9 W[9] | 18 Winv[9]
//       | 27 tdim | 28 reverse (1: channel reverse after coupling1, 0: identity) | 32 w1h[2][64] | 160 b2[64] | 224 w4[64][2] | 352 b4[2] | 384 w2[64][64]
//   cond blob (PF_COND_BLOB = 8840):  0 ws2[64][64] | 4096 wb2[64][64] | 8192 bs2[64] | 8256 bb2[64]
//       | 8320 ws4[64][4] | 8576 wb4[64][4] | 8832 bs4[4] | 8836 bb4[4]
#include "common.cuh"
#include "puflow_b200.h"

namespace {

constexpr int TP = 128;     // points per tile
constexpr int NT = 256;
constexpr int H = 64;       // hidden width of LinearA1D
constexpr float SLOPE = 0.01f;   // nn.LeakyReLU() default (interpflow.py:34,36)
constexpr int MAXB = 8;

// ---- conditioning: finish coupling2.scale_net / bias_net per input point ---------------------
__global__ void __launch_bounds__(NT, 2)
    flow_cond_kernel(const float* __restrict__ raw, const float* __restrict__ cond_w, float* __restrict__ sb,
                     int64_t M) {
    extern __shared__ __align__(16) float smem[];
    float* wsm = smem;                          // [PF_COND_BLOB]
    float* hs = wsm + PF_COND_BLOB;             // [2][H][TP]
    __shared__ __align__(8) uint64_t bar;

    const int tid = threadIdx.x;
    const int blk = blockIdx.y;
    const int64_t m0 = (int64_t)blockIdx.x * TP;
    if (tid == 0) {
        pf::mbar_init(&bar, 1);
        pf::mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {
        pf::mbar_expect_tx(&bar, PF_COND_BLOB * 4);
        pf::bulk_g2s_chunked(wsm, cond_w + (size_t)blk * PF_COND_BLOB, PF_COND_BLOB * 4, &bar);
    }
    {
        const int p = tid & (TP - 1), net = tid >> 7;
        const int64_t m = min(m0 + p, M - 1);
        const float* src = raw + ((size_t)blk * M + m) * 192 + 64 + net * 64;
        float* dst = hs + net * H * TP + p;
#pragma unroll
        for (int q = 0; q < 16; q++) {
            float4 v = __ldg(reinterpret_cast<const float4*>(src) + q);
            dst[(4 * q + 0) * TP] = pf::lrelu(v.x, SLOPE);
            dst[(4 * q + 1) * TP] = pf::lrelu(v.y, SLOPE);
            dst[(4 * q + 2) * TP] = pf::lrelu(v.z, SLOPE);
            dst[(4 * q + 3) * TP] = pf::lrelu(v.w, SLOPE);
        }
    }
    pf::mbar_wait(&bar, 0);
    __syncthreads();
    const int te = tid & 31, to = tid >> 5;
    float accs[4][8], accb[4][8];
    pf::zero_acc<4, 8>(accs);
    pf::zero_acc<4, 8>(accb);
    pf::tile_gemm<4, 8, 0>(hs + te * 4, TP, H, wsm + to * 8, H, accs);
    pf::tile_gemm<4, 8, 0>(hs + H * TP + te * 4, TP, H, wsm + 4096 + to * 8, H, accb);
    __syncthreads();
#pragma unroll
    for (int o = 0; o < 8; o++) {
        float bs = wsm[8192 + to * 8 + o], bb = wsm[8256 + to * 8 + o];
        float4 vs, vb;
        vs.x = pf::lrelu(accs[0][o] + bs, SLOPE), vs.y = pf::lrelu(accs[1][o] + bs, SLOPE);
        vs.z = pf::lrelu(accs[2][o] + bs, SLOPE), vs.w = pf::lrelu(accs[3][o] + bs, SLOPE);
        vb.x = pf::lrelu(accb[0][o] + bb, SLOPE), vb.y = pf::lrelu(accb[1][o] + bb, SLOPE);
        vb.z = pf::lrelu(accb[2][o] + bb, SLOPE), vb.w = pf::lrelu(accb[3][o] + bb, SLOPE);
        *reinterpret_cast<float4*>(hs + (to * 8 + o) * TP + te * 4) = vs;
        *reinterpret_cast<float4*>(hs + H * TP + (to * 8 + o) * TP + te * 4) = vb;
    }
    __syncthreads();
    {
        const int p = tid & (TP - 1), net = tid >> 7;
        const float* h = hs + net * H * TP + p;
        const float* w4 = wsm + 8320 + net * 256;
        const float* b4 = wsm + 8832 + net * 4;
        float o0 = b4[0], o1 = b4[1], o2 = b4[2];
#pragma unroll 8
        for (int k = 0; k < H; k++) {
            float a = h[k * TP];
            float4 w = *reinterpret_cast<const float4*>(w4 + 4 * k);
            o0 = fmaf(a, w.x, o0), o1 = fmaf(a, w.y, o1), o2 = fmaf(a, w.z, o2);
        }
        if (m0 + p < M) {
            float* dst = sb + ((size_t)blk * M + m0 + p) * 8 + net * 3;
            dst[0] = o0, dst[1] = o1, dst[2] = o2;
            if (net == 1) dst[3] = 0.f, dst[4] = 0.f;
        }
    }
}

// ---- coupling1.bias_net hidden layers on a 128-point tile ----------------------------------------
// xs[3][TP]: current coordinates (h1 = first tdim channels).  On return, threads tid < TP hold the
// (up to 2) outputs of LinearA1D for their point in o0, o1.  3 block-wide barriers inside.
__device__ __forceinline__ void coupling_mlp(const float* __restrict__ fw, const float* __restrict__ xs, float* hs,
                                             float* hs2, const float* __restrict__ rawrow, int tid, float& o0,
                                             float& o1) {
    // layer 0: a1 = lrelu(u1 + W1h . h1)          (W1h rows beyond tdim are zero)
    {
        const int p = tid & (TP - 1), c0 = (tid >> 7) * 32;
        const float x0 = xs[p], x1 = xs[TP + p];
        const float* wa = fw + 32 + c0;
        const float* wb = fw + 32 + 64 + c0;
#pragma unroll
        for (int q = 0; q < 8; q++) {
            float4 u = __ldg(reinterpret_cast<const float4*>(rawrow + c0) + q);
            float4 a = *reinterpret_cast<const float4*>(wa + 4 * q);
            float4 b = *reinterpret_cast<const float4*>(wb + 4 * q);
            hs[(c0 + 4 * q + 0) * TP + p] = pf::lrelu(fmaf(b.x, x1, fmaf(a.x, x0, u.x)), SLOPE);
            hs[(c0 + 4 * q + 1) * TP + p] = pf::lrelu(fmaf(b.y, x1, fmaf(a.y, x0, u.y)), SLOPE);
            hs[(c0 + 4 * q + 2) * TP + p] = pf::lrelu(fmaf(b.z, x1, fmaf(a.z, x0, u.z)), SLOPE);
            hs[(c0 + 4 * q + 3) * TP + p] = pf::lrelu(fmaf(b.w, x1, fmaf(a.w, x0, u.w)), SLOPE);
        }
    }
    __syncthreads();
    // layer 2: a2 = lrelu(W2 . a1 + b2)
    {
        const int te = tid & 31, to = tid >> 5;
        float acc[4][8];
        pf::zero_acc<4, 8>(acc);
        pf::tile_gemm<4, 8, 0>(hs + te * 4, TP, H, fw + 384 + to * 8, H, acc);
#pragma unroll
        for (int o = 0; o < 8; o++) {
            float b = fw[160 + to * 8 + o];
            *reinterpret_cast<float4*>(hs2 + (to * 8 + o) * TP + te * 4) =
                make_float4(pf::lrelu(acc[0][o] + b, SLOPE), pf::lrelu(acc[1][o] + b, SLOPE),
                            pf::lrelu(acc[2][o] + b, SLOPE), pf::lrelu(acc[3][o] + b, SLOPE));
        }
    }
    __syncthreads();
    // layer 4: out = W4 . a2 + b4
    o0 = fw[352], o1 = fw[353];
    if (tid < TP) {
        const float* h = hs2 + tid;
#pragma unroll 8
        for (int k = 0; k < H; k++) {
            float a = h[k * TP];
            float2 w = *reinterpret_cast<const float2*>(fw + 224 + 2 * k);
            o0 = fmaf(a, w.x, o0), o1 = fmaf(a, w.y, o1);
        }
    }
}

template <bool INVERSE>
__global__ void __launch_bounds__(NT, 1)
    flow_kernel(const float* __restrict__ xin, const float* __restrict__ raw, const float* __restrict__ sb,
                const float* __restrict__ flow_w, float* __restrict__ xout, float* __restrict__ ssum, int64_t M,
                int R, int nblocks) {
    extern __shared__ __align__(16) float smem[];
    float* fws = smem;                                   // [nblocks][PF_FLOW_BLOB]
    float* hs = fws + (size_t)nblocks * PF_FLOW_BLOB;    // [H][TP]
    float* hs2 = hs + H * TP;                            // [H][TP]
    float* xs = hs2 + H * TP;                            // [3][TP]
    __shared__ __align__(8) uint64_t bar;

    const int tid = threadIdx.x;
    const int64_t rows = INVERSE ? M * R : M;            // rows of xin / xout
    const int64_t r0 = (int64_t)blockIdx.x * TP;
    if (tid == 0) {
        pf::mbar_init(&bar, 1);
        pf::mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t bytes = (uint32_t)nblocks * PF_FLOW_BLOB * 4;
        pf::mbar_expect_tx(&bar, bytes);
        pf::bulk_g2s_chunked(fws, flow_w, bytes, &bar);
    }
    const int p = tid & (TP - 1);
    const int64_t row = min(r0 + p, rows - 1);
    const int64_t m = INVERSE ? row / R : row;           // input point that conditions this row
    float x0 = 0.f, x1 = 0.f, x2 = 0.f, ss = 0.f;
    if (tid < TP) {
        const float* s = xin + (size_t)row * 3;
        x0 = s[0], x1 = s[1], x2 = s[2];
    }
    pf::mbar_wait(&bar, 0);

    for (int it = 0; it < nblocks; it++) {
        const int blk = INVERSE ? nblocks - 1 - it : it;
        const float* fw = fws + (size_t)blk * PF_FLOW_BLOB;
        const int tdim = (int)fw[27];
        const bool rev = fw[28] != 0.f;
        const float* rawrow = raw + ((size_t)blk * M + m) * 192;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, b0 = 0.f, b1 = 0.f, b2 = 0.f;
        if (tid < TP) {
            const float4 v0 = __ldg(reinterpret_cast<const float4*>(sb + ((size_t)blk * M + m) * 8));
            const float4 v1 = __ldg(reinterpret_cast<const float4*>(sb + ((size_t)blk * M + m) * 8) + 1);
            s0 = v0.x, s1 = v0.y, s2 = v0.z, b0 = v0.w, b1 = v1.x, b2 = v1.y;
            if (!INVERSE) {
                // ActNorm.forward: x * exp(logs) + bias ; inv1x1.forward: z_i = sum_j W[i][j] x_j
                float a0 = fmaf(x0, fw[0], fw[3]), a1 = fmaf(x1, fw[1], fw[4]), a2 = fmaf(x2, fw[2], fw[5]);
                x0 = fw[9] * a0 + fw[10] * a1 + fw[11] * a2;
                x1 = fw[12] * a0 + fw[13] * a1 + fw[14] * a2;
                x2 = fw[15] * a0 + fw[16] * a1 + fw[17] * a2;
            } else {
                // coupling2.inverse: z * exp(s) + b ; permutate2.inverse: reverse
                float a0 = fmaf(x0, expf(s0), b0), a1 = fmaf(x1, expf(s1), b1), a2 = fmaf(x2, expf(s2), b2);
                x0 = rev ? a2 : a0, x1 = a1, x2 = rev ? a0 : a2;
            }
            xs[p] = x0;
            xs[TP + p] = (tdim == 2) ? x1 : 0.f;
        }
        __syncthreads();
        float o0, o1;
        coupling_mlp(fw, xs, hs, hs2, rawrow, tid, o0, o1);
        if (tid < TP) {
            if (!INVERSE) {
                // coupling1.forward (additive): h2 -= bias_net ; reverse ; coupling2.forward
                if (tdim == 1) x1 -= o0, x2 -= o1;
                else x2 -= o0;
                float a0 = rev ? x2 : x0, a1 = x1, a2 = rev ? x0 : x2;
                x0 = (a0 - b0) * expf(-s0);
                x1 = (a1 - b1) * expf(-s1);
                x2 = (a2 - b2) * expf(-s2);
                ss += (s0 + s1) + s2;
            } else {
                // coupling1.inverse: h2 += bias_net ; inv1x1.inverse ; ActNorm.inverse
                if (tdim == 1) x1 += o0, x2 += o1;
                else x2 += o0;
                float a0 = fw[18] * x0 + fw[19] * x1 + fw[20] * x2;
                float a1 = fw[21] * x0 + fw[22] * x1 + fw[23] * x2;
                float a2 = fw[24] * x0 + fw[25] * x1 + fw[26] * x2;
                x0 = (a0 - fw[3]) * fw[6];
                x1 = (a1 - fw[4]) * fw[7];
                x2 = (a2 - fw[5]) * fw[8];
            }
        }
        // xs is rewritten next iteration only after every thread passed coupling_mlp's barriers
    }
    if (tid < TP && r0 + p < rows) {
        float* d = xout + (size_t)(r0 + p) * 3;
        d[0] = x0, d[1] = x1, d[2] = x2;
        if (!INVERSE && ssum) ssum[r0 + p] = ss;
    }
}

}  // namespace

extern "C" int pf_flow_cond(const float* raw, const float* cond_w, float* sb, int64_t M, int nblocks, void* stream) {
    PF_CHECK_ARG(raw && cond_w && sb, "pf_flow_cond: null pointer");
    PF_CHECK_ARG(M >= 0 && nblocks >= 1 && nblocks <= MAXB, "pf_flow_cond: bad sizes M=%lld nblocks=%d", (long long)M,
                 nblocks);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(raw) | reinterpret_cast<uintptr_t>(cond_w) |
                   reinterpret_cast<uintptr_t>(sb)) & 15) == 0, "pf_flow_cond: pointers must be 16-byte aligned");
    if (M == 0) return 0;
    constexpr size_t smem = (size_t)(PF_COND_BLOB + 2 * H * TP) * sizeof(float);
    if (int rc = pf_smem_optin(flow_cond_kernel, smem)) return rc;
    dim3 grid((unsigned)((M + TP - 1) / TP), nblocks);
    flow_cond_kernel<<<grid, NT, smem, (cudaStream_t)stream>>>(raw, cond_w, sb, M);
    return pf_launch_status("flow_cond_kernel");
}

static int flow_launch(bool inverse, const float* xin, const float* raw, const float* sb, const float* flow_w,
                       float* xout, float* ssum, int64_t M, int R, int nblocks, void* stream) {
    PF_CHECK_ARG(xin && raw && sb && flow_w && xout, "pf_flow: null pointer");
    PF_CHECK_ARG(M >= 0 && R >= 1 && nblocks >= 1 && nblocks <= MAXB, "pf_flow: bad sizes M=%lld R=%d nblocks=%d",
                 (long long)M, R, nblocks);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(raw) | reinterpret_cast<uintptr_t>(flow_w) |
                   reinterpret_cast<uintptr_t>(sb)) & 15) == 0, "pf_flow: pointers must be 16-byte aligned");
    if (M == 0) return 0;
    const size_t smem = ((size_t)nblocks * PF_FLOW_BLOB + 2 * H * TP + 3 * TP) * sizeof(float);
    const int64_t rows = inverse ? M * R : M;
    dim3 grid((unsigned)((rows + TP - 1) / TP));
    if (inverse) {
        if (int rc = pf_smem_optin(flow_kernel<true>, smem)) return rc;
        flow_kernel<true><<<grid, NT, smem, (cudaStream_t)stream>>>(xin, raw, sb, flow_w, xout, nullptr, M, R, nblocks);
    } else {
        if (int rc = pf_smem_optin(flow_kernel<false>, smem)) return rc;
        flow_kernel<false><<<grid, NT, smem, (cudaStream_t)stream>>>(xin, raw, sb, flow_w, xout, ssum, M, 1, nblocks);
    }
    return pf_launch_status("flow_kernel");
}

extern "C" int pf_flow_fwd(const float* x, const float* raw, const float* sb, const float* flow_w, float* z,
                           float* ssum, int64_t M, int nblocks, void* stream) {
    return flow_launch(false, x, raw, sb, flow_w, z, ssum, M, 1, nblocks, stream);
}

extern "C" int pf_flow_inv(const float* zin, const float* raw, const float* sb, const float* flow_w, float* x,
                           int64_t M, int R, int nblocks, void* stream) {
    return flow_launch(true, zin, raw, sb, flow_w, x, nullptr, M, R, nblocks, stream);
}
