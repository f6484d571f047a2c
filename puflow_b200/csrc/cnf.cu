// Continuous-flow (CNF) block of the continuous variant: y(t1) of dy/dt = ODEnet([t, c], y), one thread per point.
//
// Replaces CNF.forward + torchdiffeq.odeint(method='dopri5') + ODEfunc / ODEnet / ConcatSquashLinear for the inference path
// (/root/reference/modules/continuous/cnf.py:51-112, odefunc.py:66-111,121-143, diffeq_layers.py:72-86; the block around it
// is modules/continuous/interpflow.py:19-50 with batch_norm=False, i.e. identity norms).  The conditioner in front of it is
// the discrete path's (kNN, EdgeConv, merge unit: the same kernels).  SURVEY.md section 8(f) rank 4, inference half only:
// log p (a Hutchinson trace estimate with a random probe, odefunc.py:8-32) and the backward pass are not built.
//
// Dynamics (hidden width 64, tanh):  layer_k(x) = (W_k x + b_k) * sigmoid(Wg_k [t, c] + bg_k) + Wb_k [t, c].
// The c part of every gate / bias is constant along the trajectory, so it is computed ONCE per input point by the per-point
// tensor-core linear kernel (puflow_b200/packing.py pack_cnf_block: ctx[M_c][264] = c . Wctx + bctx); the kernel adds the
// t column.  For the inverse the R up-sampled points of an input point share its row (rep = R; the reference materialises
// repeat_interleave(c, R)).
//
// Integrator: adaptive Dormand-Prince 5(4) with FSAL, torchdiffeq's controller constants (safety 0.9, growth <= 10, shrink >=
// 0.2, exponent 1/5, Hairer's starting step), atol / rtol arguments (the reference: 1e-5 each).  Two stated differences from
// torchdiffeq (oracle/cnf_oracle.py): every point controls its own step (torchdiffeq: one RMS error norm over the
// whole batch, so a point's result depends on its batch-mates and every step needs a grid-wide reduction), and the last step
// is clipped to end at t1 (torchdiffeq steps past it and interpolates).  Same ODE, same tolerance.
//
// FP32 FFMA kernel: 4,480 MACs + 131 sigmoids + 128 tanh per evaluation, 14-130 evaluations per block and point (the later
// blocks are stiff under real conditioning vectors); the 64 x 64 layer reads its weights as warp-broadcast LDS.128.  The
// tensor-core version -- 128 points = one M tile, the hidden layer as a split-fp16 MMA per stage -- is the next step for this
// row (measured times in profiles/r02_cnf.md).
#include "common.cuh"
#include "puflow_b200.h"

namespace {

constexpr int NT = 128;                       // points per CTA
constexpr int CTXW = PF_CNF_CTX;              // 264 floats per conditioning row
constexpr int CTXS = CTXW + 1;                // row stride in shared memory (odd: conflict-free column reads)
// weight blob (floats): see packing.pack_cnf_block
constexpr int O_W0 = 0, O_TG0 = 256, O_TB0 = 320, O_W1 = 384, O_B1 = 4480, O_TG1 = 4544, O_TB1 = 4608, O_W2 = 4672,
              O_L2 = 4864, WBLOB = PF_CNF_BLOB / 4;
static_assert(WBLOB == 4880, "blob layout");

// sigmoid and tanh through ex2.approx + rcp.approx (__expf, __fdividef): 2-3 ulp each plus |x| 2^-24 in the exponent, i.e. a
// relative error below 2e-6 for the |x| < 20 that occur -- an order of magnitude under the integrator's 1e-5 step tolerance
// (the accurate expf / tanhf cost 5,000 of an evaluation's 9,500 instructions; these 1,300).  The rest of the library is built
// without fast-math: there the results are compared at 1e-4 with a reference that has no solver tolerance to hide behind.
__device__ __forceinline__ float sigmoidf(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float tanh_fast(float x) {
    const float e = __expf(2.f * fminf(x, 20.f));                  // e^40 is finite; tanh(20) == 1 in fp32
    return 1.f - __fdividef(2.f, e + 1.f);
}

// dy = ODEnet([t, c], y).  w: the block's weights in shared memory; cx: this point's conditioning row in shared memory.
__device__ __noinline__ void dynamics(const float* __restrict__ w, const float* __restrict__ cx, float t, float y0, float y1,
                                      float y2, float& d0, float& d1, float& d2) {
    float h0[64], h1[64];
#pragma unroll
    for (int j = 0; j < 64; j++) {
        const float4 wj = *reinterpret_cast<const float4*>(w + O_W0 + 4 * j);                 // W0[j][0..2], b0[j]
        const float pre = fmaf(wj.x, y0, fmaf(wj.y, y1, fmaf(wj.z, y2, wj.w)));
        const float gate = sigmoidf(fmaf(w[O_TG0 + j], t, cx[j]));
        h0[j] = tanh_fast(fmaf(pre, gate, fmaf(w[O_TB0 + j], t, cx[64 + j])));
    }
#pragma unroll
    for (int j = 0; j < 64; j++) {
        float acc = w[O_B1 + j];
        const float4* wr = reinterpret_cast<const float4*>(w + O_W1 + 64 * j);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const float4 v = wr[i];
            acc = fmaf(v.x, h0[4 * i], fmaf(v.y, h0[4 * i + 1], fmaf(v.z, h0[4 * i + 2], fmaf(v.w, h0[4 * i + 3], acc))));
        }
        const float gate = sigmoidf(fmaf(w[O_TG1 + j], t, cx[128 + j]));
        h1[j] = tanh_fast(fmaf(acc, gate, fmaf(w[O_TB1 + j], t, cx[192 + j])));
    }
    float o[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        float acc = w[O_L2 + a];                                                               // b2
        const float4* wr = reinterpret_cast<const float4*>(w + O_W2 + 64 * a);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const float4 v = wr[i];
            acc = fmaf(v.x, h1[4 * i], fmaf(v.y, h1[4 * i + 1], fmaf(v.z, h1[4 * i + 2], fmaf(v.w, h1[4 * i + 3], acc))));
        }
        const float gate = sigmoidf(fmaf(w[O_L2 + 3 + a], t, cx[256 + a]));
        o[a] = fmaf(acc, gate, fmaf(w[O_L2 + 6 + a], t, cx[259 + a]));
    }
    d0 = o[0], d1 = o[1], d2 = o[2];
}

__device__ __forceinline__ float rms3(float a, float b, float c) { return sqrtf((a * a + b * b + c * c) * (1.f / 3.f)); }

__global__ void __launch_bounds__(NT, 2)
    cnf_flow_kernel(const float* __restrict__ y_in, const float* __restrict__ ctx, int ld_ctx, const float* __restrict__ wblob,
                    float* __restrict__ y_out, int32_t* __restrict__ nfe_out, int M, int rep, float t0, float t1, float atol,
                    float rtol) {
    extern __shared__ __align__(16) float sm[];
    float* w = sm;
    float* cs = sm + WBLOB;
    const int tid = threadIdx.x;
    const long long p = (long long)blockIdx.x * NT + tid;
    const long long row0 = ((long long)blockIdx.x * NT) / rep;
    long long rowl = ((long long)blockIdx.x * NT + NT - 1);
    if (rowl > M - 1) rowl = M - 1;
    const int nrows = (int)(rowl / rep - row0) + 1;
    for (int i = tid; i < WBLOB; i += NT) w[i] = __ldg(wblob + i);
    for (int i = tid; i < nrows * (CTXW / 4); i += NT) {          // 66 float4 per row
        const int r = i / (CTXW / 4), c4 = i % (CTXW / 4);
        const float4 v = __ldg(reinterpret_cast<const float4*>(ctx + (row0 + r) * (long long)ld_ctx) + c4);
        float* d = cs + r * CTXS + 4 * c4;
        d[0] = v.x, d[1] = v.y, d[2] = v.z, d[3] = v.w;
    }
    __syncthreads();
    if (p >= M) return;
    const float* cx = cs + (int)(p / rep - row0) * CTXS;

    // forward variable s = |t - t0| in [0, span]; the dynamics enter with sign sgn
    const float sgn = t1 >= t0 ? 1.f : -1.f, span = fabsf(t1 - t0);
    float y[3] = {__ldg(y_in + p * 3), __ldg(y_in + p * 3 + 1), __ldg(y_in + p * 3 + 2)};
    float k[7][3];
    auto f = [&](float s, float a0, float a1, float a2, float (&out)[3]) {
        float d0, d1, d2;
        dynamics(w, cx, fmaf(sgn, s, t0), a0, a1, a2, d0, d1, d2);
        out[0] = sgn * d0, out[1] = sgn * d1, out[2] = sgn * d2;
    };
    int nfe = 2;
    f(0.f, y[0], y[1], y[2], k[0]);
    // starting step (Hairer, Norsett, Wanner II.4; order 4 as torchdiffeq passes it)
    float h;
    {
        const float sc0 = fmaf(fabsf(y[0]), rtol, atol), sc1 = fmaf(fabsf(y[1]), rtol, atol), sc2 = fmaf(fabsf(y[2]), rtol, atol);
        const float dd0 = rms3(y[0] / sc0, y[1] / sc1, y[2] / sc2), dd1 = rms3(k[0][0] / sc0, k[0][1] / sc1, k[0][2] / sc2);
        const float h0 = (dd0 < 1e-5f || dd1 < 1e-5f) ? 1e-6f : 0.01f * dd0 / dd1;
        float f1[3];
        f(h0, fmaf(h0, k[0][0], y[0]), fmaf(h0, k[0][1], y[1]), fmaf(h0, k[0][2], y[2]), f1);
        const float dd2 = rms3((f1[0] - k[0][0]) / sc0, (f1[1] - k[0][1]) / sc1, (f1[2] - k[0][2]) / sc2) / h0;
        const float h1 = (dd1 <= 1e-15f && dd2 <= 1e-15f) ? fmaxf(1e-6f, h0 * 1e-3f) : powf(0.01f / fmaxf(dd1, dd2), 0.2f);
        h = fminf(100.f * h0, h1);
    }
    float s = 0.f;
    for (int it = 0; it < 100000 && s < span; it++) {
        const float hs = fminf(h, span - s);
        float yi[3];
#define PF_STAGE(ST, ...)                                                                    \
    {                                                                                        \
        const float b[ST] = {__VA_ARGS__};                                                   \
        _Pragma("unroll") for (int a = 0; a < 3; a++) {                                      \
            float acc = 0.f;                                                                 \
            _Pragma("unroll") for (int j = 0; j < ST; j++) acc = fmaf(b[j], k[j][a], acc);   \
            yi[a] = fmaf(hs, acc, y[a]);                                                     \
        }                                                                                    \
    }
        PF_STAGE(1, 1.f / 5.f)
        f(s + hs * (1.f / 5.f), yi[0], yi[1], yi[2], k[1]);
        PF_STAGE(2, 3.f / 40.f, 9.f / 40.f)
        f(s + hs * (3.f / 10.f), yi[0], yi[1], yi[2], k[2]);
        PF_STAGE(3, 44.f / 45.f, -56.f / 15.f, 32.f / 9.f)
        f(s + hs * (4.f / 5.f), yi[0], yi[1], yi[2], k[3]);
        PF_STAGE(4, 19372.f / 6561.f, -25360.f / 2187.f, 64448.f / 6561.f, -212.f / 729.f)
        f(s + hs * (8.f / 9.f), yi[0], yi[1], yi[2], k[4]);
        PF_STAGE(5, 9017.f / 3168.f, -355.f / 33.f, 46732.f / 5247.f, 49.f / 176.f, -5103.f / 18656.f)
        f(s + hs, yi[0], yi[1], yi[2], k[5]);
        PF_STAGE(6, 35.f / 384.f, 0.f, 500.f / 1113.f, 125.f / 192.f, -2187.f / 6784.f, 11.f / 84.f)       // = the 5th-order solution
        f(s + hs, yi[0], yi[1], yi[2], k[6]);
#undef PF_STAGE
        nfe += 6;
        const float ce[7] = {35.f / 384.f - 1951.f / 21600.f, 0.f, 500.f / 1113.f - 22642.f / 50085.f, 125.f / 192.f - 451.f / 720.f,
                             -2187.f / 6784.f + 12231.f / 42400.f, 11.f / 84.f - 649.f / 6300.f, -1.f / 60.f};
        float er[3];
#pragma unroll
        for (int a = 0; a < 3; a++) {
            float acc = 0.f;
#pragma unroll
            for (int j = 0; j < 7; j++) acc = fmaf(ce[j], k[j][a], acc);
            er[a] = hs * acc / fmaf(rtol, fmaxf(fabsf(y[a]), fabsf(yi[a])), atol);
        }
        const float ratio = rms3(er[0], er[1], er[2]);
        if (ratio <= 1.f) {                                     // accept; FSAL: k[6] = f(s + hs, y_new)
            s += hs;
#pragma unroll
            for (int a = 0; a < 3; a++) y[a] = yi[a], k[0][a] = k[6][a];
            if (span - s <= 1e-7f * span) s = span;
        }
        const float fac = ratio == 0.f ? 10.f : fminf(10.f, fmaxf(0.9f / powf(ratio, 0.2f), ratio < 1.f ? 1.f : 0.2f));
        h = hs * fac;
    }
    y_out[p * 3] = y[0], y_out[p * 3 + 1] = y[1], y_out[p * 3 + 2] = y[2];
    if (nfe_out) nfe_out[p] = s >= span ? nfe : -nfe;          // negative: the step budget ran out (never seen; tests assert > 0)
}

}  // namespace

extern "C" int pf_cnf_flow(const float* y_in, const float* ctx, int ld_ctx, const float* wblob, float* y_out, int32_t* nfe, int M,
                           int rep, float t0, float t1, float atol, float rtol, void* stream) {
    PF_CHECK_ARG(M >= 0 && rep >= 1, "pf_cnf_flow: bad sizes M=%d rep=%d", M, rep);
    if (M == 0) return 0;
    PF_CHECK_ARG(y_in && ctx && wblob && y_out, "pf_cnf_flow: null pointer");
    PF_CHECK_ARG(ld_ctx >= CTXW && ld_ctx % 4 == 0, "pf_cnf_flow: ld_ctx=%d (>= %d, multiple of 4)", ld_ctx, CTXW);
    PF_CHECK_ARG(((reinterpret_cast<uintptr_t>(ctx) | reinterpret_cast<uintptr_t>(wblob)) & 15) == 0,
                 "pf_cnf_flow: ctx and wblob must be 16-byte aligned");
    PF_CHECK_ARG(atol > 0.f && rtol >= 0.f && t0 == t0 && t1 == t1, "pf_cnf_flow: atol=%g rtol=%g t0=%g t1=%g", (double)atol,
                 (double)rtol, (double)t0, (double)t1);
    // conditioning rows a CTA's 128 points can touch: 128 for rep = 1, 33 for rep = 4 (then two CTAs share an SM)
    const int rows_max = (NT + rep - 1) / rep + 1;
    const size_t smem = (size_t)(WBLOB + (rows_max < NT ? rows_max : NT) * CTXS) * 4;
    if (int rc = pf_smem_optin(cnf_flow_kernel, smem)) return rc;
    const int grid = (M + NT - 1) / NT;
    cnf_flow_kernel<<<grid, NT, smem, (cudaStream_t)stream>>>(y_in, ctx, ld_ctx, wblob, y_out, nfe, M, rep, t0, t1, atol, rtol);
    return pf_launch_status("cnf_flow_kernel");
}
