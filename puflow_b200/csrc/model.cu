// Whole network in one C call: pf_upsample_patches = PointInterpFlow.forward
// (/root/reference/modules/discrete/interpflow.py:327-337: knn -> feat_extract -> log_prob -> interp -> g), the entry
// point SURVEY.md section 8(b) asks for.  Host orchestration only: it launches the kernels of this library on ONE stream,
// allocates nothing (the caller passes the workspace) and never synchronises, so the call can be captured in a CUDA graph.
//
// The packed weights travel as ONE device blob whose layout this file owns (pf_model_entry enumerates name / offset /
// size; puflow_b200/packing.py pack_model_blob fills it from a state dict; a C or C++ host can read the same blob from a
// file).  Only the tensor-core kernels are used here -- the FP32 FFMA variants stay reachable one by one for A/B tests.
#include <stdio.h>
#include <string.h>

#include <vector>

#include "common.cuh"
#include "prof.h"
#include "puflow_b200.h"

namespace {

constexpr int NB = 6, KNN = 16;
struct BlockCfg { int c_in, g, c_out; };
constexpr BlockCfg CFG[NB] = {{3, 8, 32}, {32, 16, 64}, {64, 32, 128}, {128, 32, 128}, {128, 32, 128}, {128, 32, 128}};
constexpr int64_t ALIGN = 256;

constexpr int64_t align_up(int64_t v) { return (v + ALIGN - 1) / ALIGN * ALIGN; }
constexpr int r2_of(int i) { return 2 * (4 * CFG[i].g + CFG[i].c_out); }                 // width of P|Q
constexpr int64_t edgeconv_tc_bytes(int g) { return 24LL * (g < 16 ? 16 : g) * (g < 16 ? 16 : g) + 2048LL * (g < 16 ? 16 : g); }
constexpr int64_t linear_tc_bytes(int K, int N) { return (int64_t)((N + 127) / 128) * 128 * K * 4; }

struct Entry { const char* name; int64_t bytes; };

// Entry order = blob order.  Per block: wpq (fp32 [3][2R], block 0 only: K = 3 has no tensor-core form), bpq (fp32 [2R]),
// wpq_tc (blocks 1..5, pf_linear_tc_split layout), wtc (pf_edgeconv_tc), chain (pf_cond_chain_tc).
struct Layout {
    Entry e[PF_MODEL_ENTRIES];
    int64_t off[PF_MODEL_ENTRIES];
    int64_t total;
    int feat_wpq0, feat_bpq[NB], feat_wpq_tc[NB], feat_wtc[NB], chain[NB], flow, interp, logdet;
};

const Layout& layout() {
    static const Layout L = [] {
        Layout l;
        memset(&l, 0, sizeof(l));
        static char names[PF_MODEL_ENTRIES][32];
        int n = 0;
        auto add = [&](const char* fmt, int i, int64_t bytes) {
            snprintf(names[n], sizeof(names[n]), fmt, i);
            l.e[n] = Entry{names[n], bytes};
            return n++;
        };
        for (int i = 0; i < NB; i++) {
            if (i == 0) l.feat_wpq0 = add("feat%d.wpq", i, 4LL * CFG[0].c_in * r2_of(0));
            l.feat_bpq[i] = add("feat%d.bpq", i, 4LL * r2_of(i));
            l.feat_wpq_tc[i] = i ? add("feat%d.wpq_tc", i, linear_tc_bytes(CFG[i].c_in, r2_of(i))) : -1;
            l.feat_wtc[i] = add("feat%d.wtc", i, edgeconv_tc_bytes(CFG[i].g));
            l.chain[i] = add("chain%d", i, (int64_t)PF_CHAIN_TC_BYTES(CFG[i].c_out));
        }
        l.flow = add("flow_tc", 0, (int64_t)NB * PF_FLOW_TC_BLOCK);
        l.interp = add("interp_tc", 0, PF_INTERP_TC_BLOB);
        l.logdet = add("logdet_const", 0, 4);
        int64_t o = 0;
        for (int i = 0; i < n; i++) {
            l.off[i] = o;
            o += align_up(l.e[i].bytes);
        }
        l.total = o;
        return l;
    }();
    return L;
}

// Workspace carve-up (all 256-byte aligned)
struct Work {
    int64_t idx, pq, c, split, raw1, sb, z, ssum, zup, pv, total;
};
Work work_layout(int64_t Bp, int64_t N, int64_t R) {
    const int64_t M = Bp * N, tiles = (M + 127) / 128;
    Work w;
    int64_t o = 0;
    auto take = [&](int64_t bytes) { const int64_t at = o; o += align_up(bytes); return at; };
    w.idx = take(M * KNN * 4);
    w.pq = take(M * 512 * 4);                   // widest P|Q (blocks 2..5)
    w.c = take(M * 128 * 4);                    // block output, fp32
    w.split = take(tiles * 512 * 128);          // block output, pre-split operand tiles (widest: 128 channels)
    w.raw1 = take((int64_t)NB * M * 64 * 4);
    w.sb = take((int64_t)NB * M * 8 * 4);
    w.z = take(M * 3 * 4);
    w.ssum = take(M * 4);
    w.zup = take(M * R * 3 * 4);
    w.pv = take(Bp * 8);
    w.total = o;
    return w;
}

// logp_x = -mean_b( sum_n,c -0.5 (z^2 + log 2 pi) + logdet_const N - sum_n ssum )   (interpflow.py:339-345, probs.py:87-93)
__global__ void logp_patch_kernel(const float* __restrict__ z, const float* __restrict__ ssum,
                                  const float* __restrict__ logdet_const, int N, double* __restrict__ pv) {
    const int b = blockIdx.x;
    double acc = 0.0;
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        const float* p = z + ((size_t)b * N + n) * 3;
        const float LOG_2PI = 1.8378770664093453f;
        float s = -0.5f * (p[0] * p[0] + LOG_2PI);
        s += -0.5f * (p[1] * p[1] + LOG_2PI);
        s += -0.5f * (p[2] * p[2] + LOG_2PI);
        acc += (double)s - (double)ssum[(size_t)b * N + n];
    }
    __shared__ double sh[256];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) pv[b] = sh[0] + (double)logdet_const[0] * N;
}
__global__ void logp_mean_kernel(const double* __restrict__ pv, int Bp, float* __restrict__ logp) {
    __shared__ double sh[256];
    double acc = 0.0;
    for (int b = threadIdx.x; b < Bp; b += blockDim.x) acc += pv[b];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) logp[0] = (float)(-sh[0] / Bp);
}

}  // namespace

pf::Recorder& pf::recorder() {
    thread_local Recorder r;
    return r;
}

extern "C" int pf_profile_enable(int on) {
    pf::Recorder& r = pf::recorder();
    r.on = on != 0;
    r.reset();
    return 0;
}
extern "C" int pf_profile_intervals(void) { return (int)pf::recorder().iv.size(); }
extern "C" int pf_profile_interval(int i, char* name40, float* ms) {
    pf::Recorder& r = pf::recorder();
    PF_CHECK_ARG(i >= 0 && (size_t)i < r.iv.size() && name40 && ms, "pf_profile_interval: index %d outside [0, %d)", i,
                 (int)r.iv.size());
    const pf::Recorder::Interval& v = r.iv[i];
    PF_CUDA(cudaEventSynchronize(r.ev[v.e1]));
    PF_CUDA(cudaEventElapsedTime(ms, r.ev[v.e0], r.ev[v.e1]));
    strncpy(name40, v.name.c_str(), 39);
    name40[39] = 0;
    return 0;
}
extern "C" int pf_profile_clear(void) {
    pf::recorder().reset();
    return 0;
}

extern "C" int64_t pf_model_bytes(void) { return layout().total; }

extern "C" int pf_model_entry(int i, char* name32, int64_t* offset, int64_t* bytes) {
    PF_CHECK_ARG(i >= 0 && i < PF_MODEL_ENTRIES && name32 && offset && bytes, "pf_model_entry: entry %d outside [0, %d)", i,
                 PF_MODEL_ENTRIES);
    const Layout& l = layout();
    strncpy(name32, l.e[i].name, 31);
    name32[31] = 0;
    *offset = l.off[i];
    *bytes = l.e[i].bytes;
    return 0;
}

extern "C" int64_t pf_upsample_workspace_bytes(int Bp, int N, int R) {
    if (Bp < 0 || N < 1 || R < 1) return -1;
    return work_layout(Bp, N, R).total;
}

extern "C" int pf_upsample_patches(const void* model, const float* xyz, int Bp, int N, int R, void* workspace,
                                   int64_t workspace_bytes, float* x, float* logp, int precision, void* stream) {
    PF_CHECK_ARG(precision == PF_PREC_FP32_SPLIT || precision == PF_PREC_FP16,
                 "pf_upsample_patches: precision %d (PF_PREC_FP32_SPLIT or PF_PREC_FP16)", precision);
    PF_CHECK_ARG(Bp >= 0 && N >= KNN && R >= 1 && R <= 32, "pf_upsample_patches: bad sizes Bp=%d N=%d R=%d", Bp, N, R);
    PF_CHECK_ARG(N <= 65535, "pf_upsample_patches: N=%d above the 16-bit neighbour index of the EdgeConv kernel", N);
    if (Bp == 0) return 0;                     // empty batch: nothing to launch, pointers may be null
    PF_CHECK_ARG(model && xyz && workspace && x, "pf_upsample_patches: null pointer");
    PF_CHECK_ARG(((uintptr_t)model & 255) == 0 && ((uintptr_t)workspace & 255) == 0,
                 "pf_upsample_patches: model and workspace must be 256-byte aligned");
    const Work w = work_layout(Bp, N, R);
    PF_CHECK_ARG(workspace_bytes >= w.total, "pf_upsample_patches: workspace of %lld B, %lld needed",
                 (long long)workspace_bytes, (long long)w.total);
    const Layout& l = layout();
    const uint8_t* mb = static_cast<const uint8_t*>(model);
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    auto m = [&](int entry) { return mb + l.off[entry]; };
    const int64_t M = (int64_t)Bp * N;
    int32_t* idx = reinterpret_cast<int32_t*>(ws + w.idx);
    float* pq = reinterpret_cast<float*>(ws + w.pq);
    float* c = reinterpret_cast<float*>(ws + w.c);
    void* split = ws + w.split;
    float* raw1 = reinterpret_cast<float*>(ws + w.raw1);
    float* sb = reinterpret_cast<float*>(ws + w.sb);
    float* z = reinterpret_cast<float*>(ws + w.z);
    float* ssum = reinterpret_cast<float*>(ws + w.ssum);
    float* zup = reinterpret_cast<float*>(ws + w.zup);
    double* pv = reinterpret_cast<double*>(ws + w.pv);
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    pf::ProfScope mk(st);
    char tag[40];

    // a1: kNN-16 inside each patch (grid.y limit: 65,535 patches per launch)
    for (int b0 = 0; b0 < Bp; b0 += 65535) {
        const int nb = Bp - b0 < 65535 ? Bp - b0 : 65535;
        const float* q = xyz + (size_t)b0 * N * 3;
        if ((rc = pf_knn_topk(q, q, nb, N, N, KNN, idx + (size_t)b0 * N * KNN, nullptr, stream))) return rc;
    }
    snprintf(tag, sizeof(tag), "knn_topk_k%d", KNN);
    mk.mark(tag);
    // a2-a4 + first layers of the conditioning nets: six EdgeConv blocks, each followed by its fused conditioning chain
    for (int i = 0; i < NB; i++) {
        const int r2 = r2_of(i);
        const float* bpq = reinterpret_cast<const float*>(m(l.feat_bpq[i]));
        if (i == 0)
            rc = pf_linear(xyz, 3, reinterpret_cast<const float*>(m(l.feat_wpq0)), bpq, pq, r2, M, 3, r2, PF_ACT_NONE, 0.f, stream);
        else
            rc = pf_linear_tc_split(split, m(l.feat_wpq_tc[i]), bpq, pq, r2, M, CFG[i].c_in, r2, PF_ACT_NONE, 0.f, precision, stream);
        if (rc) return rc;
        if (i == 0) snprintf(tag, sizeof(tag), "linear_k3_n%d", r2);
        else snprintf(tag, sizeof(tag), "linear_split_k%d_n%d", CFG[i].c_in, r2);
        mk.mark(tag);
        if ((rc = pf_edgeconv_tc_split(pq, idx, m(l.feat_wtc[i]), c, i + 1 < NB ? split : nullptr, Bp, N, KNN, CFG[i].g,
                                       CFG[i].c_out, 0.05f, precision, stream)))
            return rc;
        snprintf(tag, sizeof(tag), "edgeconv_tc_g%d_c%d_k%d", CFG[i].g, CFG[i].c_out, KNN);
        mk.mark(tag);
        if ((rc = pf_cond_chain_tc(c, CFG[i].c_out, m(l.chain[i]), raw1 + (size_t)i * M * 64, sb + (size_t)i * M * 8, M, precision, stream)))
            return rc;
        snprintf(tag, sizeof(tag), "cond_chain_tc_c%d", CFG[i].c_out);
        mk.mark(tag);
    }
    // a5-a11: flow forward -> z, log-likelihood
    if ((rc = pf_flow_fwd_tc(xyz, raw1, 64, sb, m(l.flow), z, ssum, M, NB, precision, stream))) return rc;
    mk.mark("flow_fwd_tc");
    if (logp) {
        logp_patch_kernel<<<Bp, 256, 0, st>>>(z, ssum, reinterpret_cast<const float*>(m(l.logdet)), N, pv);
        logp_mean_kernel<<<1, 256, 0, st>>>(pv, Bp, logp);
        if ((rc = pf_launch_status("logp kernels"))) return rc;
        mk.mark("logp");
    }
    // a12-a15: latent interpolation (the k = 8 list is the head of the sorted K = 16 list), a10: flow inverse
    if ((rc = pf_interp_tc(xyz, z, idx, KNN, m(l.interp), zup, Bp, N, R, precision, stream))) return rc;
    mk.mark("interp_tc");
    rc = pf_flow_inv_tc(zup, raw1, 64, sb, m(l.flow), x, M, R, NB, precision, stream);
    mk.mark("flow_inv_tc");
    return rc;
}
