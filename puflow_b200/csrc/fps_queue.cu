// Farthest point sampling as a threshold queue: one CTA per cloud, ~70-100 samples per block-wide pass instead of one.
//
// Same semantics as pf_fps / the reference's own farthest_point_sampling
// (/root/reference/modules/utils/fps.py:40-73; replaces pointnet2_ops.furthest_point_sample at
// /root/reference/modules/utils/patch.py:102,156): start at index 0, running min squared distance initialised to
// 1e10, arg-max with the lowest index winning ties.  Results are bit-identical to the brute-force kernel.
//
// FPS is a chain of npoint dependent arg-max rounds; fps_bucket.cu made a round cheap (spatial pruning) but still pays
// one block-wide reduction + barrier per sample (0.93 us).  Here the chain runs inside ONE warp on a short list:
//
//   * running minima only ever decrease.  So for any threshold theta, the set L = { p : md[p] > theta } can only
//     shrink, and while max(L) > theta the global arg-max is the arg-max of L -- whatever the stale md of the points
//     outside L are (they are <= theta).
//   * pass (all 20 warps): apply the samples drawn so far to the bucketed cloud (exact box pruning, all pending
//     samples of a bucket in one visit), take the block maximum M, collect L for theta = M * (1 - gap) into shared
//     memory (gap is steered so that |L| stays near 128 of at most 256);
//   * chain (warp 0): L lives in registers, 8 entries per lane; a sample is one warp arg-max + 8 distance updates per
//     lane, repeated until max(L) <= theta -- typically ~100 samples in the late phase, where the points above the
//     threshold are isolated local maxima of the distance field.
//
// Exactness: the same sqdist() arithmetic everywhere, min is order-independent, a pruned bucket provably cannot
// change, ties go to the lowest original index (all points that tie with the maximum are in L).  If more than 256
// points are within rounding of the maximum (duplicated points), one sample is drawn by a block-wide tie search.
#include <cooperative_groups.h>

#include <mutex>

#include "common.cuh"
#include "puflow_b200.h"

// Scratch of pf_fps_queued_static: PF_FPS_STATIC_MAX_POINTS sorted points x 5 floats = 5 MiB, held by every GPU that loads
// the library until the process exits.  In HOST code this name is only a stand-in for the symbol (not a device address):
// the entry point below asks the runtime for the address, once per device.
static __device__ __align__(256) float g_fps_static_scratch[(size_t)PF_FPS_STATIC_MAX_POINTS * 5];

namespace cg = cooperative_groups;

namespace {

constexpr int FT = 640, NW = FT / 32;
constexpr int CELLS = 4096;                  // 16^3 grid
constexpr float INIT_D = 1e10f;              // fps.py:53
constexpr float SAFE = 0.9999f;              // box lower bound is scaled down: never above any computed distance
constexpr int CAP = 256, EPL = CAP / 32;     // queue capacity, entries per lane of the chain warp
constexpr int JMAX = 256;                    // samples per pass at most
constexpr float GAP0 = 0.1f, GAP_MIN = 1e-6f;
#ifndef PF_STEER_LO
#define PF_STEER_LO 32
#define PF_STEER_HI 64
#endif
constexpr int STEER_LO = PF_STEER_LO, STEER_HI = PF_STEER_HI;   // the gap is steered to keep the queue between these lengths
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ unsigned morton4(unsigned x, unsigned y, unsigned z) {
    auto spread = [](unsigned v) { return (v & 1u) | ((v & 2u) << 2) | ((v & 4u) << 4) | ((v & 8u) << 6); };
    return spread(x) | (spread(y) << 1) | (spread(z) << 2);
}
__device__ __forceinline__ float sqdist(float px, float py, float pz, float cx, float cy, float cz) {
    const float dx = __fsub_rn(px, cx), dy = __fsub_rn(py, cy), dz = __fsub_rn(pz, cz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
// order-preserving float <-> uint map, so that REDUX can take float minima / maxima of either sign
__device__ __forceinline__ unsigned fkey(float f) {
    const unsigned b = __float_as_uint(f);
    return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float funkey(unsigned k) {
    return __uint_as_float(k ^ ((k >> 31) ? 0x80000000u : 0xffffffffu));
}
__device__ __forceinline__ float warp_minf(float v) { return funkey(__reduce_min_sync(FULL, fkey(v))); }
__device__ __forceinline__ float warp_maxf(float v) { return funkey(__reduce_max_sync(FULL, fkey(v))); }
// running minima are >= +0: their bit patterns order like the values
__device__ __forceinline__ float warp_max_pos(float v) { return __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(v))); }
// squared distance from c to the box [lo, hi]
__device__ __forceinline__ float boxd2(float cx, float cy, float cz, float lx, float ly, float lz, float hx, float hy, float hz) {
    const float ex = fmaxf(fmaxf(lx - cx, cx - hx), 0.f);
    const float ey = fmaxf(fmaxf(ly - cy, cy - hy), 0.f);
    const float ez = fmaxf(fmaxf(lz - cz, cz - hz), 0.f);
    return ex * ex + ey * ey + ez * ez;
}

// control words in shared memory
enum { C_CNT0 = 0, C_S = 3, C_NCENT = 4, C_WORDS = 8 };

// The chain (one warp): the queue's `cnt` entries live in registers, E per lane (slot 32 e + lane); a sample is one warp arg-max
// (ties: lowest original index), one broadcast of its coordinates, E distance updates per lane -- repeated until the queue's
// maximum falls to theta.  E is chosen from the queue's length (1 / 2 / 4 / 8): an empty register slot costs as much as a full
// one (440 cycles per sample with E = 8), and the gap is steered to keep the queue between 32 and 64 entries (measured best:
// a shorter queue ends its chain sooner and pays more passes, a longer one pays per sample).  emit(j, c, s): publish sample j.
template <int E, class Emit>
__device__ __forceinline__ void run_chain(int cnt, const int* Lp, const float* Ld, float4* Lc, const float* sx, const float* sy,
                                          const float* sz, const unsigned* so, float theta, int npoint, int lane, int& s, int& j,
                                          Emit&& emit) {
    float d[E], x[E], y[E], z[E];
    unsigned o[E];
#pragma unroll
    for (int e = 0; e < E; e++) {
        const int slot = 32 * e + lane;
        const bool ok = slot < cnt;
        const int p = ok ? Lp[slot] : 0;
        d[e] = ok ? Ld[slot] : 0.f;
        x[e] = sx[p], y[e] = sy[p], z[e] = sz[p], o[e] = ok ? so[p] : 0xffffffffu;
    }
#pragma unroll
    for (int e = 0; e < E; e++) Lc[32 * e + lane] = make_float4(x[e], y[e], z[e], __uint_as_float(o[e]));
    __syncwarp();
    while (s < npoint && j < JMAX) {
        float lm = d[0];
#pragma unroll
        for (int e = 1; e < E; e++) lm = fmaxf(lm, d[e]);
        const float Mq = warp_max_pos(lm);
        if (!(Mq > theta)) break;
        unsigned myo = 0xffffffffu;
        int mye = 0;
#pragma unroll
        for (int e = 0; e < E; e++)
            if (d[e] == Mq && o[e] < myo) myo = o[e], mye = e;
        const unsigned bal = __ballot_sync(FULL, lm == Mq);
        int src = __ffs(bal) - 1;
        if (bal & (bal - 1)) {      // several lanes attain the maximum: lowest original index
            const unsigned om = __reduce_min_sync(FULL, lm == Mq ? myo : 0xffffffffu);
            src = __ffs(__ballot_sync(FULL, lm == Mq && myo == om)) - 1;
        }
        const int slot = __shfl_sync(FULL, 32 * mye + lane, src);
        const float4 c = Lc[slot];
        emit(j, c, s);
        j++, s++;
#pragma unroll
        for (int e = 0; e < E; e++) d[e] = fminf(d[e], sqdist(x[e], y[e], z[e], c.x, c.y, c.z));
    }
}
template <class Emit>
__device__ __forceinline__ void run_chain_any(int cnt, const int* Lp, const float* Ld, float4* Lc, const float* sx, const float* sy,
                                              const float* sz, const unsigned* so, float theta, int npoint, int lane, int& s, int& j,
                                              Emit&& emit) {
    if (cnt <= 32) run_chain<1>(cnt, Lp, Ld, Lc, sx, sy, sz, so, theta, npoint, lane, s, j, emit);
    else if (cnt <= 64) run_chain<2>(cnt, Lp, Ld, Lc, sx, sy, sz, so, theta, npoint, lane, s, j, emit);
    else if (cnt <= 128) run_chain<4>(cnt, Lp, Ld, Lc, sx, sy, sz, so, theta, npoint, lane, s, j, emit);
    else run_chain<EPL>(cnt, Lp, Ld, Lc, sx, sy, sz, so, theta, npoint, lane, s, j, emit);
}

// work (per cloud, npad = 32 * nb entries each): sx | sy | sz | oidx | md (md only used when !MD_SMEM)
// Bucket b (32 sorted points) belongs to warp b / KB, lane (b % KB) % 32, register slot (b % KB) / 32.
#ifndef PF_FPS_PREFETCH
#define PF_FPS_PREFETCH 1
#endif
// Dev-only phase timers (make EXTRA=-DPF_PROF after touching this file; tools/dev_fps_prof.py).  A tick placed right after a
// __syncthreads() is scheduled BEFORE the barrier wait by ptxas, so a barrier's wait shows up in the phase that follows it.
#ifdef PF_PROF
__device__ long long g_fps_prof[32];
#define FPROF_TICK(k) { const long long t1_ = clock64(); facc[k] += t1_ - ft0; ft0 = t1_; }
#define FPROF_ADD(k, v) facc[k] += (v);
#else
#define FPROF_TICK(k)
#define FPROF_ADD(k, v)
#endif
template <bool MD_SMEM, int BPL>
__global__ void __launch_bounds__(FT, 1)
    fps_queue_kernel(const float* __restrict__ xyz, int n, int npoint, int32_t* __restrict__ out, float* work, int nb, int KB) {
    extern __shared__ __align__(16) float sm[];
    const int npad = nb * 32;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int cloud = blockIdx.x;
    const float* pb = xyz + (size_t)cloud * n * 3;
    int32_t* outc = out + (size_t)cloud * npoint;
    float* sx = work + (size_t)cloud * npad * 5;
    float* sy = sx + npad;
    float* sz = sy + npad;
    unsigned* so = reinterpret_cast<unsigned*>(sz + npad);
    float* mdg = sz + 2 * (size_t)npad;

    // ---- shared memory carve-up ----
    const int mdwords = MD_SMEM ? (npad > CELLS ? npad : CELLS) : CELLS;   // md (passes) aliased by the histogram (setup)
    float* md_s = sm;
    int* hist = reinterpret_cast<int*>(sm);
    float4* cent = reinterpret_cast<float4*>(sm + mdwords);         // [JMAX] samples of the last chain (x, y, z, oidx bits)
    float4* Lc = cent + JMAX;                                       // [CAP] queue entries: coordinates + original index
    float* Ld = reinterpret_cast<float*>(Lc + CAP);                 // [CAP] their running minima when collected
    int* Lp = reinterpret_cast<int*>(Ld + CAP);                     // [CAP] their sorted positions
    float* red = reinterpret_cast<float*>(Lp + CAP);                // [6 * NW] setup reductions
    float* wmax = red + 6 * NW;                                     // [NW]
    int* scan_tmp = reinterpret_cast<int*>(wmax + NW);              // [NW]
    int* ctrl = scan_tmp + NW;                                      // [C_WORDS]
    unsigned long long* tie = reinterpret_cast<unsigned long long*>(ctrl + C_WORDS);   // [1]

    // ---- setup 1: bounding box of the cloud ----
    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (int i = tid; i < n; i += FT)
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float v = pb[(size_t)i * 3 + a];
            lo[a] = fminf(lo[a], v), hi[a] = fmaxf(hi[a], v);
        }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        lo[a] = warp_minf(lo[a]), hi[a] = warp_maxf(hi[a]);
        if (lane == 0) red[a * NW + w] = lo[a], red[(3 + a) * NW + w] = hi[a];
    }
    for (int i = tid; i < CELLS; i += FT) hist[i] = 0;
    __syncthreads();
    float inv[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        lo[a] = warp_minf(red[a * NW + min(lane, NW - 1)]);
        hi[a] = warp_maxf(red[(3 + a) * NW + min(lane, NW - 1)]);
        const float ext = hi[a] - lo[a];
        inv[a] = ext > 0.f ? 15.999f / ext : 0.f;
    }
    auto cell_of = [&](float x, float y, float z) {
        const unsigned cx = min(15u, (unsigned)((x - lo[0]) * inv[0])), cy = min(15u, (unsigned)((y - lo[1]) * inv[1])),
                       cz = min(15u, (unsigned)((z - lo[2]) * inv[2]));
        return morton4(cx, cy, cz);
    };
    // ---- setup 2: histogram, exclusive scan, scatter (counting sort by cell) ----
    for (int i = tid; i < n; i += FT) atomicAdd(&hist[cell_of(pb[(size_t)i * 3], pb[(size_t)i * 3 + 1], pb[(size_t)i * 3 + 2])], 1);
    __syncthreads();
    {
        constexpr int CPT = (CELLS + FT - 1) / FT;            // cells per thread
        int v[CPT], sum = 0;
#pragma unroll
        for (int u = 0; u < CPT; u++) v[u] = (CPT * tid + u < CELLS) ? hist[CPT * tid + u] : 0, sum += v[u];
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) scan_tmp[w] = incl;
        __syncthreads();
        int wsum = lane < NW ? scan_tmp[lane] : 0;
        int wincl = wsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(FULL, wincl, o);
            if (lane >= o) wincl += t;
        }
        const int wbase = __shfl_sync(FULL, wincl - wsum, w);
        int run = wbase + incl - sum;
#pragma unroll
        for (int u = 0; u < CPT; u++)
            if (CPT * tid + u < CELLS) hist[CPT * tid + u] = run, run += v[u];
    }
    __syncthreads();
    for (int i = tid; i < n; i += FT) {
        const float x = pb[(size_t)i * 3], y = pb[(size_t)i * 3 + 1], z = pb[(size_t)i * 3 + 2];
        const int pos = atomicAdd(&hist[cell_of(x, y, z)], 1);
        sx[pos] = x, sy[pos] = y, sz[pos] = z, so[pos] = (unsigned)i;
    }
    for (int i = n + tid; i < npad; i += FT) sx[i] = pb[0], sy[i] = pb[1], sz[i] = pb[2], so[i] = 0x7fffffffu;
    __syncthreads();
    // ---- setup 3: bucket boxes into the owner lane's registers, running minima ----
    float blx[BPL], bly[BPL], blz[BPL], bhx[BPL], bhy[BPL], bhz[BPL], bmr[BPL];
#pragma unroll
    for (int h = 0; h < BPL; h++) blx[h] = bly[h] = blz[h] = bhx[h] = bhy[h] = bhz[h] = bmr[h] = 0.f;   // 0: never active
    float wlx = 3.4e38f, wly = 3.4e38f, wlz = 3.4e38f, whx = -3.4e38f, why = -3.4e38f, whz = -3.4e38f;   // box of the warp's buckets
#pragma unroll
    for (int h = 0; h < BPL; h++)
        for (int k = 32 * h; k < min(KB, 32 * h + 32); k++) {
            const int b = w * KB + k;
            if (b >= nb) break;
            const int p = 32 * b + lane;
            const float x = sx[p], y = sy[p], z = sz[p];
            const float v0 = warp_minf(x), v1 = warp_minf(y), v2 = warp_minf(z), v3 = warp_maxf(x), v4 = warp_maxf(y), v5 = warp_maxf(z);
            if (lane == (k & 31)) blx[h] = v0, bly[h] = v1, blz[h] = v2, bhx[h] = v3, bhy[h] = v4, bhz[h] = v5, bmr[h] = INIT_D;
            wlx = fminf(wlx, v0), wly = fminf(wly, v1), wlz = fminf(wlz, v2);
            whx = fmaxf(whx, v3), why = fmaxf(why, v4), whz = fmaxf(whz, v5);
            const float m0 = p < n ? INIT_D : 0.f;
            if (MD_SMEM) md_s[p] = m0; else mdg[p] = m0;
        }
    if (tid == 0) {
        cent[0] = make_float4(pb[0], pb[1], pb[2], 0.f);
        outc[0] = 0;
        ctrl[C_CNT0] = ctrl[C_CNT0 + 1] = ctrl[C_CNT0 + 2] = 0;
        *tie = ~0ull;
    }
    __syncthreads();
    if (npoint <= 1) return;

    int s = 1, ncent = 1, attempt = 0;
    float gap = GAP0, wbm = INIT_D;
    const int pbase = 32 * w * KB + lane;        // sorted position of this lane's point in the warp's bucket 0
#ifdef PF_PROF
    long long facc[12] = {0}, ft0 = clock64();
#endif

    for (;;) {
        FPROF_ADD(11, 1)
        // ---- (1) apply the pending samples to the buckets: per chunk of 32 samples, lanes first test one sample each
        //      against the box of the warp's buckets, then their own buckets against the samples that passed ----
        for (int base = 0; base < ncent; base += 32) {
            bool sa = false;
            if (base + lane < ncent) {
                const float4 c = cent[base + lane];
                sa = boxd2(c.x, c.y, c.z, wlx, wly, wlz, whx, why, whz) * SAFE < wbm;
            }
            unsigned am = __ballot_sync(FULL, sa);
            if (!am) continue;
            // A touched bucket's coordinates come from L2 (~800 cycles; shared memory holds the running minima).  Buckets are
            // updated four at a time; while a batch is processed the lines of the next one (and, up front, of the first batch
            // of the later register slots) are requested with prefetch.global.L1 -- one 128-byte line per bucket and array,
            // lanes 0..11 -- so that their loads hit L1.  (Splitting the box tests per slot to prefetch slot 0 as well, and a
            // rotating window of register loads, were measured slower: 3.82 and 4.45 ms against 3.63.)
            auto prefetch4 = [&](int h, unsigned mask) {
#if PF_FPS_PREFETCH
                int kn = -1;
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int k = mask ? __ffs(mask) - 1 : -1;
                    mask &= mask - 1;
                    if (lane / 3 == u) kn = k;
                }
                if (lane < 12 && kn >= 0) {
                    const float* a = (lane % 3 == 0 ? sx : (lane % 3 == 1 ? sy : sz)) + (pbase - lane) + 32 * (32 * h + kn);
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(a));
                }
#endif
            };
            unsigned cm[BPL];
#pragma unroll
            for (int h = 0; h < BPL; h++) cm[h] = 0u;
            while (am) {
                const int i = __ffs(am) - 1;
                am &= am - 1;
                const float4 c = cent[base + i];
#pragma unroll
                for (int h = 0; h < BPL; h++)
                    if (boxd2(c.x, c.y, c.z, blx[h], bly[h], blz[h], bhx[h], bhy[h], bhz[h]) * SAFE < bmr[h]) cm[h] |= 1u << i;
            }
#pragma unroll
            for (int h = 1; h < BPL; h++) prefetch4(h, __ballot_sync(FULL, cm[h] != 0u));
#pragma unroll
            for (int h = 0; h < BPL; h++) {
                unsigned mask = __ballot_sync(FULL, cm[h] != 0u);
                FPROF_ADD(8, __popc(mask))
                while (mask) {
                    FPROF_ADD(9, 1)
                    int kk[4];
                    unsigned cmk[4];
                    float px[4], py[4], pz[4], m[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        kk[u] = -1;
                        if (mask) {
                            kk[u] = __ffs(mask) - 1;
                            mask &= mask - 1;
                            const int p = pbase + 32 * (32 * h + kk[u]);
                            px[u] = sx[p], py[u] = sy[p], pz[u] = sz[p];
                            m[u] = MD_SMEM ? md_s[p] : mdg[p];
                            cmk[u] = __shfl_sync(FULL, cm[h], kk[u]);
                        }
                    }
                    prefetch4(h, mask);                      // the next batch
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        if (kk[u] >= 0) {
                            float mm = m[u];
                            unsigned cb = cmk[u];
                            while (cb) {
                                const int i = __ffs(cb) - 1;
                                cb &= cb - 1;
                                const float4 c = cent[base + i];
                                mm = fminf(mm, sqdist(px[u], py[u], pz[u], c.x, c.y, c.z));
                            }
                            const int p = pbase + 32 * (32 * h + kk[u]);
                            if (mm < m[u]) {
                                if (MD_SMEM) md_s[p] = mm; else mdg[p] = mm;
                            }
                            const float nm = warp_max_pos(mm);
                            if (lane == kk[u]) bmr[h] = nm;
                        }
                    }
                }
            }
        }
        {
            float lmax = bmr[0];
#pragma unroll
            for (int h = 1; h < BPL; h++) lmax = fmaxf(lmax, bmr[h]);
            wbm = warp_max_pos(lmax);
            if (lane == 0) wmax[w] = wbm;
        }
        FPROF_TICK(0)
        __syncthreads();                                                           // B1
        FPROF_TICK(1)
        const float M = warp_max_pos(wmax[min(lane, NW - 1)]);
        if (!(M > 0.f)) {            // every remaining point coincides with a sample: arg-max of all zeros is index 0
            for (int i = s + tid; i < npoint; i += FT) outc[i] = 0;
            break;
        }
        // ---- (2) collect the queue: every point whose running minimum exceeds theta ----
        float theta;
        int cnt;
        bool tie_mode = false;
        FPROF_TICK(6)
        for (;;) {
            theta = tie_mode ? M : M * (1.f - gap);
            int* cw = ctrl + C_CNT0 + attempt % 3;
            if (tid == 0) ctrl[C_CNT0 + (attempt + 1) % 3] = 0;
            if (tie_mode) {
#pragma unroll
                for (int h = 0; h < BPL; h++) {
                    unsigned mask = __ballot_sync(FULL, bmr[h] == M);
                    while (mask) {
                        const int k = __ffs(mask) - 1;
                        mask &= mask - 1;
                        const int p = pbase + 32 * (32 * h + k);
                        const float m = MD_SMEM ? md_s[p] : mdg[p];
                        if (m == M) atomicMin(tie, ((unsigned long long)so[p] << 32) | (unsigned)p);
                    }
                }
            } else {
                // two sweeps over the warp's buckets above theta: count, ONE atomic per warp for its slots, then write
                unsigned bm[BPL];
                int total = 0;
#pragma unroll
                for (int h = 0; h < BPL; h++) {
                    bm[h] = __ballot_sync(FULL, bmr[h] > theta);
                    unsigned mask = bm[h];
                    while (mask) {
                        float m[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            m[u] = 0.f;
                            if (mask) {
                                const int k = __ffs(mask) - 1;
                                mask &= mask - 1;
                                const int p = pbase + 32 * (32 * h + k);
                                m[u] = MD_SMEM ? md_s[p] : mdg[p];
                            }
                        }
#pragma unroll
                        for (int u = 0; u < 4; u++) total += __popc(__ballot_sync(FULL, m[u] > theta));   // theta >= 0
                    }
                }
                FPROF_TICK(7)
                if (total) {
                    int slot0 = 0;
                    if (lane == 0) slot0 = atomicAdd(cw, total);
                    slot0 = __shfl_sync(FULL, slot0, 0);
#pragma unroll
                    for (int h = 0; h < BPL; h++) {
                        unsigned mask = bm[h];
                        while (mask && slot0 < CAP) {
                            const int k = __ffs(mask) - 1;
                            mask &= mask - 1;
                            const int p = pbase + 32 * (32 * h + k);
                            const float m = MD_SMEM ? md_s[p] : mdg[p];
                            const bool in = m > theta;
                            const unsigned bal = __ballot_sync(FULL, in);
                            const int slot = slot0 + __popc(bal & ((1u << lane) - 1u));
                            if (in && slot < CAP) Lp[slot] = p, Ld[slot] = m;
                            slot0 += __popc(bal);
                        }
                    }
                }
            }
            FPROF_TICK(2)
            __syncthreads();                                                       // B2
            FPROF_TICK(3)
            cnt = *reinterpret_cast<volatile int*>(cw);
            attempt++;
            if (tie_mode) break;
            if (cnt > CAP) {
                gap *= 0.5f;
                if (gap < GAP_MIN) gap = GAP_MIN, tie_mode = true;
                continue;
            }
            if (cnt == 0) {          // theta rounded up to M
                if (gap >= 1.f) tie_mode = true;
                gap = fminf(1.f, gap * 4.f);
                continue;
            }
            break;
        }
        if (tie_mode) {
            // more than CAP points within rounding of the maximum: draw ONE sample, the lowest original index among the
            // points that attain M
            const unsigned long long t = *reinterpret_cast<volatile unsigned long long*>(tie);
            const int p = (int)(unsigned)t;
            if (tid == 0) {
                cent[0] = make_float4(sx[p], sy[p], sz[p], __uint_as_float((unsigned)(t >> 32)));
                outc[s] = (int32_t)(unsigned)(t >> 32);
                ctrl[C_S] = s + 1, ctrl[C_NCENT] = 1;
            }
            __syncthreads();
            if (tid == 0) *tie = ~0ull;
            gap = GAP_MIN * 4.f;
        } else {
            if (cnt < STEER_LO) gap = fminf(1.f, gap * 1.5f);
            else if (cnt > STEER_HI) gap *= 0.85f;
            // ---- (3) the chain: warp 0 draws samples from the queue until its maximum falls to theta ----
            if (w == 0) {
                int j = 0;
                run_chain_any(cnt, Lp, Ld, Lc, sx, sy, sz, so, theta, npoint, lane, s, j, [&](int jj, const float4& c, int ss) {
                    if (lane == 0) cent[jj] = c, outc[ss] = (int32_t)__float_as_uint(c.w);
                });
                if (lane == 0) ctrl[C_S] = s, ctrl[C_NCENT] = j;
            }
            FPROF_TICK(4)
            __syncthreads();                                                       // B3
            FPROF_TICK(5)
        }
        s = ctrl[C_S], ncent = ctrl[C_NCENT];
        if (s >= npoint) break;
    }
#ifdef PF_PROF
    if (blockIdx.x == 0 && (tid == 0 || tid == 32)) {
        for (int i = 0; i < 12; i++) g_fps_prof[(tid ? 16 : 0) + i] = facc[i];
        g_fps_prof[(tid ? 16 : 0) + 12] = attempt;
    }
#endif
}
#ifdef PF_PROF
}  // namespace
extern "C" __attribute__((visibility("default"))) int pf_debug_fps_prof(long long* host32) {
    return cudaMemcpyFromSymbol(host32, g_fps_prof, sizeof(long long) * 32) == cudaSuccess ? 0 : 1;
}
namespace {
#endif

// ---- the same sampler with a cloud's buckets spread over a thread-block CLUSTER (few large clouds) -----------------------
// One CTA per cloud leaves most of the GPU idle when there are 1..16 clouds, and for ~100,000 candidates two thirds of a pass
// are the update / collect sweeps over the buckets (config 3: 59 us per pass, 241 passes).  Here CTA r of an NC-CTA cluster
// owns the buckets [r nbc, (r + 1) nbc): it applies the pending samples to them and collects its points above the threshold
// into CTA 0's queue through distributed shared memory (one DSMEM atomic per warp for the slots); CTA 0's warp 0 runs the chain
// exactly as above and writes its samples into every CTA's `cent` array.  Three cluster barriers per pass replace the three
// block barriers.  Same arithmetic, same tie rule, same threshold logic evaluated identically in every CTA: bit-identical
// indices.  Set-up (counting sort into `work`) is done by CTA 0; running minima live in each CTA's shared memory.
template <int BPL>
__global__ void __launch_bounds__(FT, 1)
    fps_cqueue_kernel(const float* __restrict__ xyz, int n, int npoint, int32_t* __restrict__ out, float* work, int nb, int nbc,
                      int KB) {
    extern __shared__ __align__(16) float sm[];
    cg::cluster_group cluster = cg::this_cluster();
    const int NC = (int)cluster.num_blocks(), r = (int)cluster.block_rank();
    const int npad = nb * 32;
    const int poff = 32 * r * nbc;               // sorted position of this CTA's first point
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int cloud = blockIdx.x / NC;
    const float* pb = xyz + (size_t)cloud * n * 3;
    int32_t* outc = out + (size_t)cloud * npoint;
    float* sx = work + (size_t)cloud * npad * 5;
    float* sy = sx + npad;
    float* sz = sy + npad;
    unsigned* so = reinterpret_cast<unsigned*>(sz + npad);
    float* mdg = sz + 2 * (size_t)npad;

    // ---- shared memory carve-up ----
    const int mdwords = 32 * nbc > CELLS ? 32 * nbc : CELLS;               // md (passes) aliased by the histogram (setup, CTA 0)
    float* md_s = sm;
    int* hist = reinterpret_cast<int*>(sm);
    float4* cent = reinterpret_cast<float4*>(sm + mdwords);         // [JMAX] samples of the last chain (x, y, z, oidx bits)
    float4* Lc = cent + JMAX;                                       // [CAP] queue entries: coordinates + original index
    float* Ld = reinterpret_cast<float*>(Lc + CAP);                 // [CAP] their running minima when collected
    int* Lp = reinterpret_cast<int*>(Ld + CAP);                     // [CAP] their sorted positions
    float* red = reinterpret_cast<float*>(Lp + CAP);                // [6 * NW] setup reductions
    float* wmax = red + 6 * NW;                                     // [8 * NW] every warp of the cluster
    int* scan_tmp = reinterpret_cast<int*>(wmax + 8 * NW);          // [NW]
    int* ctrl = scan_tmp + NW;                                      // [C_WORDS]
    unsigned long long* tie = reinterpret_cast<unsigned long long*>(ctrl + C_WORDS);   // [1] this CTA's tie candidate
    unsigned long long* tiev = tie + 1;                                                 // [8] every CTA's, in CTA 0

    if (r == 0) {                                // ---- set-up 1-2 by CTA 0 alone (the sorted arrays live in `work`) ----
    // ---- setup 1: bounding box of the cloud ----
    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (int i = tid; i < n; i += FT)
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float v = pb[(size_t)i * 3 + a];
            lo[a] = fminf(lo[a], v), hi[a] = fmaxf(hi[a], v);
        }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        lo[a] = warp_minf(lo[a]), hi[a] = warp_maxf(hi[a]);
        if (lane == 0) red[a * NW + w] = lo[a], red[(3 + a) * NW + w] = hi[a];
    }
    for (int i = tid; i < CELLS; i += FT) hist[i] = 0;
    __syncthreads();
    float inv[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        lo[a] = warp_minf(red[a * NW + min(lane, NW - 1)]);
        hi[a] = warp_maxf(red[(3 + a) * NW + min(lane, NW - 1)]);
        const float ext = hi[a] - lo[a];
        inv[a] = ext > 0.f ? 15.999f / ext : 0.f;
    }
    auto cell_of = [&](float x, float y, float z) {
        const unsigned cx = min(15u, (unsigned)((x - lo[0]) * inv[0])), cy = min(15u, (unsigned)((y - lo[1]) * inv[1])),
                       cz = min(15u, (unsigned)((z - lo[2]) * inv[2]));
        return morton4(cx, cy, cz);
    };
    // ---- setup 2: histogram, exclusive scan, scatter (counting sort by cell) ----
    for (int i = tid; i < n; i += FT) atomicAdd(&hist[cell_of(pb[(size_t)i * 3], pb[(size_t)i * 3 + 1], pb[(size_t)i * 3 + 2])], 1);
    __syncthreads();
    {
        constexpr int CPT = (CELLS + FT - 1) / FT;            // cells per thread
        int v[CPT], sum = 0;
#pragma unroll
        for (int u = 0; u < CPT; u++) v[u] = (CPT * tid + u < CELLS) ? hist[CPT * tid + u] : 0, sum += v[u];
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) scan_tmp[w] = incl;
        __syncthreads();
        int wsum = lane < NW ? scan_tmp[lane] : 0;
        int wincl = wsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(FULL, wincl, o);
            if (lane >= o) wincl += t;
        }
        const int wbase = __shfl_sync(FULL, wincl - wsum, w);
        int run = wbase + incl - sum;
#pragma unroll
        for (int u = 0; u < CPT; u++)
            if (CPT * tid + u < CELLS) hist[CPT * tid + u] = run, run += v[u];
    }
    __syncthreads();
    for (int i = tid; i < n; i += FT) {
        const float x = pb[(size_t)i * 3], y = pb[(size_t)i * 3 + 1], z = pb[(size_t)i * 3 + 2];
        const int pos = atomicAdd(&hist[cell_of(x, y, z)], 1);
        sx[pos] = x, sy[pos] = y, sz[pos] = z, so[pos] = (unsigned)i;
    }
    for (int i = n + tid; i < npad; i += FT) sx[i] = pb[0], sy[i] = pb[1], sz[i] = pb[2], so[i] = 0x7fffffffu;
    __threadfence();
    }
    cluster.sync();                              // the sorted cloud is visible to every CTA of the cluster
    // ---- setup 3: bucket boxes into the owner lane's registers, running minima ----
    float blx[BPL], bly[BPL], blz[BPL], bhx[BPL], bhy[BPL], bhz[BPL], bmr[BPL];
#pragma unroll
    for (int h = 0; h < BPL; h++) blx[h] = bly[h] = blz[h] = bhx[h] = bhy[h] = bhz[h] = bmr[h] = 0.f;   // 0: never active
    float wlx = 3.4e38f, wly = 3.4e38f, wlz = 3.4e38f, whx = -3.4e38f, why = -3.4e38f, whz = -3.4e38f;   // box of the warp's buckets
#pragma unroll
    for (int h = 0; h < BPL; h++)
        for (int k = 32 * h; k < min(KB, 32 * h + 32); k++) {
            if (w * KB + k >= nbc) break;
            const int b = r * nbc + w * KB + k;
            if (b >= nb) break;
            const int p = 32 * b + lane;
            const float x = sx[p], y = sy[p], z = sz[p];
            const float v0 = warp_minf(x), v1 = warp_minf(y), v2 = warp_minf(z), v3 = warp_maxf(x), v4 = warp_maxf(y), v5 = warp_maxf(z);
            if (lane == (k & 31)) blx[h] = v0, bly[h] = v1, blz[h] = v2, bhx[h] = v3, bhy[h] = v4, bhz[h] = v5, bmr[h] = INIT_D;
            wlx = fminf(wlx, v0), wly = fminf(wly, v1), wlz = fminf(wlz, v2);
            whx = fmaxf(whx, v3), why = fmaxf(why, v4), whz = fmaxf(whz, v5);
            const float m0 = p < n ? INIT_D : 0.f;
            md_s[p - poff] = m0;
        }
    if (tid == 0) {
        cent[0] = make_float4(pb[0], pb[1], pb[2], 0.f);
        if (r == 0) outc[0] = 0;
        ctrl[C_CNT0] = ctrl[C_CNT0 + 1] = ctrl[C_CNT0 + 2] = 0;
        *tie = ~0ull;
    }
    cluster.sync();                              // every CTA is initialised before any distributed-shared-memory traffic
    if (npoint <= 1) return;
    // CTA 0's copies of the queue and the control words (distributed shared memory)
    int* ctrl0 = cluster.map_shared_rank(ctrl, 0);
    int* Lp0 = cluster.map_shared_rank(Lp, 0);
    float* Ld0 = cluster.map_shared_rank(Ld, 0);

    int s = 1, ncent = 1, attempt = 0;
    float gap = GAP0, wbm = INIT_D;
    const int pbase = 32 * (r * nbc + w * KB) + lane;   // sorted position of this lane's point in the warp's bucket 0

    for (;;) {
        // ---- (1) apply the pending samples to the buckets: per chunk of 32 samples, lanes first test one sample each
        //      against the box of the warp's buckets, then their own buckets against the samples that passed ----
        for (int base = 0; base < ncent; base += 32) {
            bool sa = false;
            if (base + lane < ncent) {
                const float4 c = cent[base + lane];
                sa = boxd2(c.x, c.y, c.z, wlx, wly, wlz, whx, why, whz) * SAFE < wbm;
            }
            unsigned am = __ballot_sync(FULL, sa);
            if (!am) continue;
            unsigned cm[BPL];
#pragma unroll
            for (int h = 0; h < BPL; h++) cm[h] = 0u;
            while (am) {
                const int i = __ffs(am) - 1;
                am &= am - 1;
                const float4 c = cent[base + i];
#pragma unroll
                for (int h = 0; h < BPL; h++)
                    if (boxd2(c.x, c.y, c.z, blx[h], bly[h], blz[h], bhx[h], bhy[h], bhz[h]) * SAFE < bmr[h]) cm[h] |= 1u << i;
            }
#pragma unroll
            for (int h = 0; h < BPL; h++) {
                unsigned mask = __ballot_sync(FULL, cm[h] != 0u);
                while (mask) {
                    int kk[4];
                    unsigned cmk[4];
                    float px[4], py[4], pz[4], m[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        kk[u] = -1;
                        if (mask) {
                            kk[u] = __ffs(mask) - 1;
                            mask &= mask - 1;
                            const int p = pbase + 32 * (32 * h + kk[u]);
                            px[u] = sx[p], py[u] = sy[p], pz[u] = sz[p];
                            m[u] = md_s[p - poff];
                            cmk[u] = __shfl_sync(FULL, cm[h], kk[u]);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        if (kk[u] >= 0) {
                            float mm = m[u];
                            unsigned cb = cmk[u];
                            while (cb) {
                                const int i = __ffs(cb) - 1;
                                cb &= cb - 1;
                                const float4 c = cent[base + i];
                                mm = fminf(mm, sqdist(px[u], py[u], pz[u], c.x, c.y, c.z));
                            }
                            const int p = pbase + 32 * (32 * h + kk[u]);
                            if (mm < m[u]) {
                                md_s[p - poff] = mm;
                            }
                            const float nm = warp_max_pos(mm);
                            if (lane == kk[u]) bmr[h] = nm;
                        }
                    }
                }
            }
        }
        {
            float lmax = bmr[0];
#pragma unroll
            for (int h = 1; h < BPL; h++) lmax = fmaxf(lmax, bmr[h]);
            wbm = warp_max_pos(lmax);
            if (lane < NC) cluster.map_shared_rank(wmax, lane)[r * NW + w] = wbm;  // every CTA gets every warp's maximum
        }
        cluster.sync();                                                            // B1
        float M;
        {
            float v = 0.f;
            for (int i = lane; i < NC * NW; i += 32) v = fmaxf(v, wmax[i]);
            M = warp_max_pos(v);
        }
        if (!(M > 0.f)) {            // every remaining point coincides with a sample: arg-max of all zeros is index 0
            if (r == 0)
                for (int i = s + tid; i < npoint; i += FT) outc[i] = 0;
            break;
        }
        // ---- (2) collect the queue: every point whose running minimum exceeds theta ----
        float theta;
        int cnt;
        bool tie_mode = false;
        for (;;) {
            theta = tie_mode ? M : M * (1.f - gap);
            int* cw = ctrl0 + C_CNT0 + attempt % 3;                                // CTA 0's counter
            if (r == 0 && tid == 0) ctrl[C_CNT0 + (attempt + 1) % 3] = 0;
            if (tie_mode) {
#pragma unroll
                for (int h = 0; h < BPL; h++) {
                    unsigned mask = __ballot_sync(FULL, bmr[h] == M);
                    while (mask) {
                        const int k = __ffs(mask) - 1;
                        mask &= mask - 1;
                        const int p = pbase + 32 * (32 * h + k);
                        const float m = md_s[p - poff];
                        if (m == M) atomicMin(tie, ((unsigned long long)so[p] << 32) | (unsigned)p);
                    }
                }
                // this CTA's lowest (original index, position) goes to CTA 0 as a plain store (64-bit atomics on distributed
                // shared memory lost candidates in testing)
                __syncthreads();
                if (tid == 0) {
                    cluster.map_shared_rank(tiev, 0)[r] = *tie;
                    *tie = ~0ull;
                }
            } else {
                // two sweeps over the warp's buckets above theta: count, ONE atomic per warp for its slots, then write
                unsigned bm[BPL];
                int total = 0;
#pragma unroll
                for (int h = 0; h < BPL; h++) {
                    bm[h] = __ballot_sync(FULL, bmr[h] > theta);
                    unsigned mask = bm[h];
                    while (mask) {
                        float m[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            m[u] = 0.f;
                            if (mask) {
                                const int k = __ffs(mask) - 1;
                                mask &= mask - 1;
                                const int p = pbase + 32 * (32 * h + k);
                                m[u] = md_s[p - poff];
                            }
                        }
#pragma unroll
                        for (int u = 0; u < 4; u++) total += __popc(__ballot_sync(FULL, m[u] > theta));   // theta >= 0
                    }
                }
                if (total) {
                    int slot0 = 0;
                    if (lane == 0) slot0 = atomicAdd(cw, total);
                    slot0 = __shfl_sync(FULL, slot0, 0);
#pragma unroll
                    for (int h = 0; h < BPL; h++) {
                        unsigned mask = bm[h];
                        while (mask && slot0 < CAP) {
                            const int k = __ffs(mask) - 1;
                            mask &= mask - 1;
                            const int p = pbase + 32 * (32 * h + k);
                            const float m = md_s[p - poff];
                            const bool in = m > theta;
                            const unsigned bal = __ballot_sync(FULL, in);
                            const int slot = slot0 + __popc(bal & ((1u << lane) - 1u));
                            if (in && slot < CAP) Lp0[slot] = p, Ld0[slot] = m;
                            slot0 += __popc(bal);
                        }
                    }
                }
            }
            cluster.sync();                                                        // B2
            cnt = *reinterpret_cast<volatile int*>(cw);
            attempt++;
            if (tie_mode) break;
            if (cnt > CAP) {
                gap *= 0.5f;
                if (gap < GAP_MIN) gap = GAP_MIN, tie_mode = true;
                continue;
            }
            if (cnt == 0) {          // theta rounded up to M
                if (gap >= 1.f) tie_mode = true;
                gap = fminf(1.f, gap * 4.f);
                continue;
            }
            break;
        }
        if (tie_mode) {
            // more than CAP points within rounding of the maximum: draw ONE sample, the lowest original index among the
            // points that attain M
            if (r == 0 && w == 0) {
                unsigned long long t = ~0ull;
                for (int i = 0; i < NC; i++) {
                    const unsigned long long v = reinterpret_cast<volatile unsigned long long*>(tiev)[i];
                    t = v < t ? v : t;
                }
                const int p = (int)(unsigned)t;
                if (lane < NC) {
                    cluster.map_shared_rank(cent, lane)[0] = make_float4(sx[p], sy[p], sz[p], __uint_as_float((unsigned)(t >> 32)));
                    int* c = cluster.map_shared_rank(ctrl, lane);
                    c[C_S] = s + 1, c[C_NCENT] = 1;
                }
                if (lane == 0) outc[s] = (int32_t)(unsigned)(t >> 32);
            }
            cluster.sync();
            gap = GAP_MIN * 4.f;
        } else {
            if (cnt < STEER_LO) gap = fminf(1.f, gap * 1.5f);
            else if (cnt > STEER_HI) gap *= 0.85f;
            // ---- (3) the chain: warp 0 draws samples from the queue until its maximum falls to theta ----
            if (r == 0 && w == 0) {
                int j = 0;
                run_chain_any(cnt, Lp, Ld, Lc, sx, sy, sz, so, theta, npoint, lane, s, j, [&](int jj, const float4& c, int ss) {
                    if (lane < NC) cluster.map_shared_rank(cent, lane)[jj] = c;    // the sample, into every CTA
                    if (lane == 0) outc[ss] = (int32_t)__float_as_uint(c.w);
                });
                if (lane < NC) {
                    int* c = cluster.map_shared_rank(ctrl, lane);
                    c[C_S] = s, c[C_NCENT] = j;
                }
            }
            cluster.sync();                                                        // B3
        }
        s = ctrl[C_S], ncent = ctrl[C_NCENT];
        if (s >= npoint) break;
    }
    cluster.sync();                              // no CTA exits while a sibling may still touch its shared memory
}

size_t queue_smem(int n, bool md_smem) {
    const size_t npad = ((size_t)n + 31) / 32 * 32;
    const size_t md = md_smem ? (npad > (size_t)CELLS ? npad : (size_t)CELLS) : (size_t)CELLS;
    return md * 4 + (size_t)JMAX * 16 + (size_t)CAP * 24 + (6 * NW + NW + NW + C_WORDS) * 4 + 16;
}

template <bool MD_SMEM, int BPL>
int launch_queue(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, int nb, int KB, cudaStream_t st) {
    const size_t smem = queue_smem(n, MD_SMEM);
    if (int rc = pf_smem_optin(fps_queue_kernel<MD_SMEM, BPL>, smem)) return rc;
    fps_queue_kernel<MD_SMEM, BPL><<<B, FT, smem, st>>>(xyz, n, npoint, idx, work, nb, KB);
    return pf_launch_status("fps_queue_kernel");
}

size_t cqueue_smem(int nbc) {
    const size_t md = (size_t)32 * nbc > (size_t)CELLS ? (size_t)32 * nbc : (size_t)CELLS;
    return md * 4 + (size_t)JMAX * 16 + (size_t)CAP * 24 + (6 * NW + 8 * NW + NW + C_WORDS) * 4 + 16 + 64;
}

template <int BPL>
int launch_cqueue(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, int nb, int nbc, int KB, int NC,
                  cudaStream_t st) {
    const size_t smem = cqueue_smem(nbc);
    if (int rc = pf_smem_optin(fps_cqueue_kernel<BPL>, smem)) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(B * NC);
    cfg.blockDim = dim3(FT);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = NC, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    PF_CUDA(cudaLaunchKernelEx(&cfg, fps_cqueue_kernel<BPL>, xyz, n, npoint, idx, work, nb, nbc, KB));
    return pf_launch_status("fps_cqueue_kernel");
}

}  // namespace

// largest cloud the register-resident bucket boxes cover: 5 buckets of 32 points per lane
extern "C" int pf_fps_queued_max_points(void) { return 32 * NW * 32 * 5; }

// pf_fps_queued with its scratch in the library's own static device array: no allocation anywhere (no cudaMalloc /
// cudaFree, no allocating library call), for callers that cannot allocate (CUDA-graph capture, serving loops).
// All calls on a device share the one array: calls that overlap on the GPU (two streams, two threads) race.  Nothing in
// it has to start at zero -- the kernel writes every entry it later reads (sorted coordinates, indices, running minima)
// before the first round -- so the previous call's contents are harmless.
extern "C" int pf_fps_queued_static(const float* xyz, int B, int n, int npoint, int32_t* idx, void* stream) {
    PF_CHECK_ARG(xyz && idx, "pf_fps_queued_static: null pointer");
    PF_CHECK_ARG(B >= 0 && n > 0 && npoint > 0, "pf_fps_queued_static: bad sizes B=%d n=%d npoint=%d", B, n, npoint);
    const long long need = (long long)B * (((long long)n + 31) / 32 * 32);
    PF_CHECK_ARG(need <= (long long)PF_FPS_STATIC_MAX_POINTS,
                 "pf_fps_queued_static: B=%d clouds of n=%d candidates need %lld points of scratch (32-padded), the static array holds "
                 "PF_FPS_STATIC_MAX_POINTS=%d; use pf_fps_queued with caller-owned scratch", B, n, need, PF_FPS_STATIC_MAX_POINTS);
    if (B == 0) return 0;
    constexpr int MAXDEV = 64;
    static float* dev_ptr[MAXDEV] = {nullptr};        // address of the array on each device, looked up once per device
    static std::mutex mu;
    int dev = -1;
    PF_CUDA(cudaGetDevice(&dev));
    PF_CHECK_ARG(dev >= 0 && dev < MAXDEV, "pf_fps_queued_static: device ordinal %d outside [0, %d)", dev, MAXDEV);
    float* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!dev_ptr[dev]) {
            void* p = nullptr;
            PF_CUDA(cudaGetSymbolAddress(&p, g_fps_static_scratch));      // with `dev` current: this device's copy
            dev_ptr[dev] = static_cast<float*>(p);
        }
        scratch = dev_ptr[dev];
    }
    return pf_fps_queued(xyz, B, n, npoint, idx, scratch, stream);
}

extern "C" int pf_fps_queued(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, void* stream) {
    PF_CHECK_ARG(xyz && idx && work, "pf_fps_queued: null pointer (work must hold pf_fps_work_floats(B, n) floats)");
    PF_CHECK_ARG(B >= 0 && n > 0 && npoint > 0, "pf_fps_queued: bad sizes B=%d n=%d npoint=%d", B, n, npoint);
    if (B == 0) return 0;
    if (n > pf_fps_queued_max_points()) return pf_fps_bucketed(xyz, B, n, npoint, idx, work, stream);
    const int nb = (n + 31) / 32, KB = (nb + NW - 1) / NW;
    cudaStream_t st = (cudaStream_t)stream;
    // FEW clouds: a cloud's buckets over a cluster of 8 CTAs (fps_cqueue_kernel), when all clusters are resident at once (clusters
    // cannot span GPCs: B = 16 clusters of 8 ran in two waves, 7.1 ms against 3.9).  One 40,960-candidate cloud 3.42 -> 2.39 ms,
    // eight 99,840-candidate clouds 14.4 -> 6.5 ms.
#ifndef PF_CQ_MIN
#define PF_CQ_MIN 16384
#endif
    if (n >= PF_CQ_MIN) {
        const int NC = 8;
        const int nbc = (nb + NC - 1) / NC, kbc = (nbc + NW - 1) / NW, bplc = (kbc + 31) / 32;
        if (bplc <= 2 && cqueue_smem(nbc) <= 227 * 1024) {
            auto kern = bplc <= 1 ? fps_cqueue_kernel<1> : fps_cqueue_kernel<2>;
            const size_t smem = cqueue_smem(nbc);
            if (int rc = pf_smem_optin(kern, smem)) return rc;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(NC), cfg.blockDim = dim3(FT), cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = NC, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr, cfg.numAttrs = 1;
            int resident = 0;
            if (cudaOccupancyMaxActiveClusters(&resident, kern, &cfg) != cudaSuccess) resident = 0, (void)cudaGetLastError();
            if (B <= resident) {
                if (bplc <= 1) return launch_cqueue<1>(xyz, B, n, npoint, idx, work, nb, nbc, kbc, NC, st);
                return launch_cqueue<2>(xyz, B, n, npoint, idx, work, nb, nbc, kbc, NC, st);
            }
        }
    }
    const bool md_smem = queue_smem(n, true) <= 227 * 1024;
    const int bpl = (KB + 31) / 32;
    if (md_smem && bpl <= 1) return launch_queue<true, 1>(xyz, B, n, npoint, idx, work, nb, KB, st);
    if (md_smem && bpl <= 2) return launch_queue<true, 2>(xyz, B, n, npoint, idx, work, nb, KB, st);
    if (bpl <= 3) return launch_queue<false, 3>(xyz, B, n, npoint, idx, work, nb, KB, st);
    return launch_queue<false, 5>(xyz, B, n, npoint, idx, work, nb, KB, st);
}
