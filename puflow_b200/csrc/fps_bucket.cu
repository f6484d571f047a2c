// Farthest point sampling with exact spatial pruning: one CTA per cloud, all clouds of a batch in one wave.
//
// Same semantics as pf_fps / the reference's own farthest_point_sampling
// (/root/reference/modules/utils/fps.py:40-73; replaces pointnet2_ops.furthest_point_sample at
// /root/reference/modules/utils/patch.py:102,156): start at index 0, running min squared distance initialised to
// 1e10, arg-max with the lowest index winning ties.  Results are bit-identical to the brute-force kernel.
//
// The merge FPS of the x4 path draws 8,216 of 40,960 candidates: 8,216 dependent rounds that each touch every
// point when done by brute force.  But a round only changes the running minimum of points closer to the new
// sample than their current minimum, which after the first few rounds is a small neighbourhood.  So:
//   setup   counting-sort the cloud into a 16^3 grid (Morton order), cut the sorted sequence into buckets of 32
//           points, keep per bucket a bounding box and the maximum running-min distance (shared memory);
//   round   every warp tests the boxes of the buckets it owns against the new sample and only updates buckets
//           whose box is closer than their current maximum (exact: a skipped bucket cannot change);
//           arg-max = max over bucket maxima, then a look inside the bucket(s) holding the maximum.
// Running minima live in shared memory (164 KB for 40,960 points), sorted coordinates in an L2-resident scratch.
// Every bucket caches the point that attains its maximum (lowest original index first), so choosing the next sample is a
// reduction over per-warp candidates: one __syncthreads per round; no clusters, no cross-CTA traffic.
#include "common.cuh"
#include "puflow_b200.h"

namespace {

#ifndef PF_FPS_THREADS
#define PF_FPS_THREADS 640                   // 20 warps: the 1,280 buckets of a 40,960-point cloud are 2 full warp-iterations each
#endif
constexpr int FT = PF_FPS_THREADS, NW = FT / 32;
#ifndef PF_FPS_BPL
#define PF_FPS_BPL 2
#endif
constexpr int BPL = PF_FPS_BPL;              // buckets per lane held in registers (REGBOX path)
static_assert(FT % 32 == 0 && NW <= 32, "per-warp partial results are reduced by one warp");
constexpr int CELLS = 4096;                  // 16^3 grid
constexpr float INIT_D = 1e10f;              // fps.py:53
constexpr float SAFE = 0.9999f;              // box lower bound is scaled down: never above any computed distance

__device__ __forceinline__ unsigned morton4(unsigned x, unsigned y, unsigned z) {
    auto spread = [](unsigned v) { return (v & 1u) | ((v & 2u) << 2) | ((v & 4u) << 4) | ((v & 8u) << 6); };
    return spread(x) | (spread(y) << 1) | (spread(z) << 2);
}
__device__ __forceinline__ float sqdist(float px, float py, float pz, float cx, float cy, float cz) {
    const float dx = __fsub_rn(px, cx), dy = __fsub_rn(py, cy), dz = __fsub_rn(pz, cz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

#ifdef PF_FPS_TIMING
__device__ long long g_fps_timing[8];
#define FSTAMP(k)                                            \
    do {                                                     \
        long long now_ = clock64();                          \
        tacc_[k] += now_ - tprev_;                           \
        tprev_ = now_;                                       \
    } while (0)
#else
#define FSTAMP(k)
#endif

struct Slot {
    float m;
    unsigned oidx;
    float x, y, z;
    float pad[3];
};

// work (per cloud, npad = 32 * nb entries each): sx | sy | sz | oidx | md (md only used when !MD_SMEM)
// REGBOX (KB <= 32 BPL): every lane keeps the box, the running maximum and the cached winner of its own (at most BPL)
// buckets in registers for the whole kernel -- the bucket -> lane map never changes -- so a round's scan of the 1,280
// boxes touches no shared memory at all (the scan, not the distance updates, is what a round spends its issue slots on).
template <bool MD_SMEM, bool REGBOX>
__global__ void __launch_bounds__(FT, 1)
    fps_bucket_kernel(const float* __restrict__ xyz, int n, int npoint, int32_t* __restrict__ out, float* work, int nb,
                      int KB, const int* __restrict__ run_flag) {
    if (run_flag && *run_flag == 0) return;      // conditional launch (pf_fps_clustered's fallback): nothing to do
    extern __shared__ __align__(16) float sm[];
    const int npad = nb * 32;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int cloud = blockIdx.x;
    const float* pb = xyz + (size_t)cloud * n * 3;
    float* sx = work + (size_t)cloud * npad * 5;
    float* sy = sx + npad;
    float* sz = sy + npad;
    unsigned* so = reinterpret_cast<unsigned*>(sz + npad);
    float* mdg = sz + 2 * (size_t)npad;

    // ---- shared memory carve-up ----
    const int mdwords = MD_SMEM ? npad : CELLS;            // region 0: md (rounds) aliased by the histogram (setup)
    float* md_s = sm;
    int* hist = reinterpret_cast<int*>(sm);
    // per bucket slot two float4: (lo.x, lo.y, lo.z, max running-min of the bucket) and (hi.x, hi.y, hi.z, -)
    float4* bbq = reinterpret_cast<float4*>(sm + (mdwords > CELLS ? mdwords : CELLS));
    const int nslot = NW * KB;
    unsigned* wc_o = reinterpret_cast<unsigned*>(bbq + 2 * nslot);   // per bucket: the point that attains bmax with the lowest
    float* wc_x = reinterpret_cast<float*>(wc_o + nslot);         // original index (oidx, x, y, z), refreshed whenever the
    float* wc_y = wc_x + nslot;                                   // bucket is updated, so that picking the next sample
    float* wc_z = wc_y + nslot;                                   // needs no trip to L2
    float* red = wc_z + nslot;                             // [6*NW] setup reductions, then wmax[NW]
    Slot* slots = reinterpret_cast<Slot*>(red + 6 * NW);   // [NW]
    int* scan_tmp = reinterpret_cast<int*>(slots + 2 * NW);    // [NW]

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
        lo[a] = warp_min(lo[a]), hi[a] = warp_max(hi[a]);
        if (lane == 0) red[a * NW + w] = lo[a], red[(3 + a) * NW + w] = hi[a];
    }
    for (int i = tid; i < CELLS; i += FT) hist[i] = 0;
    __syncthreads();
    float inv[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        lo[a] = warp_min(red[a * NW + min(lane, NW - 1)]);
        hi[a] = warp_max(red[(3 + a) * NW + min(lane, NW - 1)]);
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
        int v[CPT], s = 0;
#pragma unroll
        for (int u = 0; u < CPT; u++) v[u] = (CPT * tid + u < CELLS) ? hist[CPT * tid + u] : 0, s += v[u];
        int incl = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) scan_tmp[w] = incl;
        __syncthreads();
        int wsum = lane < NW ? scan_tmp[lane] : 0;
        int wincl = wsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, wincl, o);
            if (lane >= o) wincl += t;
        }
        const int wbase = __shfl_sync(0xffffffffu, wincl - wsum, w);
        int run = wbase + incl - s;
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
    // ---- setup 3: bucket boxes, running minima ----
    for (int k = 0; k < KB; k++) {
        const int b = w + NW * k, slot = w * KB + k;
        if (b < nb) {
            const float x = sx[32 * b + lane], y = sy[32 * b + lane], z = sz[32 * b + lane];
            const float v[6] = {warp_min(x), warp_min(y), warp_min(z), warp_max(x), warp_max(y), warp_max(z)};
            if (lane == 0) {
                bbq[2 * slot] = make_float4(v[0], v[1], v[2], INIT_D);
                bbq[2 * slot + 1] = make_float4(v[3], v[4], v[5], 0.f);
            }
            const float m0 = (32 * b + lane < n) ? INIT_D : 0.f;
            if (MD_SMEM) md_s[32 * b + lane] = m0; else mdg[32 * b + lane] = m0;
            const unsigned o = (32 * b + lane < n) ? so[32 * b + lane] : 0xffffffffu;
            if (o == __reduce_min_sync(0xffffffffu, o)) wc_o[slot] = o, wc_x[slot] = x, wc_y[slot] = y, wc_z[slot] = z;
        } else if (lane == 0) {
            bbq[2 * slot] = bbq[2 * slot + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
            wc_o[slot] = 0xffffffffu, wc_x[slot] = 0.f, wc_y[slot] = 0.f, wc_z[slot] = 0.f;
        }
    }
    float cx = pb[0], cy = pb[1], cz = pb[2];
    if (tid == 0) out[(size_t)cloud * npoint] = 0;
    __syncthreads();

    // ---- rounds ----
#ifdef PF_FPS_TIMING
    long long tacc_[5] = {0, 0, 0, 0, 0};
    long long tprev_ = clock64();
#endif
    if constexpr (REGBOX) {
        // per-lane copy of its buckets' metadata (h = 0: bucket k = lane, h = 1: k = lane + 32 of this warp)
        float blx[BPL], bly[BPL], blz[BPL], bhx[BPL], bhy[BPL], bhz[BPL], bmr[BPL], wx[BPL], wy[BPL], wz[BPL];
        unsigned wo[BPL];
#pragma unroll
        for (int h = 0; h < BPL; h++) {
            const int k = lane + 32 * h, slot = w * KB + min(k, KB - 1);
            const float4 blo = bbq[2 * slot], bhi = bbq[2 * slot + 1];
            const bool ok = k < KB;
            blx[h] = blo.x, bly[h] = blo.y, blz[h] = blo.z, bhx[h] = bhi.x, bhy[h] = bhi.y, bhz[h] = bhi.z;
            bmr[h] = ok ? blo.w : 0.f;                                  // 0: never active, never the maximum
            wo[h] = ok ? wc_o[slot] : 0xffffffffu;
            wx[h] = wc_x[slot], wy[h] = wc_y[slot], wz[h] = wc_z[slot];
        }
        for (int s = 0; s + 1 < npoint; s++) {
            bool act[BPL];
#pragma unroll
            for (int h = 0; h < BPL; h++) {
                const float ex = fmaxf(fmaxf(blx[h] - cx, cx - bhx[h]), 0.f);
                const float ey = fmaxf(fmaxf(bly[h] - cy, cy - bhy[h]), 0.f);
                const float ez = fmaxf(fmaxf(blz[h] - cz, cz - bhz[h]), 0.f);
                act[h] = (ex * ex + ey * ey + ez * ez) * SAFE < bmr[h];
            }
#pragma unroll
            for (int h = 0; h < BPL; h++) {
                unsigned mask = __ballot_sync(0xffffffffu, act[h]);
                while (mask) {
                    int kk[4];
                    float px[4], py[4], pz[4];
                    unsigned po[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        kk[u] = -1;
                        if (mask) {
                            kk[u] = __ffs(mask) - 1;
                            mask &= mask - 1;
                            const int p = 32 * (w + NW * (32 * h + kk[u])) + lane;
                            px[u] = sx[p], py[u] = sy[p], pz[u] = sz[p], po[u] = so[p];
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        if (kk[u] >= 0) {
                            const int p = 32 * (w + NW * (32 * h + kk[u])) + lane;
                            const float d = sqdist(px[u], py[u], pz[u], cx, cy, cz);
                            float m = MD_SMEM ? md_s[p] : mdg[p];
                            if (d < m) {
                                m = d;
                                if (MD_SMEM) md_s[p] = d; else mdg[p] = d;
                            }
                            const float nm = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(m)));
                            const unsigned o = (m == nm) ? po[u] : 0xffffffffu;
                            const unsigned omin = __reduce_min_sync(0xffffffffu, o);
                            const int srcl = __ffs(__ballot_sync(0xffffffffu, o == omin)) - 1;   // lane holding the winner
                            const float tx = __shfl_sync(0xffffffffu, px[u], srcl), ty = __shfl_sync(0xffffffffu, py[u], srcl),
                                        tz = __shfl_sync(0xffffffffu, pz[u], srcl);
                            if (lane == kk[u]) bmr[h] = nm, wo[h] = omin, wx[h] = tx, wy[h] = ty, wz[h] = tz;
                        }
                    }
                }
            }
            // this lane's better bucket, then the warp's, then (after the barrier) the block's
            float lmax = bmr[0], bx = wx[0], by = wy[0], bz = wz[0];
            unsigned lo_ = wo[0];
#pragma unroll
            for (int h = 1; h < BPL; h++)
                if (bmr[h] > lmax || (bmr[h] == lmax && wo[h] < lo_)) lmax = bmr[h], lo_ = wo[h], bx = wx[h], by = wy[h], bz = wz[h];
            Slot* sl = slots + (s & 1) * NW;
            {
                const float Mw = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(lmax)));
                const unsigned cand = (lmax == Mw) ? lo_ : 0xffffffffu;
                const unsigned omin = __reduce_min_sync(0xffffffffu, cand);
                const int src = __ffs(__ballot_sync(0xffffffffu, cand == omin && lmax == Mw)) - 1;
                if (lane == src) sl[w].m = Mw, sl[w].oidx = omin, sl[w].x = bx, sl[w].y = by, sl[w].z = bz;
            }
            __syncthreads();
            {
                const int l = min(lane, NW - 1);
                const float mw = sl[l].m;
                const unsigned ow = sl[l].oidx;
                const float M = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(mw)));
                const unsigned cand = (mw == M) ? ow : 0xffffffffu;
                const unsigned omin = __reduce_min_sync(0xffffffffu, cand);
                const int src = min(__ffs(__ballot_sync(0xffffffffu, cand == omin && mw == M)) - 1, NW - 1);
                cx = sl[src].x, cy = sl[src].y, cz = sl[src].z;
                if (tid == 0) out[(size_t)cloud * npoint + s + 1] = (int32_t)omin;
            }
        }
        return;
    }
    for (int s = 0; s + 1 < npoint; s++) {
        FSTAMP(0);
        // (1) update the buckets whose box is closer to the new sample than their current maximum
        // every lane also tracks the best bucket it has seen: (max running-min, lowest original index attaining it)
        float lmax = 0.f;
        unsigned lo_ = 0xffffffffu;
        int ls = w * KB;
        for (int kbase = 0; kbase < KB; kbase += 32) {
            const int k = kbase + lane, slot = w * KB + k;
            bool act = false;
            float bm = 0.f;
            if (k < KB) {
                const float4 blo = bbq[2 * slot], bhi = bbq[2 * slot + 1];
                bm = blo.w;
                const float ex = fmaxf(fmaxf(blo.x - cx, cx - bhi.x), 0.f);
                const float ey = fmaxf(fmaxf(blo.y - cy, cy - bhi.y), 0.f);
                const float ez = fmaxf(fmaxf(blo.z - cz, cz - bhi.z), 0.f);
                act = (ex * ex + ey * ey + ez * ez) * SAFE < bm;
            }
            unsigned mask = __ballot_sync(0xffffffffu, act);
            const bool any_act = mask != 0;
            while (mask) {
                int kk[4];
                float px[4], py[4], pz[4];
                unsigned po[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    kk[u] = -1;
                    if (mask) {
                        kk[u] = __ffs(mask) - 1;
                        mask &= mask - 1;
                        const int p = 32 * (w + NW * (kbase + kk[u])) + lane;
                        px[u] = sx[p], py[u] = sy[p], pz[u] = sz[p], po[u] = so[p];
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (kk[u] >= 0) {
                        const int p = 32 * (w + NW * (kbase + kk[u])) + lane;
                        const float d = sqdist(px[u], py[u], pz[u], cx, cy, cz);
                        float m = MD_SMEM ? md_s[p] : mdg[p];
                        if (d < m) {
                            m = d;
                            if (MD_SMEM) md_s[p] = d; else mdg[p] = d;
                        }
                        const float nm = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(m)));
                        if (lane == kk[u]) bm = nm;
                        const unsigned o = (m == nm) ? po[u] : 0xffffffffu;
                        if (o == __reduce_min_sync(0xffffffffu, o)) {
                            const int slot_u = w * KB + kbase + kk[u];
                            wc_o[slot_u] = o, wc_x[slot_u] = px[u], wc_y[slot_u] = py[u], wc_z[slot_u] = pz[u];
                        }
                    }
                }
            }
            if (any_act) __syncwarp();                  // winner-cache entries written by other lanes of this warp
            if (k < KB) {
                if (act) bbq[2 * slot].w = bm;
                const unsigned o = wc_o[slot];
                if (bm > lmax || (bm == lmax && o < lo_)) lmax = bm, lo_ = o, ls = slot;
            }
        }
        // (2) this warp's candidate -> slots[parity][w]
        Slot* sl = slots + (s & 1) * NW;
        {
            const float Mw = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(lmax)));
            const unsigned cand = (lmax == Mw) ? lo_ : 0xffffffffu;
            const unsigned omin = __reduce_min_sync(0xffffffffu, cand);
            const int src = __ffs(__ballot_sync(0xffffffffu, cand == omin && lmax == Mw)) - 1;
            if (lane == src) sl[w].m = Mw, sl[w].oidx = omin, sl[w].x = wc_x[ls], sl[w].y = wc_y[ls], sl[w].z = wc_z[ls];
        }
        FSTAMP(1);
        __syncthreads();
        FSTAMP(2);
        // (3) global maximum; the lowest original index among the points attaining it wins (fps.py: torch.max returns
        //     the first maximum).  slots are double-buffered by round parity, so one barrier per round suffices.
        {
            const int l = min(lane, NW - 1);
            const float mw = sl[l].m;
            const unsigned ow = sl[l].oidx;
            const float M = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(mw)));
            const unsigned cand = (mw == M) ? ow : 0xffffffffu;
            const unsigned omin = __reduce_min_sync(0xffffffffu, cand);
            const int src = min(__ffs(__ballot_sync(0xffffffffu, cand == omin && mw == M)) - 1, NW - 1);
            cx = sl[src].x, cy = sl[src].y, cz = sl[src].z;
            if (tid == 0) out[(size_t)cloud * npoint + s + 1] = (int32_t)omin;
        }
    }
#ifdef PF_FPS_TIMING
    if (blockIdx.x == 0 && tid == 0)
        for (int k = 0; k < 5; k++) g_fps_timing[k] += tacc_[k];
#endif
}

}  // namespace

#ifdef PF_FPS_TIMING
extern "C" __attribute__((visibility("default"))) int pf_fps_timing(long long* host8) {
    cudaDeviceSynchronize();
    cudaError_t e = cudaMemcpyFromSymbol(host8, g_fps_timing, sizeof(long long) * 8);
    long long z[8] = {0};
    cudaMemcpyToSymbol(g_fps_timing, z, sizeof(z));
    return (int)e;
}
#endif

static int fps_bucket_geometry(int n, int* nb, int* KB, size_t* smem_md, size_t* smem_nomd) {
    *nb = (n + 31) / 32;
    *KB = (*nb + NW - 1) / NW;
    const size_t meta = (size_t)(12 * NW * *KB + 6 * NW) * 4 + 2 * NW * sizeof(Slot) + NW * 4;
    const size_t npad = (size_t)*nb * 32;
    *smem_md = ((npad > (size_t)CELLS ? npad : (size_t)CELLS) * 4) + meta;
    *smem_nomd = (size_t)CELLS * 4 + meta;
    return 0;
}

extern "C" long long pf_fps_work_floats(int B, int n) {
    if (B <= 0 || n <= 0) return 0;
    return (long long)B * (((long long)n + 31) / 32 * 32) * 5;
}

extern "C" int pf_fps_bucketed(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, void* stream) {
    return pf_fps_bucketed_if(xyz, B, n, npoint, idx, work, nullptr, stream);
}

// run_flag: NULL = always run; otherwise a device int, the kernel returns at once when it reads 0 (no host round trip)
int pf_fps_bucketed_if(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, const int* run_flag, void* stream) {
    PF_CHECK_ARG(xyz && idx && work, "pf_fps_bucketed: null pointer (work must hold pf_fps_work_floats(B, n) floats)");
    PF_CHECK_ARG(B >= 0 && n > 0 && npoint > 0, "pf_fps_bucketed: bad sizes B=%d n=%d npoint=%d", B, n, npoint);
    if (B == 0) return 0;
    int nb, KB;
    size_t s_md, s_nomd;
    fps_bucket_geometry(n, &nb, &KB, &s_md, &s_nomd);
    cudaStream_t st = (cudaStream_t)stream;
    if (s_md <= 227 * 1024) {
        if (KB <= 32 * BPL) {    // bucket metadata in registers (BPL buckets per lane)
            if (int rc = pf_smem_optin(fps_bucket_kernel<true, true>, s_md)) return rc;
            fps_bucket_kernel<true, true><<<B, FT, s_md, st>>>(xyz, n, npoint, idx, work, nb, KB, run_flag);
        } else {
            if (int rc = pf_smem_optin(fps_bucket_kernel<true, false>, s_md)) return rc;
            fps_bucket_kernel<true, false><<<B, FT, s_md, st>>>(xyz, n, npoint, idx, work, nb, KB, run_flag);
        }
    } else {
        PF_CHECK_ARG(s_nomd <= 227 * 1024, "pf_fps_bucketed: n=%d too large (bucket metadata exceeds shared memory)", n);
        if (int rc = pf_smem_optin(fps_bucket_kernel<false, false>, s_nomd)) return rc;
        fps_bucket_kernel<false, false><<<B, FT, s_nomd, st>>>(xyz, n, npoint, idx, work, nb, KB, run_flag);
    }
    return pf_launch_status("fps_bucket_kernel");
}
