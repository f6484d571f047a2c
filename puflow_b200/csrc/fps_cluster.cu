// Farthest point sampling with exact spatial pruning, one thread-block CLUSTER per cloud.
//
// Same semantics and the same result, bit for bit, as pf_fps / pf_fps_bucketed (the reference's
// farthest_point_sampling, /root/reference/modules/utils/fps.py:40-73, standing in for
// pointnet2_ops.furthest_point_sample at /root/reference/modules/utils/patch.py:102,156): start at index 0, running min
// squared distance initialised to 1e10, arg-max with the lowest index winning ties.
//
// pf_fps_bucketed gives a whole cloud to ONE CTA: right for a batch of 128 clouds (one per SM), wrong for a few large
// clouds -- config 3's merge FPS draws 20,024 of 99,840 candidates in 20,024 dependent rounds of 2.3 us on one SM
// (46 ms, 99 % of the cloud's time) while 140 SMs idle, and a single 2,048-pt cloud waits 7.6 ms for its 8,216 rounds.
// Here the cloud's grid-sorted buckets are spread over the 2..16 CTAs of a cluster:
//   * a CTA owns a contiguous range of Morton cells (at most CAP <= 8,192 points), keeps THEIR sorted coordinates,
//     original indices and running minima in its shared memory (no L2 trip in a round) and one bucket per lane in
//     registers (box, running maximum, cached winner);
//   * a round: box test -> update of the active buckets -> warp arg-max -> CTA arg-max -> the CTA's candidate (value, index,
//     coordinates) is stored into every CTA's shared memory through DSMEM as two 16-byte stores that each carry the round
//     number -> every CTA polls its OWN shared memory until all CS candidates show the current round and picks the winner
//     from its local copy: one DSMEM store latency per round; no cluster barrier, no release / acquire fence (a remote
//     mbarrier arrive with release semantics was measured first: 1.1 us per round, no better than one CTA alone).
// Every CTA reads the whole cloud during set-up (bounding box, cell histogram, scan: identical in all CTAs, so the
// ownership of cells is consistent without communication).  If a CTA's share exceeds CAP (degenerate clouds: thousands of
// points in one cell) the kernel raises `*fallback` and returns; pf_fps_clustered then runs the single-CTA kernel.
#include <cooperative_groups.h>

#include "common.cuh"
#include "puflow_b200.h"

namespace cg = cooperative_groups;

namespace {

constexpr int CT = 256, NW = CT / 32;        // threads / warps per CTA; bucket b <-> (warp b % NW, lane b / NW)
constexpr int CELLS = 4096;                  // 16^3 grid
constexpr int MAXCS = 16;
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
__device__ __forceinline__ uint32_t mapa(uint32_t local_smem_addr, uint32_t cta) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(cta));
    return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

struct __align__(16) Cand {     // two 16-byte halves, each stored whole and tagged with the round it belongs to
    float m, x, y;
    unsigned tag_a;
    float z;
    unsigned oidx, tag_b, pad;
};
__device__ __forceinline__ uint4 ld_volatile_v4(const void* p) {
    uint4 v;
    asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(pf::smem_u32(p)) : "memory");
    return v;
}

__global__ void __launch_bounds__(CT, 1)
    fps_cluster_kernel(const float* __restrict__ xyz, int n, int npoint, int32_t* __restrict__ out, int cap, int* __restrict__ fallback) {
    cg::cluster_group cluster = cg::this_cluster();
    const int CS = (int)cluster.num_blocks(), cr = (int)cluster.block_rank();
    const int cloud = blockIdx.x / CS;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const float* pb = xyz + (size_t)cloud * n * 3;

    extern __shared__ __align__(16) float sm[];
    float* sx = sm;                                   // [cap] sorted coordinates of the points this CTA owns
    float* sy = sx + cap;
    float* sz = sy + cap;
    unsigned* so = reinterpret_cast<unsigned*>(sz + cap);   // [cap] original indices
    float* md = reinterpret_cast<float*>(so + cap);   // [cap] running minima; aliased by the cell histogram during set-up
    int* hist = reinterpret_cast<int*>(md);           // [CELLS] (cap >= CELLS)
    __shared__ Cand exch[2][MAXCS];                   // candidates of all CTAs of the cluster, by round parity
    __shared__ Cand wcand[NW];
    __shared__ float red[6 * NW];
    __shared__ int scan_tmp[NW], s_first, s_cnt, s_over;

    // ---- set-up 1: bounding box of the cloud (every CTA, whole cloud) ----
    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (int i = tid; i < n; i += CT)
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
    for (int i = tid; i < CELLS; i += CT) hist[i] = 0;
    if (tid < 2 * MAXCS) (&exch[0][0])[tid].tag_a = 0u, (&exch[0][0])[tid].tag_b = 0u;      // round tags start at 1
    if (tid == 0) s_over = 0;
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
    // ---- set-up 2: cell histogram of the whole cloud, exclusive scan, ownership ----
    for (int i = tid; i < n; i += CT) atomicAdd(&hist[cell_of(pb[(size_t)i * 3], pb[(size_t)i * 3 + 1], pb[(size_t)i * 3 + 2])], 1);
    __syncthreads();
    constexpr int CPT = CELLS / CT;                   // 16 cells per thread
    // owner(cell) = floor(start(cell) * CS / n): contiguous cell ranges of about n / CS points, identical in every CTA
    {
        int v[CPT], s = 0;
#pragma unroll
        for (int u = 0; u < CPT; u++) v[u] = hist[CPT * tid + u], s += v[u];
        int incl = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) scan_tmp[w] = incl;
        if (tid == 0) s_first = 0x7fffffff, s_cnt = 0;
        __syncthreads();
        const int wsum = lane < NW ? scan_tmp[lane] : 0;
        int wincl = wsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, wincl, o);
            if (lane >= o) wincl += t;
        }
        const int wbase = __shfl_sync(0xffffffffu, wincl - wsum, w);
        int run = wbase + incl - s, mine = 0, first = 0x7fffffff;
        int over = 0;
        // per-owner totals are needed by everybody to agree on an overflow: track the running owner segment locally
#pragma unroll
        for (int u = 0; u < CPT; u++) {
            const int start = run;
            const int owner = (int)(((long long)start * CS) / n);
            run += v[u];
            // the cursor of an owned cell = its position inside this CTA's range (set below once s_first is known)
            hist[CPT * tid + u] = (owner == cr && v[u] > 0) ? start : -1;
            if (owner == cr && v[u] > 0) mine += v[u], first = min(first, start);
            // a cell that alone is larger than the capacity, or that starts too late in its owner's range, overflows it
            const long long seg_lo = ((long long)owner * n + CS - 1) / CS;         // first position the owner can start at
            if (v[u] > 0 && (long long)start + v[u] - seg_lo > cap) over = 1;
        }
        if (mine) atomicAdd(&s_cnt, mine), atomicMin(&s_first, first);
        if (over) s_over = 1;
    }
    __syncthreads();
    if (s_over) {                                     // identical decision in every CTA of the cluster (same histogram)
        if (cr == 0 && tid == 0) *fallback = 1;
        return;
    }
    const int cnt = s_cnt, base = cnt ? s_first : 0;
    const int nbl = (cnt + 31) / 32, npad = nbl * 32;
    // ---- set-up 3: scatter the owned points (order inside a cell is arbitrary: buckets are sets, ties go by original index) ----
    for (int i = tid; i < n; i += CT) {
        const float x = pb[(size_t)i * 3], y = pb[(size_t)i * 3 + 1], z = pb[(size_t)i * 3 + 2];
        const unsigned c = cell_of(x, y, z);
        if (hist[c] >= 0) {
            const int pos = atomicAdd(&hist[c], 1) - base;
            sx[pos] = x, sy[pos] = y, sz[pos] = z, so[pos] = (unsigned)i;
        }
    }
    __syncthreads();                                  // the histogram (aliasing md) is dead from here on
    for (int i = cnt + tid; i < npad; i += CT) sx[i] = pb[0], sy[i] = pb[1], sz[i] = pb[2], so[i] = 0x7fffffffu;
    __syncthreads();
    // ---- set-up 4: one bucket per lane: box, running maximum, cached winner (lowest original index attaining it) ----
    float blx = 0.f, bly = 0.f, blz = 0.f, bhx = 0.f, bhy = 0.f, bhz = 0.f, bmr = 0.f, wx = 0.f, wy = 0.f, wz = 0.f;
    unsigned wo = 0xffffffffu;
    for (int k = 0; k < 32; k++) {
        const int b = w + NW * k;                     // warp-uniform
        if (b >= nbl) break;
        const int p = 32 * b + lane;
        const float x = sx[p], y = sy[p], z = sz[p];
        const float v0 = warp_min(x), v1 = warp_min(y), v2 = warp_min(z), v3 = warp_max(x), v4 = warp_max(y), v5 = warp_max(z);
        const bool real = p < cnt;
        md[p] = real ? INIT_D : 0.f;
        const unsigned o = real ? so[p] : 0xffffffffu;
        const unsigned omin = __reduce_min_sync(0xffffffffu, o);
        const int srcl = __ffs(__ballot_sync(0xffffffffu, o == omin)) - 1;
        const float tx = __shfl_sync(0xffffffffu, x, srcl), ty = __shfl_sync(0xffffffffu, y, srcl), tz = __shfl_sync(0xffffffffu, z, srcl);
        if (lane == k) blx = v0, bly = v1, blz = v2, bhx = v3, bhy = v4, bhz = v5, bmr = INIT_D, wo = omin, wx = tx, wy = ty, wz = tz;
    }
    float cx = pb[0], cy = pb[1], cz = pb[2];
    if (cr == 0 && tid == 0) out[(size_t)cloud * npoint] = 0;
    cluster.sync();                                   // every CTA's exchange slots are cleared before any remote store
    const uint32_t exch_addr = pf::smem_u32(&exch[0][0]);

    // ---- rounds ----
    for (int s = 0; s + 1 < npoint; s++) {
        // (1) update the buckets whose box is closer to the new sample than their current maximum
        {
            const float ex = fmaxf(fmaxf(blx - cx, cx - bhx), 0.f), ey = fmaxf(fmaxf(bly - cy, cy - bhy), 0.f),
                        ez = fmaxf(fmaxf(blz - cz, cz - bhz), 0.f);
            const bool act = (ex * ex + ey * ey + ez * ez) * SAFE < bmr;
            unsigned mask = __ballot_sync(0xffffffffu, act);
            while (mask) {
                const int kk = __ffs(mask) - 1;
                mask &= mask - 1;
                const int p = 32 * (w + NW * kk) + lane;
                const float px = sx[p], py = sy[p], pz = sz[p];
                const unsigned po = so[p];
                const float d = sqdist(px, py, pz, cx, cy, cz);
                float m = md[p];
                if (d < m) m = d, md[p] = d;
                const float nm = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(m)));
                const unsigned o = (m == nm) ? po : 0xffffffffu;
                const unsigned omin = __reduce_min_sync(0xffffffffu, o);
                const int srcl = __ffs(__ballot_sync(0xffffffffu, o == omin)) - 1;
                const float tx = __shfl_sync(0xffffffffu, px, srcl), ty = __shfl_sync(0xffffffffu, py, srcl),
                            tz = __shfl_sync(0xffffffffu, pz, srcl);
                if (lane == kk) bmr = nm, wo = omin, wx = tx, wy = ty, wz = tz;
            }
        }
        // (2) warp candidate -> shared memory
        {
            const float Mw = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(bmr)));
            const unsigned cand = (bmr == Mw) ? wo : 0xffffffffu;
            const unsigned omin = __reduce_min_sync(0xffffffffu, cand);
            const int src = __ffs(__ballot_sync(0xffffffffu, cand == omin && bmr == Mw)) - 1;
            if (lane == src) wcand[w].m = Mw, wcand[w].x = wx, wcand[w].y = wy, wcand[w].z = wz, wcand[w].oidx = omin;
        }
        __syncthreads();
        // (3) warp 0: CTA candidate, posted into every CTA of the cluster (lane r -> CTA r) and signalled
        const int par = s & 1;
        if (w == 0) {
            const int l = min(lane, NW - 1);
            const float mw = wcand[l].m;
            const unsigned ow = wcand[l].oidx;
            const float M = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(mw)));
            const unsigned cand = (mw == M) ? ow : 0xffffffffu;
            const unsigned omin = __reduce_min_sync(0xffffffffu, cand);
            const int src = min(__ffs(__ballot_sync(0xffffffffu, cand == omin && mw == M)) - 1, NW - 1);
            const float bx = wcand[src].x, by = wcand[src].y, bz = wcand[src].z;
            if (lane < CS) {
                const uint32_t dst = mapa(exch_addr + (uint32_t)((par * MAXCS + cr) * sizeof(Cand)), (uint32_t)lane);
                const float tag = __uint_as_float((unsigned)(s + 1));
                st_cluster_v4(dst, M, bx, by, tag);
                st_cluster_v4(dst + 16, bz, __uint_as_float(omin), tag, 0.f);
            }
        }
        // (4) everybody: wait for the CS candidates of this round, pick the winner (max value, lowest original index)
        {
            const int l = min(lane, CS - 1);
            const unsigned want = (unsigned)(s + 1);
            uint4 a, b;
            uint32_t spins = 0;
            for (;;) {
                a = ld_volatile_v4(&exch[par][l]);
                b = ld_volatile_v4(reinterpret_cast<const uint8_t*>(&exch[par][l]) + 16);
                if (__all_sync(0xffffffffu, a.w == want && b.z == want)) break;
                if (++spins > (1u << 26)) __trap();
            }
            const float mw = __uint_as_float(a.x);
            const unsigned ow = b.y;
            const float M = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(mw)));
            const unsigned cand = (mw == M) ? ow : 0xffffffffu;
            const unsigned omin = __reduce_min_sync(0xffffffffu, cand);
            const int src = min(__ffs(__ballot_sync(0xffffffffu, cand == omin && mw == M)) - 1, CS - 1);
            cx = __uint_as_float(__shfl_sync(0xffffffffu, a.y, src)), cy = __uint_as_float(__shfl_sync(0xffffffffu, a.z, src));
            cz = __uint_as_float(__shfl_sync(0xffffffffu, b.x, src));
            if (cr == 0 && tid == 0) out[(size_t)cloud * npoint + s + 1] = (int32_t)omin;
        }
        // wcand is rewritten only after the next round's updates; warp 0 read it before posting, and nobody passes the
        // wait above before every CTA (hence this CTA's warp 0) has posted: no second block barrier is needed.
    }
    cluster.sync();                                   // no CTA exits while a sibling may still store into its shared memory
}

__global__ void zero_flag_kernel(int* f) { *f = 0; }

}  // namespace

// Launch configuration of the cluster kernel for one cloud per cluster of cs CTAs
static void cluster_cfg(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, int B, int cs, size_t smem, cudaStream_t st) {
    *cfg = cudaLaunchConfig_t{};
    cfg->gridDim = dim3((unsigned)(B * cs));
    cfg->blockDim = dim3(CT);
    cfg->dynamicSmemBytes = smem;
    cfg->stream = st;
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg->attrs = attr;
    cfg->numAttrs = 1;
}

// Cluster size for B clouds of n points: the smallest power of two whose share fits the per-CTA capacity with headroom.
// The clusters must be (nearly) all resident at once -- clusters cannot span GPCs, so far fewer fit than SMs / cs
// (measured on B200: 4 clusters of 16, 8 of 8) -- because a second wave doubles the time: measured 0.75-0.83 us per round
// in a cluster against 0.92 us (n <= 40,960) / 2.35 us (larger clouds: metadata in shared memory) in one CTA, so two waves
// still pay for large clouds, never for small ones.  0 = use the single-CTA kernel.
static int pick_cluster(int B, int n, int* cap) {
    for (int cs = 2; cs <= MAXCS; cs <<= 1) {
        const long long share = ((long long)n + cs - 1) / cs;
        const long long want = share + share / 4 + 256;                    // 25 % + a few cells of imbalance
        if (want > 8192) continue;
        *cap = (int)((want + 31) / 32 * 32);
        if (*cap < CELLS) *cap = CELLS;                                     // the set-up histogram aliases the running minima
        const size_t smem = (size_t)*cap * 20;
        if (pf_smem_optin(fps_cluster_kernel, smem)) return 0;
        if (cs > 8 && cudaFuncSetAttribute(fps_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
            cudaGetLastError();
            return 0;
        }
        cudaLaunchConfig_t cfg;
        cudaLaunchAttribute attr[1];
        cluster_cfg(&cfg, attr, B, cs, smem, nullptr);
        int active = 0;
        if (cudaOccupancyMaxActiveClusters(&active, fps_cluster_kernel, &cfg) != cudaSuccess || active < 1) {
            cudaGetLastError();
            return 0;
        }
        const int waves = (B + active - 1) / active;
        return waves <= (n > 40960 ? 2 : 1) ? cs : 0;
    }
    return 0;
}

extern "C" long long pf_fps_clustered_work_floats(int B, int n) { return pf_fps_work_floats(B, n) + 64; }

extern "C" int pf_fps_clustered(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, void* stream) {
    PF_CHECK_ARG(xyz && idx && work, "pf_fps_clustered: null pointer (work must hold pf_fps_clustered_work_floats(B, n) floats)");
    PF_CHECK_ARG(B >= 0 && n > 0 && npoint > 0, "pf_fps_clustered: bad sizes B=%d n=%d npoint=%d", B, n, npoint);
    if (B == 0) return 0;
    int cap = 0;
    const int cs = n >= 4096 ? pick_cluster(B, n, &cap) : 0;
    if (cs == 0) return pf_fps_bucketed(xyz, B, n, npoint, idx, work, stream);
    cudaStream_t st = (cudaStream_t)stream;
    int* flag = reinterpret_cast<int*>(work + pf_fps_work_floats(B, n));   // after the single-CTA kernel's scratch
    zero_flag_kernel<<<1, 1, 0, st>>>(flag);
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    cluster_cfg(&cfg, attr, B, cs, (size_t)cap * 20, st);
    PF_CUDA(cudaLaunchKernelEx(&cfg, fps_cluster_kernel, xyz, n, npoint, idx, cap, flag));
    if (int rc = pf_launch_status("fps_cluster_kernel")) return rc;
    // degenerate clouds (a CTA's cells exceed its capacity): the single-CTA kernel runs instead -- it starts, reads the flag
    // and returns at once otherwise (no host round trip, CUDA-graph capturable)
    return pf_fps_bucketed_if(xyz, B, n, npoint, idx, work, flag, stream);
}
