// Brute-force nearest-neighbour kernels of the PU-Flow x4 hot path (bit-exact index work).
//
//   pf_knn_topk    replaces pytorch3d.ops.knn_points at
//                  /root/reference/modules/discrete/interpflow.py:104,228,328
//   pf_chamfer_nn  replaces NmDistanceKernel + launcher,
//                  /root/reference/metric/PyTorchCD/chamfer3D/chamfer3D.cu:12-154
//
// Semantics: squared L2 by direct differences, ((dx*dx + dy*dy) + dz*dz) with every operation
// individually rounded (no FMA contraction), the K smallest by lexicographic (distance, index), ascending --
// bit-identical to oracle/nn_oracle.c.
//
// One thread per query, candidate tiles staged in shared memory by TMA bulk copies (cp.async.bulk + mbarrier, double
// buffered), transposed once per tile to x[] / y[] / z[] so that a warp-broadcast LDS.128 brings four candidates and the
// distance arithmetic runs on the packed-fp32 pipe (FADD2 / FFMA2 with a zero addend: two candidates per instruction,
// each lane rounded like the scalar operation).
//
// Top-K (K > 1) is register-resident and branch-free, in three phases (round 1 kept K (distance, index) pairs sorted by
// serial insertion, which the whole warp executes whenever any lane inserts: 88 instructions per candidate):
//   A  distances only: K candidates at a time are sorted by a Batcher odd-even merge network (FMNMX pairs) and merged
//      into the running sorted top-K with a bitonic half-cleaner: 13 instructions per candidate, no indices carried.
//      Result: the K smallest distances, exactly, as a sorted multiset; T = the K-th.
//   B  second sweep: candidates with d <= T (K of them plus ties at T) are appended in index order to a per-thread list in
//      shared memory (predicated stores, no divergence).
//   C  every list entry finds its output slot: #(top[k] < d) plus the number of equal distances already placed (index
//      order = tie order); ties at T beyond slot K - 1 drop out.  Entries below T always fit.
// A list that overflows (more than K + 16 candidates at distance <= T: lattices) takes the exact serial path for that query.
#include "common.cuh"
#include "puflow_b200.h"

namespace {

constexpr int QT = 128;     // queries (threads) per CTA
constexpr int CT = 512;     // candidates per shared-memory tile (6 KB raw x2 buffers + 6 KB transposed)
constexpr float INF = __builtin_huge_valf();
#ifndef PF_NN_GRID_MIN
#define PF_NN_GRID_MIN 256              // candidates per cloud from which pf_chamfer_nn prunes with a grid
#endif

// ---- packed fp32 (two candidates per instruction) ----
__device__ __forceinline__ uint64_t pk(float a, float b) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ void upk(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// v * v rounded once: an explicit FMA with a zero addend (a bare mul.rn.f32x2 next to an add is contracted by ptxas)
__device__ __forceinline__ uint64_t sq2(uint64_t v) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(r) : "l"(v), "l"(0ull));
    return r;
}
// squared distances of the query (qx2 = (qx, qx), ...) to two candidates (x2 = (x_c, x_c+1), ...)
__device__ __forceinline__ void dist2(uint64_t qx2, uint64_t qy2, uint64_t qz2, uint64_t x2, uint64_t y2, uint64_t z2, float& d0,
                                      float& d1) {
    const uint64_t dx = sub2(qx2, x2), dy = sub2(qy2, y2), dz = sub2(qz2, z2);
    upk(add2(add2(sq2(dx), sq2(dy)), sq2(dz)), d0, d1);
}

__device__ __forceinline__ void ce(float& a, float& b) {
    const float lo = fminf(a, b), hi = fmaxf(a, b);
    a = lo, b = hi;
}
// Batcher odd-even merge sort, ascending (63 compare-exchanges for 16 inputs).  One flat, fully unrolled loop per
// stage (P, KK) -- nested unrolled loops left a dynamically indexed remainder (the array went to local memory).
template <int K, int P, int KK>
struct OddEvenSort {
    static __device__ __forceinline__ void run(float (&v)[K]) {
#pragma unroll
        for (int a = 0; a < K; a++) {
            constexpr int J0 = KK % P;
            if (a >= J0 && (a - J0) % (2 * KK) < KK && a + KK < K && a / (2 * P) == (a + KK) / (2 * P)) ce(v[a], v[a + KK]);
        }
        if constexpr (KK > 1) OddEvenSort<K, P, KK / 2>::run(v);
        else if constexpr (2 * P < K) OddEvenSort<K, 2 * P, 2 * P>::run(v);
    }
};
template <int K>
__device__ __forceinline__ void sort_net(float (&v)[K]) {
    if constexpr (K > 1) OddEvenSort<K, 1, 1>::run(v);
}
template <int K, int S>
struct BitonicMerge {
    static __device__ __forceinline__ void run(float (&v)[K]) {
#pragma unroll
        for (int i = 0; i < K; i++)
            if ((i & S) == 0) ce(v[i], v[i + S]);
        if constexpr (S > 1) BitonicMerge<K, S / 2>::run(v);
    }
};
// top (sorted) <- the K smallest of top U batch (sorted), sorted: half-cleaner + bitonic merge
template <int K>
__device__ __forceinline__ void merge_low(float (&top)[K], const float (&batch)[K]) {
#pragma unroll
    for (int k = 0; k < K; k++) top[k] = fminf(top[k], batch[K - 1 - k]);
    if constexpr (K > 1) BitonicMerge<K, K / 2>::run(top);
}

// serial exact insertion (round-1 algorithm), used for K = 1 bookkeeping-free cases and as the overflow path
template <int K>
__device__ __forceinline__ void topk_insert(float (&bd)[K], int (&bi)[K], float d, int c) {
    // bd ascending.  new[k] = median(bd[k-1], d, bd[k]); equal distances keep the older entry first.
#pragma unroll
    for (int k = K - 1; k > 0; k--) {
        bool shift = d < bd[k - 1];
        bool here = d < bd[k];
        bi[k] = shift ? bi[k - 1] : (here ? c : bi[k]);
        bd[k] = shift ? bd[k - 1] : (here ? d : bd[k]);
    }
    if (d < bd[0]) {
        bd[0] = d;
        bi[0] = c;
    }
}

template <int K>
struct Smem {
    static constexpr int CAP = K == 1 ? 0 : K + 16;                       // list entries per query (phase B)
    static constexpr size_t RAW = 2 * CT * 3 * 4, SOA = 3 * CT * 4, LST = (size_t)CAP * QT * 8;
    static constexpr size_t BYTES = RAW + SOA + LST + 16;
};

// Tile pipeline shared by the sweeps: TMA bulk copy of tile t+1 while tile t is used; every tile is transposed into
// sx / sy / sz (padded with +inf-distance sentinels up to a multiple of 32).
struct Tiles {
    float* raw;                 // [2][CT * 3]
    float *sx, *sy, *sz;        // [CT]
    uint64_t* bar;              // [2]
    const float* pb;
    int np, ntiles;
    bool bulk_ok;
    uint32_t uses;              // tiles consumed so far (mbarrier parity across sweeps)

    __device__ void issue(int t) {      // one thread
        const uint32_t u = uses + (uint32_t)t;     // position of this tile in the sequence of all sweeps
        const uint32_t bytes = (uint32_t)min(CT, np - t * CT) * 12u;
        pf::mbar_expect_tx(&bar[u & 1], bytes);
        pf::bulk_g2s(raw + (u & 1) * CT * 3, pb + (size_t)t * CT * 3, bytes, &bar[u & 1]);
    }
    // makes tile t of the current sweep available in sx / sy / sz; returns its candidate count
    __device__ int acquire(int t) {
        const int c0 = t * CT, cnt = min(CT, np - c0);
        const uint32_t u = uses + (uint32_t)t;
        const float* src;
        if (bulk_ok) {
            if (threadIdx.x == 0 && t + 1 < ntiles) issue(t + 1);
            pf::mbar_wait(&bar[u & 1], (u >> 1) & 1);
            src = raw + (u & 1) * CT * 3;
        } else {
            src = pb + (size_t)c0 * 3;
        }
        __syncthreads();                                     // the previous tile's readers are done with sx / sy / sz
        const int cpad = (cnt + 31) & ~31;                 // sentinels up to the largest batch (K <= 32)
        for (int i = threadIdx.x; i < cpad; i += QT) {
            const bool ok = i < cnt;
            // sentinel: a coordinate so far away that its squared distance overflows to +inf (never below any real one)
            sx[i] = ok ? src[3 * i] : 3e38f, sy[i] = ok ? src[3 * i + 1] : 3e38f, sz[i] = ok ? src[3 * i + 2] : 3e38f;
        }
        __syncthreads();
        return cnt;
    }
    __device__ void end_sweep() { uses += (uint32_t)ntiles; }
};

template <int K>
__global__ void __launch_bounds__(QT) knn_kernel(const float* __restrict__ q, const float* __restrict__ p, int nq,
                                                 int np, int kout, int32_t* __restrict__ idx,
                                                 float* __restrict__ dist) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    using S = Smem<K>;
    __shared__ __align__(8) uint64_t bar[2];
    const int b = blockIdx.y;
    const int qi = blockIdx.x * QT + threadIdx.x;
    const bool live = qi < nq;

    Tiles T;
    T.raw = reinterpret_cast<float*>(smem_raw);
    T.sx = reinterpret_cast<float*>(smem_raw + S::RAW);
    T.sy = T.sx + CT, T.sz = T.sy + CT;
    T.bar = bar;
    T.pb = p + (size_t)b * np * 3;
    T.np = np, T.ntiles = (np + CT - 1) / CT;
    T.bulk_ok = ((reinterpret_cast<uintptr_t>(T.pb) & 15) == 0) && ((np & 3) == 0);
    T.uses = 0;

    float qx = 0.f, qy = 0.f, qz = 0.f;
    if (live) {
        const float* qq = q + ((size_t)b * nq + qi) * 3;
        qx = qq[0], qy = qq[1], qz = qq[2];
    }
    const uint64_t qx2 = pk(qx, qx), qy2 = pk(qy, qy), qz2 = pk(qz, qz);
    if (T.bulk_ok) {
        if (threadIdx.x == 0) {
            pf::mbar_init(&bar[0], 1);
            pf::mbar_init(&bar[1], 1);
            pf::mbar_fence_init();
        }
        __syncthreads();
        if (threadIdx.x == 0) T.issue(0);
    }
    // four candidates starting at c (c % 4 == 0) of the current tile
    auto dist4 = [&](int c, float (&d)[4]) {
        const float4 x = *reinterpret_cast<const float4*>(T.sx + c), y = *reinterpret_cast<const float4*>(T.sy + c),
                     z = *reinterpret_cast<const float4*>(T.sz + c);
        dist2(qx2, qy2, qz2, pk(x.x, x.y), pk(y.x, y.y), pk(z.x, z.y), d[0], d[1]);
        dist2(qx2, qy2, qz2, pk(x.z, x.w), pk(y.z, y.w), pk(z.z, z.w), d[2], d[3]);
    };
    const size_t o = ((size_t)b * nq + qi) * kout;

    if constexpr (K == 1) {
        float best = INF;
        int bi = -1;
        for (int t = 0; t < T.ntiles; t++) {
            const int cnt = T.acquire(t), c0 = t * CT;
#pragma unroll 2
            for (int c = 0; c < cnt; c += 4) {
                float d[4];
                dist4(c, d);
#pragma unroll
                for (int u = 0; u < 4; u++)
                    if (d[u] < best) best = d[u], bi = c0 + c + u;      // padding is +inf: never strictly below
            }
        }
        if (live) {
            idx[o] = bi;
            if (dist) dist[o] = best;
        }
        return;
    } else {
        constexpr int CAP = S::CAP;
        float* lst_d = reinterpret_cast<float*>(smem_raw + S::RAW + S::SOA);          // [CAP][QT]
        int* lst_i = reinterpret_cast<int*>(lst_d + CAP * QT);
        // ---- phase A: the K smallest distances ----
        float top[K];
#pragma unroll
        for (int k = 0; k < K; k++) top[k] = INF;
        for (int t = 0; t < T.ntiles; t++) {
            const int cnt = T.acquire(t);
            for (int c = 0; c < cnt; c += K) {               // the tail of the last batch reads sentinels (+inf)
                float batch[K];
#pragma unroll
                for (int u = 0; u < K; u += 4) {
                    float d[4];
                    dist4(c + u, d);
#pragma unroll
                    for (int v = 0; v < 4; v++) batch[u + v] = d[v];
                }
                sort_net<K>(batch);
                merge_low<K>(top, batch);
            }
        }
        T.end_sweep();
        const float thr = top[K - 1];
        // ---- phase B: candidates with d <= T, in index order ----
        int n = 0;
        if (T.bulk_ok && threadIdx.x == 0 && T.ntiles > 1) T.issue(0);      // (a single tile is still in place)
        const bool resident = T.ntiles == 1;
        for (int t = 0; t < T.ntiles; t++) {
            const int cnt = resident ? min(CT, np) : T.acquire(t), c0 = t * CT;
#pragma unroll 2
            for (int c = 0; c < cnt; c += 4) {
                float d[4];
                dist4(c, d);
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const bool take = d[u] <= thr && c + u < cnt;
                    if (take && n < CAP) lst_d[n * QT + threadIdx.x] = d[u], lst_i[n * QT + threadIdx.x] = c0 + c + u;
                    n += take ? 1 : 0;
                }
            }
        }
        if (!live) return;
        if (n <= CAP) {
            // ---- phase C: slots ----
            uint32_t used = 0;
            for (int j = 0; j < n; j++) {
                const float d = lst_d[j * QT + threadIdx.x];
                int r = 0;
#pragma unroll
                for (int k = 0; k < K; k++) r += top[k] < d ? 1 : 0;
                const int slot = r + __ffs(~(used >> r)) - 1;                // first free slot at or after r
                if (slot < K) {
                    used |= 1u << slot;
                    if (slot < kout) {
                        idx[o + slot] = lst_i[j * QT + threadIdx.x];
                        if (dist) dist[o + slot] = d;
                    }
                }
            }
        } else {
            // ---- overflow (more than K + 16 candidates at distance <= T): exact serial insertion straight from global ----
            float bd[K];
            int bi[K];
#pragma unroll
            for (int k = 0; k < K; k++) bd[k] = INF, bi[k] = -1;
            for (int c = 0; c < np; c++) {
                const float dx = __fsub_rn(qx, T.pb[3 * c]), dy = __fsub_rn(qy, T.pb[3 * c + 1]), dz = __fsub_rn(qz, T.pb[3 * c + 2]);
                const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                if (d < bd[K - 1]) topk_insert<K>(bd, bi, d, c);
            }
#pragma unroll
            for (int k = 0; k < K; k++)
                if (k < kout) {
                    idx[o + k] = bi[k];
                    if (dist) dist[o + k] = bd[k];
                }
        }
    }
}

template <int K>
int launch_knn(const float* q, const float* p, int B, int nq, int np, int kout, int32_t* idx, float* dist,
               cudaStream_t st) {
    dim3 grid((nq + QT - 1) / QT, B);
    auto kern = knn_kernel<K>;
    if (int rc = pf_smem_optin(kern, Smem<K>::BYTES)) return rc;
    kern<<<grid, QT, Smem<K>::BYTES, st>>>(q, p, nq, np, kout, idx, dist);
    return pf_launch_status("knn_kernel");
}

// ---------------------------------------------------------------------------------------------------------------------
// Grid-pruned exact nearest neighbour (K = 1) for the Chamfer pass: same result as knn_kernel<1>, bit for bit (squared
// distance by the same individually rounded operations, the lowest index among equal distances), from ~30 candidates per
// query instead of all np.
//
// Every CTA bins its cloud's np candidates into a G^3 grid over their bounding box in shared memory (counting sort: cell
// histogram by shared-memory atomics, block scan, scatter as (x, y, z, index) float4 -- the order inside a cell is whatever
// the atomics give and does not matter: the minimum is taken over (distance, index)).  Four adjacent lanes serve one query:
// they walk the cells of ring r = 1, 2, ... around the query's cell (Chebyshev distance <= 1, then == r) as rows of x-adjacent
// cells (contiguous in the sorted array), every lane a quarter of each row's points, and combine their (distance, index)
// minima by two shuffles per ring.  The search stops when the best
// distance is below the distance to the nearest face of the searched cube that has grid behind it (minus a rounding
// allowance: 1e-5 of the coordinates' scale on the gap and 1e-5 on its square, against ~1e-7 of rounding in the distances
// and in the binning), or when the cube covers the grid.  Everything outside the cube is provably farther than the
// incumbent, so neither the minimum nor the tie order can change.  A cloud whose bounding box is not finite takes a plain
// scan of all candidates in the same kernel.
constexpr int GT = 256;                        // threads per CTA: 64 four-lane query groups
#ifndef PF_NN_GQ
#define PF_NN_GQ 16
#endif
#ifndef PF_NN_G
#define PF_NN_G 16
#endif
constexpr int GQ = PF_NN_GQ;                   // queries per group -> 1,024 queries per CTA

template <int G>
struct GridSmem {
    static constexpr int NC = G * G * G;
    static size_t bytes(int np) { return (size_t)((np + 3) & ~3) * 16 + (size_t)(2 * NC + 4) * 4; }
};

template <int G>
__global__ void __launch_bounds__(GT) nn_grid_kernel(const float* __restrict__ q, const float* __restrict__ p, int nq, int np,
                                                     int32_t* __restrict__ idx, float* __restrict__ dist) {
    constexpr int NC = G * G * G, CPT = (NC + GT - 1) / GT;
    extern __shared__ __align__(16) uint8_t gsm[];
    float4* pts = reinterpret_cast<float4*>(gsm);
    int* cstart = reinterpret_cast<int*>(pts + ((np + 3) & ~3));      // [NC + 1]
    int* cur = cstart + NC + 4;                                      // [NC] histogram, then scatter cursors
    __shared__ float red[GT / 32][6];
    __shared__ float bb[10];                                         // min[3], cell size[3], 1 / cell size[3], rounding allowance
    __shared__ int wsum[GT / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, b = blockIdx.y;
    const float* pb = p + (size_t)b * np * 3;
    const float* qb = q + (size_t)b * nq * 3;

    // ---- bounding box ----
    float lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
    for (int i = tid; i < np; i += GT) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float v = __ldg(pb + 3 * i + a);
            lo[a] = fminf(lo[a], v), hi[a] = fmaxf(hi[a], v);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if (lane == 0) red[warp][a] = lo[a], red[warp][3 + a] = hi[a];
    }
    for (int i = tid; i < NC; i += GT) cur[i] = 0;
    __syncthreads();
    if (tid < 3) {
        float l = red[0][tid], h = red[0][3 + tid];
        for (int w = 1; w < GT / 32; w++) l = fminf(l, red[w][tid]), h = fmaxf(h, red[w][3 + tid]);
        const float ext = h - l;
        bb[tid] = l, bb[3 + tid] = ext / G, bb[6 + tid] = ext > 0.f ? G / ext : 0.f;
        red[0][tid] = ext;                                           // (read back below, after the barrier)
    }
    __syncthreads();
    const float ex = red[0][0], ey = red[0][1], ez = red[0][2];
    const bool finite_box = ex < INF && ey < INF && ez < INF && ex == ex && ey == ey && ez == ez;   // uniform
    const float mnx = bb[0], mny = bb[1], mnz = bb[2], hx = bb[3], hy = bb[4], hz = bb[5], ivx = bb[6], ivy = bb[7], ivz = bb[8];
    const float ext_max = fmaxf(ex, fmaxf(ey, ez));
    auto cell1 = [](float v, float mn, float iv) { return min(G - 1, max(0, (int)((v - mn) * iv))); };

    if (finite_box) {
        // ---- counting sort of the candidates by cell ----
        for (int i = tid; i < np; i += GT) {
            const int c = (cell1(__ldg(pb + 3 * i + 2), mnz, ivz) * G + cell1(__ldg(pb + 3 * i + 1), mny, ivy)) * G +
                          cell1(__ldg(pb + 3 * i), mnx, ivx);
            atomicAdd(cur + c, 1);
        }
        __syncthreads();
        int cnt[CPT], tot = 0;
#pragma unroll
        for (int k = 0; k < CPT; k++) {
            const int c = tid * CPT + k;
            cnt[k] = c < NC ? cur[c] : 0;
            tot += cnt[k];
        }
        int incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        int base = incl - tot;
        for (int w = 0; w < warp; w++) base += wsum[w];
#pragma unroll
        for (int k = 0; k < CPT; k++) {
            const int c = tid * CPT + k;
            if (c < NC) cstart[c] = base, cur[c] = base;
            base += cnt[k];
        }
        if (tid == GT - 1) cstart[NC] = base;
        __syncthreads();
        for (int i = tid; i < np; i += GT) {
            const float x = __ldg(pb + 3 * i), y = __ldg(pb + 3 * i + 1), z = __ldg(pb + 3 * i + 2);
            const int c = (cell1(z, mnz, ivz) * G + cell1(y, mny, ivy)) * G + cell1(x, mnx, ivx);
            pts[atomicAdd(cur + c, 1)] = make_float4(x, y, z, __int_as_float(i));
        }
        __syncthreads();
    }

    // ---- queries: four lanes each ----
    const int grp = tid >> 2, sub = tid & 3;
    const unsigned gmask = 0xFu << (lane & 28);
    auto better = [](float d, int i, float bd, int bi) { return d < bd || (d == bd && i < bi); };
#ifdef PF_NN_SKIPQ
    if (nq > 0) return;
#endif
    for (int k = 0; k < GQ; k++) {
        const int qi = (blockIdx.x * GQ + k) * (GT / 4) + grp;
        if (qi >= nq) break;
        const float qx = __ldg(qb + 3 * qi), qy = __ldg(qb + 3 * qi + 1), qz = __ldg(qb + 3 * qi + 2);
        float best = INF;
        int bi = -1;
        auto visit = [&](float x, float y, float z, int i) {
            const float dx = __fsub_rn(qx, x), dy = __fsub_rn(qy, y), dz = __fsub_rn(qz, z);
            const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            if (better(d, i, best, bi)) best = d, bi = i;
        };
        auto combine = [&]() {
#pragma unroll
            for (int o = 1; o < 4; o <<= 1) {
                const float od = __shfl_xor_sync(gmask, best, o);
                const int oi = __shfl_xor_sync(gmask, bi, o);
                if (better(od, oi, best, bi) && oi >= 0) best = od, bi = oi;
            }
        };
        if (!finite_box) {
            for (int s = sub; s < np; s += 4) visit(__ldg(pb + 3 * s), __ldg(pb + 3 * s + 1), __ldg(pb + 3 * s + 2), s);
            combine();
        } else {
            const int cx = cell1(qx, mnx, ivx), cy = cell1(qy, mny, ivy), cz = cell1(qz, mnz, ivz);
            const float eps = 1e-5f * (ext_max + fabsf(qx) + fabsf(qy) + fabsf(qz));
            auto scan = [&](int c0, int c1) {                          // the points of cells [c0, c1]: one contiguous range
                const int s1 = cstart[c1 + 1];
                for (int s = cstart[c0] + sub; s < s1; s += 4) {
                    const float4 v = pts[s];
                    visit(v.x, v.y, v.z, __float_as_int(v.w));
                }
            };
            // ring 1 first (the 3 x 3 x 3 block: the query's own cell alone almost never proves anything), as 9 rows of
            // x-adjacent cells -- a row's points are contiguous in the sorted array; ring r >= 2: full rows on the shell's
            // z / y faces, the two end cells of the others
            for (int r = 1;; r++) {
                const int x0 = max(cx - r, 0), x1 = min(cx + r, G - 1), y0 = max(cy - r, 0), y1 = min(cy + r, G - 1),
                          z0 = max(cz - r, 0), z1 = min(cz + r, G - 1);
                for (int z = z0; z <= z1; z++)
                    for (int y = y0; y <= y1; y++) {
                        const int row = (z * G + y) * G;
                        if (r == 1 || abs(z - cz) == r || abs(y - cy) == r) {
                            scan(row + x0, row + x1);
                        } else {
                            if (cx - r >= 0) scan(row + cx - r, row + cx - r);
                            if (cx + r <= G - 1) scan(row + cx + r, row + cx + r);
                        }
                    }
                combine();
                // (an axis of zero extent has all its points in cell 0 and nothing behind any face)
                const bool allx = (x0 == 0 && x1 == G - 1) || !(ex > 0.f), ally = (y0 == 0 && y1 == G - 1) || !(ey > 0.f),
                           allz = (z0 == 0 && z1 == G - 1) || !(ez > 0.f);
                if (allx && ally && allz) break;
                float gap = INF;                                       // to the nearest face with grid behind it
                if (!allx && x0 > 0) gap = fminf(gap, qx - (mnx + x0 * hx));
                if (!allx && x1 < G - 1) gap = fminf(gap, (mnx + (x1 + 1) * hx) - qx);
                if (!ally && y0 > 0) gap = fminf(gap, qy - (mny + y0 * hy));
                if (!ally && y1 < G - 1) gap = fminf(gap, (mny + (y1 + 1) * hy) - qy);
                if (!allz && z0 > 0) gap = fminf(gap, qz - (mnz + z0 * hz));
                if (!allz && z1 < G - 1) gap = fminf(gap, (mnz + (z1 + 1) * hz) - qz);
                gap -= eps;
                if (gap > 0.f && best <= gap * gap * (1.f - 1e-5f)) break;
            }
        }
        if (sub == 0) {
            idx[(size_t)b * nq + qi] = bi;
            if (dist) dist[(size_t)b * nq + qi] = best;
        }
    }
}

template <int G>
int launch_nn_grid(const float* q, const float* p, int B, int nq, int np, int32_t* idx, float* dist, cudaStream_t st) {
    const int qpc = GQ * (GT / 4);
    dim3 grid((nq + qpc - 1) / qpc, B);
    auto kern = nn_grid_kernel<G>;
    const size_t smem = GridSmem<G>::bytes(np);
    if (int rc = pf_smem_optin(kern, smem)) return rc;
    kern<<<grid, GT, smem, st>>>(q, p, nq, np, idx, dist);
    return pf_launch_status("nn_grid_kernel");
}

// nearest candidate of every query: the grid kernel where it pays (enough candidates to prune, enough queries to amortise
// the binning every CTA repeats), the brute-force kernel otherwise
int launch_nn1(const float* q, const float* p, int B, int nq, int np, int32_t* idx, float* dist, cudaStream_t st) {
    if (np >= PF_NN_GRID_MIN && np <= 8192 && nq >= 1024)
        return np >= 1024 ? launch_nn_grid<PF_NN_G>(q, p, B, nq, np, idx, dist, st) : launch_nn_grid<8>(q, p, B, nq, np, idx, dist, st);
    return launch_knn<1>(q, p, B, nq, np, 1, idx, dist, st);
}

}  // namespace

extern "C" int pf_knn_topk(const float* q, const float* p, int B, int nq, int np, int K, int32_t* idx, float* dist,
                           void* stream) {
    PF_CHECK_ARG(q && p && idx, "pf_knn_topk: null pointer");
    PF_CHECK_ARG(B >= 0 && nq >= 0 && np > 0, "pf_knn_topk: bad sizes B=%d nq=%d np=%d", B, nq, np);
    PF_CHECK_ARG(K >= 1 && K <= 32, "pf_knn_topk: K=%d outside [1, 32] (patch extraction uses pf_knn_patch)", K);
    PF_CHECK_ARG(K <= np, "pf_knn_topk: K=%d > np=%d", K, np);
    PF_CHECK_ARG(B <= 65535, "pf_knn_topk: B=%d > 65535 (split the batch)", B);
    if (B == 0 || nq == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (K == 1) return launch_knn<1>(q, p, B, nq, np, K, idx, dist, st);
    if (K <= 4) return launch_knn<4>(q, p, B, nq, np, K, idx, dist, st);
    if (K <= 8) return launch_knn<8>(q, p, B, nq, np, K, idx, dist, st);
    if (K <= 16) return launch_knn<16>(q, p, B, nq, np, K, idx, dist, st);
    return launch_knn<32>(q, p, B, nq, np, K, idx, dist, st);
}

extern "C" int pf_chamfer_nn(const float* a, int n, const float* b, int m, int B, float* dist1, int32_t* idx1,
                             float* dist2, int32_t* idx2, void* stream) {
    PF_CHECK_ARG(a && b && dist1 && idx1 && (!dist2 == !idx2), "pf_chamfer_nn: null pointer");
    PF_CHECK_ARG(B >= 0 && n > 0 && m > 0 && B <= 65535, "pf_chamfer_nn: bad sizes B=%d n=%d m=%d", B, n, m);
    if (B == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int rc = launch_nn1(a, b, B, n, m, idx1, dist1, st);
    if (rc || !dist2) return rc;          // dist2 == idx2 == NULL: a -> b direction only (what remove_outliers consumes)
    return launch_nn1(b, a, B, m, n, idx2, dist2, st);
}
