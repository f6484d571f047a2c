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
    int rc = launch_knn<1>(a, b, B, n, m, 1, idx1, dist1, st);
    if (rc || !dist2) return rc;          // dist2 == idx2 == NULL: a -> b direction only (what remove_outliers consumes)
    return launch_knn<1>(b, a, B, m, n, 1, idx2, dist2, st);
}
