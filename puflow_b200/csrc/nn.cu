// Brute-force nearest-neighbour kernels of the PU-Flow x4 hot path (bit-exact index work).
//
//   pf_knn_topk    replaces pytorch3d.ops.knn_points at
//                  /root/reference/modules/discrete/interpflow.py:104,228,328
//   pf_chamfer_nn  replaces NmDistanceKernel + launcher,
//                  /root/reference/metric/PyTorchCD/chamfer3D/chamfer3D.cu:12-154
//
// One thread per query keeps its K best (distance, index) pairs sorted in registers; candidate
// tiles are staged in shared memory by TMA bulk copies (cp.async.bulk + mbarrier, double buffered)
// and read back as warp-wide broadcasts.  The squared distance is ((dx*dx + dy*dy) + dz*dz) with
// every operation individually rounded (__f*_rn: no FMA contraction), candidates are visited in
// index order and only a strictly smaller distance displaces an entry, so the result is the
// lexicographic (distance, index) order -- bit-identical to oracle/nn_oracle.c.
#include "common.cuh"
#include "puflow_b200.h"

namespace {

constexpr int QT = 128;     // queries (threads) per CTA
constexpr int CT = 1024;    // candidates per shared-memory tile (12 KB, x2 buffers)

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
__global__ void __launch_bounds__(QT) knn_kernel(const float* __restrict__ q, const float* __restrict__ p, int nq,
                                                 int np, int kout, int32_t* __restrict__ idx,
                                                 float* __restrict__ dist) {
    __shared__ __align__(16) float tile[2][CT * 3];
    __shared__ __align__(8) uint64_t bar[2];

    const int b = blockIdx.y;
    const int qi = blockIdx.x * QT + threadIdx.x;
    const bool live = qi < nq;
    const float* pb = p + (size_t)b * np * 3;
    const bool bulk_ok = ((reinterpret_cast<uintptr_t>(pb) & 15) == 0) && ((np & 3) == 0);
    const int ntiles = (np + CT - 1) / CT;

    float qx = 0.f, qy = 0.f, qz = 0.f;
    if (live) {
        const float* qq = q + ((size_t)b * nq + qi) * 3;
        qx = qq[0], qy = qq[1], qz = qq[2];
    }
    float bd[K];
    int bi[K];
#pragma unroll
    for (int k = 0; k < K; k++) bd[k] = __int_as_float(0x7f800000), bi[k] = -1;

    if (bulk_ok) {
        if (threadIdx.x == 0) {
            pf::mbar_init(&bar[0], 1);
            pf::mbar_init(&bar[1], 1);
            pf::mbar_fence_init();
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t bytes = (uint32_t)min(CT, np) * 12u;
            pf::mbar_expect_tx(&bar[0], bytes);
            pf::bulk_g2s(tile[0], pb, bytes, &bar[0]);
        }
    }
    for (int t = 0; t < ntiles; t++) {
        const int c0 = t * CT;
        const int cnt = min(CT, np - c0);
        const float* s = tile[t & 1];
        if (bulk_ok) {
            if (threadIdx.x == 0 && t + 1 < ntiles) {      // prefetch the next tile into the other buffer
                uint32_t bytes = (uint32_t)min(CT, np - (c0 + CT)) * 12u;
                pf::mbar_expect_tx(&bar[(t + 1) & 1], bytes);
                pf::bulk_g2s(tile[(t + 1) & 1], pb + (size_t)(c0 + CT) * 3, bytes, &bar[(t + 1) & 1]);
            }
            pf::mbar_wait(&bar[t & 1], (t >> 1) & 1);
        } else {
            float* sw = tile[t & 1];
            for (int i = threadIdx.x; i < cnt * 3; i += QT) sw[i] = pb[(size_t)c0 * 3 + i];
            __syncthreads();
        }
        if (live) {
#pragma unroll 4
            for (int c = 0; c < cnt; c++) {
                float dx = __fsub_rn(qx, s[3 * c]), dy = __fsub_rn(qy, s[3 * c + 1]), dz = __fsub_rn(qz, s[3 * c + 2]);
                float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                if (d < bd[K - 1]) topk_insert<K>(bd, bi, d, c0 + c);
            }
        }
        __syncthreads();      // everyone is done with tile[t & 1] before it is refilled
    }
    if (live) {
        size_t o = ((size_t)b * nq + qi) * kout;
#pragma unroll
        for (int k = 0; k < K; k++)
            if (k < kout) {
                idx[o + k] = bi[k];
                if (dist) dist[o + k] = bd[k];
            }
    }
}

template <int K>
int launch_knn(const float* q, const float* p, int B, int nq, int np, int kout, int32_t* idx, float* dist,
               cudaStream_t st) {
    dim3 grid((nq + QT - 1) / QT, B);
    knn_kernel<K><<<grid, QT, 0, st>>>(q, p, nq, np, kout, idx, dist);
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
