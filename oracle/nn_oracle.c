/*
 * nn_oracle.c — CPU restatement of the index-producing primitives of the PU-Flow x4 hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker / the timed CPU baseline.
 * Nothing under puflow_b200/ or modules/ links or calls it.
 *
 * Compile WITHOUT floating-point contraction (see oracle/Makefile: -ffp-contract=off) so that every
 * distance is ((dx*dx + dy*dy) + dz*dz) with each operation rounded to fp32 — the expression the
 * CUDA kernels reproduce with __fmul_rn/__fadd_rn, which is what makes indices bit-exact.
 *
 * Parity status:
 *  - oracle_knn: the reference calls pytorch3d.ops.knn_points
 *    (/root/reference/modules/discrete/interpflow.py:104,228,328), an un-vendored third-party CUDA
 *    package pinned only as `pytorch3d@stable` (/root/reference/docker/Dockerfile:50).  Its published
 *    algorithm (brute force, K smallest squared L2) is restated with the tie rule SURVEY.md §8(c)
 *    fixes: lexicographic (distance, index), output sorted ascending.  Pinned against
 *    tests/golden/net_pu1k.npz (reference model run with that stand-in) — the pytorch3d kernel
 *    itself is absent, so at THAT boundary parity is unpinned.
 *  - oracle_chamfer_nn: follows nnsearch() in
 *    /root/reference/evaluation/tf_ops/nn_distance/tf_nndistance.cpp:21-43 and NmDistanceKernel in
 *    /root/reference/metric/PyTorchCD/chamfer3D/chamfer3D.cu:12-134 (first index wins ties: strict <).
 *    Pinned against tests/golden/chamfer_unit.npz = the reference's own known-answer test
 *    (/root/reference/metric/PyTorchCD/unit_test.py:14-35: MSE < 1e-8, indices identical).
 *  - oracle_fps: follows farthest_point_sampling in /root/reference/modules/utils/fps.py:40-73
 *    (start at index 0, running min of squared distance initialised to 1e10, arg-max with the
 *    first index winning ties).  Pinned against tests/golden/e2e_*.npz (seed_idx).
 *  - oracle_knn_query: knn_cuda.KNN stand-in semantics (sqrt distances ascending, ties -> lower
 *    index), call site /root/reference/modules/utils/patch.py:33,107.  Third-party, un-vendored.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

static inline float sqdist3(const float *a, const float *b) {
    float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    float xx = dx * dx, yy = dy * dy, zz = dz * dz;
    float s = xx + yy;
    return s + zz;
}

/* K nearest of p[b, np, 3] for each query q[b, nq, 3]; idx int32 [b, nq, K], dist fp32 [b, nq, K]
 * (dist may be NULL).  Requires K <= np. */
int oracle_knn(int b, int nq, int np, int K, const float *q, const float *p, int32_t *idx, float *dist) {
    if (K > np || K <= 0) return 1;
    float *bd = (float *)malloc(sizeof(float) * (size_t)K);
    int32_t *bi = (int32_t *)malloc(sizeof(int32_t) * (size_t)K);
    for (int i = 0; i < b; i++) {
        for (int j = 0; j < nq; j++) {
            const float *qq = q + ((size_t)i * nq + j) * 3;
            int cnt = 0;
            for (int c = 0; c < np; c++) {
                float d = sqdist3(qq, p + ((size_t)i * np + c) * 3);
                if (cnt == K && !(d < bd[K - 1])) continue;     /* equal distance, larger index: loses */
                int pos = cnt < K ? cnt : K - 1;
                while (pos > 0 && d < bd[pos - 1]) { bd[pos] = bd[pos - 1]; bi[pos] = bi[pos - 1]; pos--; }
                bd[pos] = d; bi[pos] = c;
                if (cnt < K) cnt++;
            }
            for (int k = 0; k < K; k++) {
                idx[((size_t)i * nq + j) * K + k] = bi[k];
                if (dist) dist[((size_t)i * nq + j) * K + k] = bd[k];
            }
        }
    }
    free(bd); free(bi);
    return 0;
}

/* nearest neighbour in xyz2[b, m, 3] of each point of xyz1[b, n, 3]: squared distance + argmin. */
int oracle_chamfer_nn(int b, int n, int m, const float *xyz1, const float *xyz2, float *dist, int32_t *idx) {
    for (int i = 0; i < b; i++)
        for (int j = 0; j < n; j++) {
            const float *a = xyz1 + ((size_t)i * n + j) * 3;
            float best = 0.f; int besti = 0;
            for (int k = 0; k < m; k++) {
                float d = sqdist3(xyz2 + ((size_t)i * m + k) * 3, a);
                if (k == 0 || d < best) { best = d; besti = k; }
            }
            dist[(size_t)i * n + j] = best; idx[(size_t)i * n + j] = besti;
        }
    return 0;
}

/* farthest point sampling, xyz[b, n, 3] -> idx int32 [b, npoint] */
int oracle_fps(int b, int n, int npoint, const float *xyz, int32_t *idx) {
    float *mind = (float *)malloc(sizeof(float) * (size_t)n);
    for (int i = 0; i < b; i++) {
        const float *p = xyz + (size_t)i * n * 3;
        for (int j = 0; j < n; j++) mind[j] = 1e10f;
        int far = 0;
        for (int s = 0; s < npoint; s++) {
            idx[(size_t)i * npoint + s] = far;
            const float *c = p + (size_t)far * 3;
            float best = -1.f; int besti = 0;
            for (int j = 0; j < n; j++) {
                float d = sqdist3(p + (size_t)j * 3, c);
                if (d < mind[j]) mind[j] = d;
                if (mind[j] > best) { best = mind[j]; besti = j; }
            }
            far = besti;
        }
    }
    free(mind);
    return 0;
}
