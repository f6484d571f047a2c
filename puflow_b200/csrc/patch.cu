// Caller-side primitives of PatchHelper ("next" rows): farthest point sampling
// and the k=256 patch-extraction kNN.
//
//   pf_fps        replaces pointnet2_ops.furthest_point_sample at
//                 /root/reference/modules/utils/patch.py:102 (seeds) and :156 (merge: 8,216 of 40,960),
//                 with the semantics of the reference's own farthest_point_sampling,
//                 /root/reference/modules/utils/fps.py:40-73: start at index 0, running min of the
//                 squared distance initialised to 1e10, arg-max, lowest index wins ties.
//   pf_knn_patch  replaces knn_cuda.KNN(k)(ref, query) at /root/reference/modules/utils/patch.py:33,107.
//
// FPS is npoint strictly dependent arg-max rounds, so the design goal is round latency: one thread
// block CLUSTER (8 CTAs x 512 threads) per cloud keeps every point and its running min distance in
// registers; a round is: register update -> warp shuffle arg-max -> CTA arg-max in shared memory ->
// each CTA posts its candidate (key + coordinates) into all 8 CTAs' shared memory through DSMEM ->
// one cluster barrier -> every thread picks the winner locally.  No global memory on the round path
// except the 4-byte index store.
#include <cooperative_groups.h>

#include "common.cuh"
#include "puflow_b200.h"

namespace cg = cooperative_groups;

namespace {

constexpr int CS = 8;        // CTAs per cluster (portable maximum)
constexpr int FT = 512;      // threads per CTA
constexpr int NW = FT / 32;

struct Cand {
    float val, x, y, z;
    int idx, pad0, pad1, pad2;
};

__device__ __forceinline__ bool better(float av, int ai, float bv, int bi) {
    return av > bv || (av == bv && ai < bi);
}

template <int PPT>
__global__ void __launch_bounds__(FT, 1) fps_kernel(const float* __restrict__ xyz, int n, int npoint,
                                                    int32_t* __restrict__ out) {
    cg::cluster_group cluster = cg::this_cluster();
    const int cr = (int)cluster.block_rank();
    const int b = blockIdx.x / CS;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

    __shared__ __align__(16) Cand exch[2][CS];      // candidates of all CTAs, double buffered by round parity
    __shared__ float wval[NW];
    __shared__ int widx[NW];
    __shared__ int blk_best;

    const float* pb = xyz + (size_t)b * n * 3;
    float px[PPT], py[PPT], pz[PPT], md[PPT];
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        int i = k * (CS * FT) + cr * FT + tid;
        bool ok = i < n;
        px[k] = ok ? pb[(size_t)i * 3] : 0.f;
        py[k] = ok ? pb[(size_t)i * 3 + 1] : 0.f;
        pz[k] = ok ? pb[(size_t)i * 3 + 2] : 0.f;
        md[k] = ok ? 1e10f : -1.f;
    }
    int far = 0;
    float cx = pb[0], cy = pb[1], cz = pb[2];
    cluster.sync();      // all CTAs of the cluster are resident before any DSMEM traffic

    for (int s = 0; s < npoint; s++) {
        if (cr == 0 && tid == 0) out[(size_t)b * npoint + s] = far;
        if (s + 1 == npoint) break;
        // (1) update running min distances, thread-local arg-max
        float bv = -2.f;
        int bi = 0x7fffffff;
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            float dx = __fsub_rn(px[k], cx), dy = __fsub_rn(py[k], cy), dz = __fsub_rn(pz[k], cz);
            float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            float m = md[k];
            m = (m >= 0.f && d < m) ? d : m;
            md[k] = m;
            int i = k * (CS * FT) + cr * FT + tid;
            if (better(m, i, bv, bi)) bv = m, bi = i;
        }
        // (2) warp arg-max
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, bv, off);
            int oi = __shfl_xor_sync(0xffffffffu, bi, off);
            if (better(ov, oi, bv, bi)) bv = ov, bi = oi;
        }
        if (lane == 0) wval[wid] = bv, widx[wid] = bi;
        __syncthreads();
        // (3) CTA arg-max (warp 0), winner index broadcast through shared memory
        if (wid == 0) {
            float v = lane < NW ? wval[lane] : -2.f;
            int i = lane < NW ? widx[lane] : 0x7fffffff;
#pragma unroll
            for (int off = 8; off > 0; off >>= 1) {
                float ov = __shfl_xor_sync(0xffffffffu, v, off);
                int oi = __shfl_xor_sync(0xffffffffu, i, off);
                if (better(ov, oi, v, i)) v = ov, i = oi;
            }
            if (lane == 0) blk_best = i;
        }
        __syncthreads();
        // (4) the owner of the CTA's winner posts (value, coordinates, index) to every CTA of the cluster
        {
            const int w = blk_best;
            if ((w % FT) == tid && ((w / FT) % CS) == cr && w < 0x7fffffff) {
                const int kk = w / (CS * FT);
                Cand c;
                c.val = -2.f, c.x = c.y = c.z = 0.f, c.idx = w, c.pad0 = c.pad1 = c.pad2 = 0;
#pragma unroll
                for (int k = 0; k < PPT; k++)
                    if (k == kk) c.val = md[k], c.x = px[k], c.y = py[k], c.z = pz[k];
#pragma unroll
                for (int r = 0; r < CS; r++) {
                    Cand* dst = cluster.map_shared_rank(&exch[s & 1][cr], r);
                    *reinterpret_cast<float4*>(dst) = make_float4(c.val, c.x, c.y, c.z);
                    *(reinterpret_cast<int4*>(dst) + 1) = make_int4(c.idx, 0, 0, 0);
                }
            }
        }
        cluster.sync();
        // (5) every thread picks the cluster-wide winner
        {
            float v = -3.f;
            int i = 0x7fffffff;
            float nx = 0.f, ny = 0.f, nz = 0.f;
#pragma unroll
            for (int r = 0; r < CS; r++) {
                float4 a = *reinterpret_cast<const float4*>(&exch[s & 1][r]);
                int ci = exch[s & 1][r].idx;
                if (better(a.x, ci, v, i)) v = a.x, i = ci, nx = a.y, ny = a.z, nz = a.w;
            }
            far = i, cx = nx, cy = ny, cz = nz;
        }
    }
    cluster.sync();      // no CTA exits while a sibling may still write into its shared memory
}

template <int PPT>
int launch_fps(const float* xyz, int B, int n, int npoint, int32_t* idx, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(B * CS);
    cfg.blockDim = dim3(FT);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CS;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    PF_CUDA(cudaLaunchKernelEx(&cfg, fps_kernel<PPT>, xyz, n, npoint, idx));
    return pf_launch_status("fps_kernel");
}

// ---- the same sampler, one CTA per cloud: batches of MANY small clouds (the 32-seed FPS of a 128-cloud batch) ----------
// 128 clusters of 8 CTAs do not fit the GPU at once (two waves, and a cluster barrier per round): 0.26 ms for 32 rounds.  With
// one 512-thread CTA per cloud the whole batch is one wave and a round is: register update -> warp arg-max -> one block
// barrier -> every warp reduces the 16 warp candidates -> the winner's coordinates from the shared-memory copy of the cloud.
// Same arithmetic, same (value, lowest index) order: bit-identical indices.
template <int PPT>
__global__ void __launch_bounds__(FT, 1) fps_cta_kernel(const float* __restrict__ xyz, int n, int npoint,
                                                        int32_t* __restrict__ out) {
    extern __shared__ __align__(16) float pts[];           // [n][3]
    __shared__ float wval[2][NW];
    __shared__ int widx[2][NW];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const float* pb = xyz + (size_t)b * n * 3;
    for (int i = tid; i < n * 3; i += FT) pts[i] = pb[i];
    __syncthreads();
    float px[PPT], py[PPT], pz[PPT], md[PPT];
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        const int i = k * FT + tid;
        const bool ok = i < n;
        px[k] = ok ? pts[i * 3] : 0.f, py[k] = ok ? pts[i * 3 + 1] : 0.f, pz[k] = ok ? pts[i * 3 + 2] : 0.f;
        md[k] = ok ? 1e10f : -1.f;
    }
    int far = 0;
    float cx = pts[0], cy = pts[1], cz = pts[2];
    for (int s = 0; s < npoint; s++) {
        if (tid == 0) out[(size_t)b * npoint + s] = far;
        if (s + 1 == npoint) break;
        float bv = -2.f;
        int bi = 0x7fffffff;
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            float dx = __fsub_rn(px[k], cx), dy = __fsub_rn(py[k], cy), dz = __fsub_rn(pz[k], cz);
            float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            float m = md[k];
            m = (m >= 0.f && d < m) ? d : m;
            md[k] = m;
            const int i = k * FT + tid;
            if (better(m, i, bv, bi)) bv = m, bi = i;
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, bv, off);
            int oi = __shfl_xor_sync(0xffffffffu, bi, off);
            if (better(ov, oi, bv, bi)) bv = ov, bi = oi;
        }
        const int par = s & 1;                              // double buffered: one barrier per round
        if (lane == 0) wval[par][wid] = bv, widx[par][wid] = bi;
        __syncthreads();
        float v = lane < NW ? wval[par][lane] : -3.f;
        int i = lane < NW ? widx[par][lane] : 0x7fffffff;
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, v, off);
            int oi = __shfl_xor_sync(0xffffffffu, i, off);
            if (better(ov, oi, v, i)) v = ov, i = oi;
        }
        far = __shfl_sync(0xffffffffu, i, 0);
        cx = pts[far * 3], cy = pts[far * 3 + 1], cz = pts[far * 3 + 2];
    }
}

template <int PPT>
int launch_fps_cta(const float* xyz, int B, int n, int npoint, int32_t* idx, cudaStream_t st) {
    const size_t smem = (size_t)n * 12;
    if (int rc = pf_smem_optin(fps_cta_kernel<PPT>, smem)) return rc;
    fps_cta_kernel<PPT><<<B, FT, smem, st>>>(xyz, n, npoint, idx);
    return pf_launch_status("fps_cta_kernel");
}

// ---- k nearest of np points for a handful of seed queries: radix select, then a sort of the selected keys only ----
// key = (float bits of the squared distance) << 32 | index : unsigned order == (distance, index) order.
// Round 1 sorted all np keys (np = 2,048: 66 block-wide compare-exchange stages of a 1,024-thread CTA, 0.42 ms for the 4,096
// seeds of config 2 -- the barriers, not the compares, were the cost).  Now: (1) four 8-bit histogram passes find T = the
// k-th smallest DISTANCE; (2) the keys with distance <= T (k of them plus ties at T) are compacted; (3) only those are sorted
// (k = 256: 36 stages of 128 pairs) and the first k written.  More than SEL_CAP keys at or below T (massively duplicated
// points) falls back to the full sort.  Same (distance, index) order: bit-identical indices.
constexpr int ST = 256, SEL_CAP = 1024;

__device__ __forceinline__ void bitonic_sort_block(unsigned long long* keys, int n) {      // n a power of two, whole CTA
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = threadIdx.x; i < n / 2; i += ST) {
                int lo = 2 * i - (i & (stride - 1));           // index with bit `stride` clear
                int hi = lo + stride;
                bool up = (lo & size) == 0;
                unsigned long long a = keys[lo], c = keys[hi];
                if ((a > c) == up) keys[lo] = c, keys[hi] = a;
            }
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(ST) knn_patch_kernel(const float* __restrict__ q, const float* __restrict__ p,
                                                       int nq, int np, int npad, int k, int32_t* __restrict__ idx) {
    extern __shared__ __align__(16) unsigned long long keys[];          // [npad] | sel[SEL_CAP]
    unsigned long long* sel = keys + npad;
    __shared__ unsigned hist[256];
    __shared__ unsigned s_bin, s_rank, s_count;
    const int b = blockIdx.y, qi = blockIdx.x, tid = threadIdx.x;
    const float* qq = q + ((size_t)b * nq + qi) * 3;
    const float qx = qq[0], qy = qq[1], qz = qq[2];
    const float* pb = p + (size_t)b * np * 3;
    for (int i = tid; i < npad; i += ST) {
        unsigned long long key = ~0ull;
        if (i < np) {
            float dx = __fsub_rn(qx, pb[(size_t)i * 3]), dy = __fsub_rn(qy, pb[(size_t)i * 3 + 1]),
                  dz = __fsub_rn(qz, pb[(size_t)i * 3 + 2]);
            float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            key = ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)i;
        }
        keys[i] = key;
    }
    // (1) T = the k-th smallest distance (as its bit pattern: non-negative floats order like unsigned integers)
    unsigned prefix = 0, mask = 0, rank = (unsigned)k;
    for (int shift = 24; shift >= 0; shift -= 8) {
        hist[tid] = 0;                                                   // ST == 256 bins
        if (tid == 0) s_count = 0;
        __syncthreads();                                                 // (also: keys[] written)
        for (int i = tid; i < np; i += ST) {
            const unsigned d = (unsigned)(keys[i] >> 32);
            if ((d & mask) == prefix) atomicAdd(&hist[(d >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (tid < 32) {                                                  // bin in which the cumulative count reaches `rank`
            unsigned c[8], tot = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) c[j] = hist[8 * tid + j], tot += c[j];
            unsigned incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
                if (tid >= o) incl += v;
            }
            unsigned before = incl - tot;
            if (before < rank && rank <= incl) {
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (before < rank && rank <= before + c[j]) s_bin = 8u * tid + j, s_rank = rank - before;
                    before += c[j];
                }
            }
        }
        __syncthreads();
        prefix |= s_bin << shift, mask |= 255u << shift, rank = s_rank;
    }
    // (2) every key at or below T, in any order
    for (int i = tid; i < np; i += ST) {
        const unsigned long long key = keys[i];
        if ((unsigned)(key >> 32) <= prefix) {
            const unsigned slot = atomicAdd(&s_count, 1u);
            if (slot < (unsigned)SEL_CAP) sel[slot] = key;
        }
    }
    __syncthreads();
    const unsigned total = s_count;
    int32_t* o = idx + ((size_t)b * nq + qi) * k;
    if (total <= (unsigned)SEL_CAP) {
        // (3) sort the selected keys alone
        int P = 2;
        while (P < (int)total) P <<= 1;
        for (int i = (int)total + tid; i < P; i += ST) sel[i] = ~0ull;
        __syncthreads();
        bitonic_sort_block(sel, P);
        for (int i = tid; i < k; i += ST) o[i] = (int32_t)(sel[i] & 0xffffffffu);
    } else {
        bitonic_sort_block(keys, npad);
        for (int i = tid; i < k; i += ST) o[i] = (int32_t)(keys[i] & 0xffffffffu);
    }
}

}  // namespace

extern "C" int pf_fps(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, void* stream) {
    if (work != nullptr && n >= 4096) return pf_fps_bucketed(xyz, B, n, npoint, idx, work, stream);
    PF_CHECK_ARG(xyz && idx, "pf_fps: null pointer");
    PF_CHECK_ARG(B >= 0 && n > 0 && npoint > 0, "pf_fps: bad sizes B=%d n=%d npoint=%d", B, n, npoint);
    PF_CHECK_ARG((long long)B * CS < 0x7fffffffLL, "pf_fps: B too large");
    if (B == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    // many small clouds: one CTA per cloud (a single wave) instead of one 8-CTA cluster per cloud (several waves)
    if (B * CS > 2 * 148 && n <= 16 * FT) {
        const int pc = (n + FT - 1) / FT;
        if (pc <= 1) return launch_fps_cta<1>(xyz, B, n, npoint, idx, st);
        if (pc <= 2) return launch_fps_cta<2>(xyz, B, n, npoint, idx, st);
        if (pc <= 4) return launch_fps_cta<4>(xyz, B, n, npoint, idx, st);
        if (pc <= 8) return launch_fps_cta<8>(xyz, B, n, npoint, idx, st);
        return launch_fps_cta<16>(xyz, B, n, npoint, idx, st);
    }
    const int per = (n + CS * FT - 1) / (CS * FT);
    if (per <= 1) return launch_fps<1>(xyz, B, n, npoint, idx, st);
    if (per <= 2) return launch_fps<2>(xyz, B, n, npoint, idx, st);
    if (per <= 4) return launch_fps<4>(xyz, B, n, npoint, idx, st);
    if (per <= 6) return launch_fps<6>(xyz, B, n, npoint, idx, st);
    if (per <= 10) return launch_fps<10>(xyz, B, n, npoint, idx, st);
    if (per <= 16) return launch_fps<16>(xyz, B, n, npoint, idx, st);
    if (per <= 25) return launch_fps<25>(xyz, B, n, npoint, idx, st);
    pf_set_error("pf_fps: n=%d exceeds the register-resident capacity (%d points per cloud)", n, 25 * CS * FT);
    return PF_EINVAL;
}

extern "C" int pf_knn_patch(const float* q, const float* p, int B, int nq, int np, int k, int32_t* idx,
                            void* stream) {
    PF_CHECK_ARG(q && p && idx, "pf_knn_patch: null pointer");
    PF_CHECK_ARG(B >= 0 && nq >= 0 && np > 0 && k >= 1 && k <= np, "pf_knn_patch: bad sizes B=%d nq=%d np=%d k=%d", B,
                 nq, np, k);
    PF_CHECK_ARG(B <= 65535, "pf_knn_patch: B=%d > 65535", B);
    if (B == 0 || nq == 0) return 0;
    int npad = 2;
    while (npad < np) npad <<= 1;
    const size_t smem = (size_t)(npad + SEL_CAP) * sizeof(unsigned long long);
    PF_CHECK_ARG((size_t)npad * 8 <= 208 * 1024, "pf_knn_patch: np=%d too large for the shared-memory sort (max 16384)", np);
    if (int rc = pf_smem_optin(knn_patch_kernel, smem)) return rc;
    dim3 grid(nq, B);
    knn_patch_kernel<<<grid, ST, smem, (cudaStream_t)stream>>>(q, p, nq, np, npad, k, idx);
    return pf_launch_status("knn_patch_kernel");
}
