// Caller-side glue of the x4 path as kernels, and the whole per-cloud pipeline in one C call.
//
// Replaces the tensor plumbing of PatchHelper (reference: /root/reference/modules/utils/patch.py):
//   pf_normalize_pc      normalize_pc, patch.py:168-178 (centroid, max radius, scale) -- one CTA per cloud
//   pf_gather_normalize  patch gather (patch.py:112-125, pc[idx]) fused with the per-patch normalize_pc of
//                        __upsampling_patches (patch.py:85) -- one CTA per patch
//   pf_gather_points     index_points / gather_operation (patch.py:106,157) with an optional per-cloud scale + shift
//                        (the de-normalisation of patch.py:76 fused into the merge gather)
//   pf_denorm_concat     torch.cat([generated, input patch]) * radius + centroid (patch.py:90-91)
//   pf_remove_outliers   patch.py:198-214 after the Chamfer pass: drop the num_outliers largest distances (ties: lowest index
//                        first, = a stable descending argsort), survivors keep their order -- radix select on the float bits
//                        + an ordered compaction, no sort, no host round trip
//   pf_upsample_clouds   PatchHelper.upsample + remove_outliers (upsample.py:46-54) for a batch of clouds: every launch of
//                        the path on ONE stream, no allocation, no synchronisation -- CUDA-graph capturable.
// Arithmetic follows the torch expressions operation by operation (x * r then + c, separately rounded; sqrt and division
// IEEE); the centroid sums are accumulated in double and rounded once (torch's fp32 tree reduction is not specified to the
// ulp either; the parity tests allow 1e-6 on the normalised coordinates).
#include <stdio.h>

#include "common.cuh"
#include "prof.h"
#include "puflow_b200.h"

namespace {

constexpr int NT = 256;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_maxf(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// One CTA per item.  GATHER: point i of item (b, p) is pc[b][idx[(b * items + p) * k + i]].
template <bool GATHER>
__global__ void __launch_bounds__(NT) normalize_kernel(const float* __restrict__ pc, const int32_t* __restrict__ idx, int n_src,
                                                       int items, int k, float* __restrict__ out, float* __restrict__ centroid,
                                                       float* __restrict__ radius) {
    __shared__ double sred[3][NT / 32];
    __shared__ float fred[NT / 32];
    __shared__ float cen[3], rad;
    const long long item = blockIdx.x;
    const int b = (int)(item / items);
    const float* src = pc + (size_t)b * n_src * 3;
    const int32_t* ii = GATHER ? idx + item * k : nullptr;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    auto point = [&](int i, float& x, float& y, float& z) {
        const int j = GATHER ? min(max(ii[i], 0), n_src - 1) : i;
        const float* p = src + (size_t)j * 3;
        x = p[0], y = p[1], z = p[2];
    };
    double s[3] = {0.0, 0.0, 0.0};
    for (int i = tid; i < k; i += NT) {
        float x, y, z;
        point(i, x, y, z);
        s[0] += x, s[1] += y, s[2] += z;
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        s[a] = warp_sum(s[a]);
        if (lane == 0) sred[a][w] = s[a];
    }
    __syncthreads();
    if (tid < 3) {
        double t = 0.0;
        for (int i = 0; i < NT / 32; i++) t += sred[tid][i];
        cen[tid] = (float)(t / (double)k);
    }
    __syncthreads();
    const float cx = cen[0], cy = cen[1], cz = cen[2];
    float m = 0.f;
    for (int i = tid; i < k; i += NT) {
        float x, y, z;
        point(i, x, y, z);
        x = __fsub_rn(x, cx), y = __fsub_rn(y, cy), z = __fsub_rn(z, cz);
        m = fmaxf(m, sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z))));
    }
    m = warp_maxf(m);
    if (lane == 0) fred[w] = m;
    __syncthreads();
    if (tid == 0) {
        float t = fred[0];
        for (int i = 1; i < NT / 32; i++) t = fmaxf(t, fred[i]);
        rad = t;
        if (centroid) centroid[item * 3] = cx, centroid[item * 3 + 1] = cy, centroid[item * 3 + 2] = cz;
        if (radius) radius[item] = t;
    }
    __syncthreads();
    const float r = rad;
    float* dst = out + (size_t)item * k * 3;
    for (int i = tid; i < k; i += NT) {
        float x, y, z;
        point(i, x, y, z);
        dst[(size_t)i * 3] = __fdiv_rn(__fsub_rn(x, cx), r);
        dst[(size_t)i * 3 + 1] = __fdiv_rn(__fsub_rn(y, cy), r);
        dst[(size_t)i * 3 + 2] = __fdiv_rn(__fsub_rn(z, cz), r);
    }
}

// out[b][i] = pc[b][idx[b][i]] (* scale[b] + shift[b][:])
__global__ void gather_points_kernel(const float* __restrict__ pc, const int32_t* __restrict__ idx, int n, int m,
                                     const float* __restrict__ scale, const float* __restrict__ shift, float* __restrict__ out,
                                     long long total) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int b = (int)(t / m);
    const int j = min(max(idx[t], 0), n - 1);
    const float* p = pc + ((size_t)b * n + j) * 3;
    float x = p[0], y = p[1], z = p[2];
    if (scale) {
        const float s = scale[b];
        x = __fadd_rn(__fmul_rn(x, s), shift[b * 3]), y = __fadd_rn(__fmul_rn(y, s), shift[b * 3 + 1]);
        z = __fadd_rn(__fmul_rn(z, s), shift[b * 3 + 2]);
    }
    out[t * 3] = x, out[t * 3 + 1] = y, out[t * 3 + 2] = z;
}

// out[item][0 .. nx) = x[item] * rad + cen, out[item][nx .. nx + np) = patches[item] * rad + cen   (rows of 3 floats)
__global__ void denorm_concat_kernel(const float* __restrict__ x, const float* __restrict__ patches,
                                     const float* __restrict__ cen, const float* __restrict__ rad, int nx, int np,
                                     float* __restrict__ out, long long total) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;      // one float each
    if (t >= total) return;
    const int per = (nx + np) * 3;
    const long long item = t / per;
    const int o = (int)(t - item * per), a = o % 3;
    const float v = o < nx * 3 ? x[item * nx * 3 + o] : patches[item * np * 3 + (o - nx * 3)];
    out[t] = __fadd_rn(__fmul_rn(v, rad[item]), cen[item * 3 + a]);
}

// One CTA per cloud: drop the `drop` largest d (ties: lowest index first), keep the rest in order.
constexpr int RT = 1024;
__global__ void __launch_bounds__(RT) remove_outliers_kernel(const float* __restrict__ sr, const float* __restrict__ dist, int n,
                                                             int drop, float* __restrict__ out) {
    __shared__ unsigned hist[256];
    __shared__ unsigned s_prefix, s_need;
    __shared__ unsigned long long s_scan[RT / 32], s_base;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const unsigned* key = reinterpret_cast<const unsigned*>(dist) + (size_t)b * n;   // d >= 0: float order = unsigned order
    // ---- radix select: the drop-th largest key, 8 bits per pass from the top ----
    if (tid == 0) s_prefix = 0, s_need = (unsigned)drop;
    __syncthreads();
    for (int pass = 0; pass < 4; pass++) {
        const int shift = 24 - 8 * pass;
        const unsigned mask = pass == 0 ? 0u : 0xffffffffu << (shift + 8), prefix = s_prefix;
        if (tid < 256) hist[tid] = 0;
        __syncthreads();
        for (int i = tid; i < n; i += RT) {
            const unsigned kx = key[i];
            if ((kx & mask) == prefix) atomicAdd(&hist[(kx >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (tid == 0) {
            unsigned need = s_need, bin = 255;
            for (;; bin--) {                                   // from the largest digit down
                if (hist[bin] >= need || bin == 0) break;
                need -= hist[bin];
            }
            s_prefix = prefix | (bin << shift), s_need = need;
        }
        __syncthreads();
    }
    const unsigned T = s_prefix, r = s_need;                   // drop every key > T and the first r keys == T
    // ---- ordered compaction ----
    if (tid == 0) s_base = 0;
    __syncthreads();
    const float* src = sr + (size_t)b * n * 3;
    float* dst = out + (size_t)b * (n - drop) * 3;
    for (int c0 = 0; c0 < n; c0 += RT) {
        const int i = c0 + tid;
        const unsigned kx = i < n ? key[i] : 0u;
        const bool eq = i < n && kx == T, gt = i < n && kx > T;
        // one scan for both counts: (equal, greater) packed as 32 + 32 bits
        const unsigned long long v = (eq ? 1ull : 0ull) | (gt ? (1ull << 32) : 0ull);
        unsigned long long incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_scan[w] = incl;
        __syncthreads();
        unsigned long long wsum = lane < RT / 32 ? s_scan[lane] : 0ull, wincl = wsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xffffffffu, wincl, o);
            if (lane >= o) wincl += t;
        }
        const unsigned long long wbase = __shfl_sync(0xffffffffu, wincl - wsum, w), tot = __shfl_sync(0xffffffffu, wincl, 31);
        const unsigned long long base = s_base;                // packed counts of all earlier chunks
        const unsigned long long excl = base + wbase + incl - v;   // packed (equal, greater) counts of all earlier elements
        const unsigned eq_before = (unsigned)(excl & 0xffffffffull), gt_before = (unsigned)(excl >> 32);
        const bool dropped = gt || (eq && eq_before < r);
        if (i < n && !dropped) {
            const unsigned dropped_before = gt_before + min(eq_before, r);
            const size_t pos = (size_t)i - dropped_before;
            dst[pos * 3] = src[(size_t)i * 3], dst[pos * 3 + 1] = src[(size_t)i * 3 + 1], dst[pos * 3 + 2] = src[(size_t)i * 3 + 2];
        }
        __syncthreads();
        if (tid == 0) s_base = base + tot;
        __syncthreads();
    }
}

constexpr int64_t ALIGN = 256;
constexpr int64_t align_up(int64_t v) { return (v + ALIGN - 1) / ALIGN * ALIGN; }
struct CloudWork {
    int64_t pcn, gcen, grad, seed_idx, seeds, patch_idx, patches, pcen, prad, net, x, cand, fps, sel, pred, d1, i1, total;
};
CloudWork cloud_layout(int64_t B, int64_t N, int64_t R, int64_t npoint, int64_t k, int64_t n_patch) {
    CloudWork w;
    int64_t o = 0;
    auto take = [&](int64_t bytes) { const int64_t at = o; o += align_up(bytes); return at; };
    const int64_t Bp = B * n_patch, ncand = n_patch * (k * R + k);
    w.pcn = take(B * N * 12), w.gcen = take(B * 12), w.grad = take(B * 4);
    w.seed_idx = take(Bp * 4), w.seeds = take(Bp * 12), w.patch_idx = take(Bp * k * 4);
    w.patches = take(Bp * k * 12), w.pcen = take(Bp * 12), w.prad = take(Bp * 4);
    w.net = take(pf_upsample_workspace_bytes((int)Bp, (int)k, (int)R));
    w.x = take(Bp * k * R * 12), w.cand = take(B * ncand * 12);
    const int64_t fa = pf_fps_clustered_work_floats((int)B, (int)N), fb = pf_fps_clustered_work_floats((int)B, (int)ncand);
    w.fps = take((fa > fb ? fa : fb) * 4);
    w.sel = take(B * npoint * 4), w.pred = take(B * npoint * 12), w.d1 = take(B * npoint * 4), w.i1 = take(B * npoint * 4);
    w.total = o;
    return w;
}

// the sampler for a batch of clouds of n candidates: brute force below 4,096, the threshold queue up to its capacity, the
// cluster kernel (which falls back to the single-CTA pruned kernel) above; work: pf_fps_clustered_work_floats(B, n) floats
int fps_any(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, void* stream) {
    if (n < 4096) return pf_fps(xyz, B, n, npoint, idx, nullptr, stream);
    if (n <= pf_fps_queued_max_points()) return pf_fps_queued(xyz, B, n, npoint, idx, work, stream);
    return pf_fps_clustered(xyz, B, n, npoint, idx, work, stream);
}

}  // namespace

extern "C" int pf_normalize_pc(const float* pc, int B, int n, float* out, float* centroid, float* radius, void* stream) {
    PF_CHECK_ARG(B >= 0 && n > 0, "pf_normalize_pc: bad sizes B=%d n=%d", B, n);
    if (B == 0) return 0;
    PF_CHECK_ARG(pc && out, "pf_normalize_pc: null pointer");
    normalize_kernel<false><<<B, NT, 0, (cudaStream_t)stream>>>(pc, nullptr, n, 1, n, out, centroid, radius);
    return pf_launch_status("normalize_kernel");
}

extern "C" int pf_gather_normalize(const float* pc, const int32_t* idx, int B, int n, int items, int k, float* out,
                                   float* centroid, float* radius, void* stream) {
    PF_CHECK_ARG(B >= 0 && n > 0 && items >= 0 && k > 0, "pf_gather_normalize: bad sizes B=%d n=%d items=%d k=%d", B, n, items, k);
    if (B == 0 || items == 0) return 0;
    PF_CHECK_ARG(pc && idx && out, "pf_gather_normalize: null pointer");
    PF_CHECK_ARG((long long)B * items < 0x7fffffffLL, "pf_gather_normalize: too many items");
    normalize_kernel<true><<<(unsigned)(B * items), NT, 0, (cudaStream_t)stream>>>(pc, idx, n, items, k, out, centroid, radius);
    return pf_launch_status("gather_normalize_kernel");
}

extern "C" int pf_gather_points(const float* pc, const int32_t* idx, int B, int n, int m, const float* scale,
                                const float* shift, float* out, void* stream) {
    PF_CHECK_ARG(B >= 0 && n > 0 && m >= 0, "pf_gather_points: bad sizes B=%d n=%d m=%d", B, n, m);
    if (B == 0 || m == 0) return 0;
    PF_CHECK_ARG(pc && idx && out && (!scale == !shift), "pf_gather_points: null pointer (scale and shift go together)");
    const long long total = (long long)B * m;
    gather_points_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pc, idx, n, m, scale, shift, out, total);
    return pf_launch_status("gather_points_kernel");
}

extern "C" int pf_denorm_concat(const float* x, const float* patches, const float* centroid, const float* radius, int items,
                                int nx, int np, float* out, void* stream) {
    PF_CHECK_ARG(items >= 0 && nx >= 0 && np >= 0, "pf_denorm_concat: bad sizes items=%d nx=%d np=%d", items, nx, np);
    const long long total = (long long)items * (nx + np) * 3;
    if (total == 0) return 0;
    PF_CHECK_ARG(x && patches && centroid && radius && out, "pf_denorm_concat: null pointer");
    denorm_concat_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, patches, centroid, radius, nx, np, out,
                                                                                            total);
    return pf_launch_status("denorm_concat_kernel");
}

namespace {
__global__ void __launch_bounds__(256) fps_park_near_origin_kernel(const float* __restrict__ xyz, int n, float thresh2,
                                                                   float* __restrict__ out) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const float* pb = xyz + (size_t)blockIdx.y * n * 3;
    float x = pb[3 * i], y = pb[3 * i + 1], z = pb[3 * i + 2];
    if (i > 0 && x * x + y * y + z * z <= thresh2) x = pb[0], y = pb[1], z = pb[2];
    float* ob = out + ((size_t)blockIdx.y * n + i) * 3;
    ob[0] = x, ob[1] = y, ob[2] = z;
}
}  // namespace

extern "C" int pf_fps_park_near_origin(const float* xyz, int B, int n, float thresh2, float* out, void* stream) {
    PF_CHECK_ARG(B >= 0 && n > 0 && B <= 65535, "pf_fps_park_near_origin: bad sizes B=%d n=%d", B, n);
    if (B == 0) return 0;
    PF_CHECK_ARG(xyz && out && xyz != out, "pf_fps_park_near_origin: null pointer or in-place call (point 0 is read by every thread)");
    PF_CHECK_ARG(thresh2 >= 0.f, "pf_fps_park_near_origin: thresh2=%g", (double)thresh2);
    fps_park_near_origin_kernel<<<dim3((n + 255) / 256, B), 256, 0, (cudaStream_t)stream>>>(xyz, n, thresh2, out);
    return pf_launch_status("fps_park_near_origin_kernel");
}

extern "C" int pf_remove_outliers(const float* sr, const float* dist1, int B, int n, int num_outliers, float* out, void* stream) {
    PF_CHECK_ARG(B >= 0 && n > 0 && num_outliers >= 0 && num_outliers < n && n < 65536 * 16,
                 "pf_remove_outliers: bad sizes B=%d n=%d num_outliers=%d", B, n, num_outliers);
    if (B == 0) return 0;
    PF_CHECK_ARG(sr && dist1 && out, "pf_remove_outliers: null pointer");
    PF_CHECK_ARG(num_outliers > 0, "pf_remove_outliers: nothing to remove (copy the input instead)");
    remove_outliers_kernel<<<B, RT, 0, (cudaStream_t)stream>>>(sr, dist1, n, num_outliers, out);
    return pf_launch_status("remove_outliers_kernel");
}

extern "C" int64_t pf_upsample_clouds_workspace_bytes(int B, int N, int R, int npoint, int patch_points, int n_patch) {
    if (B < 0 || N < 1 || R < 1 || npoint < 1 || patch_points < 16 || n_patch < 1) return -1;
    return cloud_layout(B, N, R, npoint, patch_points, n_patch).total;
}

extern "C" int pf_upsample_clouds(const void* model, const float* clouds, int B, int N, int R, int npoint, int num_outliers,
                                  int patch_points, int n_patch, void* workspace, int64_t workspace_bytes, float* out,
                                  int precision, void* stream) {
    PF_CHECK_ARG(B >= 0 && N >= patch_points && R >= 1 && R <= 32 && num_outliers >= 0 && npoint > num_outliers &&
                     patch_points >= 16 && n_patch >= 1,
                 "pf_upsample_clouds: bad sizes B=%d N=%d R=%d npoint=%d num_outliers=%d patch_points=%d n_patch=%d", B, N, R,
                 npoint, num_outliers, patch_points, n_patch);
    PF_CHECK_ARG(n_patch <= N, "pf_upsample_clouds: n_patch=%d seeds from %d points", n_patch, N);
    if (B == 0) return 0;
    PF_CHECK_ARG(model && clouds && workspace && out, "pf_upsample_clouds: null pointer");
    PF_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "pf_upsample_clouds: workspace must be 256-byte aligned");
    const int k = patch_points, Bp = B * n_patch;
    const int ncand = n_patch * (k * R + k);
    PF_CHECK_ARG(npoint <= ncand, "pf_upsample_clouds: %d points wanted from %d candidates", npoint, ncand);
    const CloudWork w = cloud_layout(B, N, R, npoint, k, n_patch);
    PF_CHECK_ARG(workspace_bytes >= w.total, "pf_upsample_clouds: workspace of %lld B, %lld needed", (long long)workspace_bytes,
                 (long long)w.total);
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    auto F = [&](int64_t off) { return reinterpret_cast<float*>(ws + off); };
    auto I = [&](int64_t off) { return reinterpret_cast<int32_t*>(ws + off); };
    int rc;
    pf::ProfScope mk((cudaStream_t)stream);
    char tag[40];
    // normalise the clouds, pick the patch seeds, cut the patches (patch.py:59-64,96-125)
    if ((rc = pf_normalize_pc(clouds, B, N, F(w.pcn), F(w.gcen), F(w.grad), stream))) return rc;
    mk.mark("normalize_pc");
    if ((rc = fps_any(F(w.pcn), B, N, n_patch, I(w.seed_idx), F(w.fps), stream)))
        return rc;
    snprintf(tag, sizeof(tag), "fps_n%d", N);
    mk.mark(tag);
    if ((rc = pf_gather_points(F(w.pcn), I(w.seed_idx), B, N, n_patch, nullptr, nullptr, F(w.seeds), stream))) return rc;
    mk.mark("gather_points");
    if ((rc = pf_knn_patch(F(w.seeds), F(w.pcn), B, n_patch, N, k, I(w.patch_idx), stream))) return rc;
    mk.mark("knn_patch");
    if ((rc = pf_gather_normalize(F(w.pcn), I(w.patch_idx), B, N, n_patch, k, F(w.patches), F(w.pcen), F(w.prad), stream))) return rc;
    mk.mark("gather_normalize");
    // the network on every patch (patch.py:82-93 -> PointInterpFlow.sample)
    if ((rc = pf_upsample_patches(model, F(w.patches), Bp, k, R, ws + w.net, pf_upsample_workspace_bytes(Bp, k, R), F(w.x), nullptr,
                                  precision, stream)))
        return rc;
    if ((rc = pf_denorm_concat(F(w.x), F(w.patches), F(w.pcen), F(w.prad), Bp, k * R, k, F(w.cand), stream))) return rc;
    mk.mark("denorm_concat");
    // merge: farthest point sampling over all candidates of a cloud, back to the input frame (patch.py:67-70,142-158)
    if ((rc = fps_any(F(w.cand), B, ncand, npoint, I(w.sel), F(w.fps), stream)))
        return rc;
    snprintf(tag, sizeof(tag), "fps_n%d", ncand);
    mk.mark(tag);
    float* pred = num_outliers ? F(w.pred) : out;
    if ((rc = pf_gather_points(F(w.cand), I(w.sel), B, ncand, npoint, F(w.grad), F(w.gcen), pred, stream))) return rc;
    mk.mark("gather_points");
    if (!num_outliers) return 0;
    // outlier removal (patch.py:198-214)
    if ((rc = pf_chamfer_nn(pred, npoint, clouds, N, B, F(w.d1), I(w.i1), nullptr, nullptr, stream))) return rc;
    mk.mark("chamfer_nn");
    rc = pf_remove_outliers(pred, F(w.d1), B, npoint, num_outliers, out, stream);
    mk.mark("remove_outliers");
    return rc;
}
