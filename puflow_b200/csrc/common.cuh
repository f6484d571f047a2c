// Shared device helpers for the puflow_b200 kernels (sm_100a only).
//
//  * mbarrier + cp.async.bulk (TMA bulk copy, SASS: UBLKCP) staging of contiguous global blocks
//    into shared memory;
//  * tile_gemm: the register-tiled FP32 FFMA micro-kernel every per-edge / per-point MLP in the
//    path is built from (activations k-major in shared memory, weights k-major in shared or
//    global memory, warp-uniform weight reads are broadcasts);
//  * status plumbing for the C-ABI (0 = ok, otherwise a cudaError_t or a PF_E* code).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define PF_EINVAL 10001   // bad argument (shape / alignment / unsupported size)
#define PF_ENOMEM 10002   // shared memory request over the 227 KB limit

void pf_set_error(const char* fmt, ...);
// fps_bucket.cu: pf_fps_bucketed with a device-side run condition (internal; pf_fps_clustered's fallback)
int pf_fps_bucketed_if(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, const int* run_flag, void* stream);

#define PF_CHECK_ARG(cond, ...)                     \
    do {                                            \
        if (!(cond)) {                              \
            pf_set_error(__VA_ARGS__);              \
            return PF_EINVAL;                       \
        }                                           \
    } while (0)

#define PF_CUDA(call)                                                                         \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            pf_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return (int)e_;                                                                   \
        }                                                                                     \
    } while (0)

// Opt a kernel in to `bytes` of dynamic shared memory (needed above 48 KB) and report launch errors.
template <typename KernelT>
static inline int pf_smem_optin(KernelT kernel, size_t bytes) {
    if (bytes > 227 * 1024) {
        pf_set_error("kernel needs %zu B of shared memory (> 227 KB)", bytes);
        return PF_ENOMEM;
    }
    PF_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return 0;
}

static inline int pf_launch_status(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        pf_set_error("%s launch failed: %s", what, cudaGetErrorString(e));
        return (int)e;
    }
    return 0;
}

namespace pf {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier + bulk async copy (TMA, 1-D) ---------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// make the initialised barrier visible to the async (TMA) proxy
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        "  .reg .pred p;\n"
        "  mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "  selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// non-blocking: has the phase with this parity completed?
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        "  .reg .pred p;\n"
        "  mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "  selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// global -> shared bulk copy; bytes % 16 == 0, both addresses 16-byte aligned.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// Issue one logical copy as <= 64 KB pieces (keeps each transaction count well inside the
// mbarrier tx-count range).  Called by ONE thread, after mbar_expect_tx(bar, bytes).
__device__ __forceinline__ void bulk_g2s_chunked(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                                 uint64_t* bar) {
    const uint32_t CH = 32768;
    uint32_t off = 0;
    while (off < bytes) {
        uint32_t n = bytes - off < CH ? bytes - off : CH;
        bulk_g2s((char*)dst_smem + off, (const char*)src_gmem + off, n, bar);
        off += n;
    }
}

// ---- small vector loads -----------------------------------------------------------------------
template <int N>
__device__ __forceinline__ void ld_vec(const float* __restrict__ p, float (&v)[N]) {
    if constexpr (N % 4 == 0) {
#pragma unroll
        for (int i = 0; i < N / 4; i++) {
            float4 t = *reinterpret_cast<const float4*>(p + 4 * i);
            v[4 * i] = t.x, v[4 * i + 1] = t.y, v[4 * i + 2] = t.z, v[4 * i + 3] = t.w;
        }
    } else if constexpr (N == 2) {
        float2 t = *reinterpret_cast<const float2*>(p);
        v[0] = t.x, v[1] = t.y;
    } else {
#pragma unroll
        for (int i = 0; i < N; i++) v[i] = p[i];
    }
}
// read-only (LDG.CI) variant for global data written by an earlier launch
template <int N>
__device__ __forceinline__ void ldg_vec(const float* __restrict__ p, float (&v)[N]) {
    if constexpr (N % 4 == 0) {
#pragma unroll
        for (int i = 0; i < N / 4; i++) {
            float4 t = __ldg(reinterpret_cast<const float4*>(p + 4 * i));
            v[4 * i] = t.x, v[4 * i + 1] = t.y, v[4 * i + 2] = t.z, v[4 * i + 3] = t.w;
        }
    } else if constexpr (N == 2) {
        float2 t = __ldg(reinterpret_cast<const float2*>(p));
        v[0] = t.x, v[1] = t.y;
    } else {
#pragma unroll
        for (int i = 0; i < N; i++) v[i] = __ldg(p + i);
    }
}

// LeakyReLU for 0 <= slope <= 1 (checked at the C-ABI): max(v, slope v) -- one FMUL + one FMNMX, same value as the select form
__device__ __forceinline__ float lrelu(float v, float slope) { return fmaxf(v, v * slope); }

// ---- register-tiled FFMA micro-kernel ----------------------------------------------------------
// acc[e][o] += sum_k hs[k*ldh + row(e)] * w[k*ldw + o]   for e < TE, o < TO, k < kdim.
//   hs : shared memory, activations k-major (row k = channel k of every row of the tile);
//        caller passes hs already offset to its first row (16-byte aligned).  TE == 4: rows
//        0..3; TE == 8: rows 0..3 and GAP..GAP+3 (two quads half a tile apart, so that the 16
//        lanes that share a weight column read 256 contiguous bytes per LDS.128: conflict-free);
//   w  : weights k-major, already offset to the caller's first output column; shared memory when
//        W_GLOBAL == false, global memory through the read-only path otherwise.
template <int TE, int TO, int GAP, bool W_GLOBAL = false, int UNROLL = 4>
__device__ __forceinline__ void tile_gemm(const float* __restrict__ hs, int ldh, int kdim,
                                          const float* __restrict__ w, int ldw, float (&acc)[TE][TO]) {
    static_assert(TE == 4 || TE == 8, "TE must be 4 or 8");
#pragma unroll UNROLL
    for (int k = 0; k < kdim; k++) {
        float a[TE], b[TO];
        {
            float4 t = *reinterpret_cast<const float4*>(hs + (size_t)k * ldh);
            a[0] = t.x, a[1] = t.y, a[2] = t.z, a[3] = t.w;
            if constexpr (TE == 8) {
                float4 u = *reinterpret_cast<const float4*>(hs + (size_t)k * ldh + GAP);
                a[4] = u.x, a[5] = u.y, a[6] = u.z, a[7] = u.w;
            }
        }
        if constexpr (W_GLOBAL)
            ldg_vec<TO>(w + (size_t)k * ldw, b);
        else
            ld_vec<TO>(w + (size_t)k * ldw, b);
#pragma unroll
        for (int e = 0; e < TE; e++)
#pragma unroll
            for (int o = 0; o < TO; o++) acc[e][o] = fmaf(a[e], b[o], acc[e][o]);
    }
}

template <int TE, int TO>
__device__ __forceinline__ void zero_acc(float (&acc)[TE][TO]) {
#pragma unroll
    for (int e = 0; e < TE; e++)
#pragma unroll
        for (int o = 0; o < TO; o++) acc[e][o] = 0.f;
}

}  // namespace pf
