// Dev probe (not product code): semantics and throughput of cp.async.bulk.tensor.2d ... tile::gather4 on sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/g4 tools/probes/gather4_probe.cu && /tmp/g4
// Part 1: layout of 4-row gathers in shared memory with SWIZZLE_{NONE,32B,64B,128B} maps whose box is {W, 1}.
// Part 2: sustained gather rate per SM of 16 KB stages (128 rows x 128 B, and 32 rows x 512 B) through a 2/3/4-deep
//         ring, rows random inside a 256-row window (= one patch), tensor 1 Mi rows x 512 fp32 (2 GiB, as at config 2).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeFn get_encode() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    if (!fn) { printf("no cuTensorMapEncodeTiled\n"); exit(1); }
    return (EncodeFn)fn;
}

static CUtensorMap make_map(EncodeFn enc, void* base, uint64_t rows, uint64_t cols, uint32_t box_w, uint32_t box_h,
                            CUtensorMapSwizzle sw) {
    CUtensorMap m;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 4};
    cuuint32_t box[2] = {box_w, box_h};
    cuuint32_t es[2] = {1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                     CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d (box %u x %u sw %d)\n", (int)r, box_w, box_h, (int)sw); exit(1); }
    return m;
}

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint64_t* b, uint32_t ph) {
    uint32_t ok;
    asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p;}" : "=r"(ok) : "r"(s32(b)), "r"(ph) : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t ph) { uint32_t n = 0; while (!mbar_try(b, ph)) if (++n > (1u << 24)) __trap(); }
__device__ __forceinline__ void gather4(uint32_t dst, const CUtensorMap* m, uint64_t* bar, int col, int r0, int r1, int r2, int r3) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                 ::"r"(dst), "l"(m), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void tile2d(uint32_t dst, const CUtensorMap* m, uint64_t* bar, int col, int row) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(m), "r"(col), "r"(row), "r"(s32(bar)) : "memory");
}

// ---- part 1: layout ----
__global__ void layout_kernel(const __grid_constant__ CUtensorMap m, int box_w, int lanes, int col, float* out, int out_floats, int tiled_rows) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    float* s = reinterpret_cast<float*>(smem);
    for (int i = threadIdx.x; i < out_floats; i += blockDim.x) s[i] = -1.f;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tiled_rows) {
        if (threadIdx.x == 0) { mbar_expect(&bar, box_w * 4 * tiled_rows); tile2d(s32(smem), &m, &bar, col, 1000); }
    } else {
        if (threadIdx.x == 0) mbar_expect(&bar, lanes * 4 * box_w * 4);
        __syncwarp();
        if (threadIdx.x < lanes) {
            const int l = threadIdx.x;
            // rows: arbitrary, not sorted, with a repeat
            const int r0 = 7 + 37 * l, r1 = 4000 - 11 * l, r2 = 7 + 37 * l, r3 = (l * 97) % 4096;
            gather4(s32(smem) + l * 4 * box_w * 4, &m, &bar, col, r0, r1, r2, r3);
        }
    }
    mbar_wait(&bar, 0);
    __syncthreads();
    for (int i = threadIdx.x; i < out_floats; i += blockDim.x) out[i] = s[i];
}

// ---- part 2: throughput ----
// Each CTA: warp 0 = producer (32 lanes x gather4 per stage), warps 1..4 = consumers reading every staged byte.
template <int DEPTH>
__global__ void __launch_bounds__(160, 1) rate_kernel(const __grid_constant__ CUtensorMap m, int box_w, int lanes, int stages, int nrows_total,
                                                      long long* cycles, float* sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[DEPTH], empty[DEPTH];
    const int stage_bytes = lanes * 4 * box_w * 4;
    if (threadIdx.x == 0) {
        for (int i = 0; i < DEPTH; i++) { mbar_init(full + i, 1); mbar_init(empty + i, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long t0 = clock64();
    if (warp == 0) {
        uint32_t rng = 1234567u + blockIdx.x * 7919u + lane * 104729u;
        for (int s = 0; s < stages; s++) {
            const int b = s % DEPTH;
            if (s >= DEPTH) mbar_wait(empty + b, ((s / DEPTH) - 1) & 1);
            const int window = (int)(((unsigned)(blockIdx.x * 131 + s / 8) * 2654435761u) % (unsigned)(nrows_total / 256)) * 256;
            int r[4];
            for (int i = 0; i < 4; i++) { rng = rng * 1664525u + 1013904223u; r[i] = window + (rng >> 24); }
            if (lane == 0) mbar_expect(full + b, stage_bytes);
            __syncwarp();
            if (lane < lanes) gather4(s32(smem) + b * stage_bytes + lane * 4 * box_w * 4, &m, full + b, box_w == 32 ? 256 + (s % 4) * 32 : (box_w == 128 ? 384 : 256), r[0], r[1], r[2], r[3]);
        }
    } else {
        float acc = 0.f;
        const int t = threadIdx.x - 32;
        for (int s = 0; s < stages; s++) {
            const int b = s % DEPTH;
            mbar_wait(full + b, (s / DEPTH) & 1);
            const float4* p = reinterpret_cast<const float4*>(smem + b * stage_bytes);
            for (int i = t; i < stage_bytes / 16; i += 128) { const float4 v = p[i]; acc += v.x + v.y + v.z + v.w; }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + b);
        }
        if (acc == 123.456f) sink[0] = acc;
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

template <int DEPTH>
static void run_rate(const CUtensorMap& m, int box_w, int lanes, int nrows, const char* what) {
    const int stages = 4000, stage_bytes = lanes * 16 * box_w;
    long long* cyc; float* sink;
    CK(cudaMalloc(&cyc, 148 * 8)); CK(cudaMalloc(&sink, 4));
    auto k = rate_kernel<DEPTH>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, DEPTH * stage_bytes + 1024));
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int it = 0; it < 2; it++) {
        cudaEventRecord(a);
        k<<<148, 160, DEPTH * stage_bytes + 1024>>>(m, box_w, lanes, stages, nrows, cyc, sink);
        cudaEventRecord(b);
        CK(cudaDeviceSynchronize());
    }
    float ms; cudaEventElapsedTime(&ms, a, b);
    long long h[148]; CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
    double mean = 0; for (int i = 0; i < 148; i++) mean += h[i]; mean /= 148;
    printf("rate %-28s depth %d: %.3f ms, %.0f cycles/stage/SM, %.2f B/cycle/SM, %.2f TB/s chip\n", what, DEPTH, ms, mean / stages,
           stage_bytes / (mean / stages), 148.0 * stages * stage_bytes / (ms * 1e-3) / 1e12);
    cudaFree(cyc); cudaFree(sink);
}

int main() {
    EncodeFn enc = get_encode();
    // ---- part 1 ----
    {
        const int rows = 4096, cols = 512;
        std::vector<float> h((size_t)rows * cols);
        for (int r = 0; r < rows; r++) for (int c = 0; c < cols; c++) h[(size_t)r * cols + c] = r * 1000.f + c;
        float* d; CK(cudaMalloc(&d, h.size() * 4)); CK(cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
        float* out; CK(cudaMalloc(&out, 65536));
        std::vector<float> ho(16384);
        struct Case { int box_w; CUtensorMapSwizzle sw; int lanes; const char* name; } cases[] = {
            {32, CU_TENSOR_MAP_SWIZZLE_128B, 32, "box 32 sw128"}, {16, CU_TENSOR_MAP_SWIZZLE_64B, 32, "box 16 sw64"},
            {8, CU_TENSOR_MAP_SWIZZLE_32B, 32, "box 8 sw32"},     {128, CU_TENSOR_MAP_SWIZZLE_NONE, 8, "box 128 none"},
            {32, CU_TENSOR_MAP_SWIZZLE_NONE, 32, "box 32 none"},  {64, CU_TENSOR_MAP_SWIZZLE_NONE, 8, "box 64 none"},
        };
        for (auto& c : cases) {
            CUtensorMap m = make_map(enc, d, rows, cols, c.box_w, 1, c.sw);
            const int nf = c.lanes * 4 * c.box_w;
            CK(cudaFuncSetAttribute(layout_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
            layout_kernel<<<1, 128, 65536>>>(m, c.box_w, c.lanes, 256 + 8, out, nf, 0);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("%s: kernel failed: %s\n", c.name, cudaGetErrorString(e)); return 1; }
            CK(cudaMemcpy(ho.data(), out, nf * 4, cudaMemcpyDeviceToHost));
            // decode: for each lane l, row slot q (0..3), 16-byte chunk p: which (row, col) landed there
            int bad_plain = 0, bad_swz = 0;
            const int cpr = c.box_w / 4;   // chunks per row
            for (int l = 0; l < c.lanes; l++) {
                const int rr[4] = {7 + 37 * l, 4000 - 11 * l, 7 + 37 * l, (l * 97) % 4096};
                for (int q = 0; q < 4; q++) for (int p = 0; p < cpr; p++) for (int w = 0; w < 4; w++) {
                    const int srow = l * 4 + q;
                    const float got = ho[(size_t)srow * c.box_w + p * 4 + w];
                    const float want_plain = rr[q] * 1000.f + 264 + p * 4 + w;
                    // swizzled expectation: physical chunk p holds logical chunk p ^ f(addr)
                    const uint32_t addr = (uint32_t)srow * c.box_w * 4 + p * 16;
                    uint32_t la = addr;
                    if (c.sw == CU_TENSOR_MAP_SWIZZLE_128B) la = addr ^ (((addr >> 7) & 7) << 4);
                    if (c.sw == CU_TENSOR_MAP_SWIZZLE_64B) la = addr ^ (((addr >> 7) & 3) << 4);
                    if (c.sw == CU_TENSOR_MAP_SWIZZLE_32B) la = addr ^ (((addr >> 7) & 1) << 4);
                    const int lrow = la / (c.box_w * 4), lp = (la % (c.box_w * 4)) / 16;
                    const int l2 = lrow / 4, q2 = lrow % 4;
                    const int rr2[4] = {7 + 37 * l2, 4000 - 11 * l2, 7 + 37 * l2, (l2 * 97) % 4096};
                    const float want_swz = rr2[q2] * 1000.f + 264 + lp * 4 + w;
                    if (got != want_plain) bad_plain++;
                    if (got != want_swz) bad_swz++;
                }
            }
            printf("layout %-14s: mismatches plain %d, address-xor model %d   first row: %.0f %.0f %.0f %.0f | %.0f ...\n", c.name, bad_plain, bad_swz,
                   ho[0], ho[1], ho[2], ho[3], ho[4]);
        }
        {   // plain tiled 2-D load of an 8-row box for comparison (box 128 x 8, 256 x 8)
            for (int bw : {128, 256}) {
                CUtensorMap m = make_map(enc, d, rows, cols, bw, 8, CU_TENSOR_MAP_SWIZZLE_NONE);
                layout_kernel<<<1, 128, 65536>>>(m, bw, 0, 0, out, bw * 8, 8);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("tiled %d: kernel failed: %s\n", bw, cudaGetErrorString(e)); return 1; }
                CK(cudaMemcpy(ho.data(), out, bw * 8 * 4, cudaMemcpyDeviceToHost));
                int bad = 0;
                for (int r = 0; r < 8; r++) for (int cc = 0; cc < bw; cc++) if (ho[r * bw + cc] != (1000 + r) * 1000.f + cc) bad++;
                printf("tiled box %d x 8: mismatches %d\n", bw, bad);
            }
        }
        cudaFree(d); cudaFree(out);
    }
    // ---- part 2 ----
    {
        const uint64_t rows = 1ull << 20, cols = 512;
        float* d; CK(cudaMalloc(&d, rows * cols * 4)); CK(cudaMemset(d, 0, rows * cols * 4));
        CUtensorMap m32 = make_map(enc, d, rows, cols, 32, 1, CU_TENSOR_MAP_SWIZZLE_128B);
        CUtensorMap m128 = make_map(enc, d, rows, cols, 128, 1, CU_TENSOR_MAP_SWIZZLE_NONE);
        CUtensorMap m256 = make_map(enc, d, rows, cols, 256, 1, CU_TENSOR_MAP_SWIZZLE_NONE);
        run_rate<2>(m32, 32, 32, (int)rows, "128 rows x 128 B (16 KB)");
        run_rate<3>(m32, 32, 32, (int)rows, "128 rows x 128 B (16 KB)");
        run_rate<4>(m32, 32, 32, (int)rows, "128 rows x 128 B (16 KB)");
        run_rate<8>(m32, 32, 32, (int)rows, "128 rows x 128 B (16 KB)");
        run_rate<2>(m128, 128, 8, (int)rows, "32 rows x 512 B (16 KB)");
        run_rate<4>(m128, 128, 8, (int)rows, "32 rows x 512 B (16 KB)");
        run_rate<8>(m128, 128, 8, (int)rows, "32 rows x 512 B (16 KB)");
        run_rate<4>(m256, 256, 4, (int)rows, "16 rows x 1 KB (16 KB)");
        run_rate<4>(m256, 256, 8, (int)rows, "32 rows x 1 KB (32 KB)");
        cudaFree(d);
    }
    return 0;
}
