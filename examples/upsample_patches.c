/* A host that is not Python: the whole PU-Flow x R network on a batch of patches through the C-ABI alone.
 *
 *   upsample_patches MODEL.bin PATCHES.f32 Bp N R OUT.f32
 *
 * MODEL.bin   : the packed-weight blob (python -c "from puflow_b200 import packing; ..." pack_model_blob(...).numpy().tofile)
 * PATCHES.f32 : Bp * N * 3 little-endian floats, normalised patch coordinates (PatchHelper.normalize_pc, patch.py:168-178)
 * OUT.f32     : Bp * N * R * 3 floats, row n * R + r of each patch; the log-likelihood is printed.
 *
 * Build (what tests/test_gpu_modules.py does):
 *   gcc -O2 -Iinclude -I/usr/local/cuda/include examples/upsample_patches.c -o upsample_patches \
 *       -Lpuflow_b200/_lib -lpuflow_b200 -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,$PWD/puflow_b200/_lib
 */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "puflow_b200.h"

static void* read_file(const char* path, size_t want) {
    FILE* f = fopen(path, "rb");
    if (!f) { perror(path); exit(2); }
    void* buf = malloc(want ? want : 1);
    if (fread(buf, 1, want, f) != want) { fprintf(stderr, "%s: shorter than %zu bytes\n", path, want); exit(2); }
    fclose(f);
    return buf;
}

#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_)); exit(3); } } while (0)

int main(int argc, char** argv) {
    if (argc != 7) { fprintf(stderr, "usage: %s MODEL.bin PATCHES.f32 Bp N R OUT.f32\n", argv[0]); return 1; }
    const int Bp = atoi(argv[3]), N = atoi(argv[4]), R = atoi(argv[5]);
    if (pf_abi_version() != PF_ABI_VERSION) { fprintf(stderr, "ABI mismatch\n"); return 1; }
    const size_t model_bytes = (size_t)pf_model_bytes();
    const size_t in_bytes = (size_t)Bp * N * 3 * sizeof(float), out_bytes = in_bytes * R;
    const int64_t ws_bytes = pf_upsample_workspace_bytes(Bp, N, R);
    if (ws_bytes < 0) { fprintf(stderr, "bad sizes\n"); return 1; }
    void* h_model = read_file(argv[1], model_bytes);
    void* h_in = read_file(argv[2], in_bytes);
    float* h_out = (float*)malloc(out_bytes ? out_bytes : 1);

    void *d_model, *d_ws;
    float *d_in, *d_out, *d_logp;
    cudaStream_t st;
    CU(cudaStreamCreate(&st));
    CU(cudaMalloc(&d_model, model_bytes));                  /* cudaMalloc is 256-byte aligned */
    CU(cudaMalloc(&d_ws, ws_bytes ? ws_bytes : 1));
    CU(cudaMalloc((void**)&d_in, in_bytes ? in_bytes : 1));
    CU(cudaMalloc((void**)&d_out, out_bytes ? out_bytes : 1));
    CU(cudaMalloc((void**)&d_logp, sizeof(float)));
    CU(cudaMemcpyAsync(d_model, h_model, model_bytes, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_in, h_in, in_bytes, cudaMemcpyHostToDevice, st));

    const int rc = pf_upsample_patches(d_model, d_in, Bp, N, R, d_ws, ws_bytes, d_out, d_logp, PF_PREC_FP32_SPLIT, st);
    if (rc) { fprintf(stderr, "pf_upsample_patches -> %d: %s\n", rc, pf_last_error()); return 4; }

    float logp = 0.f;
    CU(cudaMemcpyAsync(h_out, d_out, out_bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&logp, d_logp, sizeof(float), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    FILE* f = fopen(argv[6], "wb");
    if (!f) { perror(argv[6]); return 2; }
    fwrite(h_out, 1, out_bytes, f);
    fclose(f);
    printf("logp_x %.9g\n", logp);
    return 0;
}
