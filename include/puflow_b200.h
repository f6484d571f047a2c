/*
 * puflow_b200 -- C-ABI of the B200-native kernels for PU-Flow's discrete x4 upsampling hot path.
 *
 * Drop-in boundary.  The reference has exactly one in-tree native FFI, the
 * pybind11 module `chamfer_3D` (forward(xyz1, xyz2, dist1, dist2, idx1, idx2) -> int,
 * /root/reference/metric/PyTorchCD/chamfer3D/chamfer_cuda.cpp:17-33); everything else on the path is
 * reached through third-party Python packages (pytorch3d.ops.knn_points / knn_gather, knn_cuda.KNN,
 * pointnet2_ops furthest_point_sample / gather_operation) and ATen.  The functions below are what a
 * binding for this path calls instead.  Conventions, all functions:
 *
 *   - plain C: raw DEVICE pointers + explicit sizes, no torch types; the caller owns every buffer
 *     (outputs and workspaces are pre-allocated by the caller, like chamfer_3D.forward's out-params,
 *     /root/reference/metric/PyTorchCD/chamfer3D/dist_chamfer_3D.py:40-52); nothing is allocated,
 *     freed or synchronised inside;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream, which is what the
 *     reference launches on, chamfer3D.cu:142-143);
 *   - return value 0 = success; otherwise a cudaError_t value or PF_EINVAL / PF_ENOMEM, with a
 *     message available from pf_last_error() (the reference returns 1/0 and printf()s,
 *     chamfer3D.cu:145-151 -- the Python binding raises instead);
 *   - tensors are dense row-major fp32 unless noted; indices are int32 (the Python wrappers widen to
 *     int64 where the replaced API returns int64);
 *   - "k-major" weights: W[k][o] = the layer's weight for input channel k and output channel o
 *     (the transpose of torch's (out, in) layout), produced by puflow_b200/packing.py.
 */
#ifndef PUFLOW_B200_H
#define PUFLOW_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PF_API __attribute__((visibility("default")))
#else
#define PF_API
#endif

#define PF_ABI_VERSION 10
#define PF_EINVAL 10001
#define PF_ENOMEM 10002

/* activation codes of pf_linear */
/* Precision of the tensor-core contractions (trailing `precision` argument of every tcgen05 entry point):
 *   PF_PREC_FP32_SPLIT  operands as fp16 hi + lo pairs, three products per contraction (hi.hi + lo.hi + hi.lo), fp32
 *                       accumulation: fp32-class results (2^-22 per product); the default and the path every parity
 *                       number refers to.
 *   PF_PREC_FP16        single pass: the hi.hi product alone (operands rounded to fp16 = 11 significant bits, the mantissa
 *                       width of TF32; fp32 accumulation): a third of the MMAs.  STATED TOLERANCE (tests/test_gpu_precision.py,
 *                       measured: 3.5e-3 / 6.1e-2 / 1.3e-2 for the three released checkpoints): network output within
 *                       1e-2 (pu1k), 3e-2 (pugeo), 1e-1 (pugan) normwise of the reference -- the split path's bar is 1e-4 --
 *                       and Chamfer distance to the reference output <= 1e-4 (pu1k; the split path meets 1e-5, this
 *                       mode does not: it is an option for the precision sweep, not the product path).  (pf_interp_tc keeps the weights split in this mode and drops the activations' lo
 *                       products only: 2 of 3 MMAs.) */
#define PF_PREC_FP32_SPLIT 0
#define PF_PREC_FP16 1

#define PF_ACT_NONE 0
#define PF_ACT_RELU 1
#define PF_ACT_LRELU 2

PF_API int pf_abi_version(void);
/* message of the last failing call on this host thread ("" if none) */
PF_API const char* pf_last_error(void);

/* ---- (1) kNN top-k ---------------------------------------------------------------------------
 * Replaces pytorch3d.ops.knn_points(p1, p2, K) at
 * /root/reference/modules/discrete/interpflow.py:104 (K=8), :228, :328 (K=16).
 * q[B][nq][3], p[B][np][3] -> idx[B][nq][K] int32 (and dist[B][nq][K] squared L2 if dist != NULL),
 * sorted ascending by (distance, index).  1 <= K <= 32, K <= np. */
PF_API int pf_knn_topk(const float* q, const float* p, int B, int nq, int np, int K, int32_t* idx, float* dist,
                void* stream);

/* ---- (5) Chamfer nearest-neighbour distance ----------------------------------------------------
 * Replaces chamfer_3D.forward / NmDistanceKernel,
 * /root/reference/metric/PyTorchCD/chamfer3D/chamfer3D.cu:12-154 (both directions, one call).
 * a[B][n][3], b[B][m][3] -> dist1[B][n], idx1[B][n] (nearest in b), dist2[B][m], idx2[B][m]
 * (nearest in a); squared distances; the lowest index wins ties (strict <, chamfer3D.cu:36-69,126).  dist2 and idx2 may both be NULL: only the a -> b direction is computed.
 * A direction with 256..8,192 candidates and >= 1,024 queries per cloud is pruned with a grid over the candidates (exact:
 * the same bits as the exhaustive scan, csrc/nn.cu); a non-finite candidate coordinate is never selected (as NaN / inf
 * distances never compare below), and idx is -1 where no candidate has a finite distance. */
PF_API int pf_chamfer_nn(const float* a, int n, const float* b, int m, int B, float* dist1, int32_t* idx1, float* dist2,
                  int32_t* idx2, void* stream);

/* ---- per-point linear layer ----------------------------------------------------------------------
 * out[m][0:N] = act(A[m][0:K] . W[K][N] + bias[N]); row strides lda / ldo in floats.  Replaces the
 * ATen Linear / 1x1 Conv2d calls that act on per-POINT tensors: FeatMergeUnit
 * (/root/reference/modules/discrete/interpflow.py:251-258), the first layers of LinearA1D
 * (:22-43), and the per-point projections the fused EdgeConv kernels gather from.
 * N % 4 == 0, lda % 4 == 0 unless K < 4, ldo % 4 == 0, 16-byte aligned pointers. */
PF_API int pf_linear(const float* A, int64_t lda, const float* W, const float* bias, float* out, int64_t ldo, int64_t M,
              int K, int N, int act, float slope, void* stream);

/* Tensor-core variant of pf_linear for K in {16, 32, 64, 128}, N % 16 == 0 (tcgen05.mma on fp16 hi/lo split operands,
 * fp32 accumulation: fp32-class result, tolerance in tests/test_gpu_kernels.py).  wtc: weights pre-split and laid out
 * per 128-column chunk by puflow_b200/packing.py pack_linear_tc: ceil(N/128) chunks of 128*K*4 bytes. */
PF_API int pf_linear_tc(const float* A, int64_t lda, const void* wtc, const float* bias, float* out, int64_t ldo, int64_t M,
                 int K, int N, int act, float slope, int precision, void* stream);

/* ---- (2) fused EdgeConv block -----------------------------------------------------------------
 * Replaces FeatureExtractUnit.forward (/root/reference/modules/discrete/interpflow.py:223-248):
 * neighbour gather + 4-layer dense block (1x1 conv + eval-BN + LeakyReLU + concat) + conv_out + max
 * over the K neighbours, for one block.  The 3C-wide edge-feature part of every conv is linear in
 * (x_i, x_j) and arrives pre-projected per point:
 *   pq[Bp*N][2R], R = 4G + C_out: columns [0,R) = P (own-point term incl. all biases),
 *   [R,2R) = Q (neighbour term); column order inside R: layer0 | layer1 | layer2 | layer3 | conv_out.
 *   idx[Bp*N][K] int32 neighbour index inside the patch;  wd: hidden-to-hidden weights of layers
 *   1..3, k-major [l*G][G] concatenated;  wo: [4G][C_out] k-major;  out[Bp*N][C_out].
 * (G, C_out) in {(8,32), (16,64), (32,128)}; K in {8,16,32}; slope = LeakyReLU slope (0.05). */
PF_API int pf_edgeconv(const float* pq, const int32_t* idx, const float* wd, const float* wo, float* out, int Bp, int N,
                int K, int G, int C_out, float slope, void* stream);

/* Tensor-core variant of pf_edgeconv for K = 16, (G, C_out) in {(8,32), (16,64), (32,128)}: the hidden-channel
 * contractions run on tcgen05.mma (fp16 hi/lo split operands, 3 MMAs per product, fp32 accumulation in TMEM; results
 * are fp32-class, tolerance in tests/test_gpu_kernels.py).  Same pq / idx / out as pf_edgeconv;  wtc: pre-split fp16
 * weights in the tensor-core operand layout (puflow_b200/packing.py pack_edgeconv_tc), 24 gp^2 + 2048 gp bytes with
 * gp = max(G, 16): PF_EDGECONV_TC_BLOB for G = 32. */
#define PF_EDGECONV_TC_BLOB 90112
PF_API int pf_edgeconv_tc(const float* pq, const int32_t* idx, const void* wtc, float* out, int Bp, int N, int K, int G,
                   int C_out, float slope, int precision, void* stream);

/* Pre-split activations.  pf_edgeconv_tc_split is pf_edgeconv_tc that ALSO (or only: out may be NULL) writes its output as
 * fp16 hi/lo in the tile-blocked tensor-core operand layout: out_split holds ceil(Bp*N / 128) tiles of
 * PF_SPLIT_TILE_BYTES(C_out) bytes; within a tile hi(r, k) at (k/8)*2048 + r*16 + (k%8)*2, lo at C_out*256 + the same.
 * pf_linear_tc_split is pf_linear_tc for an input in that layout (K = its channel count, 32 / 64 / 128): a tile is one
 * TMA bulk copy straight into the MMA operand, no conversion pass (the next block's P|Q projection).  a_split must be
 * 128-byte aligned, out 32-byte aligned, ldo a multiple of 8. */
#define PF_SPLIT_TILE_BYTES(K) (512 * (K))
PF_API int pf_edgeconv_tc_split(const float* pq, const int32_t* idx, const void* wtc, float* out, void* out_split, int Bp,
                         int N, int K, int G, int C_out, float slope, int precision, void* stream);
PF_API int pf_linear_tc_split(const void* a_split, const void* wtc, const float* bias, float* out, int64_t ldo, int64_t M,
                       int K, int N, int act, float slope, int precision, void* stream);

/* ---- (3) flow stack ---------------------------------------------------------------------------------
 * FlowBlock.forward / inverse x6 (/root/reference/modules/discrete/interpflow.py:66-82,302-321):
 * ActNorm (modules/flows/normalize.py:30-43), InvertibleConv1x1_1D (permutate.py:117-126), additive
 * spatial coupling (coupling.py:55-58,82-85,108-118), channel reverse (permutate.py:75-80), affine
 * injector (coupling.py:127-151).
 *
 * Conditioning is per INPUT point and shared by the forward pass and all R inverse copies:
 *   raw[nblocks][M][192] = cs_i . [W1c | Ws0 | Wb0]  (from pf_linear; no activation):
 *        [0,64) first-layer term of coupling1.bias_net, [64,128) / [128,192) first layers of
 *        coupling2.scale_net / bias_net;
 *   pf_flow_cond finishes the two coupling2 nets: sb[nblocks][M][8] = (s0,s1,s2,b0,b1,b2,0,0);
 *   cond_w[nblocks][PF_COND_BLOB] and flow_w[nblocks][PF_FLOW_BLOB]: packed small weights
 *   (layouts in puflow_b200/packing.py).
 * pf_flow_fwd: x[M][3] -> z[M][3]; ssum[M] = sum over blocks and channels of s (for the log-det).
 * pf_flow_inv: zin[M*R][3] (row n*R + r belongs to input point n) -> x[M*R][3]. */
#define PF_COND_BLOB 8840
#define PF_FLOW_BLOB 4480
PF_API int pf_flow_cond(const float* raw, const float* cond_w, float* sb, int64_t M, int nblocks, void* stream);
PF_API int pf_flow_fwd(const float* x, const float* raw, const float* sb, const float* flow_w, float* z, float* ssum,
                int64_t M, int nblocks, void* stream);
PF_API int pf_flow_inv(const float* zin, const float* raw, const float* sb, const float* flow_w, float* x, int64_t M,
                int R, int nblocks, void* stream);

/* Tensor-core variants (tcgen05.mma on fp16 hi/lo split operands for the 64x64 hidden layers of LinearA1D, fp32
 * accumulation; everything else fp32 as above; tolerance in tests/test_gpu_kernels.py).  Same arguments, except the
 * weights: cond_tc[nblocks][PF_COND_TC_BLOCK bytes], flow_tc[nblocks][PF_FLOW_TC_BLOCK bytes] from
 * puflow_b200/packing.py (pack_cond_tc / pack_flow_tc).  nblocks <= 6 for pf_flow_fwd_tc / pf_flow_inv_tc.
 * raw, sb and the blobs must be 32-byte aligned.  ldraw: row stride of raw in floats -- 192 for the layout above, 64
 * for the compact raw1[nblocks][M][64] written by pf_cond_chain_tc (only columns [0,64) are read by the flow).
 *
 * pf_cond_chain_tc fuses, for one block, FeatMergeUnit (interpflow.py:251-258), the conditioning first layers and
 * pf_flow_cond:  c[M][C] (EdgeConv output, C in {32,64,128}) -> raw1[M][64], sb[M][8].  conv2 of the merge unit (no
 * bias, no activation) is composed with the first layers on the host (packing.pack_chain_tc, PF_CHAIN_TC_BYTES(C)). */
#define PF_COND_TC_BLOCK 35360
#define PF_FLOW_TC_BLOCK 17920
PF_API int pf_flow_cond_tc(const float* raw, const void* cond_tc, float* sb, int64_t M, int nblocks, int precision, void* stream);
#define PF_CHAIN_TC_BYTES(C) (2 * (C) * (C) + 2 * 192 * (C) + 4 * 8192 + (64 + 648) * 4)
PF_API int pf_flow_fwd_tc(const float* x, const float* raw, int ldraw, const float* sb, const void* flow_tc, float* z,
                   float* ssum, int64_t M, int nblocks, int precision, void* stream);
PF_API int pf_flow_inv_tc(const float* zin, const float* raw, int ldraw, const float* sb, const void* flow_tc, float* x,
                   int64_t M, int R, int nblocks, int precision, void* stream);
PF_API int pf_cond_chain_tc(const float* c, int C, const void* chain_tc, float* raw1, float* sb, int64_t M, int precision, void* stream);

/* ---- (4) latent interpolation -------------------------------------------------------------------------
 * Replaces InterpolationModule.forward (/root/reference/modules/discrete/interpflow.py:162-186) with
 * KnnContextEncoder (:123-134), DistanceEncoder (:85-120) and WeightEstimationUnit (:137-159) fused:
 * per edge (k = 8 neighbours) context MLPs -> R logits -> softmax over the 8 neighbours -> weighted
 * sum of the neighbours' latents.
 *   xyz[Bp*N][3], z[Bp*N][3]; idx: int32, row stride ldidx (first 8 entries of each row are used);
 *   pq[Bp*N][640] per-point projections (P | Q, 320 each: dist-enc layer 0 (64) | dense layers (128) |
 *   composed context term (128)); w: packed weights, PF_INTERP_BLOB floats (puflow_b200/packing.py);
 *   out[Bp][N*R][3], row n*R + r (the order PointInterpFlow.g expects, interpflow.py:316).
 * 1 <= R <= 32. */
#define PF_INTERP_BLOB 46304
PF_API int pf_interp(const float* xyz, const float* z, const int32_t* idx, int ldidx, const float* pq, const float* w,
              float* out, int Bp, int N, int R, void* stream);

/* Tensor-core variant of pf_interp (tcgen05.mma on fp16 hi/lo split operands, fp32 accumulation; tolerance in
 * tests/test_gpu_kernels.py).  Needs no pre-projected pq: coordinates enter the MMAs through a "geo" K-slice.
 * wtc: PF_INTERP_TC_BLOB bytes from puflow_b200/packing.py pack_interp_tc. */
#define PF_INTERP_TC_BLOB 205440
PF_API int pf_interp_tc(const float* xyz, const float* z, const int32_t* idx, int ldidx, const void* wtc, float* out,
                 int Bp, int N, int R, int precision, void* stream);

/* ---- caller side ("next" rows of SURVEY.md section 8f) ---------------------------------------------------
 * pf_fps replaces pointnet2_ops.furthest_point_sample (/root/reference/modules/utils/patch.py:102,156)
 * with the semantics of the reference's own modules/utils/fps.py:40-73: start at index 0, running
 * min squared distance initialised to 1e10, arg-max with the first index winning ties.
 *   xyz[B][n][3] -> idx[B][npoint] int32;  work: NULL (brute force) or pf_fps_work_floats(B, n) floats of scratch.
 * pf_knn_patch replaces knn_cuda.KNN(k)(ref, query) (patch.py:33,107): for each of the nq queries the
 * k nearest of the np points, ascending by (distance, index); idx[B][nq][k] int32.  k <= 1024. */
PF_API int pf_fps(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, void* stream);
/* Same result as pf_fps, bit for bit, by exact spatial pruning (grid-sorted buckets with bounding boxes; a round only
 * touches buckets the new sample can change): one CTA per cloud, built for the merge FPS of a batch (8,216 of 40,960
 * candidates per cloud, patch.py:156).  work: pf_fps_work_floats(B, n) floats of scratch (sorted copy of the clouds).
 * pf_fps dispatches here when work != NULL and n >= 4096. */
PF_API long long pf_fps_work_floats(int B, int n);
PF_API int pf_fps_bucketed(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, void* stream);
/* Same result again, for FEW LARGE clouds: the cloud's buckets are spread over a thread-block cluster of 2..16 CTAs (sorted
 * coordinates and running minima in the cluster's shared memory, one DSMEM candidate exchange + remote mbarrier arrive per
 * round) whenever B * cluster <= SM count; otherwise (a batch of many clouds) it is pf_fps_bucketed.  Config 3's merge FPS
 * (20,024 of 99,840 candidates) and a single 2,048-pt cloud's (8,216 of 40,960) are the cases it is for.
 * work: pf_fps_clustered_work_floats(B, n) floats. */
PF_API long long pf_fps_clustered_work_floats(int B, int n);
PF_API int pf_fps_clustered(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, void* stream);
/* The same indices again, bit for bit, with the dependent chain taken out of the block-wide loop ("threshold queue"):
 * running minima only decrease, so the points above a threshold theta form a list that can only shrink, and while its
 * maximum exceeds theta the next sample is the arg-max of the LIST.  One warp draws ~70-100 samples in a row from a list of
 * <= 256 points held in its registers; all warps then apply those samples to the bucketed cloud in one pruned pass and
 * collect the next list.  One CTA per cloud, any batch size; clouds above pf_fps_queued_max_points() candidates go to
 * pf_fps_bucketed.  Few large clouds (n >= 65,536 and every cloud's 8-CTA cluster resident at once) run the same algorithm with
 * a cloud's buckets spread over a thread-block cluster (fps_cqueue_kernel: config 3's merge FPS).  work: pf_fps_work_floats(B, n) floats. */
PF_API int pf_fps_queued_max_points(void);
PF_API int pf_fps_queued(const float* xyz, int B, int n, int npoint, int32_t* idx, float* work, void* stream);
/* pf_fps_queued_static: no counterpart in the reference; added for callers that
 * cannot allocate device memory (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same inputs, outputs and indices as pf_fps_queued, but the scratch is a file-scope static __device__ array inside the
 * library (PF_FPS_STATIC_MAX_POINTS points x 20 B = 5 MiB, resident on every GPU that loads the library until exit): the
 * call allocates nothing.  Limit: B * ceil32(n) <= PF_FPS_STATIC_MAX_POINTS (e.g. one cloud of 99,840 candidates or six of
 * 40,960), otherwise PF_EINVAL with a message.  All calls on a device share the array: calls that overlap on the GPU race;
 * nothing in it needs clearing between calls (the kernel rewrites every entry it reads). */
#define PF_FPS_STATIC_MAX_POINTS 262144
PF_API int pf_fps_queued_static(const float* xyz, int B, int n, int npoint, int32_t* idx, void* stream);
PF_API int pf_knn_patch(const float* q, const float* p, int B, int nq, int np, int k, int32_t* idx, void* stream);

/* ---- caller glue as kernels ------------------------------------------------------
 * pf_normalize_pc     PatchHelper.normalize_pc (/root/reference/modules/utils/patch.py:168-178): pc[B][n][3] ->
 *                     out[B][n][3] = (pc - centroid) / max radius, centroid[B][3], radius[B] (either may be NULL).
 * pf_gather_normalize the patch gather of extract_knn_patch (patch.py:112-125) fused with the per-patch normalize_pc of
 *                     __upsampling_patches (patch.py:85): item (b, p) takes the points pc[b][idx[b][p][0..k)] of cloud b
 *                     (n points); out[B*items][k][3], centroid[B*items][3], radius[B*items].
 * pf_gather_points    index_points / gather_operation (patch.py:106,157): out[b][i] = pc[b][idx[b][i]], i < m, optionally
 *                     * scale[b] + shift[b][0..3) (the de-normalisation of patch.py:76; scale and shift both or neither).
 * pf_denorm_concat    torch.cat([generated, patch], 1) * radius + centroid (patch.py:90-91): out[items][nx + np][3].
 * pf_remove_outliers  patch.py:198-214 given dist1 = the squared distance of every point of sr to its nearest input point
 *                     (pf_chamfer_nn): drops the num_outliers largest (ties: the lowest index goes first, i.e. a stable
 *                     descending argsort), survivors keep their order; sr[B][n][3], dist1[B][n] -> out[B][n - num_outliers][3]. */
PF_API int pf_normalize_pc(const float* pc, int B, int n, float* out, float* centroid, float* radius, void* stream);
PF_API int pf_gather_normalize(const float* pc, const int32_t* idx, int B, int n, int items, int k, float* out, float* centroid,
                        float* radius, void* stream);
PF_API int pf_gather_points(const float* pc, const int32_t* idx, int B, int n, int m, const float* scale, const float* shift,
                     float* out, void* stream);
PF_API int pf_denorm_concat(const float* x, const float* patches, const float* centroid, const float* radius, int items, int nx,
                     int np, float* out, void* stream);
PF_API int pf_remove_outliers(const float* sr, const float* dist1, int B, int n, int num_outliers, float* out, void* stream);
/* pf_fps_park_near_origin: the candidate rule of UPSTREAM pointnet2_ops.furthest_point_sample (what
 * /root/reference/modules/utils/patch.py:102,156 imports; not vendored: `if (mag <= 1e-3) continue;` in its kernel), as a
 * pre-pass for any pf_fps* entry point: out = xyz with every point i > 0 whose squared norm is <= thresh2 moved onto point 0.
 * Such a point's running minimum is 0 after the first round, so it is never selected while any other candidate is left, and
 * it is never a sample, so its coordinates influence nothing -- the indices are those of an FPS that skips it.  The default
 * semantics of this library (and of the oracle and the golden fixtures) are the reference's own modules/utils/fps.py:40-73,
 * which has no such rule; this is the opt-in (modules/utils/patch.py FPS_SKIP_NEAR_ORIGIN).  xyz, out: [B][n][3]. */
PF_API int pf_fps_park_near_origin(const float* xyz, int B, int n, float thresh2, float* out, void* stream);

/* ---- whole network in one call --------------------------------------------------------------------------------
 * pf_upsample_patches replaces PointInterpFlow.forward (/root/reference/modules/discrete/interpflow.py:327-337:
 * knn_points -> feat_extract -> log_prob -> interp -> g) for a batch of patches:
 *   xyz[Bp][N][3] (normalised patch coordinates) -> x[Bp][N*R][3] (row n*R + r), logp[1] (may be NULL: -mean log-likelihood,
 *   interpflow.py:339-345).  16 <= N <= 65535, 1 <= R <= 32.
 * It launches this library's tensor-core kernels on `stream`, allocates nothing and never synchronises (CUDA-graph
 * capturable).  workspace: pf_upsample_workspace_bytes(Bp, N, R) bytes of device scratch, 256-byte aligned.
 * model: the packed weights as ONE device blob of pf_model_bytes() bytes, 256-byte aligned; pf_model_entry(i, ...) for
 * i in [0, PF_MODEL_ENTRIES) gives the name (<= 31 chars), byte offset and size of every piece, in blob order:
 *   feat0.wpq (fp32 [3][128]) | per block i: feat<i>.bpq (fp32 [2R]) | feat<i>.wpq_tc (i >= 1, pf_linear_tc layout) |
 *   feat<i>.wtc (pf_edgeconv_tc) | chain<i> (pf_cond_chain_tc) | flow_tc (6 x PF_FLOW_TC_BLOCK) | interp_tc | logdet_const
 * (puflow_b200/packing.py pack_model_blob fills it from a reference state dict and can save it to a file). */
#define PF_MODEL_ENTRIES 27
PF_API int64_t pf_model_bytes(void);
PF_API int pf_model_entry(int i, char* name32, int64_t* offset, int64_t* bytes);
PF_API int64_t pf_upsample_workspace_bytes(int Bp, int N, int R);
PF_API int pf_upsample_patches(const void* model, const float* xyz, int Bp, int N, int R, void* workspace,
                        int64_t workspace_bytes, float* x, float* logp, int precision, void* stream);

/* pf_upsample_clouds replaces the body of the reference's per-file loop (/root/reference/modules/discrete/upsample.py:46-54:
 * PatchHelper.upsample(jitter=False) + remove_outliers) for a batch of clouds:
 *   clouds[B][N][3] -> out[B][npoint - num_outliers][3]:  normalise -> n_patch seeds by FPS -> patch kNN (patch_points per
 *   patch) -> per-patch normalise -> pf_upsample_patches -> de-normalise + append the input patch -> merge FPS to npoint
 *   points (the reference: N*R + num_outliers) -> back to the input frame -> Chamfer pass against the input -> drop the
 *   num_outliers farthest points.
 * n_patch is the caller's int(N / patch_points * expand_ratio) (patch.py:101).  Every launch goes to `stream`; nothing is
 * allocated and nothing synchronises (CUDA-graph capturable).  workspace: pf_upsample_clouds_workspace_bytes(...) bytes,
 * 256-byte aligned.  num_outliers = 0 skips the Chamfer pass and writes the npoint merged points directly
 * (= PatchHelper.upsample alone). */
PF_API int64_t pf_upsample_clouds_workspace_bytes(int B, int N, int R, int npoint, int patch_points, int n_patch);
PF_API int pf_upsample_clouds(const void* model, const float* clouds, int B, int N, int R, int npoint, int num_outliers,
                       int patch_points, int n_patch, void* workspace, int64_t workspace_bytes, float* out, int precision,
                       void* stream);

/* Optional per-launch timing of the one-call entry points (what bench.py's breakdown and roofline read): while enabled,
 * pf_upsample_patches / pf_upsample_clouds record a CUDA event on their stream after each of their launches; every interval
 * carries the launch's name ("knn_topk_k16", "linear_split_k128_n512", "edgeconv_tc_g32_c128_k16", "cond_chain_tc_c128",
 * "flow_fwd_tc", "logp", "interp_tc", "flow_inv_tc"; "normalize_pc", "fps_n<candidates>", "gather_points", "knn_patch",
 * "gather_normalize", "denorm_concat", "chamfer_nn", "remove_outliers").  pf_profile_intervals: how many were recorded
 * since pf_profile_enable(1) / pf_profile_clear; pf_profile_interval(i, name40, ms) waits for interval i and returns its
 * name and duration.  The record is per HOST THREAD (thread_local): threads driving different streams do not interfere.
 * Off by default; do not enable during CUDA-graph capture. */
PF_API int pf_profile_enable(int on);
PF_API int pf_profile_intervals(void);
PF_API int pf_profile_interval(int i, char* name40, float* ms);
PF_API int pf_profile_clear(void);

/* ---- continuous (CNF) variant, inference ------------------------------------------------
 * pf_cnf_flow replaces one FlowBlock of /root/reference/modules/continuous/interpflow.py:19-50 at inference time
 * (batch_norm=False: identity norms): CNF.forward (modules/continuous/cnf.py:51-112) -> torchdiffeq.odeint(dopri5) over
 * ODEfunc / ODEnet / ConcatSquashLinear (modules/continuous/odefunc.py:66-111,121-143, diffeq_layers.py:72-86).
 *   y_in[M][3] -> y_out[M][3] = y(t1) of dy/dt = ODEnet([t, c], y), y(t0) = y_in; forward pass t0 = 0, t1 = T =
 *   sqrt_end_time^2, inverse t0 = T, t1 = 0.  Point p uses conditioning row p / rep (rep = 1 forward, = the up-sampling ratio
 *   on the inverse: the reference's repeat_interleave(c, R)).
 *   ctx[ceil(M / rep)][ld_ctx >= PF_CNF_CTX]: the c-dependent part of every layer's gate and bias, one row per INPUT point:
 *   c . Wctx + bctx with (Wctx, bctx, wblob) from puflow_b200/packing.py pack_cnf_block (computed with pf_linear[_tc]).
 *   wblob: PF_CNF_BLOB bytes.  nfe: NULL or int32[M], function evaluations per point (negative: step budget exhausted).
 * Adaptive Dormand-Prince 5(4), atol / rtol as the reference passes them (1e-5); every point controls its own step and the
 * last step is clipped at t1 -- torchdiffeq (absent from the reference tree: solver parity unpinned) uses one batch-wide
 * error norm and interpolates at t1; oracle/cnf_oracle.py restates both and a solver-independent RK4 reference.
 * Not built: log p (Hutchinson trace estimate, odefunc.py:8-32) and the backward pass. */
#define PF_CNF_CTX 264
#define PF_CNF_BLOB 19520
PF_API int pf_cnf_flow(const float* y_in, const float* ctx, int ld_ctx, const float* wblob, float* y_out, int32_t* nfe, int M,
                       int rep, float t0, float t1, float atol, float rtol, void* stream);

/* ---- on-disk format (host code, no CUDA) ------------------------------------------
 * The reference reads clouds with np.loadtxt(path, dtype=np.float32) (/root/reference/modules/discrete/upsample.py:42)
 * and writes them with np.savetxt(path, pred, fmt='%.6f') (:57): whitespace-separated text, one point per line,
 * '#' comments.  Same values bit for bit (parsed as double, narrowed to float) and the same bytes on output
 * ('%.6f' of the widened value, ' ' between columns, '\n' after each row).
 *   pf_xyz_count: number of data lines and columns of the file;  pf_xyz_read: fills out[rows][cols] (HOST memory);
 *   pf_xyz_write: data[rows][cols] (HOST memory) -> file. */
PF_API int pf_xyz_count(const char* path, int64_t* rows, int* cols);
PF_API int pf_xyz_read(const char* path, float* out, int64_t rows, int cols);
PF_API int pf_xyz_write(const char* path, const float* data, int64_t rows, int cols);

#ifdef __cplusplus
}
#endif
#endif /* PUFLOW_B200_H */
