/* dpvo_b200.h — C ABI of libdpvo_b200.so (sm_100a).
 *
 * Drop-in boundary for DPVO's per-iteration update hot path.  Every entry point
 * takes plain device pointers, sizes, element strides and a CUDA stream
 * (`void* stream` == cudaStream_t).  No torch types.  No allocation happens
 * behind this ABI: outputs and workspaces are caller-provided.
 *
 * Return value: 0 on success, non-zero on failure; dpvo_last_error() gives the
 * message (thread-local).  Nothing here synchronises the stream.
 *
 * Reference interfaces replaced (paths under /root/reference/methods/dpvo/dpvo):
 *   altcorr/correlation.cpp:28-63   cuda_corr.{forward,backward,patchify_forward,patchify_backward}
 *   fastba/ba.cpp:32-97,183-189     cuda_ba.{forward,neighbors,reproject}
 *   lietorch/src/lietorch.cpp:18-316 lietorch_backends.{expm,logm,inv,mul,adj,adjT,act,act4,...}
 */
#ifndef DPVO_B200_H
#define DPVO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DPVO_B200_ABI_VERSION 5

typedef enum { DPVO_F16 = 0, DPVO_F32 = 1, DPVO_F64 = 2 } dpvo_dtype;

int dpvo_abi_version(void);
const char* dpvo_last_error(void);
/* Number of CUDA kernels this library has launched so far in this process (for bench.py's
 * gpu_launches; memsets are not counted). */
uint64_t dpvo_launch_count(void);

/* ------------------------------------------------------------------ altcorr */

/* One pyramid level of the correlation lookup.  Strides are in ELEMENTS. */
typedef struct {
  const void* fmap;      /* [B, N2, C, H2, W2] feature ring (logical shape), any strides   */
  int64_t sb, sn, sc, sy, sx;
  int32_t N2, H2, W2;
  float coord_scale;     /* coords are multiplied by this before use (1, 0.25)             */
  void* out;             /* out[(b*M+m)*out_edge_stride + o*out_elem_stride],              */
  int64_t out_edge_stride; /* o = ((xoff*(D-1) + yoff)*PH + i0)*PW + j0, D = 2*radius+2    */
  int64_t out_elem_stride;
} dpvo_corr_level;

/* Correlation forward — replaces cuda_corr.forward (altcorr/correlation.cpp:28-35,
 * correlation_kernel.cu:83-137 raw window + :222-233 bilinear blend + permute).
 * fmap1: [B, N1, C, PH, PW] patch features (gmap), element strides g_*.
 * coords: [B, M, 2, PH, PW] float32 contiguous.  us/vs: [M] int64 (patch / frame index).
 * nlev levels are evaluated in ONE launch (nlev=1 reproduces the reference call;
 * nlev=2 with out pointers interleaved gives DPVO.corr's stacked [E,882] layout,
 * dpvo.py:200-207).  dtype is the fmap/out element type. */
int dpvo_corr_forward(dpvo_dtype dtype,
                      const void* fmap1, int64_t g_sb, int64_t g_sn, int64_t g_sc, int64_t g_sy, int64_t g_sx,
                      int32_t N1, int32_t C, int32_t PH, int32_t PW,
                      const float* coords, const int64_t* us, const int64_t* vs,
                      int32_t B, int32_t M, int32_t radius,
                      const dpvo_corr_level* levels, int32_t nlev,
                      void* stream);

/* Locality order for the lookup (csrc/corr_order.cu): counting sort of the B = 1 edge list by (target-frame slot vs[e],
 * tile x tile pixel tile of the patch centre at level 1) so that the warps resident on an SM read the same part of the
 * same frame.  perm: [M] int32 out; ws: dpvo_corr_order_workspace_bytes(...) bytes.  No host synchronisation. */
int64_t dpvo_corr_order_workspace_bytes(int64_t M, int32_t N2, int32_t H1, int32_t W1, int32_t tile);
int dpvo_corr_order(const float* coords, const int64_t* vs, int64_t M, int32_t N2, int32_t H1, int32_t W1,
                    int32_t tile, int32_t* perm, void* ws, int64_t ws_bytes, void* stream);
/* dpvo_corr_forward walked in the order `perm`; each edge's result is written at its own index (bit-identical output). */
int dpvo_corr_forward_ordered(dpvo_dtype dtype,
                              const void* fmap1, int64_t g_sb, int64_t g_sn, int64_t g_sc, int64_t g_sy, int64_t g_sx,
                              int32_t N1, int32_t C, int32_t PH, int32_t PW,
                              const float* coords, const int64_t* us, const int64_t* vs,
                              int32_t B, int32_t M, int32_t radius,
                              const dpvo_corr_level* levels, int32_t nlev, const int32_t* perm, void* stream);

/* Name of the kernel the last dpvo_corr_forward call launched (static string; instrumentation for bench.py). */
const char* dpvo_corr_last_kernel(void);

/* Correlation backward — replaces cuda_corr.backward (correlation_kernel.cu:140-191,237-287).
 * corr_grad: logical [B, M, D-1(x), D-1(y), PH, PW] with element strides cg_*.
 * fmap1_grad / fmap2_grad: contiguous, same logical shape as the fmaps, PRE-ZEROED
 * by the caller; accumulated in `acc` dtype buffers (float32 when dtype is F16:
 * acc1/acc2 are float32 scratch of the same numel, else NULL and the grads are
 * accumulated in place). */
int dpvo_corr_backward(dpvo_dtype dtype,
                       const void* fmap1, int64_t g_sb, int64_t g_sn, int64_t g_sc, int64_t g_sy, int64_t g_sx,
                       const void* fmap2, int64_t sb, int64_t sn, int64_t sc, int64_t sy, int64_t sx,
                       int32_t N1, int32_t N2, int32_t C, int32_t PH, int32_t PW, int32_t H2, int32_t W2,
                       const float* coords, const int64_t* us, const int64_t* vs,
                       int32_t B, int32_t M, int32_t radius,
                       const void* corr_grad, int64_t cg_sb, int64_t cg_sm, int64_t cg_sx, int64_t cg_sy,
                       int64_t cg_si, int64_t cg_sj,
                       void* fmap1_grad, void* fmap2_grad, float* acc1, float* acc2,
                       void* stream);

/* Same operator, training fast path: fp16 or fp32 features, C = 128, 3x3 patches, radius 3 (the only
 * shape net.py:156-169 produces).  fmap1: dense patches [B][N1][128][3][3] (batch / patch strides g_sb,
 * g_sn); fmap2: CHANNELS-LAST ring (channel stride 1; strides sb, sn, sy, sx in elements) — the layout the
 * forward lookup already reads; fmap2_grad_nhwc: contiguous [B][N2][H2][W2][128] (return it to autograd as
 * a permuted view); fmap1_grad: contiguous like fmap1.
 * F16: acc1 ([B*N1*1152]) and acc2 ([B*N2*H2*W2*128]) are ZEROED float32 scratch and both gradients are
 * written in full.  F32: acc1/acc2 are ignored, the gradients must be PRE-ZEROED and are accumulated in place.
 * perm (optional): processing order from dpvo_corr_order — edges of one target frame / tile back to back keep the
 * gradient lines they reduce into resident in L2 (the sums are order-independent up to fp32 rounding). */
int dpvo_corr_backward_nhwc(
    dpvo_dtype dtype,
    const void* fmap1, int64_t g_sb, int64_t g_sn,
    const void* fmap2, int64_t sb, int64_t sn, int64_t sy, int64_t sx,
    int32_t N1, int32_t N2, int32_t H2, int32_t W2,
    const float* coords, const int64_t* us, const int64_t* vs, int32_t B, int32_t M,
    const void* corr_grad, int64_t cg_sb, int64_t cg_sm, int64_t cg_sx, int64_t cg_sy, int64_t cg_si, int64_t cg_sj,
    void* fmap1_grad, void* fmap2_grad_nhwc, float* acc1, float* acc2, const int32_t* perm /* optional, dpvo_corr_order */,
    void* stream);

/* Patch gather — replaces cuda_corr.patchify_forward (correlation_kernel.cu:17-48,289-308).
 * net: [B, C, H, W] strides n_*; coords: [B, M, 2] float32 contiguous;
 * patches: [B, M, C, D, D] contiguous (D = 2*radius+2), fully written (zeros outside). */
int dpvo_patchify_forward(dpvo_dtype dtype, const void* net,
                          int64_t n_sb, int64_t n_sc, int64_t n_sy, int64_t n_sx,
                          int32_t B, int32_t C, int32_t H, int32_t W,
                          const float* coords, int32_t M, int32_t radius,
                          void* patches, void* stream);

/* Replaces cuda_corr.patchify_backward (correlation_kernel.cu:50-81,311-334).
 * net_grad: [B, C, H, W] contiguous, PRE-ZEROED.  patch_grad: [B,M,C,D,D] contiguous. */
int dpvo_patchify_backward(dpvo_dtype dtype, const void* patch_grad,
                           const float* coords, int32_t B, int32_t M, int32_t C,
                           int32_t H, int32_t W, int32_t radius,
                           void* net_grad, void* stream);

/* Layout helper for the feature ring: src [N, C, HW] -> dst [N, HW, C] (fp16/fp32). */
int dpvo_nchw_to_nhwc(dpvo_dtype dtype, const void* src, void* dst,
                      int64_t N, int32_t C, int64_t HW, void* stream);

/* Incremental form of the above for a ring the CALLER keeps in the reference's NCHW layout
 * (dpvo/dpvo.py:72-73; its slot writes at :435-438 and :296-299 are plain torch copies): hashes
 * every slot of `src` (one read pass), and re-transposes into `dst` only the slots whose hash
 * differs from the previous call's.  `hashes` = 2*N uint64 of device memory owned by the caller,
 * filled with 0xFF bytes before the first call; `parity` alternates 0,1,0,... between calls on
 * the same ring.  No host synchronisation. */
int dpvo_ring_sync(dpvo_dtype dtype, const void* src, void* dst, int64_t N, int32_t C, int64_t HW,
                   void* hashes, int32_t parity, void* stream);

/* ------------------------------------------------------------------- fastba */

/* Bytes of device workspace dpvo_ba_forward needs for a problem of this size.
 * n_patches_total = rows of the patch buffer (patches.numel()/(3*P*P)); it bounds kk. */
int64_t dpvo_ba_workspace_bytes(int64_t n_edges, int64_t n_patches_total, int32_t n_poses_opt,
                                int32_t eff_impl, int32_t ppf, int64_t n_frames_total);

/* Bundle adjustment — replaces cuda_ba.forward (fastba/ba.cpp:32-45, ba_cuda.cu:433-582,
 * block_e.cu).  `iterations` Gauss-Newton steps on poses[t0,t1) and the inverse depths of
 * every patch referenced by kk; poses [n_frames_total,7] and patches [n_patches_total,3,P,P]
 * are updated IN PLACE.  intrinsics: row 0 is used for every frame (ba_cuda.cu:255-258).
 * target/weight: [E,2] float32.  lmbda: device pointer to one float.  ii/jj/kk: [E] int64.
 * eff_impl selects the block-E ("efficient E") Schur path used for global BA; it assumes
 * kk == ii*ppf + slot like the reference (block_e.cu:57-62).  workspace: >= the size
 * returned by dpvo_ba_workspace_bytes, 256-byte aligned, contents irrelevant. */
int dpvo_ba_forward(float* poses, float* patches, const float* intrinsics,
                    const float* target, const float* weight, const float* lmbda,
                    const int64_t* ii, const int64_t* jj, const int64_t* kk,
                    int64_t n_edges, int64_t n_frames_total, int64_t n_patches_total, int32_t P,
                    int32_t ppf, int32_t t0, int32_t t1, int32_t iterations, int32_t eff_impl,
                    void* workspace, int64_t workspace_bytes, void* stream);
/* The same call in two parts.  dpvo_ba_prepare does what only depends on the edge list and the window
 * (set of patches seen, compact numbering, per-patch edge lists, cleared accumulators) and may run ahead
 * on another stream; dpvo_ba_forward_prepared then runs the iterations on the SAME workspace. */
int dpvo_ba_prepare(const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t n_edges,
                    int64_t n_frames_total, int64_t n_patches_total, int32_t P, int32_t ppf, int32_t t0, int32_t t1,
                    int32_t eff_impl, void* workspace, int64_t workspace_bytes, void* stream);
int dpvo_ba_forward_prepared(float* poses, float* patches, const float* intrinsics,
                    const float* target, const float* weight, const float* lmbda,
                    const int64_t* ii, const int64_t* jj, const int64_t* kk,
                    int64_t n_edges, int64_t n_frames_total, int64_t n_patches_total, int32_t P,
                    int32_t ppf, int32_t t0, int32_t t1, int32_t iterations, int32_t eff_impl,
                    void* workspace, int64_t workspace_bytes, void* stream);

/* Temporal neighbours — replaces cuda_ba.neighbors (fastba/ba.cpp:59-97), entirely on the
 * device (the reference round-trips through the host).  For edge e: ix[e] / jx[e] = index of
 * the previous / next edge with the same ii value when that group is stably sorted by jj;
 * -1 at the ends.  Bit-exact.  Keys may be any int64 except INT64_MIN-like 0x8080808080808080.
 * workspace: dpvo_neighbors_workspace_bytes(n) bytes. */
int64_t dpvo_neighbors_workspace_bytes(int64_t n);
int dpvo_neighbors(const int64_t* ii, const int64_t* jj, int64_t n,
                   int64_t* ix, int64_t* jx, void* workspace, int64_t workspace_bytes, void* stream);

/* Sorted unique + inverse of a non-negative int64 key array bounded by key_bound
 * (what torch::_unique(kk, true, true) yields at ba_cuda.cu:447-449): uniq[0..*n_uniq),
 * inverse[n].  n_uniq is a device int32.  workspace: dpvo_unique_workspace_bytes(key_bound). */
int64_t dpvo_unique_workspace_bytes(int64_t key_bound);
int dpvo_unique_inverse(const int64_t* keys, int64_t n, int64_t key_bound,
                        int64_t* uniq, int64_t* inverse, int32_t* n_uniq,
                        void* workspace, int64_t workspace_bytes, void* stream);

/* Dense SPD solve S x = y with the blocked Cholesky the BA uses for 6N > 127 (csrc/ba_chol.cu) —
 * what at::linalg_cholesky_ex + torch::cholesky_solve do at ba_cuda.cu:547-548,561-562.  S: [n,n]
 * row-major float32 (lower triangle read), y: [n], x: [n] out, info: device int32, 0 = ok, 1 = a
 * pivot was not positive (x is then zero).  workspace: dpvo_chol_workspace_bytes(n). */
int64_t dpvo_chol_workspace_bytes(int32_t n);
int dpvo_chol_solve(const float* S, const float* y, int32_t n, float* x, int32_t* info,
                    void* workspace, int64_t workspace_bytes, void* stream);

/* dpvo_chol_solve without a workspace argument and without any allocation (no cudaMalloc / cudaFree, no stream or
 * event creation): for callers that cannot allocate — CUDA graph capture, serving loops, harnesses that forbid it.
 * No counterpart in the reference.  The scratch is a file-scope
 * device array of the library (about 2.6 MB, held by every GPU that loads the library until the process exits), so
 * n <= dpvo_chol_solve_noalloc_max_n() (768 = 128 free poses); larger n returns 1 with a message in
 * dpvo_last_error().  All calls on a device share that array: calls that overlap on the GPU (different streams
 * without an ordering between them) race; calls on one stream, or ordered by events, are safe.  Same arguments,
 * results and info semantics as dpvo_chol_solve otherwise. */
int32_t dpvo_chol_solve_noalloc_max_n(void);
int dpvo_chol_solve_noalloc(const float* S, const float* y, int32_t n, float* x, int32_t* info, void* stream);

/* Arithmetic of the blocked Cholesky's bulk trailing update (the 128x128 tiles used while the trailing matrix has at
 * least one tile per SM, i.e. global bundle adjustment; smaller systems never reach it): 0 (default) = split-TF32 on
 * the tensor cores (each fp32 operand as hi + lo TF32 halves, three mma products per term, fp32 accumulation — fp32
 * results to ~2^-21 per product), 1 = fp32 FMA.  Process-wide switch for A/B measurements and the accuracy test;
 * where the reference calls cuSOLVER potrf (ba_cuda.cu:547) there is no such choice. */
int dpvo_chol_bulk_mode(int32_t mode);

/* ---- Update operator (net.py:23-88, blocks.py:15-48): edge-MLP layers on the tensor cores -------
 * One nn.Linear(K, 384) over E edge rows with the row-local work that follows it in the reference
 * fused into the epilogue:
 *     v = A W^T + bias;  v *= gate;  v += res1;  v += res2[res2_index ? res2_index[row] : row];
 *     v = LayerNorm(v; ln_gamma, ln_beta, ln_eps) when ln_gamma != NULL;  v = act(v);  Y = fp16(v)
 * A: fp16 [M, K], row stride lda elements; W: fp16 [384, K] (nn.Linear weight layout), row stride
 * ldw; strides are multiples of 8 elements and pointers 16-byte aligned (TMA); K is arbitrary (the
 * tail is zero-filled).  gate / res1 / Y: fp16 [M, 384] contiguous; res2: fp16 [*, 384] rows
 * gathered through the int64 index.  bias / ln_gamma / ln_beta: float32 [384].  fp32 accumulation.
 * act: 0 none, 1 ReLU, 2 sigmoid.  Replaces the nn.Linear / nn.LayerNorm / ReLU / Sigmoid / add
 * sequences of Update.forward (net.py:69-88) and GatedResidual.forward (blocks.py:30-31). */
typedef struct dpvo_linear_params {
  const void* A; int64_t lda;
  const void* W; int64_t ldw;
  const float* bias;
  const void* gate;
  const void* res1;
  const void* res2; const int64_t* res2_index;
  const float* ln_gamma; const float* ln_beta; float ln_eps;
  int32_t act;
  void* Y;
  int64_t M; int32_t K;
  /* optional (K = 384, no LayerNorm): the operator's two 384->2 heads on ReLU(v) (net.py:57-66,88):
   * head_acc[row][h] += sum_n ReLU(v[row][n]) head_w[h][n], h = (d0, d1, w0, w1); head_acc is
   * float32 [M,4], zeroed by the caller, finished by dpvo_update_heads_finish. */
  const float* head_w; float* head_acc;
} dpvo_linear_params;
int dpvo_linear384_f16(const dpvo_linear_params* p, void* stream);
/* Diagnostics: 16 per-role cycle counters of CTA 0 of the last K=384 launch made with the
 * environment variable DPVO_LINEAR_PROFILE=1 in a library built with -DDPVO_DEBUG_HOOKS (csrc/update_gemm.cu).
 * Synchronises the device. */
int dpvo_linear_profile_read(long long* out16);

/* out[r] = index[r] >= 0 ? src[index[r]] : 0 for fp16 rows of `dim` elements (dim % 8 == 0) —
 * `mask * net[:, ix]` of Update.forward (net.py:76-80; ix from fastba.neighbors, -1 = no neighbour). */
int dpvo_gather_rows_f16(const void* src, const int64_t* index, int64_t n_rows, int32_t dim, void* out,
                         void* stream);

/* SoftAgg's scatter_softmax + scatter_sum (blocks.py:42-50, torch_scatter.py:17-132) over fp16
 * [E,384] rows f and g grouped by ids[e] in [0, group_bound) (any dense labelling of the groups, e.g.
 * the inverse of dpvo_unique_inverse):  y[grp] = sum_e f_e softmax_grp(g)_e.  y: fp16
 * [group_bound,384]; rows of empty groups are left untouched.  fp32 arithmetic, running maximum.
 * workspace: dpvo_softagg_workspace_bytes(E, group_bound).  The bucket sort of the edges by group
 * depends on ids only: dpvo_softagg_prepare does it ahead of time (e.g. on another stream while the
 * layers that produce f and g run) and dpvo_softagg_f16(prepared = 1) then reuses the workspace;
 * with prepared = 0 the sort happens inside the call (ids required). */
int64_t dpvo_softagg_workspace_bytes(int64_t n_edges, int64_t group_bound);
int dpvo_softagg_prepare(const int64_t* ids, int64_t n_edges, int64_t group_bound, void* workspace,
                         int64_t workspace_bytes, void* stream);
int dpvo_softagg_f16(const void* f, const void* g, const int64_t* ids, int64_t n_edges, int64_t group_bound,
                     void* y, int32_t prepared, void* workspace, int64_t workspace_bytes, void* stream);

/* out[e] = net[e] + hy[ids[e]] (SoftAgg's expand `h(y)[:, jx]` + the residual add, net.py:82-83),
 * followed by LayerNorm(ln_gamma, ln_beta, ln_eps) when ln_gamma != NULL (gru.0, net.py:42).
 * fp16 [*,384] rows, float32 gamma/beta. */
int dpvo_expand_add_f16(const void* net, const void* hy, const int64_t* ids, int64_t n_edges,
                        const float* ln_gamma, const float* ln_beta, float ln_eps, void* out, void* stream);

/* delta = Linear(ReLU(net)), weight = sigmoid(Linear(ReLU(net))) (net.py:57-66,88): net fp16 [E,384],
 * wd / ww float32 [2,384], bd / bw float32 [2]; delta, weight: float32 [E,2]. */
int dpvo_update_heads_f16(const void* net, const float* wd, const float* bd, const float* ww, const float* bw,
                          int64_t n_edges, float* delta, float* weight, void* stream);
/* Same result from the dot products the last layer accumulated (dpvo_linear_params.head_acc):
 * delta = acc[:, 0:2] + bd, weight = sigmoid(acc[:, 2:4] + bw). */
int dpvo_update_heads_finish(const float* acc, const float* bd, const float* bw, int64_t n_edges,
                             float* delta, float* weight, void* stream);

/* The whole operator (net.py:69-88) in one call: the same launches update.py would issue one ctypes call at a time
 * (csrc/update_forward.cu), without the host time of ~45 Python round trips.  Layer order of W / ldw / bias:
 *   c1.0 c1.2 c2.0 c2.2 agg_kk.f agg_kk.g agg_kk.h agg_ij.f agg_ij.g agg_ij.h gru.1.gate.0 gru.1.res.0 gru.1.res.2
 *   gru.3.gate.0 gru.3.res.0 gru.3.res.2 corr.0 corr.2 corr.5          (reference state-dict names, net.py:23-66)
 * LayerNorm order of ln_gamma / ln_beta: norm, gru.0, gru.2, corr.3.  W: fp16 [384, K] rows ldw apart (K = 384, corr.0:
 * corr_features = 882); bias / gamma / beta: float32 [384]; head_w: float32 [4, 384] (d.1 rows then w.1 rows);
 * head_bias_d / head_bias_w: float32 [2]. */
typedef struct {
  const void* W[19]; int64_t ldw[19]; const float* bias[19];
  const float* ln_gamma[4]; const float* ln_beta[4]; float ln_eps;
  int32_t corr_features;
  const float* head_w; const float* head_bias_d; const float* head_bias_w;
} dpvo_update_weights;
/* net: fp16 [E,384]; table: fp16 [*,384] context rows, gathered with table_index [E] (NULL: row e); corr: fp16
 * [E, corr_features] rows ld_corr elements apart (16-byte aligned rows); ii / jj / kk: int64 [E]; n_frames /
 * n_patches: exclusive upper bounds of the frame / patch indices.  Outputs: net_out fp16 [E,384] (must not alias an
 * input), delta / weight float32 [E,2].  workspace: dpvo_update_workspace_bytes(E, n_frames, n_patches). */
int64_t dpvo_update_workspace_bytes(int64_t n_edges, int64_t n_frames, int64_t n_patches);
int dpvo_update_forward(const dpvo_update_weights* w, const void* net, const void* table, const int64_t* table_index,
                        const void* corr, int64_t ld_corr, const int64_t* ii, const int64_t* jj, const int64_t* kk,
                        int64_t n_edges, int64_t n_frames, int64_t n_patches, void* net_out, float* delta, float* weight,
                        void* workspace, int64_t workspace_bytes, void* stream);

/* cuda_ba.solve_system (fastba/ba.cpp:99-180): one Gauss-Newton step of the classic loop-closure
 * pose graph over Sim(3).  J_Ginv_i / J_Ginv_j: float32 [r,7,7] blocks of edge x at poses ii[x] / jj[x]
 * (int64), res: float32 [r,7].  A = J^T J, A.diag += A.diag * lm, A.diag += ep, b = -J^T res; the first
 * `freen` poses are solved for (all n_poses when freen < 0), the others stay fixed.  delta: float32
 * [n_poses,7] (zeros for fixed poses).  info (device int32): 0 ok, 1 factorisation failed, 2 self edge
 * or negative index (the reference exit(1)s) — delta is zero then.  Normal equations in float64,
 * float32 blocked Cholesky + `refinements` rounds of float64-residual iterative refinement (0..4).
 * workspace: dpvo_pgo_workspace_bytes(n_poses, freen). */
int64_t dpvo_pgo_workspace_bytes(int32_t n_poses, int32_t freen);
int dpvo_pgo_solve(const float* J_Ginv_i, const float* J_Ginv_j, const int64_t* ii, const int64_t* jj,
                   const float* res, int64_t n_edges, int32_t n_poses, float ep, float lm, int32_t freen,
                   int32_t refinements, float* delta, int32_t* info, void* workspace, int64_t workspace_bytes,
                   void* stream);

/* Reprojection of all PxP pixels of patch kk[e] from frame ii[e] into frame jj[e] —
 * replaces cuda_ba.reproject (ba_cuda.cu:379-429,585-617).  coords: [E,2,P,P] float32. */
int dpvo_reproject(const float* poses, const float* patches, const float* intrinsics,
                   const int64_t* ii, const int64_t* jj, const int64_t* kk,
                   int64_t n_edges, int32_t P, float* coords, void* stream);

/* --------------------------------------------------------- projective_ops */

/* Fused projective transform — replaces the inference path of projective_ops.transform
 * (projective_ops.py:53-108): coords of every pixel of patch kk[e] (source frame ii[e]) in frame jj[e],
 * per-frame intrinsics, quaternions normalised on load, depth clamp 1 / max(Z, 0.1).
 * coords: [E,P,P,2] (or [E,P,P,3] with the inverse depth when depth != 0).  valid (optional):
 * [E,P,P] float, 1 where Z > 0.2.  tonly != 0 drops the rotation of G_ij.  Ji/Jj: [E,2,6],
 * Jz: [E,2], jvalid: [E] (all four or none): centre-pixel Jacobians as at :71-106.  No autograd. */
int dpvo_transform(const float* poses, const float* patches, const float* intrinsics,
                   const int64_t* ii, const int64_t* jj, const int64_t* kk,
                   int64_t n_edges, int64_t n_frames, int64_t n_patches, int32_t P,
                   int32_t depth, int32_t tonly, float* coords, float* valid,
                   float* Ji, float* Jj, float* Jz, float* jvalid, void* stream);

/* projective_ops.flow_mag (:116-126): flow [E,P,P] = beta |x_ij - x_ii| + (1-beta) |x_ij(t only) - x_ii|,
 * valid [E,P,P] uint8 = (Z_ij > 0.2). */
int dpvo_flow_mag(const float* poses, const float* patches, const float* intrinsics,
                  const int64_t* ii, const int64_t* jj, const int64_t* kk,
                  int64_t n_edges, int64_t n_frames, int64_t n_patches, int32_t P, float beta,
                  float* flow, uint8_t* valid, void* stream);

/* DPVO.motionmag(a, b) + DPVO.motionmag(b, a) (dpvo/dpvo.py:257-264 as called by keyframe(), :266-272): out[0] =
 * mean of flow_mag over the edges a -> b plus the mean over the edges b -> a, in one launch over the live edge list
 * (the reference selects each direction with a boolean mask, i.e. two device synchronisations, and runs flow_mag twice).
 * Fixed reduction order; an empty direction gives NaN like torch's mean of an empty tensor. */
int dpvo_motionmag_pair(const float* poses, const float* patches, const float* intrinsics,
                        const int64_t* ii, const int64_t* jj, const int64_t* kk,
                        int64_t n_edges, int64_t n_frames, int64_t n_patches, int32_t P, float beta,
                        int64_t frame_a, int64_t frame_b, float* out, void* stream);

/* dpvo_motionmag_pair plus a dry run of what DPVO.keyframe() (dpvo/dpvo.py:266-310) does to the edge list under both
 * outcomes of its test, in the same pass over the live edges, so that the caller learns the decision AND the edge
 * counts that follow from it in one host read and can issue the edits (dpvo_edges_mask_keyframe / _remove / _renumber /
 * _mask_window) without reading their result back.  out: 8 x 32-bit words —
 *   [0] float  motionmag(a, b) + motionmag(b, a)  (NaN when a direction has no edge)
 *   [1] int32  no removal: edges the window rule (:304-309) moves to the inactive list at frame count n
 *   [2] int32  no removal: smallest source frame among the edges that stay (0x7fffffff if none)
 *   [3] int32  frame k removed: edges touching k (dropped, :280-281)
 *   [4] int32  frame k removed: of the others, renumbered (:283-285), those the window rule moves at frame count n - 1
 *   [5] int32  frame k removed: smallest (renumbered) source frame among the edges that stay (0x7fffffff if none)
 *   [6] int32  *live_count (the edge store's device count, dpvo_edge_store.count) or 0 when live_count is NULL
 *   [7] int32  *info (the edge kernels' error flag) or 0 when info is NULL
 * so a caller that mirrors the count on the host sees in the same read whether the mirror is still true.
 * ix: patch -> source frame table [n_patches] (DPVO.ix).  n_edges: exact live count. */
int dpvo_keyframe_plan(const float* poses, const float* patches, const float* intrinsics,
                       const int64_t* ii, const int64_t* jj, const int64_t* kk, const int64_t* ix,
                       int64_t n_edges, int64_t n_frames, int64_t n_patches, int32_t P, float beta,
                       int64_t frame_a, int64_t frame_b, int64_t k, int64_t n, int64_t patches_per_frame,
                       int64_t removal_window, int32_t loop_closure, int64_t optimization_window,
                       const int64_t* live_count, const int32_t* info, int32_t* out, void* stream);

/* projective_ops.point_cloud (:111-113): points [n,P,P,4] = G_{ix[n]}^-1 * iproj(patches[n], intrinsics[ix[n]]). */
int dpvo_point_cloud(const float* poses, const float* patches, const float* intrinsics, const int64_t* ix,
                     int64_t n, int64_t n_frames, int32_t P, float* points, void* stream);

/* ----------------------------------------------------------------- lietorch */

/* Element-wise Lie-group operators — replace lietorch_backends (lietorch/src/lietorch.cpp:18-316,
 * kernels lietorch/src/lietorch_gpu.cu:20-601).  `group` uses the reference's group ids
 * (lietorch/groups.py:236,252,268,290); SO3 and SE3 are provided, the others return an error.
 * All tensors are contiguous [n, width] rows of `dt` (DPVO_F32 / DPVO_F64) on the device:
 * group element N = 4 (SO3: qx qy qz qw) / 7 (SE3: tx ty tz qx qy qz qw), tangent K = 3 / 6
 * (SE3: tau, phi).  Quaternions are normalised on load like the reference (so3.h:35-37).
 * Gradients follow lietorch_cpu.cpp: row vectors in the tangent space stored in the first K slots
 * of an N-wide row (remaining slots written as zero).  Every output element is written. */
typedef enum { DPVO_LIE_SO3 = 1, DPVO_LIE_RXSO3 = 2, DPVO_LIE_SE3 = 3, DPVO_LIE_SIM3 = 4 } dpvo_lie_group;

int dpvo_lie_expm(int32_t group, dpvo_dtype dt, const void* a /*[n,K]*/, void* X /*[n,N]*/, int64_t n, void* stream);
int dpvo_lie_expm_backward(int32_t group, dpvo_dtype dt, const void* grad /*[n,N]*/, const void* a /*[n,K]*/,
                           void* da /*[n,K]*/, int64_t n, void* stream);
int dpvo_lie_logm(int32_t group, dpvo_dtype dt, const void* X /*[n,N]*/, void* a /*[n,K]*/, int64_t n, void* stream);
int dpvo_lie_logm_backward(int32_t group, dpvo_dtype dt, const void* grad /*[n,K]*/, const void* X /*[n,N]*/,
                           void* dX /*[n,N]*/, int64_t n, void* stream);
int dpvo_lie_inv(int32_t group, dpvo_dtype dt, const void* X, void* Y, int64_t n, void* stream);
int dpvo_lie_inv_backward(int32_t group, dpvo_dtype dt, const void* grad /*[n,N]*/, const void* X, void* dX,
                          int64_t n, void* stream);
int dpvo_lie_mul(int32_t group, dpvo_dtype dt, const void* X, const void* Y, void* Z, int64_t n, void* stream);
int dpvo_lie_mul_backward(int32_t group, dpvo_dtype dt, const void* grad /*[n,N]*/, const void* X, const void* Y,
                          void* dX, void* dY, int64_t n, void* stream);
int dpvo_lie_adj(int32_t group, dpvo_dtype dt, const void* X, const void* a /*[n,K]*/, void* b /*[n,K]*/, int64_t n,
                 void* stream);
int dpvo_lie_adj_backward(int32_t group, dpvo_dtype dt, const void* grad /*[n,K]*/, const void* X, const void* a,
                          void* dX /*[n,N]*/, void* da /*[n,K]*/, int64_t n, void* stream);
int dpvo_lie_adjT(int32_t group, dpvo_dtype dt, const void* X, const void* a, void* b, int64_t n, void* stream);
int dpvo_lie_adjT_backward(int32_t group, dpvo_dtype dt, const void* grad /*[n,K]*/, const void* X, const void* a,
                           void* dX, void* da, int64_t n, void* stream);
int dpvo_lie_act(int32_t group, dpvo_dtype dt, const void* X, const void* p /*[n,3]*/, void* q /*[n,3]*/, int64_t n,
                 void* stream);
int dpvo_lie_act_backward(int32_t group, dpvo_dtype dt, const void* grad /*[n,3]*/, const void* X, const void* p,
                          void* dX /*[n,N]*/, void* dp /*[n,3]*/, int64_t n, void* stream);
int dpvo_lie_act4(int32_t group, dpvo_dtype dt, const void* X, const void* p /*[n,4]*/, void* q /*[n,4]*/, int64_t n,
                  void* stream);
int dpvo_lie_act4_backward(int32_t group, dpvo_dtype dt, const void* grad /*[n,4]*/, const void* X, const void* p,
                           void* dX, void* dp /*[n,4]*/, int64_t n, void* stream);
int dpvo_lie_as_matrix(int32_t group, dpvo_dtype dt, const void* X, void* T4x4 /*[n,4,4]*/, int64_t n, void* stream);
int dpvo_lie_projector(int32_t group, dpvo_dtype dt, const void* X, void* P /*[n,N,N]*/, int64_t n, void* stream);
int dpvo_lie_jinv(int32_t group, dpvo_dtype dt, const void* X, const void* a /*[n,K]*/, void* b /*[n,K]*/, int64_t n,
                  void* stream);

/* The same operators on HOST memory (CPU tensors): the reference module dispatches on the tensor's device
 * (lietorch/include/dispatch.h:12-43, src/lietorch_cpu.cpp).  `op` selects the operator; in0..in2 / out0..out1 are the
 * pointer arguments of the device entry point of that operator, in its order (unused ones NULL). */
typedef enum {
  DPVO_LIE_OP_EXP = 0, DPVO_LIE_OP_EXP_BWD, DPVO_LIE_OP_LOG, DPVO_LIE_OP_LOG_BWD, DPVO_LIE_OP_INV, DPVO_LIE_OP_INV_BWD,
  DPVO_LIE_OP_MUL, DPVO_LIE_OP_MUL_BWD, DPVO_LIE_OP_ADJ, DPVO_LIE_OP_ADJ_BWD, DPVO_LIE_OP_ADJT, DPVO_LIE_OP_ADJT_BWD,
  DPVO_LIE_OP_ACT, DPVO_LIE_OP_ACT_BWD, DPVO_LIE_OP_ACT4, DPVO_LIE_OP_ACT4_BWD, DPVO_LIE_OP_MATRIX, DPVO_LIE_OP_PROJECTOR,
  DPVO_LIE_OP_JINV
} dpvo_lie_op;
int dpvo_lie_host(int32_t op, int32_t group, dpvo_dtype dt, const void* in0, const void* in1, const void* in2,
                  void* out0, void* out1, int64_t n);

/* ---- patch-graph edge bookkeeping on the device -------------------------------
 * Replaces the torch.cat / boolean-mask edits of dpvo.py:215-239 (append_factors, remove_factors),
 * :262-310 (keyframe) and :362-375 (__edges_forw/__edges_back), and the numba reduce_edges of
 * loop_closure/optim_utils.py:27-64.  The edge list lives in caller-owned fixed-capacity device arrays;
 * the live count is a device scalar, so none of these calls synchronises.  `bound` is any host-known upper
 * bound on the live count (<= capacity) and only sizes the launch.  `info` is a caller-zeroed device flag:
 * 1 = a capacity was exceeded (extra edges dropped), 2 = a patch index outside ix. */
typedef struct dpvo_edge_store {
  int64_t* ii;       /* [capacity] source frame  */
  int64_t* jj;       /* [capacity] target frame  */
  int64_t* kk;       /* [capacity] patch index   */
  void* net;         /* fp16 hidden state rows [capacity, net_ld], or NULL */
  int64_t net_ld;    /* row stride in elements (even) */
  int32_t net_dim;   /* columns that are live (even; 384 in DPVO) */
  int32_t reserved;
  float* target;     /* [capacity, 2] or NULL */
  float* weight;     /* [capacity, 2] or NULL */
  int64_t* count;    /* device scalar: live edges */
  int64_t capacity;
} dpvo_edge_store;

/* append_factors(ii=kk_new, jj=jj_new): kk += kk_new, jj += jj_new, ii += ix[kk_new], zero net/target/weight rows */
int dpvo_edges_append(const dpvo_edge_store* store, const int64_t* ix, int64_t n_ix, const int64_t* kk_new,
                      const int64_t* jj_new, int64_t n_new, int32_t* info, void* stream);
/* append_factors(*flatmeshgrid(arange(kk0,kk1), arange(jj0,jj1), indexing='ij')) without materialising the grid */
int dpvo_edges_append_grid(const dpvo_edge_store* store, const int64_t* ix, int64_t n_ix, int64_t kk0, int64_t kk1,
                           int64_t jj0, int64_t jj1, int32_t* info, void* stream);
int64_t dpvo_edges_workspace_bytes(int64_t capacity, int32_t net_dim);
/* remove_factors(m, store): edges with mask != 0 leave the list (order of the rest kept); with `inactive`
 * non-NULL (store=True) their ii/jj/kk/target/weight are appended, in order, to that list. */
int dpvo_edges_remove(const dpvo_edge_store* store, const uint8_t* mask, const dpvo_edge_store* inactive,
                      int64_t bound, int32_t* info, void* workspace, int64_t workspace_bytes, void* stream);
/* keyframe(): mask = (ii == k) | (jj == k)  (dpvo.py:280) */
int dpvo_edges_mask_keyframe(const dpvo_edge_store* store, int64_t k, uint8_t* mask, int64_t bound, void* stream);
/* keyframe(): mask = ix[kk] < n - removal_window, minus loop-closure edges (jj - ii > 30 and
 * jj > n - optimization_window) when loop_closure != 0  (dpvo.py:304-309) */
int dpvo_edges_mask_window(const dpvo_edge_store* store, const int64_t* ix, int64_t n_ix, int64_t n,
                           int64_t removal_window, int32_t loop_closure, int64_t optimization_window, uint8_t* mask,
                           int64_t bound, void* stream);
/* keyframe(): kk[ii > k] -= M; ii[ii > k] -= 1; jj[jj > k] -= 1  (dpvo.py:283-285) */
int dpvo_edges_renumber(const dpvo_edge_store* store, int64_t k, int64_t patches_per_frame, int64_t bound,
                        void* stream);
/* keyframe(): for i in range(k, n-1): buf[i % mod] = buf[(i+1) % mod] over slots of slot_bytes (dpvo.py:287-297) */
int dpvo_frames_shift(void* buf, int64_t slot_bytes, int32_t mod, int32_t k, int32_t n, void* stream);
/* reduce_edges(flow_mag, ii, jj, max_num_edges, nms): greedy pick in ascending flow magnitude (ties: lower
 * index), skipping j - i < 30 and magnitude >= 1000, suppressing source frames i-nms..i+nms at the same
 * target.  ni / nj: any bounds > max(ii) / max(jj).  out_edges [max_num_edges, 2], out_count device scalar. */
int64_t dpvo_edges_reduce_workspace_bytes(int64_t n, int64_t ni, int64_t nj);
int dpvo_edges_reduce(const float* flow_mag, const int64_t* ii, const int64_t* jj, int64_t n, int64_t ni, int64_t nj,
                      int32_t max_num_edges, int32_t nms, int64_t* out_edges, int64_t* out_count, void* workspace,
                      int64_t workspace_bytes, void* stream);

/* ---- DPViewer's CUDA helpers, headless: plain device buffers, no window / GL interop ----
 * poseToMatrix (modules/DPViewer/dpviewer/viewer_cuda.cu:221-277): mat4x4[n,4,4] = matrix of the INVERSE of each
 * pose (tx ty tz qx qy qz qw). */
int dpvo_viewer_pose_to_matrix(const float* poses /*[n,7]*/, float* mat4x4 /*[n,4,4]*/, int64_t n, void* stream);
/* iproj_kernel (viewer_cuda.cu:56-164): back-projects frame `index` (disps [N,ht,wd], intrinsics [N,4] at 1/8
 * resolution, images u8 [N,3,ht,wd] BGR) into points [ht*wd,3] / colors [ht*wd,3] (RGB) and counts, per pixel, the
 * other frames whose disparity agrees within thresh percent (count -100 marks a depth edge). Border pixels: zeros. */
int dpvo_viewer_iproj(int32_t index, int32_t n_frames, float thresh, const uint8_t* images, const float* poses,
                      const float* disps, const float* intrinsics, int32_t ht, int32_t wd, float* points,
                      uint8_t* colors, float* count, void* stream);

/* ---- per-frame feature encoders: BasicEncoder4, methods/dpvo/dpvo/extractor.py:200-264 -------
 * Channels-last fp16 activations [img, H, W, C]; fp32 accumulation.  `stats` ([img, DPVO_ENCODER_STATS_REPLICAS, C, 2]
 * float64, caller-zeroed, or NULL) receives each (image, channel) plane's sum and sum of squares for the
 * InstanceNorm2d that follows, spread over replicas (different CTAs add to different cache lines; the consumer sums
 * them). */
#define DPVO_ENCODER_STATS_REPLICAS 16
/* conv1: nn.Conv2d(3, 32, 7, stride=2, padding=3) on the fp32 NCHW image; weight fp16 packed [ky=7][cout=32][32]
 * with k = kx * 3 + c for the 21 live columns and zeros in the other 11 (torch weight [32,3,7,7] permuted) */
int dpvo_encoder_conv7x7s2(const float* images, const void* weight, const float* bias, int32_t n_img, int32_t H,
                           int32_t W, int32_t relu, void* out, double* stats, void* stream);
/* 3x3 (padding 1) / 1x1 (padding 0) convolutions of the residual blocks and the output layer; weight fp16
 * [cout, ksize*ksize, cin]; out = (conv + bias) * scale.  Layers: 3x3 s1 32->32, 3x3 s2 32->64, 3x3 s1 64->64,
 * 1x1 s2 32->64, 1x1 s1 64->(multiple of 64).  Fusions: in_stats != NULL applies relu(InstanceNorm(in)) while the
 * input patch is staged (the producer's statistics, eps); relu != 0 and skip != NULL ([img, OH, OW, cout]) give
 * out = relu(conv) and out = relu(skip + relu(conv)) for the norm-free encoder (exclusive with `stats`). */
int dpvo_encoder_conv(const void* in, const void* weight, const float* bias, int32_t n_img, int32_t H, int32_t W,
                      int32_t cin, int32_t cout, int32_t ksize, int32_t stride, float scale, const double* in_stats,
                      float eps, int32_t relu, const void* skip, void* out, double* stats, void* stream);
/* out = relu(norm(raw)), or relu(skipv + relu(norm(raw))) with skipv = skip or norm(skip) (ResidualBlock.forward,
 * extractor.py:47-55); norm = InstanceNorm2d(affine=False, eps) from `stats`, identity when stats is NULL */
int dpvo_encoder_norm_act(const void* raw, const double* stats, const void* skip, const double* skip_stats,
                          int32_t n_img, int64_t hw, int32_t channels, float eps, void* out, void* stream);
/* F.avg_pool2d(x, 4, 4) on a channels-last fp16 map (pyramid level 2, dpvo.py:438) */
int dpvo_avgpool4_nhwc_f16(const void* in, int32_t n_img, int32_t H, int32_t W, int32_t channels, void* out,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DPVO_B200_H */
