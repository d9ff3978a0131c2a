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

#define DPVO_B200_ABI_VERSION 1

typedef enum { DPVO_F16 = 0, DPVO_F32 = 1, DPVO_F64 = 2 } dpvo_dtype;

int dpvo_abi_version(void);
const char* dpvo_last_error(void);

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

#ifdef __cplusplus
}
#endif
#endif /* DPVO_B200_H */
