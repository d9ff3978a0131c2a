// Correlation lookup backward (altcorr.corr, training) for sm_100a.
//
// Replaces /root/reference/methods/dpvo/dpvo/altcorr/correlation_kernel.cu:237-287 (adjoint of
// the bilinear blend: four zero-padded products rounded in the feature dtype, summed left to
// right) and :140-191 (scatter of g*fmap2 into fmap1_grad and g*fmap1 into fmap2_grad).
//
// One CTA per edge.  The adjoint window cg[p][D*D] is formed once in shared memory with the
// reference's rounding points.  fmap1_grad is a gather over the edge's window (one atomic per
// (patch pixel, channel) instead of one per window tap); fmap2_grad contributions of the 9
// overlapping windows are combined over the union bounding box before the atomic (one per
// (box pixel, channel) instead of one per (patch pixel, tap, channel)).  fp16 gradients are
// accumulated in caller-provided fp32 scratch and rounded once at the end (the reference uses
// fp16 atomics, whose result depends on the order of arrival).
#include "common.cuh"
#include <limits.h>

namespace dpvo {

template <typename T> struct BwdT;
template <> struct BwdT<float> {
  using Acc = float;
  static __device__ __forceinline__ float rnd(float v) { return v; }
  static __device__ __forceinline__ float ld(const float* p) { return *p; }
};
template <> struct BwdT<double> {
  using Acc = double;
  static __device__ __forceinline__ double rnd(double v) { return v; }
  static __device__ __forceinline__ double ld(const double* p) { return *p; }
};
template <> struct BwdT<__half> {
  using Acc = float;
  static __device__ __forceinline__ float rnd(float v) { return __half2float(__float2half_rn(v)); }
  static __device__ __forceinline__ float ld(const __half* p) { return __half2float(*p); }
};

struct CorrBwdParams {
  const void* fmap1; int64_t g_sb, g_sn, g_sc, g_sy, g_sx;
  const void* fmap2; int64_t sb, sn, sc, sy, sx;
  int N1, N2, C, PH, PW, H2, W2;
  const float* coords; const int64_t* us; const int64_t* vs;
  int B, M, R;
  const void* cg; int64_t cg_sb, cg_sm, cg_sx, cg_sy, cg_si, cg_sj;
  void* acc1; void* acc2;   // accumulation buffers (contiguous, shapes of fmap1 / fmap2), Acc dtype
};

constexpr int kBwdMaxBox = 256;

// dynamic smem: Acc cgs[PP*D*D]; int fxy[PP*2]
template <typename T>
__global__ void __launch_bounds__(128)
corr_bwd_kernel(const __grid_constant__ CorrBwdParams prm) {
  using A = typename BwdT<T>::Acc;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int PP = prm.PH * prm.PW;
  const int R = prm.R, D = 2 * R + 2, DD = D * D;
  A* cgs = reinterpret_cast<A*>(smem_raw);
  int* fxy = reinterpret_cast<int*>(cgs + PP * DD);
  __shared__ int s_box[4];   // x0, y0, bw, bh
  const T* fmap1 = reinterpret_cast<const T*>(prm.fmap1);
  const T* fmap2 = reinterpret_cast<const T*>(prm.fmap2);
  const T* cgin = reinterpret_cast<const T*>(prm.cg);
  A* acc1 = reinterpret_cast<A*>(prm.acc1);
  A* acc2 = reinterpret_cast<A*>(prm.acc2);
  const int64_t nedge = (int64_t)prm.B * prm.M;
  const int C = prm.C, H2 = prm.H2, W2 = prm.W2;

  for (int64_t e = blockIdx.x; e < nedge; e += gridDim.x) {
    const int b = (int)(e / prm.M);
    const int m = (int)(e - (int64_t)b * prm.M);
    const int64_t u = prm.us[m], v = prm.vs[m];
    if (u < 0 || u >= prm.N1 || v < 0 || v >= prm.N2) continue;   // uniform per CTA
    __syncthreads();
    // ---- adjoint of the blend, reference rounding: ((g1 + g2) + g3) + g4, each product w*grad rounded
    for (int idx = threadIdx.x; idx < PP * DD; idx += blockDim.x) {
      const int p = idx / DD, w = idx - p * DD;
      const int a = w / D, bb = w - a * D;     // a: y offset, bb: x offset
      const int i0 = p / prm.PW, j0 = p - i0 * prm.PW;
      const float x = prm.coords[e * (2 * PP) + p];
      const float y = prm.coords[e * (2 * PP) + PP + p];
      if (w == 0) { fxy[2 * p] = floor_to_int(x); fxy[2 * p + 1] = floor_to_int(y); }
      const A dx = BwdT<T>::rnd((A)(x - floorf(x))), dy = BwdT<T>::rnd((A)(y - floorf(y)));
      const A mdx = BwdT<T>::rnd((A)1 - dx), mdy = BwdT<T>::rnd((A)1 - dy);
      const T* gp = cgin + b * prm.cg_sb + m * prm.cg_sm + i0 * prm.cg_si + j0 * prm.cg_sj;
      auto G = [&](int yy, int xx) -> A { return BwdT<T>::ld(gp + xx * prm.cg_sx + yy * prm.cg_sy); };
      A s = 0;
      if (a < D - 1 && bb < D - 1) s = BwdT<T>::rnd(BwdT<T>::rnd(mdx * mdy) * G(a, bb));
      if (a < D - 1 && bb >= 1) s = BwdT<T>::rnd(s + BwdT<T>::rnd(BwdT<T>::rnd(dx * mdy) * G(a, bb - 1)));
      if (a >= 1 && bb < D - 1) s = BwdT<T>::rnd(s + BwdT<T>::rnd(BwdT<T>::rnd(mdx * dy) * G(a - 1, bb)));
      if (a >= 1 && bb >= 1) s = BwdT<T>::rnd(s + BwdT<T>::rnd(BwdT<T>::rnd(dx * dy) * G(a - 1, bb - 1)));
      cgs[idx] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int mnx = INT_MAX, mny = INT_MAX, mxx = INT_MIN, mxy = INT_MIN;
      for (int p = 0; p < PP; ++p) {
        mnx = min(mnx, fxy[2 * p]); mxx = max(mxx, fxy[2 * p]);
        mny = min(mny, fxy[2 * p + 1]); mxy = max(mxy, fxy[2 * p + 1]);
      }
      s_box[0] = mnx - R; s_box[1] = mny - R; s_box[2] = mxx - mnx + D; s_box[3] = mxy - mny + D;
    }
    __syncthreads();
    const T* f1 = fmap1 + b * prm.g_sb + u * prm.g_sn;
    const T* f2 = fmap2 + b * prm.sb + v * prm.sn;

    // ---- fmap1_grad[u][c][p] += sum_taps cg[p][tap] * fmap2[v][c][tap]   (taps fastest: contiguous reads)
    for (int idx = threadIdx.x; idx < PP * C; idx += blockDim.x) {
      const int c = idx / PP, p = idx - c * PP;
      const int x0 = fxy[2 * p] - R, y0 = fxy[2 * p + 1] - R;
      A s = 0;
      for (int a = 0; a < D; ++a) {
        const int i1 = y0 + a;
        if (i1 < 0 || i1 >= H2) continue;
        const T* row = f2 + c * prm.sc + (int64_t)i1 * prm.sy;
        for (int bb = 0; bb < D; ++bb) {
          const int j1 = x0 + bb;
          if (j1 >= 0 && j1 < W2) s += cgs[p * DD + a * D + bb] * BwdT<T>::ld(row + (int64_t)j1 * prm.sx);
        }
      }
      if (s != (A)0) {
        const int i0 = p / prm.PW, j0 = p - i0 * prm.PW;
        atomicAdd(&acc1[(((int64_t)b * prm.N1 + u) * C + c) * PP + i0 * prm.PW + j0], s);
      }
    }

    // ---- fmap2_grad[v][c][pixel] += sum_p cg[p][tap_p(pixel)] * fmap1[u][c][p]
    const int bx0 = s_box[0], by0 = s_box[1], bw = s_box[2], bh = s_box[3];
    const int64_t nbox = (int64_t)bw * bh;
    if (nbox <= kBwdMaxBox) {
      for (int64_t idx = threadIdx.x; idx < nbox * C; idx += blockDim.x) {
        const int c = (int)(idx / nbox), q = (int)(idx - (int64_t)c * nbox);
        const int qy = q / bw, qx = q - qy * bw;
        const int i1 = by0 + qy, j1 = bx0 + qx;
        if (i1 < 0 || i1 >= H2 || j1 < 0 || j1 >= W2) continue;
        A s = 0;
        for (int p = 0; p < PP; ++p) {
          const int a = i1 - (fxy[2 * p + 1] - R), bb = j1 - (fxy[2 * p] - R);
          if (a >= 0 && a < D && bb >= 0 && bb < D) {
            const int i0 = p / prm.PW, j0 = p - i0 * prm.PW;
            s += cgs[p * DD + a * D + bb] * BwdT<T>::ld(f1 + c * prm.g_sc + i0 * prm.g_sy + j0 * prm.g_sx);
          }
        }
        if (s != (A)0) atomicAdd(&acc2[((((int64_t)b * prm.N2 + v) * C + c) * H2 + i1) * W2 + j1], s);
      }
    } else {
      for (int64_t idx = threadIdx.x; idx < (int64_t)PP * DD * C; idx += blockDim.x) {
        const int c = (int)(idx / (PP * DD)), r = (int)(idx - (int64_t)c * PP * DD);
        const int p = r / DD, w = r - p * DD;
        const int i1 = fxy[2 * p + 1] - R + w / D, j1 = fxy[2 * p] - R + w % D;
        if (i1 < 0 || i1 >= H2 || j1 < 0 || j1 >= W2) continue;
        const int i0 = p / prm.PW, j0 = p - i0 * prm.PW;
        const A s = cgs[r] * BwdT<T>::ld(f1 + c * prm.g_sc + i0 * prm.g_sy + j0 * prm.g_sx);
        if (s != (A)0) atomicAdd(&acc2[((((int64_t)b * prm.N2 + v) * C + c) * H2 + i1) * W2 + j1], s);
      }
    }
  }
}

__global__ void __launch_bounds__(256)
f32_to_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = __float2half_rn(src[i]);
}

template <typename T>
static int launch_bwd(const CorrBwdParams& prm, cudaStream_t st) {
  using A = typename BwdT<T>::Acc;
  const int PP = prm.PH * prm.PW, D = 2 * prm.R + 2;
  const size_t smem = (size_t)PP * D * D * sizeof(A) + (size_t)PP * 2 * sizeof(int) + 16;
  if (smem > 200 * 1024) { set_error("corr_backward: window too large for shared memory"); return 1; }
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(corr_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int64_t nedge = (int64_t)prm.B * prm.M;
  const int grid = (int)(nedge < (int64_t)kNumSMs * 16 ? nedge : (int64_t)kNumSMs * 16);
  note_launches(1);
  corr_bwd_kernel<T><<<grid, 128, smem, st>>>(prm);
  return check_launch("corr_bwd");
}

}  // namespace dpvo

using namespace dpvo;

extern "C" int dpvo_corr_backward(dpvo_dtype dtype,
                                  const void* fmap1, int64_t g_sb, int64_t g_sn, int64_t g_sc, int64_t g_sy, int64_t g_sx,
                                  const void* fmap2, int64_t sb, int64_t sn, int64_t sc, int64_t sy, int64_t sx,
                                  int32_t N1, int32_t N2, int32_t C, int32_t PH, int32_t PW, int32_t H2, int32_t W2,
                                  const float* coords, const int64_t* us, const int64_t* vs,
                                  int32_t B, int32_t M, int32_t radius,
                                  const void* corr_grad, int64_t cg_sb, int64_t cg_sm, int64_t cg_sx, int64_t cg_sy,
                                  int64_t cg_si, int64_t cg_sj,
                                  void* fmap1_grad, void* fmap2_grad, float* acc1, float* acc2, void* stream) {
  DPVO_REQUIRE(radius >= 0 && PH > 0 && PW > 0 && C > 0 && B >= 0 && M >= 0, "corr_backward: bad sizes");
  if ((int64_t)B * M == 0) return 0;
  DPVO_REQUIRE(fmap1 && fmap2 && coords && us && vs && corr_grad && fmap1_grad && fmap2_grad,
               "corr_backward: null pointer");
  DPVO_REQUIRE(dtype != DPVO_F16 || (acc1 && acc2), "corr_backward: fp16 needs the float32 scratch buffers acc1/acc2");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  CorrBwdParams p;
  memset(&p, 0, sizeof(p));
  p.fmap1 = fmap1; p.g_sb = g_sb; p.g_sn = g_sn; p.g_sc = g_sc; p.g_sy = g_sy; p.g_sx = g_sx;
  p.fmap2 = fmap2; p.sb = sb; p.sn = sn; p.sc = sc; p.sy = sy; p.sx = sx;
  p.N1 = N1; p.N2 = N2; p.C = C; p.PH = PH; p.PW = PW; p.H2 = H2; p.W2 = W2;
  p.coords = coords; p.us = us; p.vs = vs; p.B = B; p.M = M; p.R = radius;
  p.cg = corr_grad; p.cg_sb = cg_sb; p.cg_sm = cg_sm; p.cg_sx = cg_sx; p.cg_sy = cg_sy; p.cg_si = cg_si; p.cg_sj = cg_sj;
  if (dtype == DPVO_F16) {
    p.acc1 = acc1; p.acc2 = acc2;
    if (launch_bwd<__half>(p, st)) return 1;
    const int64_t n1 = (int64_t)B * N1 * C * PH * PW, n2 = (int64_t)B * N2 * C * H2 * W2;
    note_launches(2);
    f32_to_f16_kernel<<<(int)(ceil_div64(n1, 1024) < kNumSMs * 8 ? ceil_div64(n1, 1024) : kNumSMs * 8), 256, 0, st>>>(
        acc1, reinterpret_cast<__half*>(fmap1_grad), n1);
    f32_to_f16_kernel<<<(int)(ceil_div64(n2, 1024) < kNumSMs * 8 ? ceil_div64(n2, 1024) : kNumSMs * 8), 256, 0, st>>>(
        acc2, reinterpret_cast<__half*>(fmap2_grad), n2);
    return check_launch("corr_bwd_cast");
  }
  p.acc1 = fmap1_grad; p.acc2 = fmap2_grad;
  if (dtype == DPVO_F32) return launch_bwd<float>(p, st);
  if (dtype == DPVO_F64) return launch_bwd<double>(p, st);
  set_error("corr_backward: unknown dtype %d", (int)dtype);
  return 1;
}
