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

// ------------------------------------------------------------------------------------------------
// Fast path (training shape: fp16 features, C = 128, 3x3 patches, radius 3).
//
// One WARP per edge, lane l owns channels 4l..4l+3.  fmap2 is read from the channels-last ring the
// forward already uses (a box pixel's 128 channels = one coalesced 256-byte warp load) and its
// gradient is accumulated channels-last in fp32 (one `red.global.add.v4.f32` per lane per box
// pixel: 100 vector reductions per edge instead of 12,800 scalar atomics); the patch gradient
// (9 pixels x 4 channels per lane) stays in registers for the whole edge and leaves with 9 vector
// reductions.  The adjoint of the blend is expanded once per edge into a [box pixel][patch pixel]
// weight table in shared memory (zeros outside a pixel's 8x8 window), so the inner loop is
//     w[0..8] (3 broadcast LDS.128)  ->  s[k] += w[p] f1[p][k];  a1[p][k] += w[p] f2[k]
// with no index arithmetic.  Rounding points are the reference's (each product and partial sum of
// the adjoint rounded to fp16, accumulation in fp32 — the reference's fp16 atomics round more).
constexpr int kFB = 12;                 // largest box side held in the weight table (nominal box: 10)
constexpr int kFBWarps = 4;

struct CorrBwdFastParams {
  const void* fmap1; int64_t g_sb, g_sn;                   // [B][N1][128][3][3] dense patches
  const void* fmap2; int64_t sb, sn, sy, sx;               // channels-last: channel stride 1
  int N1, N2, H2, W2;
  const float* coords; const int64_t* us; const int64_t* vs;
  int B, M;
  const void* cg; int64_t cg_sb, cg_sm, cg_sx, cg_sy, cg_si, cg_sj;
  float* acc1;                                             // [B][N1][128][9]  fp32, zeroed
  float* acc2;                                             // [B][N2][H2][W2][128] fp32, zeroed
  const int32_t* perm;                                     // optional processing order (csrc/corr_order.cu)
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ float h_rnd(float v) { return __half2float(__float2half_rn(v)); }

// adjoint of the bilinear blend at raw-window tap (a: y offset, bb: x offset), reference rounding
template <typename T>
__device__ __forceinline__ float blend_adjoint(const T* gp, int64_t sx, int64_t sy, int a, int bb,
                                               float dx, float dy, float mdx, float mdy) {
  constexpr int D = 8;
  auto G = [&](int yy, int xx) -> float { return BwdT<T>::ld(gp + xx * sx + yy * sy); };
  auto rnd = [](float v) -> float { return BwdT<T>::rnd(v); };
  float s = 0.f;
  if (a < D - 1 && bb < D - 1) s = rnd(rnd(mdx * mdy) * G(a, bb));
  if (a < D - 1 && bb >= 1) s = rnd(s + rnd(rnd(dx * mdy) * G(a, bb - 1)));
  if (a >= 1 && bb < D - 1) s = rnd(s + rnd(rnd(mdx * dy) * G(a - 1, bb)));
  if (a >= 1 && bb >= 1) s = rnd(s + rnd(rnd(dx * dy) * G(a - 1, bb - 1)));
  return s;
}

// this lane's 4 channels of one channels-last pixel
__device__ __forceinline__ void load4(const __half* p, float (&v)[4]) {
  const uint2 raw = __ldg(reinterpret_cast<const uint2*>(p));
  const __half2 h01 = *reinterpret_cast<const __half2*>(&raw.x);
  const __half2 h23 = *reinterpret_cast<const __half2*>(&raw.y);
  v[0] = __low2float(h01); v[1] = __high2float(h01); v[2] = __low2float(h23); v[3] = __high2float(h23);
}
__device__ __forceinline__ void load4(const float* p, float (&v)[4]) {
  const float4 raw = __ldg(reinterpret_cast<const float4*>(p));
  v[0] = raw.x; v[1] = raw.y; v[2] = raw.z; v[3] = raw.w;
}

template <typename T>
__global__ void __launch_bounds__(kFBWarps * 32)
corr_bwd_fast_kernel(const __grid_constant__ CorrBwdFastParams prm) {
  constexpr int R = 3, D = 8, PP = 9, C = 128;
  __shared__ __align__(16) float s_w[kFBWarps][kFB * kFB][12];
  __shared__ int s_fx[kFBWarps][PP], s_fy[kFBWarps][PP];
  __shared__ float s_fr[kFBWarps][PP][4];      // dx, dy, 1-dx, 1-dy (rounded to fp16)
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t nedge = (int64_t)prm.B * prm.M;
  const int64_t warp0 = (int64_t)blockIdx.x * kFBWarps + wid;
  const int64_t nwarps = (int64_t)gridDim.x * kFBWarps;
  const int H2 = prm.H2, W2 = prm.W2;
  float(*wtab)[12] = s_w[wid];

  for (int64_t es = warp0; es < nedge; es += nwarps) {
    const int64_t e = prm.perm ? (int64_t)prm.perm[es] : es;
    const int b = (int)(e / prm.M);
    const int m = (int)(e - (int64_t)b * prm.M);
    const int64_t u = prm.us[m], v = prm.vs[m];
    if (u < 0 || u >= prm.N1 || v < 0 || v >= prm.N2) continue;     // uniform per warp
    __syncwarp();
    if (lane < PP) {
      const float x = prm.coords[e * (2 * PP) + lane];
      const float y = prm.coords[e * (2 * PP) + PP + lane];
      s_fx[wid][lane] = floor_to_int(x);
      s_fy[wid][lane] = floor_to_int(y);
      const float dx = BwdT<T>::rnd(x - floorf(x)), dy = BwdT<T>::rnd(y - floorf(y));
      s_fr[wid][lane][0] = dx; s_fr[wid][lane][1] = dy;
      s_fr[wid][lane][2] = BwdT<T>::rnd(1.f - dx); s_fr[wid][lane][3] = BwdT<T>::rnd(1.f - dy);
    }
    __syncwarp();
    int mnx = s_fx[wid][0], mxx = mnx, mny = s_fy[wid][0], mxy = mny;
#pragma unroll
    for (int p = 1; p < PP; ++p) {
      mnx = min(mnx, s_fx[wid][p]); mxx = max(mxx, s_fx[wid][p]);
      mny = min(mny, s_fy[wid][p]); mxy = max(mxy, s_fy[wid][p]);
    }
    const int bx0 = mnx - R, by0 = mny - R, bw = mxx - mnx + D, bh = mxy - mny + D;
    // whole box outside the map: nothing to do
    if (bx0 >= W2 || by0 >= H2 || bx0 + bw <= 0 || by0 + bh <= 0) continue;

    // patch features of this lane's 4 channels: f1[p][k], 36 contiguous halfs at channel 4*lane
    const T* f1p = reinterpret_cast<const T*>(prm.fmap1) + b * prm.g_sb + u * prm.g_sn + (int64_t)lane * 4 * PP;
    float f1[PP][4];
    {
      float tmp[36];
#pragma unroll
      for (int i = 0; i < 9; ++i) {
        float q[4];
        load4(f1p + 4 * i, q);
        tmp[4 * i] = q[0]; tmp[4 * i + 1] = q[1]; tmp[4 * i + 2] = q[2]; tmp[4 * i + 3] = q[3];
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int p = 0; p < PP; ++p) f1[p][k] = tmp[k * PP + p];
    }
    float a1[PP][4];
#pragma unroll
    for (int p = 0; p < PP; ++p)
#pragma unroll
      for (int k = 0; k < 4; ++k) a1[p][k] = 0.f;

    const T* f2 = reinterpret_cast<const T*>(prm.fmap2) + b * prm.sb + v * prm.sn + lane * 4;
    float* g2 = prm.acc2 + (((int64_t)b * prm.N2 + v) * H2) * (int64_t)W2 * C + lane * 4;
    const T* cgb = reinterpret_cast<const T*>(prm.cg) + b * prm.cg_sb + m * prm.cg_sm;

    if (bw <= kFB && bh <= kFB) {
      const int nbox = bw * bh;
      // ---- weight table: wtab[q][p] = adjoint-blended gradient of (patch pixel p, tap of box pixel q)
      for (int idx = lane; idx < nbox * PP; idx += 32) {
        const int q = idx / PP, p = idx - q * PP;
        const int qy = q / bw, qx = q - qy * bw;
        const int a = qy - (s_fy[wid][p] - R - by0), bb = qx - (s_fx[wid][p] - R - bx0);
        float w = 0.f;
        if (a >= 0 && a < D && bb >= 0 && bb < D) {
          const int i0 = p / 3, j0 = p - 3 * i0;
          w = blend_adjoint(cgb + i0 * prm.cg_si + j0 * prm.cg_sj, prm.cg_sx, prm.cg_sy, a, bb,
                            s_fr[wid][p][0], s_fr[wid][p][1], s_fr[wid][p][2], s_fr[wid][p][3]);
        }
        wtab[q][p] = w;
      }
      __syncwarp();
      // ---- main loop over the box pixels inside the map
      const int qy0 = max(0, -by0), qy1 = min(bh, H2 - by0);
      const int qx0 = max(0, -bx0), qx1 = min(bw, W2 - bx0);
      for (int qy = qy0; qy < qy1; ++qy) {
        const int i1 = by0 + qy;
#pragma unroll 2
        for (int qx = qx0; qx < qx1; ++qx) {
          const int j1 = bx0 + qx;
          const int64_t pix = (int64_t)i1 * prm.sy + (int64_t)j1 * prm.sx;
          float fv[4];
          load4(f2 + pix, fv);
          const float4 w0 = *reinterpret_cast<const float4*>(&wtab[qy * bw + qx][0]);
          const float4 w1 = *reinterpret_cast<const float4*>(&wtab[qy * bw + qx][4]);
          const float w8 = wtab[qy * bw + qx][8];
          const float w[PP] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w8};
          float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int p = 0; p < PP; ++p) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              s[k] = fmaf(w[p], f1[p][k], s[k]);
              a1[p][k] = fmaf(w[p], fv[k], a1[p][k]);
            }
          }
          red_add_v4(g2 + ((int64_t)i1 * W2 + j1) * C, s[0], s[1], s[2], s[3]);
        }
      }
    } else {
      // ---- rare: a patch stretched over more than kFB pixels — tap by tap, weights recomputed by every lane
#pragma unroll
      for (int p = 0; p < PP; ++p) {
        const int i0 = p / 3, j0 = p - 3 * i0;
        const int x0 = s_fx[wid][p] - R, y0 = s_fy[wid][p] - R;
#pragma unroll 1
        for (int t = 0; t < D * D; ++t) {
          const int a = t / D, bb = t - a * D;
          const int i1 = y0 + a, j1 = x0 + bb;
          if (i1 < 0 || i1 >= H2 || j1 < 0 || j1 >= W2) continue;
          const float w = blend_adjoint(cgb + i0 * prm.cg_si + j0 * prm.cg_sj, prm.cg_sx, prm.cg_sy, a, bb,
                                        s_fr[wid][p][0], s_fr[wid][p][1], s_fr[wid][p][2], s_fr[wid][p][3]);
          if (w == 0.f) continue;
          float fv[4];
          load4(f2 + (int64_t)i1 * prm.sy + (int64_t)j1 * prm.sx, fv);
          float s[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            s[k] = w * f1[p][k];
            a1[p][k] = fmaf(w, fv[k], a1[p][k]);
          }
          red_add_v4(g2 + ((int64_t)i1 * W2 + j1) * C, s[0], s[1], s[2], s[3]);
        }
      }
    }
    // ---- patch gradient: acc1[u][c][p], this lane's 36 floats are contiguous (9 vector reductions)
    {
      float* g1 = prm.acc1 + (((int64_t)b * prm.N1 + u) * C + lane * 4) * PP;
      float flat[36];
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int p = 0; p < PP; ++p) flat[k * PP + p] = a1[p][k];
#pragma unroll
      for (int i = 0; i < 9; ++i) red_add_v4(g1 + 4 * i, flat[4 * i], flat[4 * i + 1], flat[4 * i + 2], flat[4 * i + 3]);
    }
  }
}

__global__ void __launch_bounds__(256)
f32_to_f16_vec4_kernel(const float4* __restrict__ src, uint2* __restrict__ dst, int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = __ldcs(src + i);
    const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<const uint32_t*>(&a);
    o.y = *reinterpret_cast<const uint32_t*>(&b);
    dst[i] = o;
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

// Fast path entry: see the comment above corr_bwd_fast_kernel.  fmap2 and fmap2_grad are CHANNELS-LAST
// ([B][N2][H2][W2][128], channel stride 1); fmap1_grad has fmap1's dense [B][N1][128][3][3] layout.
extern "C" int dpvo_corr_backward_nhwc(
    dpvo_dtype dtype,
    const void* fmap1, int64_t g_sb, int64_t g_sn,
    const void* fmap2, int64_t sb, int64_t sn, int64_t sy, int64_t sx,
    int32_t N1, int32_t N2, int32_t H2, int32_t W2,
    const float* coords, const int64_t* us, const int64_t* vs, int32_t B, int32_t M,
    const void* corr_grad, int64_t cg_sb, int64_t cg_sm, int64_t cg_sx, int64_t cg_sy, int64_t cg_si, int64_t cg_sj,
    void* fmap1_grad, void* fmap2_grad_nhwc, float* acc1, float* acc2, const int32_t* perm, void* stream) {
  DPVO_REQUIRE(dtype == DPVO_F16 || dtype == DPVO_F32, "corr_backward_nhwc: float16 or float32 features");
  DPVO_REQUIRE(B >= 0 && M >= 0 && N1 > 0 && N2 > 0 && H2 > 0 && W2 > 0, "corr_backward_nhwc: bad sizes");
  if ((int64_t)B * M == 0) return 0;
  DPVO_REQUIRE(fmap1 && fmap2 && coords && us && vs && corr_grad && fmap1_grad && fmap2_grad_nhwc,
               "corr_backward_nhwc: null pointer");
  const bool half = dtype == DPVO_F16;
  DPVO_REQUIRE(!half || (acc1 && acc2), "corr_backward_nhwc: float16 needs the float32 scratch buffers");
  const uintptr_t amask = half ? 7 : 15;
  DPVO_REQUIRE((reinterpret_cast<uintptr_t>(fmap1) & amask) == 0 && g_sn % 4 == 0 && g_sb % 4 == 0,
               "corr_backward_nhwc: fmap1 must be dense patches, each lane's 4-channel group vector aligned");
  DPVO_REQUIRE((reinterpret_cast<uintptr_t>(fmap2) & amask) == 0 && sn % 4 == 0 && sy % 4 == 0 && sx % 4 == 0 && sb % 4 == 0,
               "corr_backward_nhwc: fmap2 must be channels-last with vector-aligned pixels");
  float* a1 = half ? acc1 : reinterpret_cast<float*>(fmap1_grad);
  float* a2 = half ? acc2 : reinterpret_cast<float*>(fmap2_grad_nhwc);
  DPVO_REQUIRE((reinterpret_cast<uintptr_t>(a1) & 15) == 0 && (reinterpret_cast<uintptr_t>(a2) & 15) == 0,
               "corr_backward_nhwc: accumulation buffers must be 16-byte aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  CorrBwdFastParams p;
  memset(&p, 0, sizeof(p));
  p.fmap1 = fmap1; p.g_sb = g_sb; p.g_sn = g_sn;
  p.fmap2 = fmap2; p.sb = sb; p.sn = sn; p.sy = sy; p.sx = sx;
  p.N1 = N1; p.N2 = N2; p.H2 = H2; p.W2 = W2;
  p.coords = coords; p.us = us; p.vs = vs; p.B = B; p.M = M;
  p.cg = corr_grad;
  p.cg_sb = cg_sb; p.cg_sm = cg_sm; p.cg_sx = cg_sx; p.cg_sy = cg_sy; p.cg_si = cg_si; p.cg_sj = cg_sj;
  p.acc1 = a1; p.acc2 = a2;
  p.perm = perm;
  const int64_t nedge = (int64_t)B * M;
  const int64_t want = ceil_div64(nedge, kFBWarps);
  const int grid = (int)(want < (int64_t)kNumSMs * 4 ? want : (int64_t)kNumSMs * 4);
  note_launches(1);
  if (half) corr_bwd_fast_kernel<__half><<<grid, kFBWarps * 32, 0, st>>>(p);
  else corr_bwd_fast_kernel<float><<<grid, kFBWarps * 32, 0, st>>>(p);
  if (check_launch("corr_bwd_fast")) return 1;
  if (!half) return 0;
  const int64_t n1 = (int64_t)B * N1 * 128 * 9 / 4, n2 = (int64_t)B * N2 * H2 * W2 * 128 / 4;
  auto blocks = [](int64_t n4) { const int64_t b = ceil_div64(n4, 256 * 4); return (int)(b < kNumSMs * 16 ? (b > 0 ? b : 1) : kNumSMs * 16); };
  note_launches(2);
  f32_to_f16_vec4_kernel<<<blocks(n1), 256, 0, st>>>(reinterpret_cast<const float4*>(acc1),
                                                     reinterpret_cast<uint2*>(fmap1_grad), n1);
  f32_to_f16_vec4_kernel<<<blocks(n2), 256, 0, st>>>(reinterpret_cast<const float4*>(acc2),
                                                     reinterpret_cast<uint2*>(fmap2_grad_nhwc), n2);
  return check_launch("corr_bwd_fast_cast");
}
