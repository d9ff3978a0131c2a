// Blocked right-looking Cholesky solve for the reduced camera system when it does not fit one
// CTA's registers / shared memory (6N > 127: wide local windows, global BA with loop closures).
//
// Replaces at::linalg_cholesky_ex + torch::cholesky_solve of
// /root/reference/methods/dpvo/dpvo/fastba/ba_cuda.cu:547-548,561-562 (cuSOLVER potrf/potrs there).
//
// A is (n+1) x lda row-major in global memory: rows 0..n-1 hold the lower triangle of
// S = B - E Q E^T (+ damping), row n holds y^T, so the forward substitution L z = y is part of
// the factorisation (row n of L is z^T).  Panels of 64 columns:
//   panel  ONE launch per panel on the critical path (chol_panel_kernel): every CTA applies the previous panel's
//          update to the column it needs, factors the 64x64 diagonal block AND inverts the factor (redundantly, in
//          registers / shared memory), and forms its 64 rows of L21 = A21 * inv(L11)^T as a 64^3 product
//   bulk   trailing update A22 -= L21 L21^T of everything right of the next panel, on the caller's stream underneath
//          the next panel (look-ahead): 128x128 tiles on the tensor cores as split-TF32 tcgen05.mma
//          (chol_syrk128_umma_kernel) while there is a tile per SM, 64x64 fp32 tiles (chol_syrk_kernel) below that;
//          chol_syrk128_kernel is the fp32 form of the big tiles (dpvo_chol_bulk_mode)
// Back substitution L^T x = z from the bottom, four panels per launch (chol_bs_kernel).
// Structure of S is NOT used.  Measured: a bitmap of the non-zero 64x64 tiles (one warp per tile), its symbolic
// elimination (one CTA) and early exits in the panel / trailing-update kernels.  In time order the factor of a
// 1,000-keyframe system with 193 loop-closure pairs fills 85 % of its tiles (24 % before elimination), 74 % at 1,500
// keyframes: 21.2 -> 13.9 ms at 1,500 keyframes, nothing at 1,000, and the 1,000-frame sequence (S 45 % occupied, the
// factor dense) paid 0.3 - 0.7 ms per call for the bookkeeping on its critical path.  A fill-reducing order (loop
// closure targets last) is the prerequisite for this to pay; not built.
// Finally dX -> global and the SE3 retraction (ba_cuda.cu:178-206).
// A non-positive pivot sets meta[1]; every later kernel returns early and dX is zeroed.
#include "se3.cuh"
#include "ba_internal.cuh"
#include "sm100.cuh"
#include <stdlib.h>

#ifndef DPVO_CHOL_NOALLOC_MAX_N
#define DPVO_CHOL_NOALLOC_MAX_N 768
#endif

namespace dpvo {

constexpr int NB = 64;

constexpr int LDT = 68;                    // shared-memory tile row stride (floats): rows stay 16-byte aligned for
                                           // float4 reads along k, and 8 consecutive rows cover all 32 banks
constexpr int TILE = NB * LDT;
constexpr int kPanelSmemBytes = 5 * TILE * (int)sizeof(float);   // Pd, Pr, D, R, X: 85 KB (+ 129 KB of a bulk CTA fit one SM)

// Stage clocks of chol_panel_kernel (development builds with -DDPVO_CHOL_PROFILE only: tools/build_prof.sh): thread 0
// of CTA 1 adds the cycles between consecutive marks to g_chol_prof[mark]; [31] counts launches.
#ifdef DPVO_CHOL_PROFILE
__device__ long long g_chol_prof[32];
#define CHOL_MARK(i)                                                                         \
  do {                                                                                       \
    if (threadIdx.x == 0 && blockIdx.x == 1) {                                               \
      const long long now_ = clock64();                                                      \
      atomicAdd(reinterpret_cast<unsigned long long*>(&g_chol_prof[i]), (unsigned long long)(now_ - chol_t_)); \
      chol_t_ = now_;                                                                        \
    }                                                                                        \
  } while (0)
#define CHOL_MARK_DECL long long chol_t_ = clock64()
#define CHOL_MARK_ARG , long long& chol_t_
#define CHOL_MARK_PASS , chol_t_
#else
#define CHOL_MARK(i) do {} while (0)
#define CHOL_MARK_DECL do {} while (0)
#define CHOL_MARK_ARG
#define CHOL_MARK_PASS
#endif

// acc[i][j] += sum_{k < 64} As[ty + 16 i][k] * Bs[tx + 16 j][k]   (both tiles row-major with stride LDT).
// float4 along k: per 4 k-steps a thread issues 8 LDS.128 for 64 FMAs (scalar reads: 32 LDS for 64 FMAs, which made
// the old 64^3 tile products shared-memory-issue bound); the 16 lanes that share ty read one address (broadcast).
// Output block (i, j) is the 16x16 block (i, j) of the tile, k block kb its 16 columns 16 kb ..:
//   SKIP = 1: only blocks on or below the diagonal are wanted (symmetric update of a diagonal tile): j > i skipped
//   SKIP = 2: Bs is lower triangular (Bs[c][k] = 0 for k > c): products with kb > j skipped
template <int SKIP>
__device__ __forceinline__ void tile_mma(float (&acc)[4][4], const float* __restrict__ As, const float* __restrict__ Bs,
                                         int ty, int tx) {
#pragma unroll
  for (int kb = 0; kb < 4; ++kb) {
#pragma unroll 1
    for (int k = 16 * kb; k < 16 * kb + 16; k += 4) {
      float4 a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        a[i] = *reinterpret_cast<const float4*>(As + (ty + 16 * i) * LDT + k);
        if (SKIP != 2 || i >= kb) b[i] = *reinterpret_cast<const float4*>(Bs + (tx + 16 * i) * LDT + k);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (SKIP == 1 && j > i) continue;
          if (SKIP == 2 && j < kb) continue;
          acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
          acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
          acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
          acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
        }
    }
  }
}

// Rows [32 h, 32 h + 32) of  R -= Pr Pd^T  (64 columns, K = 64) by 128 threads (t = 0..127): the previous panel's update
// of a row tile, done by warps 4..7 of chol_panel_kernel UNDERNEATH the factorisation of the diagonal tile (warps 0 / 1),
// one half per trip of chol_factor_invert64.  Same k order per output as tile_mma<0>.
__device__ __forceinline__ void chol_r_update_half(float* __restrict__ R, const float* __restrict__ Pr,
                                                   const float* __restrict__ Pd, int h, int t) {
  const int ty = t >> 4, tx = t & 15;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const float* Ar = Pr + (32 * h + ty) * LDT;
  const float* Br = Pd + tx * LDT;
#pragma unroll 1
  for (int k = 0; k < NB; k += 4) {
    float4 a[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      a[i] = *reinterpret_cast<const float4*>(Ar + 8 * i * LDT + k);
      b[i] = *reinterpret_cast<const float4*>(Br + 16 * i * LDT + k);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
        acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
        acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
        acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
      }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) R[(32 * h + ty + 8 * i) * LDT + tx + 16 * j] -= acc[i][j];
}

// One warp factors the 32x32 SPD block at (o, o) of D in registers (lane l owns row l), right-looking: the column below
// pivot j is used UNSCALED in the update,  A[l][k] -= (A[l][j] / d_j) * A[k][j],  and scaled to L[l][j] = A[l][j] *
// rsqrt(d_j) off the critical path.  Column j travels between the lanes through shared memory: every lane parks its
// A[l][j] in a 32-float buffer (double-buffered by pivot parity, one __syncwarp per pivot) and reads the pivot and the
// A[k][j], k > j, back as broadcast LDS.128 — 145 vector loads per block where one shuffle per (j, k) pair took 528
// (stage clocks: 153 cycles per pivot with shuffles, shuffle-issue bound).  The scaled columns go to Lt[j][l] =
// L[l][j] for chol_invert32 / the panel substitution, 1 / L_jj to rinv; L (lower, zero upper) is written back into D.
// Returns false on a non-positive (or NaN) pivot.
__device__ __forceinline__ bool chol_factor32(float* __restrict__ D, float* __restrict__ rinv, int o, int lane,
                                              float* __restrict__ Lt, float* __restrict__ ucol) {
  constexpr int SB = 32;
  constexpr unsigned FULL = 0xffffffffu;
  float a[SB];
#pragma unroll
  for (int k = 0; k < SB; ++k) a[k] = D[(o + lane) * LDT + o + k];
  bool bad = false;
#pragma unroll
  for (int j = 0; j < SB; ++j) {
    float* uc = ucol + (j & 1) * SB;
    uc[lane] = a[j];
    __syncwarp();
    const float d = uc[j];
    if (!(d > 0.f)) bad = true;
    // 1 / d as the square of the refined rsqrt (no slow-path branch in the pivot chain)
    float rs = rsqrtf(d);
    rs = rs * (1.5f - 0.5f * d * rs * rs);
    const float t = a[j] * (rs * rs);
#pragma unroll
    for (int g = (j + 1) / 4; g < SB / 4; ++g) {
      const float4 u = *reinterpret_cast<const float4*>(uc + 4 * g);
      if (4 * g + 0 > j) a[4 * g + 0] = fmaf(-t, u.x, a[4 * g + 0]);   // lanes l < k compute don't-care values
      if (4 * g + 1 > j) a[4 * g + 1] = fmaf(-t, u.y, a[4 * g + 1]);
      if (4 * g + 2 > j) a[4 * g + 2] = fmaf(-t, u.z, a[4 * g + 2]);
      if (4 * g + 3 > j) a[4 * g + 3] = fmaf(-t, u.w, a[4 * g + 3]);
    }
    Lt[j * SB + lane] = (lane > j) ? a[j] * rs : ((lane == j) ? d * rs : 0.f);
    if (lane == 0) rinv[o + j] = rs;
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < SB; ++k) D[(o + lane) * LDT + o + k] = Lt[k * SB + lane];
  return !__any_sync(FULL, bad);
}

// inv(L) of the block chol_factor32 left in Lt / rinv: lane c solves L x = e_c forwards (right-looking); column i of
// L is row i of Lt, read as broadcast LDS.128.  Writes the block (lower, zero upper) into X.
__device__ __forceinline__ void chol_invert32(float* __restrict__ X, int o, int lane, const float* __restrict__ Lt,
                                              const float* __restrict__ rinv) {
  constexpr int SB = 32;
  float x[SB];
#pragma unroll
  for (int i = 0; i < SB; ++i) x[i] = (i == lane) ? 1.f : 0.f;
#pragma unroll
  for (int i = 0; i < SB; ++i) {
    x[i] *= rinv[o + i];
#pragma unroll
    for (int g = (i + 1) / 4; g < SB / 4; ++g) {
      const float4 l = *reinterpret_cast<const float4*>(Lt + i * SB + 4 * g);
      if (4 * g + 0 > i) x[4 * g + 0] = fmaf(-l.x, x[i], x[4 * g + 0]);
      if (4 * g + 1 > i) x[4 * g + 1] = fmaf(-l.y, x[i], x[4 * g + 1]);
      if (4 * g + 2 > i) x[4 * g + 2] = fmaf(-l.z, x[i], x[4 * g + 2]);
      if (4 * g + 3 > i) x[4 * g + 3] = fmaf(-l.w, x[i], x[4 * g + 3]);
    }
  }
#pragma unroll
  for (int k = 0; k < SB; ++k) X[(o + k) * LDT + o + lane] = (k >= lane) ? x[k] : 0.f;
}

// Factor the 64x64 SPD tile D (lower triangle; identity-padded by the caller when the panel is narrower) in place
// and build X = inv(L), by 256 threads: the two 32x32 diagonal blocks are factored and inverted by warp 0 in registers
// (above), the 32x32 panel between them by substitution on warp 1 WHILE warp 0 inverts the first block, the trailing
// update and the off-diagonal block of the inverse (X10 = -X11 L10 X00) on all warps.  Returns false on a non-positive
// pivot.
// The two diagonal blocks go through ONE copy of the factor / invert code (a two-trip loop that is not unrolled):
// those ~2,400 straight-line instructions are executed by a single warp and ncu shows them waiting for instruction
// fetch 3 cycles out of 4 the first time through (and the second block's copy, when the code was inlined twice, taking
// several times longer than the first); on the second trip they come from the instruction cache.
// R / Pr / Pd (Pr may be null): the row tile's pending update R -= Pr Pd^T, worked off by warps 4..7 while they would
// otherwise wait for warp 0 (above).
__device__ __forceinline__ bool chol_factor_invert64(float* __restrict__ D, float* __restrict__ X,
                                                     float* __restrict__ rinv, int* __restrict__ s_fail,
                                                     float* __restrict__ Lt, float* __restrict__ ucol,
                                                     float* __restrict__ R, const float* __restrict__ Pr,
                                                     const float* __restrict__ Pd CHOL_MARK_ARG) {
  constexpr int H = 32;
  const int tid = threadIdx.x, lane = tid & 31;
  // the warp index through a shuffle: tells the compiler it is warp-uniform, so the shuffles inside the per-warp
  // roles below are not each wrapped in a WARPSYNC / reconvergence sequence
  const int wid = __shfl_sync(0xffffffffu, tid >> 5, 0);
  for (int i = tid; i < NB * NB; i += 256) X[(i >> 6) * LDT + (i & 63)] = 0.f;
  if (tid == 0) *s_fail = 0;
  __syncthreads();
#pragma unroll 1
  for (int blk = 0; blk < 2; ++blk) {
    const int o = blk * H;
    if (wid == 0) {
      CHOL_MARK(3);                                         // barrier / trailing update since the last mark
      if (!chol_factor32(D, rinv, o, lane, Lt, ucol) && lane == 0) *s_fail = 1;
      CHOL_MARK(4 + 4 * blk);                               // factor32
      __threadfence_block();
      asm volatile("bar.arrive 1, 64;" ::: "memory");      // L_oo, Lt and rinv are in shared memory: warp 1 may go on
      chol_invert32(X, o, lane, Lt, rinv);
      CHOL_MARK(5 + 4 * blk);                               // invert32
    } else {
      if (wid == 1) asm volatile("bar.sync 1, 64;" ::: "memory");
      if (blk == 0) {
        if (wid == 1 && !*s_fail) {
          // panel: row r = 32 + lane of L10 = A10 * L00^-T by forward substitution (right-looking; column c of L00 is
          // row c of Lt: broadcast LDS.128)
          float a[H];
#pragma unroll
          for (int c = 0; c < H; ++c) a[c] = D[(H + lane) * LDT + c];
#pragma unroll
          for (int c = 0; c < H; ++c) {
            a[c] *= rinv[c];
#pragma unroll
            for (int g = (c + 1) / 4; g < H / 4; ++g) {
              const float4 l = *reinterpret_cast<const float4*>(Lt + c * H + 4 * g);
              if (4 * g + 0 > c) a[4 * g + 0] = fmaf(-a[c], l.x, a[4 * g + 0]);
              if (4 * g + 1 > c) a[4 * g + 1] = fmaf(-a[c], l.y, a[4 * g + 1]);
              if (4 * g + 2 > c) a[4 * g + 2] = fmaf(-a[c], l.z, a[4 * g + 2]);
              if (4 * g + 3 > c) a[4 * g + 3] = fmaf(-a[c], l.w, a[4 * g + 3]);
            }
          }
#pragma unroll
          for (int c = 0; c < H; ++c) D[(H + lane) * LDT + c] = a[c];
        }
        if (Pr && wid >= 4) chol_r_update_half(R, Pr, Pd, 0, tid - 128);
      } else {
        // Y = L10 * X00 (32x32, X00 lower) into the X10 block, by warps 1..7 underneath the second block's factorisation
        for (int idx = tid - 32; idx < 256; idx += 224) {
          const int r = idx >> 3, c0 = (idx & 7) * 4;
          float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
          for (int k = 0; k < H; ++k) {
            const float l = D[(H + r) * LDT + k];
            const float4 xv = *reinterpret_cast<const float4*>(X + k * LDT + c0);
            s[0] = fmaf(l, xv.x, s[0]); s[1] = fmaf(l, xv.y, s[1]); s[2] = fmaf(l, xv.z, s[2]); s[3] = fmaf(l, xv.w, s[3]);
          }
          *reinterpret_cast<float4*>(X + (H + r) * LDT + c0) = make_float4(s[0], s[1], s[2], s[3]);
        }
        if (Pr && wid >= 4) chol_r_update_half(R, Pr, Pd, 1, tid - 128);
      }
    }
    __syncthreads();
    CHOL_MARK(6 + 4 * blk);                                 // waiting for the other warps of this trip
    if (*s_fail) return false;
    if (blk == 0) {
      // trailing update D11 -= L10 L10^T: thread -> row r, 4 columns; blocks above the diagonal are skipped
      const int r = tid >> 3, c0 = (tid & 7) * 4;
      if (c0 <= r) {
        float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < H; k += 4) {
          const float4 av = *reinterpret_cast<const float4*>(D + (H + r) * LDT + k);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 bv = *reinterpret_cast<const float4*>(D + (H + c0 + j) * LDT + k);
            s[j] = fmaf(av.x, bv.x, s[j]); s[j] = fmaf(av.y, bv.y, s[j]);
            s[j] = fmaf(av.z, bv.z, s[j]); s[j] = fmaf(av.w, bv.w, s[j]);
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) D[(H + r) * LDT + H + c0 + j] -= s[j];
      }
      __syncthreads();
    }
  }
  {
    // X10 = -X11 * Y (X11 lower: k <= r)
    const int r = tid >> 3, c0 = (tid & 7) * 4;
    float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
    for (int k = 0; k < H; ++k) {
      const float xr = X[(H + r) * LDT + H + k];
      const float4 yv = *reinterpret_cast<const float4*>(X + (H + k) * LDT + c0);
      s[0] = fmaf(xr, yv.x, s[0]); s[1] = fmaf(xr, yv.y, s[1]); s[2] = fmaf(xr, yv.z, s[2]); s[3] = fmaf(xr, yv.w, s[3]);
    }
    __syncthreads();
    *reinterpret_cast<float4*>(X + (H + r) * LDT + c0) = make_float4(-s[0], -s[1], -s[2], -s[3]);
  }
  __syncthreads();
  return true;
}

__device__ __forceinline__ void cp_async16(float* dst_smem, const float* src, bool valid) {
  const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(dst_smem));
  const int sz = valid ? 16 : 0;             // 0 source bytes: the 16 destination bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// Full-width tiles (64 valid columns) by cp.async: all of a CTA's tile loads are in flight together and nothing waits
// in registers (stage clocks: the load-process-store loops below cost 8-10k cycles per launch, four L2 round trips
// back to back).  Rows past `nrows` are zero-filled.
__device__ __forceinline__ void chol_issue_tile(float* __restrict__ T, const float* __restrict__ A, int lda, int row0,
                                                int col0, int nrows) {
#pragma unroll
  for (int q = 0; q < NB * (NB / 4) / 256; ++q) {
    const int i = threadIdx.x + q * 256;
    const int r = i >> 4, c = (i & 15) * 4;
    const bool ok = r < nrows;
    cp_async16(T + r * LDT + c, A + (int64_t)(row0 + (ok ? r : 0)) * lda + col0 + c, ok);
  }
}
// strict upper triangle of a freshly loaded diagonal tile -> 0 (the workspace above the diagonal is never written)
__device__ __forceinline__ void chol_zero_upper(float* __restrict__ T) {
  for (int i = threadIdx.x; i < NB * (NB / 4); i += 256) {
    const int r = i >> 4, c = (i & 15) * 4;
    if (c + 3 <= r) continue;
    float4 v = *reinterpret_cast<float4*>(T + r * LDT + c);
    if (c + 0 > r) v.x = 0.f;
    if (c + 1 > r) v.y = 0.f;
    if (c + 2 > r) v.z = 0.f;
    if (c + 3 > r) v.w = 0.f;
    *reinterpret_cast<float4*>(T + r * LDT + c) = v;
  }
}

// global (row-major, leading dimension lda, 64-float aligned columns) -> shared tile, 4 floats per thread and step
//   mode 0: rectangular rows [row0, row0 + nrows) x cols [col0, col0 + ncols), zero elsewhere
//   mode 1: lower triangle of the diagonal tile, identity where r >= nrows or c >= ncols (narrow last panel)
__device__ __forceinline__ void chol_load_tile(float* __restrict__ T, const float* __restrict__ A, int lda, int row0,
                                               int col0, int nrows, int ncols, int mode) {
  for (int i = threadIdx.x; i < NB * (NB / 4); i += 256) {
    const int r = i >> 4, c = (i & 15) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < nrows) v = *reinterpret_cast<const float4*>(A + (int64_t)(row0 + r) * lda + col0 + c);
    float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int cc = c + q;
      if (mode == 0) {
        if (cc >= ncols) e[q] = 0.f;
      } else {
        if (r >= nrows || cc >= ncols) e[q] = (r == cc) ? 1.f : 0.f;
        else if (cc > r) e[q] = 0.f;
      }
    }
    *reinterpret_cast<float4*>(T + r * LDT + c) = make_float4(e[0], e[1], e[2], e[3]);
  }
}

// ONE launch per panel on the critical path.  CTA 0 owns the diagonal tile, CTA i >= 1 the i-th 64-row tile below it
// (the last one includes the y row).  Every CTA
//   1. applies the PREVIOUS panel's trailing update to what it needs of this panel's column — the diagonal tile and
//      its own row tile:  D -= Lp_d Lp_d^T,  R -= Lp_r Lp_d^T  (the bulk update of the previous panel, which runs
//      underneath this kernel, leaves this column alone); only the diagonal tile's update is on the critical path,
//      the row tile's is done by warps 4..7 underneath step 2 (two warps' work, the other six used to wait),
//   2. factors the diagonal tile and inverts the factor — redundantly in every CTA, so that there is no second
//      launch and no dependency between "diag" and "trsm" (identical inputs and code: identical results),
//   3. forms its rows of L21 = R inv(L11)^T as one 64^3 product and stores them.
// CTA 0 stores L11 and inv(L11) (for the back substitution) instead of step 3.
__global__ void __launch_bounds__(256)
chol_panel_kernel(float* __restrict__ A, int lda, int n, int p0, int has_prev, float* __restrict__ Linv,
                  int* __restrict__ meta) {
  extern __shared__ __align__(16) float chol_sm[];
  float* Pd = chol_sm;               // previous panel's L, rows of the diagonal tile
  float* Pr = chol_sm + TILE;        // previous panel's L, rows of this CTA's tile
  float* D = chol_sm + 2 * TILE;
  float* R = chol_sm + 3 * TILE;
  float* X = chol_sm + 4 * TILE;     // inv(L11)
  __shared__ float rinv[NB];
  __shared__ int s_fail;
  __shared__ __align__(16) float Lt[32 * 32];   // columns of the 32x32 block being factored, as rows
  __shared__ __align__(16) float ucol[2 * 32];  // the current pivot's unscaled column, double-buffered
  if (meta[1]) return;
  const int tid = threadIdx.x;
  const int nb = min(NB, n - p0);
  const bool diag_cta = blockIdx.x == 0;
  const int r0 = p0 + nb + ((int)blockIdx.x - 1) * NB;
  const int nr = diag_cta ? 0 : min(NB, n + 1 - r0);
  CHOL_MARK_DECL;
  if (nb == NB) {                    // every panel but a ragged last one
    chol_issue_tile(D, A, lda, p0, p0, NB);
    if (!diag_cta) chol_issue_tile(R, A, lda, r0, p0, nr);
    if (has_prev) {
      chol_issue_tile(Pd, A, lda, p0, p0 - NB, NB);
      if (!diag_cta) chol_issue_tile(Pr, A, lda, r0, p0 - NB, nr);
    }
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();
    chol_zero_upper(D);
  } else {
    chol_load_tile(D, A, lda, p0, p0, nb, nb, 1);
    if (!diag_cta) chol_load_tile(R, A, lda, r0, p0, nr, nb, 0);
    if (has_prev) {
      chol_load_tile(Pd, A, lda, p0, p0 - NB, nb, NB, 0);
      if (!diag_cta) chol_load_tile(Pr, A, lda, r0, p0 - NB, nr, NB, 0);
    }
  }
  __syncthreads();
  CHOL_MARK(1);                       // tile loads
  const int tx = tid & 15, ty = tid >> 4;
  if (has_prev) {
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    tile_mma<1>(acc, Pd, Pd, ty, tx);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = ty + 16 * i, c = tx + 16 * j;
        if (c <= r) D[r * LDT + c] -= acc[i][j];       // rows / columns >= nb of Pd are zero: the padding survives
      }
    __syncthreads();
  }
  CHOL_MARK(2);                       // previous panel's update of the diagonal tile (the row tile's follows below)
  // R -= Pr Pd^T is not needed before the L21 product: warps 4..7 do it underneath the factorisation
  const bool pending_r = has_prev && !diag_cta;
  if (!chol_factor_invert64(D, X, rinv, &s_fail, Lt, ucol, R, pending_r ? Pr : nullptr, Pd CHOL_MARK_PASS)) {
    if (tid == 0) meta[1] = 1;
    return;
  }
  if (diag_cta) {
    for (int i = tid; i < NB * NB; i += 256) {
      const int r = i >> 6, c = i & 63;
      if (r < nb && c <= r) A[(int64_t)(p0 + r) * lda + p0 + c] = D[r * LDT + c];
      Linv[i] = X[r * LDT + c];
    }
    return;
  }
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  CHOL_MARK(12);                      // X10 = -X11 Y and its barriers
  tile_mma<2>(acc, R, X, ty, tx);    // out[r][c] = sum_k R[r][k] X[c][k]  (X lower: terms with k > c are zero)
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = ty + 16 * i, c = tx + 16 * j;
      if (r < nr && c < nb) A[(int64_t)(r0 + r) * lda + p0 + c] = acc[i][j];
    }
  CHOL_MARK(13);                      // L21 product + store
#ifdef DPVO_CHOL_PROFILE
  if (threadIdx.x == 0 && blockIdx.x == 1) atomicAdd(reinterpret_cast<unsigned long long*>(&g_chol_prof[31]), 1ull);
#endif
}

// Bulk of panel p0's trailing update: tile (ti, tj), 1 <= tj <= ti, of the rows / columns starting at q0 = p0 + nb.
// The first tile column (tj = 0) is the next panel's and is updated by that panel's chol_panel_kernel.
__global__ void __launch_bounds__(256)
chol_syrk_kernel(float* __restrict__ A, int lda, int n, int p0, const int* __restrict__ meta) {
  __shared__ __align__(16) float As[TILE];
  __shared__ __align__(16) float Bs[TILE];
  if (meta[1]) return;
  const int t = blockIdx.x;
  int ti = (int)((sqrtf(8.f * t + 1.f) - 1.f) * 0.5f);
  while ((ti + 1) * (ti + 2) / 2 <= t) ++ti;
  while (ti * (ti + 1) / 2 > t) --ti;
  int tj = t - ti * (ti + 1) / 2;
  ++ti; ++tj;
  const int nb = min(NB, n - p0);
  const int q0 = p0 + nb;
  const int ri = q0 + ti * NB, rj = q0 + tj * NB;
  const int nri = min(NB, n + 1 - ri);        // rows may include the y row
  const int nrj = min(NB, n - rj);            // columns never do
  if (nrj <= 0 || nri <= 0) return;
  chol_load_tile(As, A, lda, ri, p0, nri, nb, 0);
  chol_load_tile(Bs, A, lda, rj, p0, nrj, nb, 0);
  __syncthreads();
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  tile_mma<0>(acc, As, Bs, ty, tx);
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = ty + 16 * i, c = tx + 16 * j;
      if (r < nri && c < nrj && (rj + c) <= (ri + r)) A[(int64_t)(ri + r) * lda + rj + c] -= acc[i][j];
    }
}

// The same bulk update as ONE persistent CTA per SM over 128x128 tiles, 8x8 outputs per thread (rows ty + 16 i,
// columns tx + 16 j): per 4 k-steps a thread issues 16 LDS.128 for 256 FMAs (the 64x64 kernel: 8 for 64, LDS issue and
// FMA issue neck and neck) and the L21 rows cross L2 -> SM half as often.  The next tile's rows arrive by cp.async in a
// second buffer while this one is multiplied; the result leaves as fire-and-forget reductions (RED.ADD.F32 of -acc:
// every element has exactly one writer per launch, so the sum is the same single rounding as load-subtract-store and
// no thread waits for the old value).
// Why persistent and why 136 KB of shared memory: the factorisation's critical path (chol_panel_kernel, on the
// high-priority stream) runs UNDERNEATH this kernel and needs a free slot on <= 95 SMs the moment it is launched.  Two
// buffers make a second bulk CTA impossible on an SM (2 x 136 KB > 227 KB) while one bulk CTA + one panel CTA always fit
// (136 + 68 KB, 127 + 121 registers x 256 threads) — with short-lived 2-per-SM CTAs of this size the panel kernels
// queued behind 10 us tiles and the look-ahead was lost.
// Used while the trailing matrix has at least one 128-tile per SM (ba_chol_solve picks); tile (ti, tj), tj <= ti,
// starts at row / column c0 + 128 ti / c0 + 128 tj with c0 = p0 + 128 (the 64 columns before c0 are the next panel's).
constexpr int SB = 128;
constexpr int kSyrkStage = 2 * SB * LDT;                      // floats: rows of the tile's row block + of its column block
constexpr int kSyrk128SmemBytes = 2 * kSyrkStage * (int)sizeof(float);
constexpr int kSyrk128MinTiles = kNumSMs;    // fewer 128-tiles than SMs: the 64x64 kernel spreads the work better

__device__ __forceinline__ uint32_t tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

__device__ __forceinline__ void syrk_tile_coords(int t, int& ti, int& tj) {
  ti = (int)((sqrtf(8.f * t + 1.f) - 1.f) * 0.5f);
  while ((ti + 1) * (ti + 2) / 2 <= t) ++ti;
  while (ti * (ti + 1) / 2 > t) --ti;
  tj = t - ti * (ti + 1) / 2;
}

// rows [row0, row0 + 128) x the panel's 64 columns -> T (stride LDT); rows past `nrows` are zero-filled
__device__ __forceinline__ void syrk_issue_rows(float* __restrict__ T, const float* __restrict__ A, int lda, int row0,
                                                int col0, int nrows) {
#pragma unroll
  for (int q = 0; q < SB * (NB / 4) / 256; ++q) {
    const int i = threadIdx.x + q * 256;
    const int r = i >> 4, c = (i & 15) * 4;
    const bool ok = r < nrows;
    cp_async16(T + r * LDT + c, A + (int64_t)(row0 + (ok ? r : 0)) * lda + col0 + c, ok);
  }
}

// 0: split-TF32 tensor-core tiles (default); 1: fp32 FMA tiles (dpvo_chol_bulk_mode, A/B measurements and tests)
static int g_chol_bulk_mode = 0;

__global__ void __launch_bounds__(256, 2)   // <= 128 registers: a panel CTA (121) must fit beside this one
chol_syrk128_kernel(float* __restrict__ A, int lda, int n, int p0, int ntiles, const int* __restrict__ meta) {
  extern __shared__ __align__(16) float syrk_sm[];
  if (meta[1]) return;
  const int c0 = p0 + 2 * NB;                 // only full panels have a bulk update
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  int t = blockIdx.x;
  if (t >= ntiles) return;
  {
    int ti, tj;
    syrk_tile_coords(t, ti, tj);
    const int ri = c0 + ti * SB, rj = c0 + tj * SB;
    syrk_issue_rows(syrk_sm, A, lda, ri, p0, min(SB, n + 1 - ri));
    if (ti != tj) syrk_issue_rows(syrk_sm + SB * LDT, A, lda, rj, p0, min(SB, n - rj));
    cp_async_commit();
  }
  for (int it = 0; t < ntiles; t += gridDim.x, ++it) {
    float* As = syrk_sm + (it & 1) * kSyrkStage;
    int ti, tj;
    syrk_tile_coords(t, ti, tj);
    const int ri = c0 + ti * SB, rj = c0 + tj * SB;
    const int nri = min(SB, n + 1 - ri);      // rows may include the y row
    const int nrj = min(SB, n - rj);          // columns never do (<= 0: the tile holds nothing, e.g. the y row's own column)
    const float* Bs = (ti == tj) ? As : As + SB * LDT;      // diagonal tile: the same rows on both sides
    const int tn = t + gridDim.x;
    if (tn < ntiles) {
      float* An = syrk_sm + ((it + 1) & 1) * kSyrkStage;
      int ni, nj;
      syrk_tile_coords(tn, ni, nj);
      const int rni = c0 + ni * SB, rnj = c0 + nj * SB;
      syrk_issue_rows(An, A, lda, rni, p0, min(SB, n + 1 - rni));
      if (ni != nj) syrk_issue_rows(An + SB * LDT, A, lda, rnj, p0, min(SB, n - rnj));
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (nrj > 0) {
      float acc[8][8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll 1
      for (int k = 0; k < NB; k += 4) {
        float4 a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4*>(As + (ty + 16 * i) * LDT + k);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 b = *reinterpret_cast<const float4*>(Bs + (tx + 16 * j) * LDT + k);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            acc[i][j] = fmaf(a[i].x, b.x, acc[i][j]);
            acc[i][j] = fmaf(a[i].y, b.y, acc[i][j]);
            acc[i][j] = fmaf(a[i].z, b.z, acc[i][j]);
            acc[i][j] = fmaf(a[i].w, b.w, acc[i][j]);
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = ty + 16 * i;
        if (r >= nri) continue;
        float* row = A + (int64_t)(ri + r) * lda + rj;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = tx + 16 * j;
          if (c < nrj && (rj + c) <= (ri + r)) atomicAdd(&row[c], -acc[i][j]);   // result unused: RED.ADD.F32
        }
      }
    }
    __syncthreads();                          // this buffer is the target of the next iteration's prefetch
  }
}

// ---- the same tiles on the 5th-generation tensor cores: split-TF32 ("3xTF32") with tcgen05.mma.
// Every fp32 operand x is split into hi = tf32(x) and lo = tf32(x - hi) (x - hi is exact in fp32), and
//   a b ~= a_lo b_hi + a_hi b_lo + a_hi b_hi      (the dropped a_lo b_lo term is 2^-22 relative)
// is accumulated in fp32 in tensor memory, small terms first: 24 UMMA instructions (M128 x N128 x K8) per tile.
// Measured against numpy float64 at n = 5,994: 5.8e-7 (split) vs 4.4e-7 (fp32 FMA) on the solve.
// (mma.sync m16n8k8 TF32 was tried first: HMMA.1688.F32.TF32 retires one instruction per ~17 cycles per sub-partition
//  on this part, so three products cost exactly what the FMA loop costs — the legacy path is no tensor core to speak of.)
// Pipeline per persistent CTA (one per SM; 128 KB of shared memory keeps a second one out and a panel CTA in):
//   TMA (128-byte swizzle, zero fill past the last row) brings the 128 x 64 fp32 row blocks of tile t  ->  all threads
//   split them IN PLACE (hi) and into a second buffer (lo) — element-wise, so the swizzle never matters  ->  one
//   thread issues the 24 MMAs and commits  ->  as soon as they have read shared memory the TMA of tile t + 1 starts
//   underneath the epilogue: TMEM -> registers (a thread owns a row, 2 x 32 columns) -> red.global.add.v4.f32.
// Per tile (stage clocks of a development build): split 1.8k cycles, MMAs 1.8k, epilogue 5.0k + 0.9k at its barrier,
// of 10.4k.  The epilogue runs at the L2's reduction rate (~13 bytes per clock per SM with every SM reducing):
// coalescing the reductions through shared memory (8 lanes per 128-byte row segment) changed nothing, a prefetched
// load - subtract - store of the thread's row was slower (7.9k cycles), and a warp-specialised version (control warp
// for TMA / MMA, two accumulators, epilogue of tile t - 1 underneath the MMAs of tile t) took the same time per
// launch.  Tiles are assigned with a stride of gridDim: giving a CTA a contiguous run of tiles (to keep the row
// block's operand resident) made the solve irreproducible — about one solve in ten off by 1e-2 at 6N = 4,890, with
// or without the reuse, cause not found — so it is not done; tests/test_ba_gpu.py checks that both arithmetics
// repeat bit for bit.
constexpr int kUmmaAtomBytes = SB * 128;                   // 128 rows x 32 fp32: one K-atom of an operand (16 KB)
constexpr int kUmmaOperandBytes = 2 * kUmmaAtomBytes;      // K = 64: two atoms
constexpr int kSyrkUmmaSmemBytes = 4 * kUmmaOperandBytes + 1024;   // A hi, B hi, A lo, B lo + alignment slack
constexpr uint32_t kSyrkTmemCols = 128;

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(256, 1)
chol_syrk128_umma_kernel(const __grid_constant__ CUtensorMap tmA, float* __restrict__ A, int lda, int n, int p0,
                         int ntiles, const int* __restrict__ meta) {
  using namespace sm100;
  extern __shared__ uint8_t umma_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(umma_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_hi = smem;
  uint8_t* b_hi = smem + kUmmaOperandBytes;
  uint8_t* a_lo = smem + 2 * kUmmaOperandBytes;
  uint8_t* b_lo = smem + 3 * kUmmaOperandBytes;
  __shared__ uint64_t full, mma_done;
  __shared__ uint32_t tmem_base_s;
  if (meta[1]) return;
  if ((int)blockIdx.x >= ntiles) return;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c0 = p0 + 2 * NB;

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&tmA);
      mbar_init(&full, 1);
      mbar_init(&mma_done, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(&tmem_base_s, kSyrkTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  auto issue_loads = [&](int t) {              // one thread
    int ti, tj;
    syrk_tile_coords(t, ti, tj);
    const int ri = c0 + ti * SB, rj = c0 + tj * SB;
    mbar_arrive_expect_tx(&full, (ti == tj ? 1 : 2) * kUmmaOperandBytes);
    tma_load_2d(a_hi, &tmA, &full, p0, ri);
    tma_load_2d(a_hi + kUmmaAtomBytes, &tmA, &full, p0 + 32, ri);
    if (ti != tj) {
      tma_load_2d(b_hi, &tmA, &full, p0, rj);
      tma_load_2d(b_hi + kUmmaAtomBytes, &tmA, &full, p0 + 32, rj);
    }
  };
  if (tid == 0) issue_loads(blockIdx.x);

  int it = 0;
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
    int ti, tj;
    syrk_tile_coords(t, ti, tj);
    const int ri = c0 + ti * SB, rj = c0 + tj * SB;
    const int nri = min(SB, n + 1 - ri);
    const int nrj = min(SB, n - rj);
    const bool diag = ti == tj;
    mbar_wait(&full, it & 1);
    // hi / lo split, 16 bytes per thread and step (A and, off the diagonal, B are contiguous)
    {
      const int nvec = (diag ? 1 : 2) * kUmmaOperandBytes / 16;
      float4* hi = reinterpret_cast<float4*>(a_hi);
      float4* lo = reinterpret_cast<float4*>(a_lo);
      for (int q = tid; q < nvec; q += 256) {
        const float4 x = hi[q];
        float4 h, l;
        h.x = __uint_as_float(tf32_rna(x.x)); l.x = __uint_as_float(tf32_rna(x.x - h.x));
        h.y = __uint_as_float(tf32_rna(x.y)); l.y = __uint_as_float(tf32_rna(x.y - h.y));
        h.z = __uint_as_float(tf32_rna(x.z)); l.z = __uint_as_float(tf32_rna(x.z - h.z));
        h.w = __uint_as_float(tf32_rna(x.w)); l.w = __uint_as_float(tf32_rna(x.w - h.w));
        hi[q] = h;
        lo[q] = l;
      }
    }
    fence_proxy_async_smem();                 // generic-proxy writes -> visible to the tensor core's operand reads
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      constexpr uint32_t idesc = instr_desc_tf32(SB, SB);
      const uint32_t ah = smem_u32(a_hi), al = smem_u32(a_lo);
      const uint32_t bh = diag ? ah : smem_u32(b_hi), bl = diag ? al : smem_u32(b_lo);
      uint32_t accf = 0;
#pragma unroll
      for (int prod = 0; prod < 3; ++prod) {
        const uint32_t pa = prod == 0 ? al : ah;
        const uint32_t pb = prod == 1 ? bl : bh;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t off = (ks >> 2) * kUmmaAtomBytes + (ks & 3) * 32;
          mma_tf32_ss(tmem_base, smem_desc_sw128(pa + off), smem_desc_sw128(pb + off), idesc, accf);
          accf = 1;
        }
      }
      mma_commit(&mma_done);
    }
    mbar_wait(&mma_done, it & 1);
    tc_fence_after();
    if (tid == 0 && t + (int)gridDim.x < ntiles) issue_loads(t + gridDim.x);   // operands have been read
    // epilogue: warp w reads TMEM lanes 32 (w % 4) .., columns 64 (w / 4) ..; thread = one row, 2 x 32 columns
    {
      const int q = warp & 3, hcol = (warp >> 2) * 64;
      const int r = q * 32 + lane;
      float* row = A + (int64_t)(ri + (r < nri ? r : 0)) * lda + rj;
      const bool fast = !diag && nrj == SB;
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + hcol + c * 32, v);
        tmem_ld_wait();
        const int cb = hcol + c * 32;
        if (r < nri) {
          if (fast) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              red_add_v4(row + cb + j, -__uint_as_float(v[j]), -__uint_as_float(v[j + 1]), -__uint_as_float(v[j + 2]),
                         -__uint_as_float(v[j + 3]));
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int col = cb + j;
              if (col < nrj && (rj + col) <= (ri + r)) atomicAdd(&row[col], -__uint_as_float(v[j]));
            }
          }
        }
      }
    }
    tc_fence_before();
    __syncthreads();                          // tensor memory drained before the next tile's MMAs overwrite it
  }
  if (warp == 0) tmem_dealloc(tmem_base, kSyrkTmemCols);
}

// Back substitution L^T x = z, FOUR panels per launch (the launches are a serial chain: 94 of them at 6N = 5,994 took
// 0.65 ms of a 4.1 ms solve).  Panels pl..ph, from ph down: every CTA redundantly forms
//   x_p = inv(L_pp)^T (z_p - acc_p - sum_{p' in (p, ph]} L[p', p]^T x_p')
// from shared-memory copies of the group's inv(L_pp) and of its (up to six) coupling tiles (cp.async, all in flight
// together), then CTA 0 stores the x_p and CTA q+1 folds sum_p L[p rows, q cols]^T x_p into acc_q for its 64-column
// tile q < pl (each acc segment has exactly one writer per launch: no atomics).
constexpr int BSG = 4;
constexpr int kBsSmemBytes = (BSG + BSG * (BSG - 1) / 2) * NB * NB * (int)sizeof(float);   // 160 KB

__global__ void __launch_bounds__(256)
chol_bs_kernel(const float* __restrict__ A, int lda, int n, int pl, int ph, const float* __restrict__ Linv,
               const float* __restrict__ z, float* __restrict__ acc, float* __restrict__ sol,
               const int* __restrict__ meta) {
  extern __shared__ __align__(16) float bs_sm[];
  __shared__ float xs[BSG][NB];
  __shared__ float rhs[NB];
  __shared__ float part[4][NB];
  if (meta[1]) return;
  const int tid = threadIdx.x;
  const int G = ph - pl + 1;
  float* sLinv = bs_sm;                              // [G][64][64]
  float* sL = bs_sm + BSG * NB * NB;                 // [pair(b, a), a < b][64 rows of panel pl + b][64 columns of panel pl + a]
  for (int g = 0; g < G; ++g) {
    const float* src = Linv + (int64_t)(pl + g) * NB * NB;
    for (int i = tid; i < NB * NB / 4; i += 256) cp_async16(sLinv + g * NB * NB + 4 * i, src + 4 * i, true);
  }
  for (int b = 1; b < G; ++b) {
    const int nbb = min(NB, n - (pl + b) * NB);      // valid rows of panel pl + b
    for (int a = 0; a < b; ++a) {
      float* dst = sL + (b * (b - 1) / 2 + a) * NB * NB;
      for (int i = tid; i < NB * NB / 4; i += 256) {
        const int r = i >> 4, c = (i & 15) * 4;
        const bool ok = r < nbb;
        cp_async16(dst + r * NB + c, A + (int64_t)((pl + b) * NB + (ok ? r : 0)) * lda + (pl + a) * NB + c, ok);
      }
    }
  }
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  const int c = tid & 63, rg = tid >> 6;
  for (int g = G - 1; g >= 0; --g) {
    const int p0 = (pl + g) * NB;
    float s = 0.f;
    for (int b = g + 1; b < G; ++b) {
      const float* t = sL + (b * (b - 1) / 2 + g) * NB * NB;
#pragma unroll 4
      for (int r = rg; r < NB; r += 4) s = fmaf(t[r * NB + c], xs[b][r], s);
    }
    part[rg][c] = s;
    __syncthreads();
    if (tid < NB) {
      const int i = p0 + tid;
      rhs[tid] = (i < n) ? z[i] - acc[i] - (part[0][tid] + part[1][tid] + part[2][tid] + part[3][tid]) : 0.f;
    }
    __syncthreads();
    {
      // x[c] = sum_{k >= c} inv(L)[k][c] rhs[k], four partial sums per column
      const float* li = sLinv + g * NB * NB;
      float v = 0.f;
      for (int k = c + rg; k < NB; k += 4) v = fmaf(li[k * NB + c], rhs[k], v);
      __syncthreads();                               // part[] of the coupling sums has been consumed
      part[rg][c] = v;
    }
    __syncthreads();
    if (tid < NB) xs[g][tid] = part[0][tid] + part[1][tid] + part[2][tid] + part[3][tid];
    __syncthreads();
  }
  if (blockIdx.x == 0) {
    for (int i = tid; i < G * NB; i += 256) {
      const int g = i >> 6, k = i & 63;
      if ((pl + g) * NB + k < n) sol[(pl + g) * NB + k] = xs[g][k];
    }
    return;
  }
  const int c0 = (blockIdx.x - 1) * NB;
  float s = 0.f;
  for (int g = 0; g < G; ++g) {
    const int p0 = (pl + g) * NB;
    const int nb = min(NB, n - p0);
    // 16 independent loads in flight per thread (a rolled load-FMA loop exposed one L2 round trip per row)
    float v[NB / 4];
#pragma unroll
    for (int i = 0; i < NB / 4; ++i) {
      const int r = rg + 4 * i;
      v[i] = (r < nb) ? __ldg(A + (int64_t)(p0 + r) * lda + c0 + c) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < NB / 4; ++i) s = fmaf(v[i], xs[g][rg + 4 * i], s);
  }
  part[rg][c] = s;
  __syncthreads();
  if (tid < NB) acc[c0 + tid] += part[0][tid] + part[1][tid] + part[2][tid] + part[3][tid];
}

__global__ void __launch_bounds__(256)
chol_init_x_kernel(const float* __restrict__ A, int lda, int n, float* __restrict__ z, float* __restrict__ acc,
                   const int* __restrict__ meta) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { z[i] = meta[1] ? 0.f : A[(int64_t)n * lda + i]; acc[i] = 0.f; }
}

__global__ void __launch_bounds__(256)
chol_finish_kernel(const __grid_constant__ BADev p, const float* __restrict__ x) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool failed = p.meta[1] != 0;
  if (i < p.n) p.dX[i] = failed ? 0.f : x[i];
  if (i < p.N && !failed) {
    float* ps = p.poses + (int64_t)(p.t0 + i) * 7;
    const Vec3 t = {ps[0], ps[1], ps[2]};
    const Quat q = {ps[3], ps[4], ps[5], ps[6]};
    const float xi[6] = {x[6 * i + 0], x[6 * i + 1], x[6 * i + 2], x[6 * i + 3], x[6 * i + 4], x[6 * i + 5]};
    Vec3 t1;
    Quat q1;
    retr_se3(xi, t, q, t1, q1);
    ps[0] = t1.x; ps[1] = t1.y; ps[2] = t1.z;
    ps[3] = q1.x; ps[4] = q1.y; ps[5] = q1.z; ps[6] = q1.w;
  }
}

// scratch floats ba_chol_solve needs behind the (n+1) x lda matrix: z, acc, sol (3n) + inv(L11) per panel
int64_t chol_scratch_floats(int n) { return 3 * (int64_t)n + 64 + (int64_t)((n + NB - 1) / NB) * NB * NB; }

// A: (n+1) x lda, lower triangle + y row already in place.  xbuf: chol_scratch_floats(n) floats.
// DPVO_DEBUG_SYNC=1: synchronise after every launch and name the kernel that faulted
#define CHOL_DBG(what)                                                                              \
  do {                                                                                              \
    if (dbg) {                                                                                      \
      cudaError_t e_ = cudaStreamSynchronize(st);                                                   \
      if (e_ == cudaSuccess) e_ = cudaGetLastError();                                               \
      if (e_ != cudaSuccess) { set_error("ba_chol_solve: %s (p0=%d): %s", what, p0, cudaGetErrorString(e_)); return 1; } \
    }                                                                                               \
  } while (0)

// Look-ahead resources (one set per device): a HIGHEST-priority internal stream for the critical path and a small ring
// of events.  Panel p's trailing update is split: its first tile column (all the NEXT panel needs) follows diag(p) /
// trsm(p) on the critical-path stream; the bulk (thousands of short CTAs) runs on the caller's stream underneath
// diag(p+1) / trsm(p+1), whose few CTAs win every SM slot that frees up because of the stream priority (the caller's
// stream — torch's default — has the lowest priority there is, so the roles cannot be the other way round).
struct CholLookahead {
  cudaStream_t side = nullptr;
  cudaEvent_t trsm_done[2] = {nullptr, nullptr};
  cudaEvent_t rest_done[2] = {nullptr, nullptr};
  cudaEvent_t fork = nullptr;
  bool ok = false;
};
static CholLookahead* chol_lookahead() {
  static CholLookahead per_dev[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  CholLookahead& L = per_dev[dev];
  if (!L.ok) {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);             // hi = most urgent (numerically smallest)
    if (cudaStreamCreateWithPriority(&L.side, cudaStreamNonBlocking, hi) != cudaSuccess) return nullptr;
    bool good = cudaEventCreateWithFlags(&L.fork, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; i < 2 && good; ++i)
      good = cudaEventCreateWithFlags(&L.trsm_done[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&L.rest_done[i], cudaEventDisableTiming) == cudaSuccess;
    if (!good) return nullptr;
    L.ok = true;
  }
  return &L;
}

int ba_chol_solve(const BADev& p, float* A, int lda, float* xbuf, cudaStream_t st) {
  const int n = p.n;
  int launches = 0;
  static const bool dbg = debug_env("DPVO_DEBUG_SYNC") != nullptr;
  static const bool no_la = debug_env("DPVO_CHOL_NO_LOOKAHEAD") != nullptr;
  int p0 = -1;
  float* z = xbuf;
  float* acc = z + n;
  float* sol = acc + n;
  float* Linv = sol + n + (64 - (3 * n) % 64) % 64;     // 256-byte aligned panels
  {
    // per device and cheap: once per solve, not per panel
    cudaError_t e = cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPanelSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(chol_bs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBsSmemBytes);
    if (e != cudaSuccess) {
      set_error("ba_chol_solve: cannot reserve %d bytes of shared memory: %s", kPanelSmemBytes, cudaGetErrorString(e));
      return 1;
    }
  }
  CUtensorMap tmA;
  bool have_map = false;
  {
    const int rows0 = n + 1 - NB, t128 = (rows0 - NB + SB - 1) / SB;      // the first panel's bulk region
    if (n >= 16 * NB && t128 * (t128 + 1) / 2 >= kSyrk128MinTiles) {
      cudaError_t e = cudaFuncSetAttribute(chol_syrk128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSyrk128SmemBytes);
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(chol_syrk128_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSyrkUmmaSmemBytes);
      if (e != cudaSuccess) {
        set_error("ba_chol_solve: cannot reserve shared memory for the bulk update: %s", cudaGetErrorString(e));
        return 1;
      }
      // fp32 matrix as a [n + 1][lda] tensor, boxes of 32 columns x 128 rows (rows past n + 1 read as zero)
      have_map = sm100::make_tensor_map_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, A, (uint64_t)lda, (uint64_t)n + 1,
                                           (uint64_t)lda, 32, SB);
    }
  }
  // look-ahead pays once the trailing update is big enough to hide the next panel under it (>= 16 panels)
  CholLookahead* la = (!dbg && !no_la && n >= 16 * NB) ? chol_lookahead() : nullptr;
  cudaStream_t crit = st;                          // the panel kernels
  if (la) {
    crit = la->side;
    cudaEventRecord(la->fork, st);
    cudaStreamWaitEvent(crit, la->fork, 0);
  }
  // panel(p) = previous panel's update of column p + factor + L21;  bulk(p) = trailing update of columns >= p + 2.
  // bulk(q) writes column p for every q <= p - 2, so panel(p) waits for bulk(p - 2) and runs underneath bulk(p - 1);
  // bulk(p) needs panel(p).  Without look-ahead everything is on one stream in that order.
  int np = 0;
  for (p0 = 0; p0 < n; p0 += NB, ++np) {
    const int nb = n - p0 < NB ? n - p0 : NB;
    const int rows = n + 1 - p0 - nb;            // rows below the panel, incl. the y row
    const int T = (rows + NB - 1) / NB;
    float* Li = Linv + (int64_t)(p0 / NB) * NB * NB;
    const int slot = np & 1;
    if (la && np >= 2) cudaStreamWaitEvent(crit, la->rest_done[slot], 0);     // bulk(np - 2): same slot, recorded earlier
    chol_panel_kernel<<<1 + T, 256, kPanelSmemBytes, crit>>>(A, lda, n, p0, p0 > 0 ? 1 : 0, Li, p.meta);
    ++launches;
    CHOL_DBG("panel");
    const int Tc = (n - p0 - nb + NB - 1) / NB;  // column tiles of the trailing matrix
    if (la) {
      cudaEventRecord(la->trsm_done[slot], crit);
      cudaStreamWaitEvent(st, la->trsm_done[slot], 0);
    }
    if (Tc > 1 && T > 1) {
      const int T128 = (rows - NB + SB - 1) / SB;          // 128-row tiles of the bulk region (rows >= p0 + nb + 64)
      const int nt128 = T128 * (T128 + 1) / 2;
      if (nb == NB && nt128 >= kSyrk128MinTiles)
      {
        if (g_chol_bulk_mode == 0 && have_map)
          chol_syrk128_umma_kernel<<<kNumSMs, 256, kSyrkUmmaSmemBytes, st>>>(tmA, A, lda, n, p0, nt128, p.meta);
        else
          chol_syrk128_kernel<<<kNumSMs, 256, kSyrk128SmemBytes, st>>>(A, lda, n, p0, nt128, p.meta);
      }
      else
        chol_syrk_kernel<<<(T - 1) * T / 2, 256, 0, st>>>(A, lda, n, p0, p.meta);
      ++launches;
      CHOL_DBG("syrk");
    }
    if (la) cudaEventRecord(la->rest_done[slot], st);
  }
  if (la) {                                       // join: the critical-path stream's work is visible to the caller's stream
    cudaEventRecord(la->fork, crit);
    cudaStreamWaitEvent(st, la->fork, 0);
  }
  chol_init_x_kernel<<<(n + 255) / 256, 256, 0, st>>>(A, lda, n, z, acc, p.meta);
  ++launches;
  CHOL_DBG("init_x");
  for (int ph = (n - 1) / NB; ph >= 0;) {
    const int pl = ph - BSG + 1 > 0 ? ph - BSG + 1 : 0;
    p0 = pl * NB;
    chol_bs_kernel<<<1 + pl, 256, kBsSmemBytes, st>>>(A, lda, n, pl, ph, Linv, z, acc, sol, p.meta);
    ++launches;
    CHOL_DBG("bs");
    ph = pl - 1;
  }
  const int m = n > p.N ? n : p.N;
  chol_finish_kernel<<<(m + 255) / 256, 256, 0, st>>>(p, sol);
  CHOL_DBG("finish");
  note_launches(launches + 1);
  return check_launch("ba_chol_solve");
}

}  // namespace dpvo

// ------------------------------------------------------------------ stand-alone entry (tests, tools)
namespace dpvo {
__global__ void __launch_bounds__(256)
chol_pack_kernel(const float* __restrict__ S, const float* __restrict__ y, int n, float* __restrict__ A, int lda,
                 int* __restrict__ meta) {
  const int64_t total = (int64_t)(n + 1) * n;
  if (blockIdx.x == 0 && threadIdx.x == 0) meta[1] = 0;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int a = (int)(idx / n), b = (int)(idx - (int64_t)a * n);
    if (a == n) A[(int64_t)n * lda + b] = y[b];
    else if (b <= a) A[(int64_t)a * lda + b] = S[(int64_t)a * n + b];
  }
}
}  // namespace dpvo

#ifdef DPVO_CHOL_PROFILE
extern "C" int dpvo_chol_profile_read(long long* out32, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out32, dpvo::g_chol_prof, sizeof(long long) * 32);
  if (reset) {
    long long z[32] = {};
    cudaMemcpyToSymbol(dpvo::g_chol_prof, z, sizeof(z));
  }
  return 0;
}
#endif

extern "C" int dpvo_chol_bulk_mode(int32_t mode) {
  DPVO_REQUIRE(mode == 0 || mode == 1, "chol_bulk_mode: 0 (split-TF32 tensor-core tiles) or 1 (fp32 FMA tiles)");
  dpvo::g_chol_bulk_mode = mode;
  return 0;
}

extern "C" int64_t dpvo_chol_workspace_bytes(int32_t n) {
  if (n < 0) n = 0;
  return 256 + ((int64_t)(n + 1) * dpvo::chol_lda(n) + dpvo::chol_scratch_floats(n) + 128) * 4;
}

extern "C" int dpvo_chol_solve(const float* S, const float* y, int32_t n, float* x, int32_t* info,
                               void* workspace, int64_t workspace_bytes, void* stream) {
  using namespace dpvo;
  DPVO_REQUIRE(n >= 0, "chol_solve: negative size");
  if (n == 0) return 0;
  DPVO_REQUIRE(S && y && x && info && workspace, "chol_solve: null pointer");
  DPVO_REQUIRE(workspace_bytes >= dpvo_chol_workspace_bytes(n), "chol_solve: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  BADev p;
  memset(&p, 0, sizeof(p));
  p.n = n; p.N = 0; p.dX = x;
  p.meta = reinterpret_cast<int*>(base);          // 64 ints
  float* A = reinterpret_cast<float*>(base + 256);
  const int lda = chol_lda(n);
  const int64_t tot = (int64_t)(n + 1) * n;
  const int grid = (int)(ceil_div64(tot, 256) < kNumSMs * 8 ? ceil_div64(tot, 256) : kNumSMs * 8);
  chol_pack_kernel<<<grid, 256, 0, st>>>(S, y, n, A, lda, p.meta);
  note_launches(1);
  if (debug_env("DPVO_DEBUG_SYNC")) {
    cudaError_t e_ = cudaStreamSynchronize(st);
    if (e_ != cudaSuccess) { set_error("chol_solve: pack: %s", cudaGetErrorString(e_)); return 1; }
  }
  if (ba_chol_solve(p, A, lda, A + (int64_t)(n + 1) * lda, st)) return 1;
  cudaMemcpyAsync(info, p.meta + 1, sizeof(int), cudaMemcpyDeviceToDevice, st);
  return check_launch("chol_solve");
}

// ------------------------------------------------------------------ allocation-free variant
// dpvo_chol_solve for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc): the workspace is a file-scope device array, so every GPU that loads the library holds
// kNoAllocBytes (about 2.6 MB) until the process exits.  No counterpart in the reference.
namespace dpvo {
constexpr int kNoAllocMaxN = DPVO_CHOL_NOALLOC_MAX_N;      // 768 = 128 free poses; below the look-ahead threshold
                                                           // (16 panels), so no internal stream / events either
constexpr int64_t kNoAllocFloats =
    64 /* meta */ + (int64_t)(kNoAllocMaxN + 1) * ((kNoAllocMaxN + 63) & ~63) /* A + y row */ +
    3 * (int64_t)kNoAllocMaxN + 64 + (int64_t)((kNoAllocMaxN + NB - 1) / NB) * NB * NB /* chol_scratch_floats */ + 128;
static_assert(kNoAllocMaxN < 16 * NB, "the static-scratch solve must stay below the look-ahead threshold (stream / event creation)");
static __device__ __align__(256) float g_chol_noalloc_scratch[kNoAllocFloats];

// In host code `g_chol_noalloc_scratch` is only a handle: the device address comes from cudaGetSymbolAddress,
// once per device with that device current.
static float* chol_noalloc_scratch() {
  static float* per_dev[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    set_error("chol_solve_noalloc: no current device");
    return nullptr;
  }
  float* ptr = __atomic_load_n(&per_dev[dev], __ATOMIC_ACQUIRE);
  if (!ptr) {
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_chol_noalloc_scratch);
    if (e != cudaSuccess || !ptr) {
      set_error("chol_solve_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return nullptr;
    }
    __atomic_store_n(&per_dev[dev], ptr, __ATOMIC_RELEASE);  // every thread resolves the same address: benign race
  }
  return ptr;
}
}  // namespace dpvo

extern "C" int32_t dpvo_chol_solve_noalloc_max_n(void) { return dpvo::kNoAllocMaxN; }

extern "C" int dpvo_chol_solve_noalloc(const float* S, const float* y, int32_t n, float* x, int32_t* info,
                                       void* stream) {
  using namespace dpvo;
  DPVO_REQUIRE(n >= 0, "chol_solve_noalloc: negative size");
  DPVO_REQUIRE(n <= kNoAllocMaxN,
               "chol_solve_noalloc: n = %d exceeds the static scratch (max %d); use dpvo_chol_solve with a workspace",
               n, kNoAllocMaxN);
  if (n == 0) return 0;
  DPVO_REQUIRE(S && y && x && info, "chol_solve_noalloc: null pointer");
  float* base = chol_noalloc_scratch();
  if (!base) return 1;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  BADev p;
  memset(&p, 0, sizeof(p));
  p.n = n; p.N = 0; p.dX = x;
  p.meta = reinterpret_cast<int*>(base);          // 64 ints; meta[1] (failure flag) is cleared by the pack kernel
  float* A = base + 64;
  const int lda = chol_lda(n);
  const int64_t tot = (int64_t)(n + 1) * n;
  const int grid = (int)(ceil_div64(tot, 256) < kNumSMs * 8 ? ceil_div64(tot, 256) : kNumSMs * 8);
  // everything the factorisation reads is rewritten per call: lower triangle + y row (pack), z / acc (init_x),
  // inv(L11) per panel (diag); the rest of the array keeps the previous call's bytes and is never read
  chol_pack_kernel<<<grid, 256, 0, st>>>(S, y, n, A, lda, p.meta);
  note_launches(1);
  if (ba_chol_solve(p, A, lda, A + (int64_t)(n + 1) * lda, st)) return 1;
  cudaMemcpyAsync(info, p.meta + 1, sizeof(int), cudaMemcpyDeviceToDevice, st);
  return check_launch("chol_solve_noalloc");
}
