// Blocked right-looking Cholesky solve for the reduced camera system when it does not fit one
// CTA's registers / shared memory (6N > 127: wide local windows, global BA with loop closures).
//
// Replaces at::linalg_cholesky_ex + torch::cholesky_solve of
// /root/reference/methods/dpvo/dpvo/fastba/ba_cuda.cu:547-548,561-562 (cuSOLVER potrf/potrs there).
//
// A is (n+1) x lda row-major in global memory: rows 0..n-1 hold the lower triangle of
// S = B - E Q E^T (+ damping), row n holds y^T, so the forward substitution L z = y is part of
// the factorisation (row n of L is z^T).  Panels of 64 columns, three launches per panel:
//   diag   one CTA factors the 64x64 diagonal block AND inverts the factor (blocked 4 x 16 inside the
//          tile, 16x16 blocks in one warp's registers with shuffles: ~22 barriers, not 64 + a serial solve)
//   trsm   one CTA per 64-row tile below the panel: L21 = A21 * inv(L11)^T as a 64x64x64 product
//          (inv(L11) was left behind by the CTA that factored the diagonal block), no sequential loop
//   syrk   one CTA per 64x64 lower tile of the trailing matrix: A22 -= L21 L21^T (fp32 SIMT; the
//          update is K=64 deep per panel — not worth a tensor-core pipeline at these sizes).
// Back substitution L^T x = z from the bottom, ONE launch per panel: every CTA forms
// x_p = inv(L11)^T (z_p - acc_p) (a 64x64 mat-vec), CTA q then adds L[p, q]^T x_p to acc_q for its
// 64-column tile q < p (each acc segment has exactly one writer per launch: no atomics).
// Finally dX -> global and the SE3 retraction (ba_cuda.cu:178-206).
// A non-positive pivot sets meta[1]; every later kernel returns early and dX is zeroed.
#include "se3.cuh"
#include "ba_internal.cuh"
#include <stdlib.h>

#ifndef DPVO_CHOL_NOALLOC_MAX_N
#define DPVO_CHOL_NOALLOC_MAX_N 768
#endif

namespace dpvo {

constexpr int NB = 64;

// Factor the nb x nb SPD block held in L (lower triangle, row-major, leading dimension NB+1) in place
// and invert the factor into X (lower).  Blocked 4 x 16 inside the tile so that the serial chain is
// short: each 16x16 diagonal block is factored AND inverted by one warp in registers (row l in lane l,
// pivots / multipliers broadcast with shuffles: no barriers), the panel below is a product with that
// inverse, the trailing update and the assembly of inv(L) from the block inverses use all 256 threads.
// ~22 barriers instead of 64 + a 2,000-step serial substitution.  Returns false on a non-positive pivot.
__device__ __forceinline__ bool chol_factor_invert(float (*L)[NB + 1], float (*X)[NB + 1], int nb) {
  constexpr int SB = 16;
  __shared__ float Tm[3][SB][SB + 1];
  __shared__ int s_fail;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) s_fail = 0;
  for (int i = tid; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    X[r][c] = 0.f;
    if (r >= nb || c > r) L[r][c] = (r == c) ? 1.f : 0.f;     // identity padding, clean upper triangle
  }
  __syncthreads();
#pragma unroll 1
  for (int b = 0; b < NB / SB; ++b) {
    const int o = b * SB;
    if (wid == 0) {
      // ---- 16x16 diagonal block: lane l < 16 owns row l
      const int l = lane & 15;
      float a[SB];
#pragma unroll
      for (int k = 0; k < SB; ++k) a[k] = L[o + l][o + k];      // upper triangle is zero (kept so below)
      bool bad = false;
#pragma unroll
      for (int j = 0; j < SB; ++j) {
        const float d = __shfl_sync(0xffffffffu, a[j], j);
        if (!(d > 0.f)) bad = true;
        float rs = rsqrtf(d);
        rs = rs * (1.5f - 0.5f * d * rs * rs);
        a[j] = (l == j) ? d * rs : a[j] * rs;                    // L[l][j] for l >= j
#pragma unroll
        for (int k = j + 1; k < SB; ++k) {
          const float lkj = __shfl_sync(0xffffffffu, a[j], k);   // L[k][j]
          a[k] -= a[j] * lkj;      // unconditional: entries above the diagonal (l < k) become don't-care values
                                  // (never read by another lane, masked at the store) and a[] stays in registers
        }
      }
      // ---- inverse: lane c < 16 owns column c of Xd; L[i][k] comes from lane i
      float x[SB];
#pragma unroll
      for (int i = 0; i < SB; ++i) {
        float sacc = (i == l) ? 1.f : 0.f;
#pragma unroll
        for (int k = 0; k < i; ++k) sacc -= __shfl_sync(0xffffffffu, a[k], i) * x[k];
        x[i] = sacc / __shfl_sync(0xffffffffu, a[i], i);
      }
      if (lane < SB) {
#pragma unroll
        for (int k = 0; k < SB; ++k) {                           // unconditional stores: selects, not branches,
          L[o + l][o + k] = (k <= l) ? a[k] : 0.f;               // so the arrays stay in registers
          X[o + k][o + l] = (k >= l) ? x[k] : 0.f;
        }
        if (bad) s_fail = 1;
      }
    }
    __syncthreads();
    if (s_fail) return false;
    const int below = NB - o - SB;           // rows under the diagonal block
    if (below > 0) {
      // ---- panel: L21 = A21 * Xd^T  (below x 16 outputs, <= 3 per thread)
      float outv[3];
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const int idx = tid + 256 * q;
        float sacc = 0.f;
        if (idx < below * SB) {
          const int r = o + SB + idx / SB, c = idx % SB;
          for (int k = 0; k <= c; ++k) sacc += L[r][o + k] * X[o + c][o + k];
        }
        outv[q] = sacc;
      }
      __syncthreads();
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const int idx = tid + 256 * q;
        if (idx < below * SB) L[o + SB + idx / SB][o + idx % SB] = outv[q];
      }
      __syncthreads();
      // ---- trailing update of the lower triangle: A22[r][c] -= L21[r][:] . L21[c][:]
#pragma unroll 1
      for (int idx = tid; idx < below * below; idx += 256) {
        const int r = idx / below, c = idx - r * below;
        if (c > r) continue;
        float sacc = 0.f;
#pragma unroll
        for (int k = 0; k < SB; ++k) sacc += L[o + SB + r][o + k] * L[o + SB + c][o + k];
        L[o + SB + r][o + SB + c] -= sacc;
      }
      __syncthreads();
    }
  }
  // ---- inv(L): X_ij = -X_ii * sum_{k=j}^{i-1} L_ik X_kj for block distance 1, 2, 3
#pragma unroll 1
  for (int dist = 1; dist < NB / SB; ++dist) {
    const int nblk = NB / SB - dist;
    const int r = tid / SB, c = tid % SB;
#pragma unroll 1
    for (int q = 0; q < nblk; ++q) {         // block (i = q + dist, j = q): one output per thread
      const int i = q + dist, j = q;
      float sacc = 0.f;
#pragma unroll 1
      for (int k = j; k < i; ++k)
#pragma unroll
        for (int t = 0; t < SB; ++t) sacc += L[i * SB + r][k * SB + t] * X[k * SB + t][j * SB + c];
      Tm[q][r][c] = sacc;
    }
    __syncthreads();
#pragma unroll 1
    for (int q = 0; q < nblk; ++q) {
      const int i = q + dist, j = q;
      float sacc = 0.f;
#pragma unroll 4
      for (int t = 0; t <= r; ++t) sacc += X[i * SB + r][i * SB + t] * Tm[q][t][c];
      X[i * SB + r][j * SB + c] = -sacc;
    }
    __syncthreads();
  }
  return true;
}

// write L11 (lower incl. diagonal) into A and inv(L11) into Linv (dense NB x NB, zero upper)
__device__ __forceinline__ void chol_store_diag(float* __restrict__ A, int lda, int p0, int nb, float (*L)[NB + 1],
                                                float (*X)[NB + 1], float* __restrict__ Linv) {
  for (int i = threadIdx.x; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    if (r < nb && c <= r) A[(int64_t)(p0 + r) * lda + p0 + c] = L[r][c];
    Linv[i] = X[r][c];
  }
}

// diagonal block of panel p0: factor + invert (one CTA)
__global__ void __launch_bounds__(256)
chol_diag_kernel(float* __restrict__ A, int lda, int n, int p0, float* __restrict__ Linv, int* __restrict__ meta) {
  __shared__ float L[NB][NB + 1];
  __shared__ float X[NB][NB + 1];
  if (meta[1]) return;
  const int nb = min(NB, n - p0);
  for (int i = threadIdx.x; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    L[r][c] = (r < nb && c <= r) ? A[(int64_t)(p0 + r) * lda + p0 + c] : 0.f;
  }
  __syncthreads();
  if (!chol_factor_invert(L, X, nb)) { if (threadIdx.x == 0) meta[1] = 1; return; }
  chol_store_diag(A, lda, p0, nb, L, X, Linv);
}

// rows [p0+nb, n] (incl. the y row) in tiles of 64: L21 = A21 * inv(L11)^T
__global__ void __launch_bounds__(256)
chol_trsm_kernel(float* __restrict__ A, int lda, int n, int p0, const float* __restrict__ Linv,
                 const int* __restrict__ meta) {
  __shared__ float Xi[NB][NB + 1];   // inv(L11), lower
  __shared__ float T[NB][NB + 1];    // A21 tile
  if (meta[1]) return;
  const int nb = min(NB, n - p0);
  const int r0 = p0 + nb + blockIdx.x * NB;
  const int nr = min(NB, n + 1 - r0);
  const int tid = threadIdx.x;
  for (int i = tid; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    Xi[r][c] = Linv[i];
    T[r][c] = (r < nr && c < nb) ? A[(int64_t)(r0 + r) * lda + p0 + c] : 0.f;
  }
  __syncthreads();
  // out[r][c] = sum_{k <= c} T[r][k] * Xi[c][k];  thread: 4 rows x 4 cols
  const int tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
  for (int k = 0; k < NB; ++k) {
    float av[4], bv[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { av[i] = T[ty + 16 * i][k]; bv[i] = Xi[tx + 16 * i][k]; }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] += av[i] * bv[j];
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = ty + 16 * i, c = tx + 16 * j;
      if (r < nr && c < nb) A[(int64_t)(r0 + r) * lda + p0 + c] = acc[i][j];
    }
}

// trailing update: tile (ti, tj), tj <= ti, of rows/cols starting at q0 = p0 + nb.
// mode 0: every tile; mode 1: the first tile column only (tj = 0: what the NEXT panel's factorisation needs);
// mode 2: every tile with tj >= 1 (runs on the look-ahead stream under the next panel's diag + trsm).
__global__ void __launch_bounds__(256)
chol_syrk_kernel(float* __restrict__ A, int lda, int n, int p0, const int* __restrict__ meta, int mode) {
  __shared__ float As[NB][NB + 1];
  __shared__ float Bs[NB][NB + 1];
  if (meta[1]) return;
  // linear tile index -> (ti, tj) with tj <= ti
  int ti, tj;
  if (mode == 1) {
    ti = blockIdx.x;
    tj = 0;
  } else {
    const int t = blockIdx.x;
    ti = (int)((sqrtf(8.f * t + 1.f) - 1.f) * 0.5f);
    while ((ti + 1) * (ti + 2) / 2 <= t) ++ti;
    while (ti * (ti + 1) / 2 > t) --ti;
    tj = t - ti * (ti + 1) / 2;
    if (mode == 2) { ++ti; ++tj; }
  }
  const int nb = min(NB, n - p0);
  const int q0 = p0 + nb;
  const int ri = q0 + ti * NB, rj = q0 + tj * NB;
  const int nri = min(NB, n + 1 - ri);        // rows may include the y row
  const int nrj = min(NB, n - rj);            // columns never do
  if (nrj <= 0) return;
  const int tid = threadIdx.x;
  for (int i = tid; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    As[r][c] = (r < nri && c < nb) ? A[(int64_t)(ri + r) * lda + p0 + c] : 0.f;
    Bs[r][c] = (r < nrj && c < nb) ? A[(int64_t)(rj + r) * lda + p0 + c] : 0.f;
  }
  __syncthreads();
  const int tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
  for (int k = 0; k < NB; ++k) {
    float av[4], bv[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { av[i] = As[ty + 16 * i][k]; bv[i] = Bs[tx + 16 * i][k]; }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] += av[i] * bv[j];
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = ty + 16 * i, c = tx + 16 * j;
      if (r < nri && c < nrj && (rj + c) <= (ri + r)) A[(int64_t)(ri + r) * lda + rj + c] -= acc[i][j];
    }
}

// Back substitution, panel p0: x_p = inv(L11)^T (z_p - acc_p); CTA 0 stores it, CTA q+1 folds it into
// acc for column tile q (< p):  acc_q += L[p rows, q cols]^T x_p.
__global__ void __launch_bounds__(256)
chol_bs_kernel(const float* __restrict__ A, int lda, int n, int p0, const float* __restrict__ Linv,
               const float* __restrict__ z, float* __restrict__ acc, float* __restrict__ sol,
               const int* __restrict__ meta) {
  __shared__ float rhs[NB];
  __shared__ float xs[NB];
  __shared__ float part[4][NB];
  if (meta[1]) return;
  const int nb = min(NB, n - p0);
  const int tid = threadIdx.x;
  if (tid < NB) rhs[tid] = (tid < nb) ? z[p0 + tid] - acc[p0 + tid] : 0.f;
  __syncthreads();
  if (tid < NB) {
    float s = 0.f;
    for (int k = tid; k < nb; ++k) s += Linv[k * NB + tid] * rhs[k];
    xs[tid] = s;
  }
  __syncthreads();
  if (blockIdx.x == 0) {
    if (tid < nb) sol[p0 + tid] = xs[tid];
    return;
  }
  const int c0 = (blockIdx.x - 1) * NB;
  const int c = tid & 63, rg = tid >> 6;
  float s = 0.f;
  for (int r = rg; r < nb; r += 4) s += A[(int64_t)(p0 + r) * lda + c0 + c] * xs[r];
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
  // look-ahead pays once the trailing update is big enough to hide a diag + trsm under it (>= 16 panels)
  CholLookahead* la = (!dbg && !no_la && n >= 16 * NB) ? chol_lookahead() : nullptr;
  bool rest_pending = false;
  int slot = 0;
  cudaStream_t crit = st;                          // diag / trsm / first trailing column
  if (la) {
    crit = la->side;
    cudaEventRecord(la->fork, st);
    cudaStreamWaitEvent(crit, la->fork, 0);
  }
  for (p0 = 0; p0 < n; p0 += NB) {
    const int nb = n - p0 < NB ? n - p0 : NB;
    const int rows = n + 1 - p0 - nb;            // rows below the panel, incl. the y row
    const int T = (rows + NB - 1) / NB;
    float* Li = Linv + (int64_t)(p0 / NB) * NB * NB;
    chol_diag_kernel<<<1, 256, 0, crit>>>(A, lda, n, p0, Li, p.meta);
    ++launches;
    CHOL_DBG("diag");
    chol_trsm_kernel<<<T, 256, 0, crit>>>(A, lda, n, p0, Li, p.meta);
    ++launches;
    CHOL_DBG("trsm");
    const int Tc = (n - p0 - nb + NB - 1) / NB;  // column tiles of the trailing matrix
    if (Tc > 0) {
      if (!la) {
        chol_syrk_kernel<<<T * (T + 1) / 2, 256, 0, st>>>(A, lda, n, p0, p.meta, 0);
        ++launches;
        CHOL_DBG("syrk");
      } else {
        cudaEventRecord(la->trsm_done[slot], crit);
        // the previous panel's bulk update also wrote this panel's first trailing column: order the two writers
        if (rest_pending) cudaStreamWaitEvent(crit, la->rest_done[slot ^ 1], 0);
        chol_syrk_kernel<<<T, 256, 0, crit>>>(A, lda, n, p0, p.meta, 1);
        ++launches;
        rest_pending = false;
        if (T > 1) {
          cudaStreamWaitEvent(st, la->trsm_done[slot], 0);
          chol_syrk_kernel<<<(T - 1) * T / 2, 256, 0, st>>>(A, lda, n, p0, p.meta, 2);
          ++launches;
          cudaEventRecord(la->rest_done[slot], st);
          rest_pending = true;
        }
        slot ^= 1;
      }
    }
  }
  if (la) {                                       // join: the critical-path stream's work is visible to the caller's stream
    cudaEventRecord(la->fork, crit);
    cudaStreamWaitEvent(st, la->fork, 0);
  }
  chol_init_x_kernel<<<(n + 255) / 256, 256, 0, st>>>(A, lda, n, z, acc, p.meta);
  ++launches;
  CHOL_DBG("init_x");
  const int last = ((n - 1) / NB) * NB;
  for (p0 = last; p0 >= 0; p0 -= NB) {
    chol_bs_kernel<<<1 + p0 / NB, 256, 0, st>>>(A, lda, n, p0, Linv + (int64_t)(p0 / NB) * NB * NB, z, acc, sol, p.meta);
    ++launches;
    CHOL_DBG("bs");
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
