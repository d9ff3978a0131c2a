// Blocked right-looking Cholesky solve for the reduced camera system when it does not fit one
// CTA's registers / shared memory (6N > 127: wide local windows, global BA with loop closures).
//
// Replaces at::linalg_cholesky_ex + torch::cholesky_solve of
// /root/reference/methods/dpvo/dpvo/fastba/ba_cuda.cu:547-548,561-562 (cuSOLVER potrf/potrs there).
//
// A is (n+1) x lda row-major in global memory: rows 0..n-1 hold the lower triangle of
// S = B - E Q E^T (+ damping), row n holds y^T, so the forward substitution L z = y is part of
// the factorisation (row n of L is z^T).  Panels of 64 columns, TWO launches per panel:
//   trsm   one CTA per 64-row tile below the panel: L21 = A21 * inv(L11)^T as a 64x64x64 product
//          (inv(L11) was left behind by the CTA that factored the diagonal block), no sequential loop
//   syrk   one CTA per 64x64 lower tile of the trailing matrix: A22 -= L21 L21^T (fp32 SIMT; the
//          update is K=64 deep per panel — not worth a tensor-core pipeline at these sizes).  The CTA
//          that owns tile (0,0) — the NEXT panel's diagonal block, touched by nobody else — goes on to
//          factor it in shared memory (one barrier per column) and to invert the factor, so the
//          serial 64-column factorisation hides behind the other tiles' updates.
// Back substitution L^T x = z from the bottom, ONE launch per panel: every CTA forms
// x_p = inv(L11)^T (z_p - acc_p) (a 64x64 mat-vec), CTA q then adds L[p, q]^T x_p to acc_q for its
// 64-column tile q < p (each acc segment has exactly one writer per launch: no atomics).
// Finally dX -> global and the SE3 retraction (ba_cuda.cu:178-206).
// A non-positive pivot sets meta[1]; every later kernel returns early and dX is zeroed.
#include "se3.cuh"
#include "ba_internal.cuh"

namespace dpvo {

constexpr int NB = 64;

// Factor the nb x nb SPD block held in L (lower triangle, row-major, leading dimension NB+1) and
// invert the factor: on return the caller writes L11 = {strictly lower: L[c][r] (scaled copies in the
// upper triangle), diagonal: dg[r]} and X = inv(L11) (lower).  256 threads.  Returns false on a
// non-positive pivot (uniform across the CTA).
__device__ __forceinline__ bool chol_factor_invert(float (*L)[NB + 1], float (*X)[NB + 1], float* dg, int nb) {
  const int tid = threadIdx.x;
  const int cl = tid & 63, rg = tid >> 6;
  bool fail = false;
  for (int j = 0; j < nb; ++j) {
    const float d = L[j][j];
    if (!(d > 0.f)) { fail = true; break; }   // uniform
    float rs = rsqrtf(d);
    rs = rs * (1.5f - 0.5f * d * rs * rs);
    const float invd = rs * rs;
    // trailing update reads the unscaled column j; scaled values go to the upper triangle
    for (int i = j + 1 + tid; i < nb; i += 256) L[j][i] = L[i][j] * rs;
    if (tid == 0) dg[j] = d * rs;             // sqrt(d); L[j][j] itself is still being read as d by slower warps
    for (int k = j + 1 + cl; k < nb; k += 64) {
      const float f = L[k][j] * invd;
      for (int i = k + ((rg - k) & 3); i < nb; i += 4) L[i][k] -= L[i][j] * f;
    }
    __syncthreads();
  }
  if (fail) return false;
  // X = inv(L11): thread c solves L11 X[:,c] = e_c by forward substitution (L[i][k], i>k, is at L[k][i])
  for (int i = tid; i < NB * NB; i += 256) X[i / NB][i % NB] = 0.f;
  __syncthreads();
  if (tid < nb) {
    const int c = tid;
    X[c][c] = 1.0f / dg[c];
    for (int i = c + 1; i < nb; ++i) {
      float s = 0.f;
      for (int k = c; k < i; ++k) s += L[k][i] * X[k][c];
      X[i][c] = -s / dg[i];
    }
  }
  __syncthreads();
  return true;
}

// write L11 (lower incl. diagonal) into A and inv(L11) into Linv (dense NB x NB, zero upper)
__device__ __forceinline__ void chol_store_diag(float* __restrict__ A, int lda, int p0, int nb, float (*L)[NB + 1],
                                                float (*X)[NB + 1], const float* dg, float* __restrict__ Linv) {
  for (int i = threadIdx.x; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    if (r < nb && c <= r) A[(int64_t)(p0 + r) * lda + p0 + c] = (c == r) ? dg[r] : L[c][r];
    Linv[i] = X[r][c];
  }
}

// panel 0 only: the first diagonal block
__global__ void __launch_bounds__(256)
chol_diag0_kernel(float* __restrict__ A, int lda, int n, float* __restrict__ Linv, int* __restrict__ meta) {
  __shared__ float L[NB][NB + 1];
  __shared__ float X[NB][NB + 1];
  __shared__ float dg[NB];
  if (meta[1]) return;
  const int nb = min(NB, n);
  for (int i = threadIdx.x; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    L[r][c] = (r < nb && c <= r) ? A[(int64_t)r * lda + c] : 0.f;
  }
  __syncthreads();
  if (!chol_factor_invert(L, X, dg, nb)) { if (threadIdx.x == 0) meta[1] = 1; return; }
  chol_store_diag(A, lda, 0, nb, L, X, dg, Linv);
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

// trailing update: tile (ti, tj), tj <= ti, of rows/cols starting at q0 = p0 + nb; the CTA of tile
// (0,0) then factors + inverts that tile (the next panel's diagonal block).
__global__ void __launch_bounds__(256)
chol_syrk_kernel(float* __restrict__ A, int lda, int n, int p0, float* __restrict__ Linv_next,
                 int* __restrict__ meta) {
  __shared__ float As[NB][NB + 1];
  __shared__ float Bs[NB][NB + 1];
  __shared__ float dg[NB];
  if (meta[1]) return;
  // linear tile index -> (ti, tj) with tj <= ti
  const int t = blockIdx.x;
  int ti = (int)((sqrtf(8.f * t + 1.f) - 1.f) * 0.5f);
  while ((ti + 1) * (ti + 2) / 2 <= t) ++ti;
  while (ti * (ti + 1) / 2 > t) --ti;
  const int tj = t - ti * (ti + 1) / 2;
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
  const bool next_diag = (t == 0);
  if (next_diag) __syncthreads();             // As / Bs are about to be reused
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = ty + 16 * i, c = tx + 16 * j;
      const bool in = r < nri && c < nrj && (rj + c) <= (ri + r);
      float v = 0.f;
      if (in) {
        v = A[(int64_t)(ri + r) * lda + rj + c] - acc[i][j];
        if (!next_diag || r >= nrj) A[(int64_t)(ri + r) * lda + rj + c] = v;   // the y row (r == nrj) if present
      }
      if (next_diag) As[r][c] = (in && r < nrj) ? v : 0.f;
    }
  if (!next_diag) return;
  __syncthreads();
  const int nbn = nrj;                        // size of the next diagonal block
  if (!chol_factor_invert(As, Bs, dg, nbn)) { if (tid == 0) meta[1] = 1; return; }
  chol_store_diag(A, lda, q0, nbn, As, Bs, dg, Linv_next);
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
int ba_chol_solve(const BADev& p, float* A, int lda, float* xbuf, cudaStream_t st) {
  const int n = p.n;
  int launches = 0;
  float* z = xbuf;
  float* acc = z + n;
  float* sol = acc + n;
  float* Linv = sol + n + (64 - (3 * n) % 64) % 64;     // 256-byte aligned panels
  chol_diag0_kernel<<<1, 256, 0, st>>>(A, lda, n, Linv, p.meta);
  ++launches;
  for (int p0 = 0; p0 < n; p0 += NB) {
    const int nb = n - p0 < NB ? n - p0 : NB;
    const int rows = n + 1 - p0 - nb;            // rows below the panel, incl. the y row
    const int T = (rows + NB - 1) / NB;
    const float* Li = Linv + (int64_t)(p0 / NB) * NB * NB;
    chol_trsm_kernel<<<T, 256, 0, st>>>(A, lda, n, p0, Li, p.meta);
    ++launches;
    const int Tc = (n - p0 - nb + NB - 1) / NB;  // column tiles of the trailing matrix
    if (Tc > 0) {
      chol_syrk_kernel<<<T * (T + 1) / 2, 256, 0, st>>>(A, lda, n, p0, Linv + (int64_t)(p0 / NB + 1) * NB * NB, p.meta);
      ++launches;
    }
  }
  chol_init_x_kernel<<<(n + 255) / 256, 256, 0, st>>>(A, lda, n, z, acc, p.meta);
  ++launches;
  const int last = ((n - 1) / NB) * NB;
  for (int p0 = last; p0 >= 0; p0 -= NB) {
    chol_bs_kernel<<<1 + p0 / NB, 256, 0, st>>>(A, lda, n, p0, Linv + (int64_t)(p0 / NB) * NB * NB, z, acc, sol, p.meta);
    ++launches;
  }
  const int m = n > p.N ? n : p.N;
  chol_finish_kernel<<<(m + 255) / 256, 256, 0, st>>>(p, sol);
  note_launches(launches + 1);
  return check_launch("ba_chol_solve");
}

}  // namespace dpvo
