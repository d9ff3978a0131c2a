// Blocked right-looking Cholesky solve for the reduced camera system when it does not fit one
// CTA's registers / shared memory (6N > 127: wide local windows, global BA with loop closures).
//
// Replaces at::linalg_cholesky_ex + torch::cholesky_solve of
// /root/reference/methods/dpvo/dpvo/fastba/ba_cuda.cu:547-548,561-562 (cuSOLVER potrf/potrs there).
//
// A is (n+1) x lda row-major in global memory: rows 0..n-1 hold the lower triangle of
// S = B - E Q E^T (+ damping), row n holds y^T, so the forward substitution L z = y is part of
// the factorisation (row n of L is z^T).  Panels of 64 columns:
//   diag   one CTA factors the 64x64 diagonal block in shared memory (one barrier per column)
//   trsm   one CTA per 64-row tile below: X L11^T = A21
//   syrk   one CTA per 64x64 lower tile of the trailing matrix: A22 -= L21 L21^T (fp32 SIMT; the
//          update is K=64 deep per panel — not worth a tensor-core pipeline at these sizes)
// then L^T x = z panel by panel from the bottom (gemv over the rows below with atomics + a
// 64x64 transposed triangular solve), dX -> global and the SE3 retraction (ba_cuda.cu:178-206).
// A non-positive pivot sets meta[1]; every later kernel returns early and dX is zeroed.
#include "se3.cuh"
#include "ba_internal.cuh"

namespace dpvo {

constexpr int NB = 64;

__global__ void __launch_bounds__(256)
chol_diag_kernel(float* __restrict__ A, int lda, int n, int p0, int* __restrict__ meta) {
  __shared__ float L[NB][NB + 1];
  __shared__ float dg[NB];
  if (meta[1]) return;
  const int nb = min(NB, n - p0);
  const int tid = threadIdx.x;
  for (int i = tid; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    L[r][c] = (r < nb && c <= r) ? A[(int64_t)(p0 + r) * lda + p0 + c] : 0.f;
  }
  __syncthreads();
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
  if (fail) { if (tid == 0) meta[1] = 1; return; }
  // write back L11: strictly-lower from the (scaled) upper copies, diagonal as is
  for (int i = tid; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    if (r < nb && c <= r) A[(int64_t)(p0 + r) * lda + p0 + c] = (c == r) ? dg[r] : L[c][r];
  }
}

// rows [p0+nb, n] (incl. the y row) in tiles of 64: X L11^T = A21, column by column.
__global__ void __launch_bounds__(256)
chol_trsm_kernel(float* __restrict__ A, int lda, int n, int p0, const int* __restrict__ meta) {
  __shared__ float L[NB][NB + 1];
  __shared__ float X[NB][NB + 1];
  if (meta[1]) return;
  const int nb = min(NB, n - p0);
  const int r0 = p0 + nb + blockIdx.x * NB;
  const int nr = min(NB, n + 1 - r0);
  const int tid = threadIdx.x;
  for (int i = tid; i < NB * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    L[r][c] = (r < nb && c <= r) ? A[(int64_t)(p0 + r) * lda + p0 + c] : 0.f;
    X[r][c] = (r < nr && c < nb) ? A[(int64_t)(r0 + r) * lda + p0 + c] : 0.f;
  }
  __syncthreads();
  const int r = tid & 63, cg = tid >> 6;   // thread: row r, columns c == cg (mod 4)
  for (int c = 0; c < nb; ++c) {
    if (cg == 0) X[r][c] = X[r][c] / L[c][c];
    __syncthreads();
    const float x = X[r][c];
    for (int k = c + 1 + ((cg - (c + 1)) & 3); k < nb; k += 4) X[r][k] -= x * L[k][c];
    __syncthreads();
  }
  for (int i = tid; i < NB * NB; i += 256) {
    const int rr = i / NB, c = i % NB;
    if (rr < nr && c < nb) A[(int64_t)(r0 + rr) * lda + p0 + c] = X[rr][c];
  }
}

// trailing update: tile (ti, tj), tj <= ti, of rows/cols starting at q0 = p0 + nb.
__global__ void __launch_bounds__(256)
chol_syrk_kernel(float* __restrict__ A, int lda, int n, int p0, const int* __restrict__ meta) {
  __shared__ float As[NB][NB + 1];
  __shared__ float Bs[NB][NB + 1];
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
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = ty + 16 * i, c = tx + 16 * j;
      if (r < nri && c < nrj && (rj + c) <= (ri + r)) A[(int64_t)(ri + r) * lda + rj + c] -= acc[i][j];
    }
}

// x (n floats, global) starts as z = row n of L.  For panel p0: acc[c] += sum_{i >= p0+nb} L[i][p0+c] x[i]
__global__ void __launch_bounds__(256)
chol_bs_gemv_kernel(const float* __restrict__ A, int lda, int n, int p0, const float* __restrict__ x,
                    float* __restrict__ acc, const int* __restrict__ meta) {
  __shared__ float part[4][NB];
  if (meta[1]) return;
  const int nb = min(NB, n - p0);
  const int c = threadIdx.x & 63, rg = threadIdx.x >> 6;
  const int r0 = p0 + nb + blockIdx.x * 256;
  float s = 0.f;
  if (c < nb)
    for (int i = r0 + rg; i < min(r0 + 256, n); i += 4) s += A[(int64_t)i * lda + p0 + c] * x[i];
  part[rg][c] = s;
  __syncthreads();
  if (threadIdx.x < nb) {
    const float v = part[0][c] + part[1][c] + part[2][c] + part[3][c];
    if (v != 0.f) atomicAdd(&acc[p0 + c], v);
  }
}

// x_p = L11^-T (z_p - acc_p), in place in x
__global__ void __launch_bounds__(64)
chol_bs_trsv_kernel(const float* __restrict__ A, int lda, int n, int p0, float* __restrict__ x,
                    const float* __restrict__ acc, const int* __restrict__ meta) {
  __shared__ float L[NB][NB + 1];
  __shared__ float xs[NB];
  if (meta[1]) return;
  const int nb = min(NB, n - p0);
  const int tid = threadIdx.x;
  for (int i = tid; i < NB * NB; i += 64) {
    const int r = i / NB, c = i % NB;
    L[r][c] = (r < nb && c <= r) ? A[(int64_t)(p0 + r) * lda + p0 + c] : 0.f;
  }
  if (tid < nb) xs[tid] = x[p0 + tid] - acc[p0 + tid];
  __syncthreads();
  for (int j = nb - 1; j >= 0; --j) {
    if (tid == 0) xs[j] = xs[j] / L[j][j];
    __syncthreads();
    if (tid < j) xs[tid] -= L[j][tid] * xs[j];
    __syncthreads();
  }
  if (tid < nb) x[p0 + tid] = xs[tid];
}

__global__ void __launch_bounds__(256)
chol_init_x_kernel(const float* __restrict__ A, int lda, int n, float* __restrict__ x, float* __restrict__ acc,
                   const int* __restrict__ meta) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { x[i] = meta[1] ? 0.f : A[(int64_t)n * lda + i]; acc[i] = 0.f; }
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

// A: (n+1) x lda, lower triangle + y row already in place.  xbuf: 2*n floats of scratch.
int ba_chol_solve(const BADev& p, float* A, int lda, float* xbuf, cudaStream_t st) {
  const int n = p.n;
  int launches = 0;
  for (int p0 = 0; p0 < n; p0 += NB) {
    const int nb = n - p0 < NB ? n - p0 : NB;
    chol_diag_kernel<<<1, 256, 0, st>>>(A, lda, n, p0, p.meta);
    const int rows = n + 1 - p0 - nb;            // rows below the panel, incl. the y row
    const int T = (rows + NB - 1) / NB;
    chol_trsm_kernel<<<T, 256, 0, st>>>(A, lda, n, p0, p.meta);
    launches += 2;
    const int Tc = (n - p0 - nb + NB - 1) / NB;  // column tiles of the trailing matrix
    if (Tc > 0) {
      chol_syrk_kernel<<<T * (T + 1) / 2, 256, 0, st>>>(A, lda, n, p0, p.meta);
      ++launches;
    }
  }
  float* x = xbuf;
  float* acc = xbuf + n;
  chol_init_x_kernel<<<(n + 255) / 256, 256, 0, st>>>(A, lda, n, x, acc, p.meta);
  ++launches;
  const int last = ((n - 1) / NB) * NB;
  for (int p0 = last; p0 >= 0; p0 -= NB) {
    const int nb = n - p0 < NB ? n - p0 : NB;
    const int below = n - p0 - nb;
    if (below > 0) {
      chol_bs_gemv_kernel<<<(below + 255) / 256, 256, 0, st>>>(A, lda, n, p0, x, acc, p.meta);
      ++launches;
    }
    chol_bs_trsv_kernel<<<1, 64, 0, st>>>(A, lda, n, p0, x, acc, p.meta);
    ++launches;
  }
  const int m = n > p.N ? n : p.N;
  chol_finish_kernel<<<(m + 255) / 256, 256, 0, st>>>(p, x);
  note_launches(launches + 1);
  return check_launch("ba_chol_solve");
}

}  // namespace dpvo
