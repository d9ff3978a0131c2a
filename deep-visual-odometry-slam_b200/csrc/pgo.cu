// cuda_ba.solve_system on the device (fastba/ba.cpp:99-180): one Gauss-Newton step of the classic
// loop-closure pose-graph optimisation over Sim(3) (7 parameters per pose).
//
//   J  : r x n block rows, two 7x7 blocks per edge x at block columns ii[x], jj[x]
//   A  = J^T J;  A.diag += A.diag * lm;  A.diag += ep;  b = -J^T res
//   delta[:freen] = A[:F,:F]^-1 b[:F]  (F = 7 freen; the remaining poses stay fixed: zeros)
//
// The reference builds J as an Eigen sparse matrix on the host and solves in double with
// SimplicialCholesky.  Here the F x F normal equations are accumulated densely in float64 on the
// device (49 x 4 atomics per edge), factorised in float32 by the blocked Cholesky the BA uses
// (ba_chol.cu) and brought back to float64 accuracy by iterative refinement: the residual b - A x
// is formed in float64 and solved for a correction with the same float32 solver.
#include "common.cuh"

namespace dpvo {
namespace {

constexpr int kSim = 7;

// one 64-thread group per edge: thread t < 49 owns entry (k, l) of each 7x7 product
__global__ void __launch_bounds__(256)
pgo_assemble_kernel(const float* __restrict__ Ji, const float* __restrict__ Jj, const int64_t* __restrict__ ii,
                    const int64_t* __restrict__ jj, const float* __restrict__ res, int64_t r, int nf, int F,
                    double* __restrict__ A, double* __restrict__ b, int* __restrict__ info) {
  const int t = threadIdx.x & 63;
  const int64_t x = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 6);
  if (x >= r || t >= kSim * kSim) return;
  const int64_t i = ii[x], j = jj[x];
  if (i == j || i < 0 || j < 0) {          // the reference exit(1)s on a self edge (ba.cpp:151-152)
    if (t == 0) atomicExch(info, 2);
    return;
  }
  const int k = t / kSim, l = t % kSim;
  const float* a = Ji + x * 49;
  const float* c = Jj + x * 49;
  double sii = 0, sij = 0, sji = 0, sjj = 0;
#pragma unroll
  for (int m = 0; m < kSim; ++m) {
    const double aik = a[m * 7 + k], ail = a[m * 7 + l], cjk = c[m * 7 + k], cjl = c[m * 7 + l];
    sii += aik * ail;
    sij += aik * cjl;
    sji += cjk * ail;
    sjj += cjk * cjl;
  }
  const bool fi = i < nf, fj = j < nf;
  if (fi) atomicAdd(A + (i * 7 + k) * F + i * 7 + l, sii);
  if (fi && fj) {
    atomicAdd(A + (i * 7 + k) * F + j * 7 + l, sij);
    atomicAdd(A + (j * 7 + k) * F + i * 7 + l, sji);
  }
  if (fj) atomicAdd(A + (j * 7 + k) * F + j * 7 + l, sjj);
  if (l == 0) {
    double bi = 0, bj = 0;
#pragma unroll
    for (int m = 0; m < kSim; ++m) {
      const double rm = res[x * 7 + m];
      bi += (double)a[m * 7 + k] * rm;
      bj += (double)c[m * 7 + k] * rm;
    }
    if (fi) atomicAdd(b + i * 7 + k, -bi);
    if (fj) atomicAdd(b + j * 7 + k, -bj);
  }
}

// damping on the diagonal, then the float32 copies the factorisation reads
__global__ void __launch_bounds__(256)
pgo_damp_cast_kernel(double* __restrict__ A, const double* __restrict__ b, int F, double lm, double ep,
                     float* __restrict__ S, float* __restrict__ y) {
  const int64_t total = (int64_t)F * F;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(idx / F), col = (int)(idx - (int64_t)row * F);
    double v = A[idx];
    if (row == col) {
      v = v + v * lm;
      v = v + ep;
      A[idx] = v;
      y[row] = (float)b[row];
    }
    S[idx] = (float)v;
  }
}

// rr = b - A x (float64, one warp per row), as the float32 right-hand side of the correction solve
__global__ void __launch_bounds__(256)
pgo_residual_kernel(const double* __restrict__ A, const double* __restrict__ b, const double* __restrict__ x, int F,
                    float* __restrict__ rr) {
  const int lane = threadIdx.x & 31;
  const int row = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (row >= F) return;
  const double* ar = A + (int64_t)row * F;
  double s = 0;
  for (int c = lane; c < F; c += 32) s += ar[c] * x[c];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) rr[row] = (float)(b[row] - s);
}

// x64 (+)= x32
__global__ void __launch_bounds__(256)
pgo_accumulate_kernel(double* __restrict__ x64, const float* __restrict__ x32, int F, int first) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < F) x64[i] = first ? (double)x32[i] : x64[i] + (double)x32[i];
}

// a failed factorisation (info 1) or a self edge (info 2) leaves delta at zero
__global__ void __launch_bounds__(256)
pgo_write_kernel(const double* __restrict__ x64, int F, int total, const int* __restrict__ chol_flag,
                 int* __restrict__ info, float* __restrict__ delta) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool bad = info[0] != 0 || chol_flag[0] != 0;
  if (i < total) delta[i] = (i < F && !bad) ? (float)x64[i] : 0.f;
  if (i == 0 && info[0] == 0 && chol_flag[0] != 0) info[0] = 1;
}

int64_t a256(int64_t x) { return (x + 255) & ~(int64_t)255; }

}  // namespace
}  // namespace dpvo

using namespace dpvo;

extern "C" int64_t dpvo_pgo_workspace_bytes(int32_t n_poses, int32_t freen) {
  if (n_poses < 0) return -1;
  const int nf = freen < 0 ? n_poses : (freen < n_poses ? freen : n_poses);
  const int64_t F = 7 * (int64_t)nf;
  // flags | A64 | b64 | x64 | S32 | y32 | x32 | chol workspace
  return 256 + a256(F * F * 8) + 2 * a256(F * 8) + a256(F * F * 4) + 2 * a256(F * 4) + dpvo_chol_workspace_bytes((int32_t)F) + 512;
}

extern "C" int dpvo_pgo_solve(const float* J_Ginv_i, const float* J_Ginv_j, const int64_t* ii, const int64_t* jj,
                              const float* res, int64_t n_edges, int32_t n_poses, float ep, float lm, int32_t freen,
                              int32_t refinements, float* delta, int32_t* info, void* workspace,
                              int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && n_poses >= 0 && refinements >= 0 && refinements <= 4, "pgo_solve: bad sizes");
  DPVO_REQUIRE(delta && info, "pgo_solve: null output");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  cudaMemsetAsync(info, 0, sizeof(int), st);
  cudaMemsetAsync(delta, 0, sizeof(float) * 7 * (size_t)n_poses, st);
  const int nf = freen < 0 ? n_poses : (freen < n_poses ? freen : n_poses);
  const int F = 7 * nf;
  if (F == 0 || n_edges == 0) return check_launch("pgo_solve");
  DPVO_REQUIRE(7 * (int64_t)nf < (int64_t(1) << 24), "pgo_solve: too many free poses");
  DPVO_REQUIRE(J_Ginv_i && J_Ginv_j && ii && jj && res && workspace, "pgo_solve: null pointer");
  DPVO_REQUIRE(workspace_bytes >= dpvo_pgo_workspace_bytes(n_poses, freen), "pgo_solve: workspace too small");
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  int* chol_flag = reinterpret_cast<int*>(base); base += 256;
  double* A = reinterpret_cast<double*>(base); base += a256((int64_t)F * F * 8);
  double* b = reinterpret_cast<double*>(base); base += a256((int64_t)F * 8);
  double* x64 = reinterpret_cast<double*>(base); base += a256((int64_t)F * 8);
  float* S = reinterpret_cast<float*>(base); base += a256((int64_t)F * F * 4);
  float* y = reinterpret_cast<float*>(base); base += a256((int64_t)F * 4);
  float* x32 = reinterpret_cast<float*>(base); base += a256((int64_t)F * 4);
  void* chol_ws = base;
  const int64_t chol_bytes = dpvo_chol_workspace_bytes(F);

  cudaMemsetAsync(A, 0, sizeof(double) * (size_t)F * F, st);
  cudaMemsetAsync(b, 0, sizeof(double) * F, st);
  pgo_assemble_kernel<<<(unsigned)ceil_div64(n_edges, 4), 256, 0, st>>>(J_Ginv_i, J_Ginv_j, ii, jj, res, n_edges, nf, F,
                                                                        A, b, info);
  const int64_t tot = (int64_t)F * F;
  const int grid = (int)(ceil_div64(tot, 256) < kNumSMs * 8 ? ceil_div64(tot, 256) : kNumSMs * 8);
  pgo_damp_cast_kernel<<<grid, 256, 0, st>>>(A, b, F, (double)lm, (double)ep, S, y);
  note_launches(2);
  for (int it = 0; it <= refinements; ++it) {
    if (dpvo_chol_solve(S, y, F, x32, chol_flag, chol_ws, chol_bytes, stream)) return 1;
    pgo_accumulate_kernel<<<(F + 255) / 256, 256, 0, st>>>(x64, x32, F, it == 0);
    note_launches(1);
    if (it < refinements) {
      pgo_residual_kernel<<<(unsigned)ceil_div64((int64_t)F * 32, 256), 256, 0, st>>>(A, b, x64, F, y);
      note_launches(1);
    }
  }
  pgo_write_kernel<<<(7 * n_poses + 255) / 256, 256, 0, st>>>(x64, F, 7 * n_poses, chol_flag, info, delta);
  note_launches(1);
  return check_launch("pgo_solve");
}
