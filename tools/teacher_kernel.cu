// HARNESS code, not product: the "teacher" of tools/refdpvo.py (the stand-in for a trained Update operator on the
// synthetic sequence: it pulls the predicted patch centres towards the ground-truth reprojection) as ONE launch instead
// of ~70 float64 torch kernels per update.  Both arms of every sequence measurement (reference DPVO class / this
// repo's class) call the same function, so the comparison is unchanged; only the harness' own host overhead shrinks,
// which matters for the arm that is bound by host time (see profiles/r02_seq_host_profile_*.txt).
// Arithmetic: float64 like the torch formulation it replaces (refdpvo.py Teacher.gt_centres / forward).
//   nvcc -shared -Xcompiler -fPIC -gencode arch=compute_100a,code=sm_100a -o tools/libteacher.so tools/teacher_kernel.cu
#include <cuda_runtime.h>
#include <cstdint>

namespace {

struct D3 { double x, y, z; };
__device__ __forceinline__ D3 cross(D3 a, D3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
// v + w t + u x t with t = 2 u x v  (unit quaternion x y z w)
__device__ __forceinline__ D3 qrot(const double* q, double sgn, D3 v) {
  const D3 u = {sgn * q[0], sgn * q[1], sgn * q[2]};
  const double w = q[3];
  D3 t = cross(u, v);
  t = {2.0 * t.x, 2.0 * t.y, 2.0 * t.z};
  const D3 c = cross(u, t);
  return {v.x + w * t.x + c.x, v.y + w * t.y + c.y, v.z + w * t.z + c.z};
}

__global__ void __launch_bounds__(256)
teacher_kernel(const double* __restrict__ gt, int64_t n_gt, const int64_t* __restrict__ ts, int64_t n_ts,
               const float* __restrict__ patches, int64_t n_patches, const float* __restrict__ intr,
               const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, const int64_t* __restrict__ kk, int64_t E,
               const float* __restrict__ coords, int64_t cs_e, int64_t cs_k, int64_t c_off,
               const float* __restrict__ delta, const float* __restrict__ weight, float alpha,
               float* __restrict__ d_out, float* __restrict__ w_out) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t fi = ii[e], fj = jj[e], kf = kk[e];
    bool ok = fi >= 0 && fi < n_ts && fj >= 0 && fj < n_ts && kf >= 0 && kf < n_patches;
    float du = 0.f, dv = 0.f;
    if (ok) {
      const int64_t ti = ts[fi], tj = ts[fj];
      ok = ti >= 0 && ti < n_gt && tj >= 0 && tj < n_gt;
      if (ok) {
        const double x = patches[kf * 27 + 4], y = patches[kf * 27 + 9 + 4];
        const float* Ki = intr + fi * 4;
        const float* Kj = intr + fj * 4;
        const double d = 0.4 + 0.1 * sin(0.05 * x + 0.3 * (double)ti) * cos(0.07 * y);
        const double* Gi = gt + ti * 7;
        const double* Gj = gt + tj * 7;
        D3 X = {(x - Ki[2]) / Ki[0] / d, (y - Ki[3]) / Ki[1] / d, 1.0 / d};
        X = {X.x - Gi[0], X.y - Gi[1], X.z - Gi[2]};
        const D3 Xw = qrot(Gi + 3, -1.0, X);
        D3 Xj = qrot(Gj + 3, 1.0, Xw);
        Xj = {Xj.x + Gj[0], Xj.y + Gj[1], Xj.z + Gj[2]};
        ok = Xj.z > 0.2;
        const double zc = ok ? Xj.z : 1.0;
        const float u = (float)(Kj[0] * Xj.x / zc + Kj[2]);
        const float v = (float)(Kj[1] * Xj.y / zc + Kj[3]);
        const float* c = coords + e * cs_e + c_off;
        du = (u - c[0]) + alpha * tanhf(delta[2 * e + 0]);
        dv = (v - c[cs_k]) + alpha * tanhf(delta[2 * e + 1]);
      }
    }
    d_out[2 * e + 0] = ok ? du : 0.f;
    d_out[2 * e + 1] = ok ? dv : 0.f;
    w_out[2 * e + 0] = ok ? weight[2 * e + 0] : 0.f;
    w_out[2 * e + 1] = ok ? weight[2 * e + 1] : 0.f;
  }
}

}  // namespace

// gt [n_gt,7] float64 (t, quaternion x y z w: world -> camera), ts [n_ts] frame -> time stamp, patches [n_patches,3,3,3],
// intr [n_ts,4], coords: element (e, k) of the predicted centres at coords[e * cs_e + k * cs_k + c_off],
// delta / weight / d_out / w_out [E,2] float32.  Returns the CUDA error code of the launch.
extern "C" int teacher_targets(const double* gt, int64_t n_gt, const int64_t* ts, int64_t n_ts, const float* patches,
                               int64_t n_patches, const float* intr, const int64_t* ii, const int64_t* jj,
                               const int64_t* kk, int64_t E, const float* coords, int64_t cs_e, int64_t cs_k,
                               int64_t c_off, const float* delta, const float* weight, float alpha, float* d_out,
                               float* w_out, void* stream) {
  if (E <= 0) return 0;
  const int grid = (int)((E + 255) / 256 < 148 * 8 ? (E + 255) / 256 : 148 * 8);
  teacher_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(gt, n_gt, ts, n_ts, patches, n_patches, intr, ii,
                                                                          jj, kk, E, coords, cs_e, cs_k, c_off, delta,
                                                                          weight, alpha, d_out, w_out);
  return (int)cudaGetLastError();
}
