// Device-side SE3 helpers for the BA / reprojection kernels (float32).
// Math follows /root/reference/methods/dpvo/dpvo/fastba/ba_cuda.cu:36-174
// (quaternion action, relative pose, adjoint pull-back, Exp and retraction),
// including its thresholds (theta_sq < 1e-8, theta > 1e-4) and the absence of
// quaternion renormalisation after retraction.
#pragma once
#include "common.cuh"

namespace dpvo {

struct Vec3 { float x, y, z; };
struct Quat { float x, y, z, w; };

__device__ __forceinline__ Vec3 rot(const Quat& q, const Vec3& X) {
  Vec3 uv;
  uv.x = 2.0f * (q.y * X.z - q.z * X.y);
  uv.y = 2.0f * (q.z * X.x - q.x * X.z);
  uv.z = 2.0f * (q.x * X.y - q.y * X.x);
  Vec3 Y;
  Y.x = X.x + q.w * uv.x + (q.y * uv.z - q.z * uv.y);
  Y.y = X.y + q.w * uv.y + (q.z * uv.x - q.x * uv.z);
  Y.z = X.z + q.w * uv.z + (q.x * uv.y - q.y * uv.x);
  return Y;
}

// Gij = Gj * Gi^-1
__device__ __forceinline__ void rel_se3(const Vec3& ti, const Quat& qi, const Vec3& tj, const Quat& qj,
                                        Vec3& tij, Quat& qij) {
  qij.x = -qj.w * qi.x + qj.x * qi.w - qj.y * qi.z + qj.z * qi.y;
  qij.y = -qj.w * qi.y + qj.y * qi.w - qj.z * qi.x + qj.x * qi.z;
  qij.z = -qj.w * qi.z + qj.z * qi.w - qj.x * qi.y + qj.y * qi.x;
  qij.w = qj.w * qi.w + qj.x * qi.x + qj.y * qi.y + qj.z * qi.z;
  const Vec3 r = rot(qij, ti);
  tij.x = tj.x - r.x;
  tij.y = tj.y - r.y;
  tij.z = tj.z - r.z;
}

// Y = pull-back of the 6-vector X through (t,q): Ji = adj(Gij, Jj)
__device__ __forceinline__ void adj_se3(const Vec3& t, const Quat& q, const float (&X)[6], float (&Y)[6]) {
  const Quat qi = {-q.x, -q.y, -q.z, q.w};
  const Vec3 a = {X[0], X[1], X[2]};
  const Vec3 b = {X[3], X[4], X[5]};
  const Vec3 ya = rot(qi, a);
  const Vec3 yb = rot(qi, b);
  Vec3 u;
  u.x = t.z * X[1] - t.y * X[2];
  u.y = t.x * X[2] - t.z * X[0];
  u.z = t.y * X[0] - t.x * X[1];
  const Vec3 v = rot(qi, u);
  Y[0] = ya.x; Y[1] = ya.y; Y[2] = ya.z;
  Y[3] = yb.x + v.x; Y[4] = yb.y + v.y; Y[5] = yb.z + v.z;
}

__device__ __forceinline__ Quat exp_so3(const Vec3& phi) {
  const float th2 = phi.x * phi.x + phi.y * phi.y + phi.z * phi.z;
  const float th4 = th2 * th2;
  const float th = sqrtf(th2);
  float imag, real;
  if (th2 < 1e-8f) {
    imag = 0.5f - (1.0f / 48.0f) * th2 + (1.0f / 3840.0f) * th4;
    real = 1.0f - (1.0f / 8.0f) * th2 + (1.0f / 384.0f) * th4;
  } else {
    imag = sinf(0.5f * th) / th;
    real = cosf(0.5f * th);
  }
  return Quat{imag * phi.x, imag * phi.y, imag * phi.z, real};
}

__device__ __forceinline__ Vec3 cross3(const Vec3& a, const Vec3& b) {
  return Vec3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

__device__ __forceinline__ void exp_se3(const float (&xi)[6], Vec3& t, Quat& q) {
  const Vec3 phi = {xi[3], xi[4], xi[5]};
  Vec3 tau = {xi[0], xi[1], xi[2]};
  q = exp_so3(phi);
  const float th2 = phi.x * phi.x + phi.y * phi.y + phi.z * phi.z;
  const float th = sqrtf(th2);
  t = tau;
  if (th > 1e-4f) {
    const float a = (1.0f - cosf(th)) / th2;
    tau = cross3(phi, tau);
    t.x += a * tau.x; t.y += a * tau.y; t.z += a * tau.z;
    const float b = (th - sinf(th)) / (th * th2);
    tau = cross3(phi, tau);
    t.x += b * tau.x; t.y += b * tau.y; t.z += b * tau.z;
  }
}

// T <- Exp(xi) * T
__device__ __forceinline__ void retr_se3(const float (&xi)[6], const Vec3& t, const Quat& q, Vec3& t1, Quat& q1) {
  Vec3 dt;
  Quat dq;
  exp_se3(xi, dt, dq);
  q1.x = dq.w * q.x + dq.x * q.w + dq.y * q.z - dq.z * q.y;
  q1.y = dq.w * q.y + dq.y * q.w + dq.z * q.x - dq.x * q.z;
  q1.z = dq.w * q.z + dq.z * q.w + dq.x * q.y - dq.y * q.x;
  q1.w = dq.w * q.w - dq.x * q.x - dq.y * q.y - dq.z * q.z;
  const Vec3 r = rot(dq, t);
  t1.x = r.x + dt.x; t1.y = r.y + dt.y; t1.z = r.z + dt.z;
}

}  // namespace dpvo
