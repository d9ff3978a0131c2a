// Fixed-size Lie-group algebra for the lietorch_backends replacement (SO3, RxSO3, SE3, Sim3; float / double;
// device AND host: the reference module dispatches on the tensor's device, lietorch/src/lietorch.cpp:18-316).
//
// Semantics follow /root/reference/methods/dpvo/dpvo/lietorch/include/so3.h:29-37 (quaternion is
// normalised on every load), :43-60 (product, action), :115-170 (Log / Exp with EPS = 1e-6 branches),
// :172-208 (left Jacobian and inverse) and include/se3.h:36-56 (inv, mul, act, act4), :58-86 (Adj,
// matrix), :113-142 (projector, Log, Exp), :144-201 (Q, left Jacobian and inverse), :203-217
// (action Jacobians); include/rxso3.h:41-96 (inv, mul, act, Adj, matrix, projector), :119-190 (Log / Exp), :192-286
// (W and its use), :288-320 (Jacobians); include/sim3.h:41-108 (inv, mul, act, matrix, projector, Adj), :114-170 (hat,
// adj, Log / Exp through W), :172-196 (series left Jacobian and inverse), :198-214 (action Jacobians).
// Written from the formulae, without Eigen: everything is plain registers.
#pragma once
#include "common.cuh"
#include <math.h>

#define LIE_HD __host__ __device__ __forceinline__

namespace dpvo {
namespace lie {

template <typename T> struct Eps { static LIE_HD T v() { return T(1e-6); } };
// Deliberate deviation: the reference switches the Jacobian coefficients ((1-cos t)/t^2, (t-sin t)/t^3, the
// Q coefficients of se3.h:155-165, ...) from their Taylor series to the closed forms at t = EPS = 1e-6.
// Just above that the closed forms cancel catastrophically (in float32 the reference's own values are
// noise for t in ~[1e-6, 1e-2]); here the series (three terms, truncation error < 1e-14) is used below
// t = 1e-2, so results follow the exact value instead of the rounding noise.
template <typename T> struct SeriesTh { static LIE_HD T v() { return T(1e-2); } };

template <typename T> struct V3 { T x, y, z; };
template <typename T> struct M3 { T a[3][3]; };

template <typename T> LIE_HD V3<T> operator+(V3<T> a, V3<T> b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <typename T> LIE_HD V3<T> operator-(V3<T> a, V3<T> b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <typename T> LIE_HD V3<T> operator-(V3<T> a) { return {-a.x, -a.y, -a.z}; }
template <typename T> LIE_HD V3<T> operator*(T s, V3<T> a) { return {s * a.x, s * a.y, s * a.z}; }
template <typename T> LIE_HD T dot(V3<T> a, V3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> LIE_HD V3<T> cross(V3<T> a, V3<T> b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

template <typename T> LIE_HD M3<T> hat(V3<T> v) {
  M3<T> m;
  m.a[0][0] = T(0); m.a[0][1] = -v.z; m.a[0][2] = v.y;
  m.a[1][0] = v.z;  m.a[1][1] = T(0); m.a[1][2] = -v.x;
  m.a[2][0] = -v.y; m.a[2][1] = v.x;  m.a[2][2] = T(0);
  return m;
}
template <typename T> LIE_HD M3<T> eye3() {
  M3<T> m;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) m.a[i][j] = (i == j) ? T(1) : T(0);
  return m;
}
template <typename T> LIE_HD M3<T> operator*(const M3<T>& A, const M3<T>& B) {
  M3<T> C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[i][j] = A.a[i][0] * B.a[0][j] + A.a[i][1] * B.a[1][j] + A.a[i][2] * B.a[2][j];
  return C;
}
template <typename T> LIE_HD M3<T> operator+(const M3<T>& A, const M3<T>& B) {
  M3<T> C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[i][j] = A.a[i][j] + B.a[i][j];
  return C;
}
template <typename T> LIE_HD M3<T> operator-(const M3<T>& A, const M3<T>& B) {
  M3<T> C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[i][j] = A.a[i][j] - B.a[i][j];
  return C;
}
template <typename T> LIE_HD M3<T> operator*(T s, const M3<T>& A) {
  M3<T> C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[i][j] = s * A.a[i][j];
  return C;
}
template <typename T> LIE_HD V3<T> operator*(const M3<T>& A, V3<T> v) {
  return {A.a[0][0] * v.x + A.a[0][1] * v.y + A.a[0][2] * v.z, A.a[1][0] * v.x + A.a[1][1] * v.y + A.a[1][2] * v.z,
          A.a[2][0] * v.x + A.a[2][1] * v.y + A.a[2][2] * v.z};
}
// row vector times matrix: (v^T A)^T
template <typename T> LIE_HD V3<T> rowmul(V3<T> v, const M3<T>& A) {
  return {v.x * A.a[0][0] + v.y * A.a[1][0] + v.z * A.a[2][0], v.x * A.a[0][1] + v.y * A.a[1][1] + v.z * A.a[2][1],
          v.x * A.a[0][2] + v.y * A.a[1][2] + v.z * A.a[2][2]};
}

// ----------------------------------------------------------------------------- SO3
template <typename T> struct SO3 {
  static constexpr int K = 3, N = 4;
  T x, y, z, w;  // unit quaternion

  static LIE_HD SO3 from_quat(T qx, T qy, T qz, T qw) {
    const T n = sqrt(qx * qx + qy * qy + qz * qz + qw * qw);
    return SO3{qx / n, qy / n, qz / n, qw / n};
  }
  static LIE_HD SO3 load(const T* d) { return from_quat(d[0], d[1], d[2], d[3]); }
  LIE_HD void store(T* d) const { d[0] = x; d[1] = y; d[2] = z; d[3] = w; }
  LIE_HD V3<T> vec() const { return {x, y, z}; }

  LIE_HD SO3 inv() const { return from_quat(-x, -y, -z, w); }
  LIE_HD SO3 mul(const SO3& o) const {
    return from_quat(w * o.x + x * o.w + y * o.z - z * o.y, w * o.y - x * o.z + y * o.w + z * o.x,
                     w * o.z + x * o.y - y * o.x + z * o.w, w * o.w - x * o.x - y * o.y - z * o.z);
  }
  LIE_HD V3<T> rot(V3<T> p) const {
    V3<T> uv = cross(vec(), p);
    uv = uv + uv;
    return p + w * uv + cross(vec(), uv);
  }
  LIE_HD void act(const T* p, T* q) const {
    const V3<T> r = rot(V3<T>{p[0], p[1], p[2]});
    q[0] = r.x; q[1] = r.y; q[2] = r.z;
  }
  LIE_HD void act4(const T* p, T* q) const {
    act(p, q);
    q[3] = p[3];
  }
  LIE_HD M3<T> R() const {
    M3<T> m;
    const T tx = T(2) * x, ty = T(2) * y, tz = T(2) * z;
    const T twx = tx * w, twy = ty * w, twz = tz * w;
    const T txx = tx * x, txy = ty * x, txz = tz * x;
    const T tyy = ty * y, tyz = tz * y, tzz = tz * z;
    m.a[0][0] = T(1) - (tyy + tzz); m.a[0][1] = txy - twz;          m.a[0][2] = txz + twy;
    m.a[1][0] = txy + twz;          m.a[1][1] = T(1) - (txx + tzz); m.a[1][2] = tyz - twx;
    m.a[2][0] = txz - twy;          m.a[2][1] = tyz + twx;          m.a[2][2] = T(1) - (txx + tyy);
    return m;
  }
  LIE_HD void Adj(T (&A)[K][K]) const {
    const M3<T> r = R();
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) A[i][j] = r.a[i][j];
  }
  static LIE_HD void ad(const T* a, T (&A)[K][K]) {
    const M3<T> h = hat(V3<T>{a[0], a[1], a[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) A[i][j] = h.a[i][j];
  }
  LIE_HD void matrix4(T (&M)[4][4]) const {
    const M3<T> r = R();
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) M[i][j] = (i < 3 && j < 3) ? r.a[i][j] : ((i == j) ? T(1) : T(0));
  }
  // 4x4 block of the embedding projector (so3.h:81-91)
  LIE_HD void projector(T (&P)[N][N]) const {
    const M3<T> h = hat(-vec());
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) P[i][j] = T(0);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) P[i][j] = T(0.5) * (((i == j) ? w : T(0)) + h.a[i][j]);
    P[3][0] = T(-0.5) * x; P[3][1] = T(-0.5) * y; P[3][2] = T(-0.5) * z;
  }

  LIE_HD V3<T> log_vec() const {
    const T n2 = x * x + y * y + z * z;
    T f;
    if (n2 < Eps<T>::v() * Eps<T>::v()) {
      f = T(2) / w - T(2.0 / 3.0) * n2 / (w * w * w);
    } else {
      const T n = sqrt(n2);
      if (fabs(w) < Eps<T>::v()) {
        f = (w > T(0) ? T(3.14159265358979323846) : -T(3.14159265358979323846)) / n;
      } else {
        f = T(2) * atan(n / w) / n;
      }
    }
    return {f * x, f * y, f * z};
  }
  LIE_HD void Log(T* a) const {
    const V3<T> v = log_vec();
    a[0] = v.x; a[1] = v.y; a[2] = v.z;
  }
  static LIE_HD SO3 exp_vec(V3<T> phi) {
    const T th2 = dot(phi, phi);
    const T th = sqrt(th2);
    T im, re;
    if (th < Eps<T>::v()) {
      const T th4 = th2 * th2;
      im = T(0.5) - T(1.0 / 48.0) * th2 + T(1.0 / 3840.0) * th4;
      re = T(1) - T(1.0 / 8.0) * th2 + T(1.0 / 384.0) * th4;
    } else {
      im = sin(T(0.5) * th) / th;
      re = cos(T(0.5) * th);
    }
    return from_quat(im * phi.x, im * phi.y, im * phi.z, re);
  }
  static LIE_HD SO3 Exp(const T* a) { return exp_vec(V3<T>{a[0], a[1], a[2]}); }

  static LIE_HD M3<T> left_jac3(V3<T> phi) {
    const M3<T> Phi = hat(phi);
    const T th2 = dot(phi, phi);
    const T th = sqrt(th2);
    const bool small = th < SeriesTh<T>::v();
    const T c1 = small ? T(0.5) - T(1.0 / 24.0) * th2 + T(1.0 / 720.0) * th2 * th2 : (T(1) - cos(th)) / th2;
    const T c2 = small ? T(1.0 / 6.0) - T(1.0 / 120.0) * th2 + T(1.0 / 5040.0) * th2 * th2 : (th - sin(th)) / (th2 * th);
    return eye3<T>() + c1 * Phi + c2 * (Phi * Phi);
  }
  static LIE_HD M3<T> left_jac_inv3(V3<T> phi) {
    const M3<T> Phi = hat(phi);
    const T th2 = dot(phi, phi);
    const T th = sqrt(th2);
    const T h = T(0.5) * th;
    const T c2 = (th < SeriesTh<T>::v()) ? T(1.0 / 12.0) + T(1.0 / 720.0) * th2 + T(1.0 / 30240.0) * th2 * th2
                                         : (T(1) - th * cos(h) / (T(2) * sin(h))) / (th * th);
    return eye3<T>() + T(-0.5) * Phi + c2 * (Phi * Phi);
  }
  static LIE_HD void left_jacobian(const T* a, T (&J)[K][K]) {
    const M3<T> m = left_jac3(V3<T>{a[0], a[1], a[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] = m.a[i][j];
  }
  static LIE_HD void left_jacobian_inverse(const T* a, T (&J)[K][K]) {
    const M3<T> m = left_jac_inv3(V3<T>{a[0], a[1], a[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] = m.a[i][j];
  }
  // d(X p)/d(xi) evaluated at the TRANSFORMED point q (so3.h:210-221)
  static LIE_HD void act_jacobian(const T* q, T (&J)[3][K]) {
    const M3<T> h = hat(V3<T>{-q[0], -q[1], -q[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] = h.a[i][j];
  }
  static LIE_HD void act4_jacobian(const T* q, T (&J)[4][K]) {
    const M3<T> h = hat(V3<T>{-q[0], -q[1], -q[2]});
#pragma unroll
    for (int j = 0; j < 3; ++j) {
#pragma unroll
      for (int i = 0; i < 3; ++i) J[i][j] = h.a[i][j];
      J[3][j] = T(0);
    }
  }
};

// ----------------------------------------------------------------------------- SE3
template <typename T> struct SE3 {
  static constexpr int K = 6, N = 7;
  V3<T> t;
  SO3<T> q;

  static LIE_HD SE3 load(const T* d) { return SE3{V3<T>{d[0], d[1], d[2]}, SO3<T>::load(d + 3)}; }
  LIE_HD void store(T* d) const { d[0] = t.x; d[1] = t.y; d[2] = t.z; q.store(d + 3); }

  LIE_HD SE3 inv() const {
    const SO3<T> qi = q.inv();
    return SE3{-(qi.rot(t)), qi};
  }
  LIE_HD SE3 mul(const SE3& o) const { return SE3{t + q.rot(o.t), q.mul(o.q)}; }
  LIE_HD void act(const T* p, T* r) const {
    const V3<T> v = q.rot(V3<T>{p[0], p[1], p[2]}) + t;
    r[0] = v.x; r[1] = v.y; r[2] = v.z;
  }
  LIE_HD void act4(const T* p, T* r) const {
    const V3<T> v = q.rot(V3<T>{p[0], p[1], p[2]}) + p[3] * t;
    r[0] = v.x; r[1] = v.y; r[2] = v.z; r[3] = p[3];
  }
  LIE_HD void Adj(T (&A)[K][K]) const {
    const M3<T> R = q.R();
    const M3<T> tR = hat(t) * R;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        A[i][j] = R.a[i][j];
        A[i][j + 3] = tR.a[i][j];
        A[i + 3][j] = T(0);
        A[i + 3][j + 3] = R.a[i][j];
      }
  }
  static LIE_HD void ad(const T* a, T (&A)[K][K]) {
    const M3<T> Tau = hat(V3<T>{a[0], a[1], a[2]});
    const M3<T> Phi = hat(V3<T>{a[3], a[4], a[5]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        A[i][j] = Phi.a[i][j];
        A[i][j + 3] = Tau.a[i][j];
        A[i + 3][j] = T(0);
        A[i + 3][j + 3] = Phi.a[i][j];
      }
  }
  LIE_HD void matrix4(T (&M)[4][4]) const {
    q.matrix4(M);
    M[0][3] = t.x; M[1][3] = t.y; M[2][3] = t.z;
  }
  LIE_HD void projector(T (&P)[N][N]) const {
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
      for (int j = 0; j < N; ++j) P[i][j] = (i == j && i < 3) ? T(1) : T(0);
    const M3<T> h = hat(-t);
    T S[4][4];
    q.projector(S);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) P[i][j + 3] = h.a[i][j];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) P[i + 3][j + 3] = S[i][j];
  }
  LIE_HD void Log(T* a) const {
    const V3<T> phi = q.log_vec();
    const V3<T> tau = SO3<T>::left_jac_inv3(phi) * t;
    a[0] = tau.x; a[1] = tau.y; a[2] = tau.z; a[3] = phi.x; a[4] = phi.y; a[5] = phi.z;
  }
  static LIE_HD SE3 Exp(const T* a) {
    const V3<T> tau{a[0], a[1], a[2]}, phi{a[3], a[4], a[5]};
    return SE3{SO3<T>::left_jac3(phi) * tau, SO3<T>::exp_vec(phi)};
  }
  static LIE_HD M3<T> calcQ(const T* a) {
    const V3<T> tau{a[0], a[1], a[2]}, phi{a[3], a[4], a[5]};
    const M3<T> Tau = hat(tau), Phi = hat(phi);
    const T th = sqrt(dot(phi, phi));
    const T th2 = th * th, th4 = th2 * th2;
    const bool small = th < SeriesTh<T>::v();
    const T c1 = small ? T(1.0 / 6.0) - T(1.0 / 120.0) * th2 + T(1.0 / 5040.0) * th4 : (th - sin(th)) / (th2 * th);
    const T c2 = small ? T(1.0 / 24.0) - T(1.0 / 720.0) * th2 + T(1.0 / 40320.0) * th4
                       : (th2 + T(2) * cos(th) - T(2)) / (T(2) * th4);
    const T c3 = small ? T(1.0 / 120.0) - T(1.0 / 2520.0) * th2 + T(1.0 / 120960.0) * th4
                       : (T(2) * th - T(3) * sin(th) + th * cos(th)) / (T(2) * th4 * th);
    const M3<T> PT = Phi * Tau, TP = Tau * Phi, PTP = PT * Phi, PP = Phi * Phi;
    return T(0.5) * Tau + c1 * (PT + TP + PTP) + c2 * (PP * Tau + TP * Phi - T(3) * PTP) + c3 * (PTP * Phi + Phi * PTP);
  }
  static LIE_HD void blocks(const M3<T>& D, const M3<T>& O, T (&J)[K][K]) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        J[i][j] = D.a[i][j];
        J[i][j + 3] = O.a[i][j];
        J[i + 3][j] = T(0);
        J[i + 3][j + 3] = D.a[i][j];
      }
  }
  static LIE_HD void left_jacobian(const T* a, T (&J)[K][K]) {
    blocks(SO3<T>::left_jac3(V3<T>{a[3], a[4], a[5]}), calcQ(a), J);
  }
  static LIE_HD void left_jacobian_inverse(const T* a, T (&J)[K][K]) {
    const M3<T> Ji = SO3<T>::left_jac_inv3(V3<T>{a[3], a[4], a[5]});
    const M3<T> Q = calcQ(a);
    blocks(Ji, T(-1) * (Ji * Q * Ji), J);
  }
  static LIE_HD void act_jacobian(const T* r, T (&J)[3][K]) {
    const M3<T> h = hat(V3<T>{-r[0], -r[1], -r[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        J[i][j] = (i == j) ? T(1) : T(0);
        J[i][j + 3] = h.a[i][j];
      }
  }
  static LIE_HD void act4_jacobian(const T* r, T (&J)[4][K]) {
    const M3<T> h = hat(V3<T>{-r[0], -r[1], -r[2]});
#pragma unroll
    for (int j = 0; j < 3; ++j) {
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        J[i][j] = (i == j) ? r[3] : T(0);
        J[i][j + 3] = h.a[i][j];
      }
      J[3][j] = T(0);
      J[3][j + 3] = T(0);
    }
  }
};

// generic small dense helpers (Sim3's 7x7 series Jacobians)
template <typename T, int K> LIE_HD void matmulKK(const T (&A)[K][K], const T (&B)[K][K], T (&C)[K][K]) {
#pragma unroll
  for (int i = 0; i < K; ++i)
#pragma unroll
    for (int j = 0; j < K; ++j) {
      T s = T(0);
#pragma unroll
      for (int k = 0; k < K; ++k) s += A[i][k] * B[k][j];
      C[i][j] = s;
    }
}
template <typename T> LIE_HD M3<T> inverse3(const M3<T>& m) {
  const T (&a)[3][3] = m.a;
  const T c00 = a[1][1] * a[2][2] - a[1][2] * a[2][1];
  const T c01 = a[1][2] * a[2][0] - a[1][0] * a[2][2];
  const T c02 = a[1][0] * a[2][1] - a[1][1] * a[2][0];
  const T det = a[0][0] * c00 + a[0][1] * c01 + a[0][2] * c02;
  const T id = T(1) / det;
  M3<T> r;
  r.a[0][0] = c00 * id; r.a[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) * id; r.a[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) * id;
  r.a[1][0] = c01 * id; r.a[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) * id; r.a[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) * id;
  r.a[2][0] = c02 * id; r.a[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) * id; r.a[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) * id;
  return r;
}

// --------------------------------------------------------------------------- RxSO3
// data: [qx qy qz qw s]; tangent: [phi(3) sigma]   (rxso3.h)
template <typename T> struct RxSO3 {
  static constexpr int K = 4, N = 5;
  SO3<T> q;
  T s;

  static LIE_HD RxSO3 load(const T* d) { return RxSO3{SO3<T>::load(d), d[4]}; }
  LIE_HD void store(T* d) const { q.store(d); d[4] = s; }
  LIE_HD RxSO3 inv() const { return RxSO3{q.inv(), T(1) / s}; }
  LIE_HD RxSO3 mul(const RxSO3& o) const { return RxSO3{q.mul(o.q), s * o.s}; }
  LIE_HD V3<T> rot(V3<T> p) const { return s * q.rot(p); }
  LIE_HD void act(const T* p, T* r) const {
    const V3<T> v = rot(V3<T>{p[0], p[1], p[2]});
    r[0] = v.x; r[1] = v.y; r[2] = v.z;
  }
  LIE_HD void act4(const T* p, T* r) const { act(p, r); r[3] = p[3]; }
  LIE_HD void Adj(T (&A)[K][K]) const {
    const M3<T> R = q.R();
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) A[i][j] = (i < 3 && j < 3) ? R.a[i][j] : ((i == j) ? T(1) : T(0));
  }
  static LIE_HD void ad(const T* a, T (&A)[K][K]) {
    const M3<T> h = hat(V3<T>{a[0], a[1], a[2]});
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) A[i][j] = (i < 3 && j < 3) ? h.a[i][j] : T(0);
  }
  LIE_HD void matrix4(T (&M)[4][4]) const {
    const M3<T> R = q.R();
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) M[i][j] = (i < 3 && j < 3) ? s * R.a[i][j] : ((i == j) ? T(1) : T(0));
  }
  LIE_HD void projector(T (&P)[N][N]) const {
    T S4[4][4];
    q.projector(S4);
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
      for (int j = 0; j < N; ++j) P[i][j] = (i < 4 && j < 3) ? S4[i][j] : T(0);
    P[4][3] = s;
  }
  LIE_HD void Log(T* a) const {
    const V3<T> v = q.log_vec();
    a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = log(s);
  }
  static LIE_HD RxSO3 Exp(const T* a) { return RxSO3{SO3<T>::exp_vec(V3<T>{a[0], a[1], a[2]}), exp(a[3])}; }
  // W of rxso3.h:192-236 (the "V" matrix of Sim3's exponential), reference branches at EPS
  static LIE_HD M3<T> calcW(const T* ps) {
    const V3<T> phi{ps[0], ps[1], ps[2]};
    const T sigma = ps[3];
    const T theta = sqrt(dot(phi, phi));
    const M3<T> Phi = hat(phi);
    const T scale = exp(sigma);
    const T one = T(1), half = T(0.5), eps = Eps<T>::v();
    T A, B, C;
    if (fabs(sigma) < eps) {
      C = one;
      if (fabs(theta) < eps) {
        A = half;
        B = T(1.0 / 6.0);
      } else {
        const T t2 = theta * theta;
        A = (one - cos(theta)) / t2;
        B = (theta - sin(theta)) / (t2 * theta);
      }
    } else {
      C = (scale - one) / sigma;
      if (fabs(theta) < eps) {
        const T s2 = sigma * sigma;
        A = ((sigma - one) * scale + one) / s2;
        B = (scale * half * s2 + scale - one - sigma * scale) / (s2 * sigma);
      } else {
        const T t2 = theta * theta;
        const T a = scale * sin(theta), b = scale * cos(theta), c = t2 + sigma * sigma;
        A = (a * sigma + (one - b) * theta) / (theta * c);
        B = (C - ((b - one) * sigma + a * theta) / c) * one / t2;
      }
    }
    return A * Phi + B * (Phi * Phi) + C * eye3<T>();
  }
  static LIE_HD void embed(const M3<T>& m, T (&J)[K][K]) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) J[i][j] = (i < 3 && j < 3) ? m.a[i][j] : ((i == j) ? T(1) : T(0));
  }
  static LIE_HD void left_jacobian(const T* a, T (&J)[K][K]) { embed(SO3<T>::left_jac3(V3<T>{a[0], a[1], a[2]}), J); }
  static LIE_HD void left_jacobian_inverse(const T* a, T (&J)[K][K]) {
    embed(SO3<T>::left_jac_inv3(V3<T>{a[0], a[1], a[2]}), J);
  }
  static LIE_HD void act_jacobian(const T* r, T (&J)[3][K]) {
    const M3<T> h = hat(V3<T>{-r[0], -r[1], -r[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] = h.a[i][j];
      J[i][3] = r[i];
    }
  }
  static LIE_HD void act4_jacobian(const T* r, T (&J)[4][K]) {
    const M3<T> h = hat(V3<T>{-r[0], -r[1], -r[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] = h.a[i][j];
      J[i][3] = r[i];
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) J[3][j] = T(0);
  }
};

// ---------------------------------------------------------------------------- Sim3
// data: [tx ty tz qx qy qz qw s]; tangent: [tau(3) phi(3) sigma]   (sim3.h)
template <typename T> struct Sim3 {
  static constexpr int K = 7, N = 8;
  V3<T> t;
  RxSO3<T> r;

  static LIE_HD Sim3 load(const T* d) { return Sim3{V3<T>{d[0], d[1], d[2]}, RxSO3<T>::load(d + 3)}; }
  LIE_HD void store(T* d) const { d[0] = t.x; d[1] = t.y; d[2] = t.z; r.store(d + 3); }
  LIE_HD Sim3 inv() const {
    const RxSO3<T> ri = r.inv();
    return Sim3{-(ri.rot(t)), ri};
  }
  LIE_HD Sim3 mul(const Sim3& o) const { return Sim3{t + r.rot(o.t), r.mul(o.r)}; }
  LIE_HD void act(const T* p, T* q) const {
    const V3<T> v = r.rot(V3<T>{p[0], p[1], p[2]}) + t;
    q[0] = v.x; q[1] = v.y; q[2] = v.z;
  }
  LIE_HD void act4(const T* p, T* q) const {
    const V3<T> v = r.rot(V3<T>{p[0], p[1], p[2]}) + p[3] * t;
    q[0] = v.x; q[1] = v.y; q[2] = v.z; q[3] = p[3];
  }
  LIE_HD void Adj(T (&A)[K][K]) const {
    const M3<T> R = r.q.R();
    const M3<T> tR = hat(t) * R;
    const T tv[3] = {t.x, t.y, t.z};
#pragma unroll
    for (int i = 0; i < K; ++i)
#pragma unroll
      for (int j = 0; j < K; ++j) A[i][j] = (i == j) ? T(1) : T(0);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        A[i][j] = r.s * R.a[i][j];
        A[i][j + 3] = tR.a[i][j];
        A[i + 3][j + 3] = R.a[i][j];
      }
      A[i][6] = -tv[i];
    }
  }
  static LIE_HD void ad(const T* a, T (&A)[K][K]) {
    const M3<T> Tau = hat(V3<T>{a[0], a[1], a[2]});
    const M3<T> Phi = hat(V3<T>{a[3], a[4], a[5]});
#pragma unroll
    for (int i = 0; i < K; ++i)
#pragma unroll
      for (int j = 0; j < K; ++j) A[i][j] = T(0);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        A[i][j] = Phi.a[i][j] + ((i == j) ? a[6] : T(0));
        A[i][j + 3] = Tau.a[i][j];
        A[i + 3][j + 3] = Phi.a[i][j];
      }
      A[i][6] = -a[i];
    }
  }
  LIE_HD void matrix4(T (&M)[4][4]) const {
    r.matrix4(M);
    M[0][3] = t.x; M[1][3] = t.y; M[2][3] = t.z;
  }
  LIE_HD void projector(T (&P)[N][N]) const {
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
      for (int j = 0; j < N; ++j) P[i][j] = (i == j && i < 3) ? T(1) : T(0);
    const M3<T> h = hat(-t);
    T S5[5][5];
    r.projector(S5);
    const T tv[3] = {t.x, t.y, t.z};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
      for (int j = 0; j < 3; ++j) P[i][j + 3] = h.a[i][j];
      P[i][6] = tv[i];
    }
#pragma unroll
    for (int i = 0; i < 5; ++i)
#pragma unroll
      for (int j = 0; j < 5; ++j) P[i + 3][j + 3] = S5[i][j];
  }
  LIE_HD void Log(T* a) const {
    T ps[4];
    r.Log(ps);
    const V3<T> tau = inverse3(RxSO3<T>::calcW(ps)) * t;
    a[0] = tau.x; a[1] = tau.y; a[2] = tau.z; a[3] = ps[0]; a[4] = ps[1]; a[5] = ps[2]; a[6] = ps[3];
  }
  static LIE_HD Sim3 Exp(const T* a) {
    const V3<T> tau{a[0], a[1], a[2]};
    return Sim3{RxSO3<T>::calcW(a + 3) * tau, RxSO3<T>::Exp(a + 3)};
  }
  // sim3.h:172-184: I + Xi/2 + Xi^2/6 + Xi^3/24 + Xi^4/120 (the statement that would add Xi^5/720 is dead code there)
  static LIE_HD void left_jacobian(const T* a, T (&J)[K][K]) {
    T X[K][K], X2[K][K], X3[K][K], X4[K][K];
    ad(a, X);
    matmulKK<T, K>(X, X, X2);
    matmulKK<T, K>(X, X2, X3);
    matmulKK<T, K>(X2, X2, X4);
#pragma unroll
    for (int i = 0; i < K; ++i)
#pragma unroll
      for (int j = 0; j < K; ++j)
        J[i][j] = ((i == j) ? T(1) : T(0)) + T(1.0 / 2.0) * X[i][j] + T(1.0 / 6.0) * X2[i][j] + T(1.0 / 24.0) * X3[i][j] +
                  T(1.0 / 120.0) * X4[i][j];
  }
  // sim3.h:186-196: I - Xi/2 + Xi^2/12 - Xi^4/720
  static LIE_HD void left_jacobian_inverse(const T* a, T (&J)[K][K]) {
    T X[K][K], X2[K][K], X4[K][K];
    ad(a, X);
    matmulKK<T, K>(X, X, X2);
    matmulKK<T, K>(X2, X2, X4);
#pragma unroll
    for (int i = 0; i < K; ++i)
#pragma unroll
      for (int j = 0; j < K; ++j)
        J[i][j] = ((i == j) ? T(1) : T(0)) - T(1.0 / 2.0) * X[i][j] + T(1.0 / 12.0) * X2[i][j] - T(1.0 / 720.0) * X4[i][j];
  }
  static LIE_HD void act_jacobian(const T* q, T (&J)[3][K]) {
    const M3<T> h = hat(V3<T>{-q[0], -q[1], -q[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        J[i][j] = (i == j) ? T(1) : T(0);
        J[i][j + 3] = h.a[i][j];
      }
      J[i][6] = q[i];
    }
  }
  static LIE_HD void act4_jacobian(const T* q, T (&J)[4][K]) {
    const M3<T> h = hat(V3<T>{-q[0], -q[1], -q[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        J[i][j] = (i == j) ? q[3] : T(0);
        J[i][j + 3] = h.a[i][j];
      }
      J[i][6] = q[i];
    }
#pragma unroll
    for (int j = 0; j < K; ++j) J[3][j] = T(0);
  }
};

}  // namespace lie
}  // namespace dpvo
