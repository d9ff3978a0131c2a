// Fixed-size Lie-group algebra for the lietorch_backends replacement (SO3 and SE3, float / double).
//
// Semantics follow /root/reference/methods/dpvo/dpvo/lietorch/include/so3.h:29-37 (quaternion is
// normalised on every load), :43-60 (product, action), :115-170 (Log / Exp with EPS = 1e-6 branches),
// :172-208 (left Jacobian and inverse) and include/se3.h:36-56 (inv, mul, act, act4), :58-86 (Adj,
// matrix), :113-142 (projector, Log, Exp), :144-201 (Q, left Jacobian and inverse), :203-217
// (action Jacobians).  Written from the formulae, without Eigen: everything is plain registers.
#pragma once
#include "common.cuh"

namespace dpvo {
namespace lie {

template <typename T> struct Eps { static __device__ __forceinline__ T v() { return T(1e-6); } };
// Deliberate deviation: the reference switches the Jacobian coefficients ((1-cos t)/t^2, (t-sin t)/t^3, the
// Q coefficients of se3.h:155-165, ...) from their Taylor series to the closed forms at t = EPS = 1e-6.
// Just above that the closed forms cancel catastrophically (in float32 the reference's own values are
// noise for t in ~[1e-6, 1e-2]); here the series (three terms, truncation error < 1e-14) is used below
// t = 1e-2, so results follow the exact value instead of the rounding noise.
template <typename T> struct SeriesTh { static __device__ __forceinline__ T v() { return T(1e-2); } };

template <typename T> struct V3 { T x, y, z; };
template <typename T> struct M3 { T a[3][3]; };

template <typename T> __device__ __forceinline__ V3<T> operator+(V3<T> a, V3<T> b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <typename T> __device__ __forceinline__ V3<T> operator-(V3<T> a, V3<T> b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <typename T> __device__ __forceinline__ V3<T> operator-(V3<T> a) { return {-a.x, -a.y, -a.z}; }
template <typename T> __device__ __forceinline__ V3<T> operator*(T s, V3<T> a) { return {s * a.x, s * a.y, s * a.z}; }
template <typename T> __device__ __forceinline__ T dot(V3<T> a, V3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> __device__ __forceinline__ V3<T> cross(V3<T> a, V3<T> b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

template <typename T> __device__ __forceinline__ M3<T> hat(V3<T> v) {
  M3<T> m;
  m.a[0][0] = T(0); m.a[0][1] = -v.z; m.a[0][2] = v.y;
  m.a[1][0] = v.z;  m.a[1][1] = T(0); m.a[1][2] = -v.x;
  m.a[2][0] = -v.y; m.a[2][1] = v.x;  m.a[2][2] = T(0);
  return m;
}
template <typename T> __device__ __forceinline__ M3<T> eye3() {
  M3<T> m;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) m.a[i][j] = (i == j) ? T(1) : T(0);
  return m;
}
template <typename T> __device__ __forceinline__ M3<T> operator*(const M3<T>& A, const M3<T>& B) {
  M3<T> C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[i][j] = A.a[i][0] * B.a[0][j] + A.a[i][1] * B.a[1][j] + A.a[i][2] * B.a[2][j];
  return C;
}
template <typename T> __device__ __forceinline__ M3<T> operator+(const M3<T>& A, const M3<T>& B) {
  M3<T> C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[i][j] = A.a[i][j] + B.a[i][j];
  return C;
}
template <typename T> __device__ __forceinline__ M3<T> operator-(const M3<T>& A, const M3<T>& B) {
  M3<T> C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[i][j] = A.a[i][j] - B.a[i][j];
  return C;
}
template <typename T> __device__ __forceinline__ M3<T> operator*(T s, const M3<T>& A) {
  M3<T> C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[i][j] = s * A.a[i][j];
  return C;
}
template <typename T> __device__ __forceinline__ V3<T> operator*(const M3<T>& A, V3<T> v) {
  return {A.a[0][0] * v.x + A.a[0][1] * v.y + A.a[0][2] * v.z, A.a[1][0] * v.x + A.a[1][1] * v.y + A.a[1][2] * v.z,
          A.a[2][0] * v.x + A.a[2][1] * v.y + A.a[2][2] * v.z};
}
// row vector times matrix: (v^T A)^T
template <typename T> __device__ __forceinline__ V3<T> rowmul(V3<T> v, const M3<T>& A) {
  return {v.x * A.a[0][0] + v.y * A.a[1][0] + v.z * A.a[2][0], v.x * A.a[0][1] + v.y * A.a[1][1] + v.z * A.a[2][1],
          v.x * A.a[0][2] + v.y * A.a[1][2] + v.z * A.a[2][2]};
}

// ----------------------------------------------------------------------------- SO3
template <typename T> struct SO3 {
  static constexpr int K = 3, N = 4;
  T x, y, z, w;  // unit quaternion

  static __device__ __forceinline__ SO3 from_quat(T qx, T qy, T qz, T qw) {
    const T n = sqrt(qx * qx + qy * qy + qz * qz + qw * qw);
    return SO3{qx / n, qy / n, qz / n, qw / n};
  }
  static __device__ __forceinline__ SO3 load(const T* d) { return from_quat(d[0], d[1], d[2], d[3]); }
  __device__ __forceinline__ void store(T* d) const { d[0] = x; d[1] = y; d[2] = z; d[3] = w; }
  __device__ __forceinline__ V3<T> vec() const { return {x, y, z}; }

  __device__ __forceinline__ SO3 inv() const { return from_quat(-x, -y, -z, w); }
  __device__ __forceinline__ SO3 mul(const SO3& o) const {
    return from_quat(w * o.x + x * o.w + y * o.z - z * o.y, w * o.y - x * o.z + y * o.w + z * o.x,
                     w * o.z + x * o.y - y * o.x + z * o.w, w * o.w - x * o.x - y * o.y - z * o.z);
  }
  __device__ __forceinline__ V3<T> rot(V3<T> p) const {
    V3<T> uv = cross(vec(), p);
    uv = uv + uv;
    return p + w * uv + cross(vec(), uv);
  }
  __device__ __forceinline__ void act(const T* p, T* q) const {
    const V3<T> r = rot(V3<T>{p[0], p[1], p[2]});
    q[0] = r.x; q[1] = r.y; q[2] = r.z;
  }
  __device__ __forceinline__ void act4(const T* p, T* q) const {
    act(p, q);
    q[3] = p[3];
  }
  __device__ __forceinline__ M3<T> R() const {
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
  __device__ __forceinline__ void Adj(T (&A)[K][K]) const {
    const M3<T> r = R();
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) A[i][j] = r.a[i][j];
  }
  static __device__ __forceinline__ void ad(const T* a, T (&A)[K][K]) {
    const M3<T> h = hat(V3<T>{a[0], a[1], a[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) A[i][j] = h.a[i][j];
  }
  __device__ __forceinline__ void matrix4(T (&M)[4][4]) const {
    const M3<T> r = R();
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) M[i][j] = (i < 3 && j < 3) ? r.a[i][j] : ((i == j) ? T(1) : T(0));
  }
  // 4x4 block of the embedding projector (so3.h:81-91)
  __device__ __forceinline__ void projector(T (&P)[N][N]) const {
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

  __device__ __forceinline__ V3<T> log_vec() const {
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
  __device__ __forceinline__ void Log(T* a) const {
    const V3<T> v = log_vec();
    a[0] = v.x; a[1] = v.y; a[2] = v.z;
  }
  static __device__ __forceinline__ SO3 exp_vec(V3<T> phi) {
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
  static __device__ __forceinline__ SO3 Exp(const T* a) { return exp_vec(V3<T>{a[0], a[1], a[2]}); }

  static __device__ __forceinline__ M3<T> left_jac3(V3<T> phi) {
    const M3<T> Phi = hat(phi);
    const T th2 = dot(phi, phi);
    const T th = sqrt(th2);
    const bool small = th < SeriesTh<T>::v();
    const T c1 = small ? T(0.5) - T(1.0 / 24.0) * th2 + T(1.0 / 720.0) * th2 * th2 : (T(1) - cos(th)) / th2;
    const T c2 = small ? T(1.0 / 6.0) - T(1.0 / 120.0) * th2 + T(1.0 / 5040.0) * th2 * th2 : (th - sin(th)) / (th2 * th);
    return eye3<T>() + c1 * Phi + c2 * (Phi * Phi);
  }
  static __device__ __forceinline__ M3<T> left_jac_inv3(V3<T> phi) {
    const M3<T> Phi = hat(phi);
    const T th2 = dot(phi, phi);
    const T th = sqrt(th2);
    const T h = T(0.5) * th;
    const T c2 = (th < SeriesTh<T>::v()) ? T(1.0 / 12.0) + T(1.0 / 720.0) * th2 + T(1.0 / 30240.0) * th2 * th2
                                         : (T(1) - th * cos(h) / (T(2) * sin(h))) / (th * th);
    return eye3<T>() + T(-0.5) * Phi + c2 * (Phi * Phi);
  }
  static __device__ __forceinline__ void left_jacobian(const T* a, T (&J)[K][K]) {
    const M3<T> m = left_jac3(V3<T>{a[0], a[1], a[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] = m.a[i][j];
  }
  static __device__ __forceinline__ void left_jacobian_inverse(const T* a, T (&J)[K][K]) {
    const M3<T> m = left_jac_inv3(V3<T>{a[0], a[1], a[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] = m.a[i][j];
  }
  // d(X p)/d(xi) evaluated at the TRANSFORMED point q (so3.h:210-221)
  static __device__ __forceinline__ void act_jacobian(const T* q, T (&J)[3][K]) {
    const M3<T> h = hat(V3<T>{-q[0], -q[1], -q[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] = h.a[i][j];
  }
  static __device__ __forceinline__ void act4_jacobian(const T* q, T (&J)[4][K]) {
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

  static __device__ __forceinline__ SE3 load(const T* d) { return SE3{V3<T>{d[0], d[1], d[2]}, SO3<T>::load(d + 3)}; }
  __device__ __forceinline__ void store(T* d) const { d[0] = t.x; d[1] = t.y; d[2] = t.z; q.store(d + 3); }

  __device__ __forceinline__ SE3 inv() const {
    const SO3<T> qi = q.inv();
    return SE3{-(qi.rot(t)), qi};
  }
  __device__ __forceinline__ SE3 mul(const SE3& o) const { return SE3{t + q.rot(o.t), q.mul(o.q)}; }
  __device__ __forceinline__ void act(const T* p, T* r) const {
    const V3<T> v = q.rot(V3<T>{p[0], p[1], p[2]}) + t;
    r[0] = v.x; r[1] = v.y; r[2] = v.z;
  }
  __device__ __forceinline__ void act4(const T* p, T* r) const {
    const V3<T> v = q.rot(V3<T>{p[0], p[1], p[2]}) + p[3] * t;
    r[0] = v.x; r[1] = v.y; r[2] = v.z; r[3] = p[3];
  }
  __device__ __forceinline__ void Adj(T (&A)[K][K]) const {
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
  static __device__ __forceinline__ void ad(const T* a, T (&A)[K][K]) {
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
  __device__ __forceinline__ void matrix4(T (&M)[4][4]) const {
    q.matrix4(M);
    M[0][3] = t.x; M[1][3] = t.y; M[2][3] = t.z;
  }
  __device__ __forceinline__ void projector(T (&P)[N][N]) const {
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
  __device__ __forceinline__ void Log(T* a) const {
    const V3<T> phi = q.log_vec();
    const V3<T> tau = SO3<T>::left_jac_inv3(phi) * t;
    a[0] = tau.x; a[1] = tau.y; a[2] = tau.z; a[3] = phi.x; a[4] = phi.y; a[5] = phi.z;
  }
  static __device__ __forceinline__ SE3 Exp(const T* a) {
    const V3<T> tau{a[0], a[1], a[2]}, phi{a[3], a[4], a[5]};
    return SE3{SO3<T>::left_jac3(phi) * tau, SO3<T>::exp_vec(phi)};
  }
  static __device__ __forceinline__ M3<T> calcQ(const T* a) {
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
  static __device__ __forceinline__ void blocks(const M3<T>& D, const M3<T>& O, T (&J)[K][K]) {
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
  static __device__ __forceinline__ void left_jacobian(const T* a, T (&J)[K][K]) {
    blocks(SO3<T>::left_jac3(V3<T>{a[3], a[4], a[5]}), calcQ(a), J);
  }
  static __device__ __forceinline__ void left_jacobian_inverse(const T* a, T (&J)[K][K]) {
    const M3<T> Ji = SO3<T>::left_jac_inv3(V3<T>{a[3], a[4], a[5]});
    const M3<T> Q = calcQ(a);
    blocks(Ji, T(-1) * (Ji * Q * Ji), J);
  }
  static __device__ __forceinline__ void act_jacobian(const T* r, T (&J)[3][K]) {
    const M3<T> h = hat(V3<T>{-r[0], -r[1], -r[2]});
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        J[i][j] = (i == j) ? T(1) : T(0);
        J[i][j + 3] = h.a[i][j];
      }
  }
  static __device__ __forceinline__ void act4_jacobian(const T* r, T (&J)[4][K]) {
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

}  // namespace lie
}  // namespace dpvo
