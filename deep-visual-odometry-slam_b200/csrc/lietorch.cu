// lietorch_backends replacement: element-wise Lie-group operators (SO3, RxSO3, SE3, Sim3) and their
// vector-Jacobian products, float32 / float64, on the device (kernels below) and on the host
// (dpvo_lie_host: the same templates in a plain loop — the reference module dispatches on the tensor's device,
// lietorch/include/dispatch.h:12-43, src/lietorch_cpu.cpp).
//
// Replaces /root/reference/methods/dpvo/dpvo/lietorch/src/lietorch_gpu.cu (kernels :20-290,
// launchers :296-601) behind the C ABI in include/dpvo_b200.h.  Gradient conventions are the
// reference's (src/lietorch_cpu.cpp:15-340): gradients are row vectors in the tangent space stored
// in the first K slots of an N-wide row, remaining slots zero.
//
// One thread owns one batch element.  Rows are 3..7 scalars wide, so a thread-per-row global access
// would touch 7 cache lines per warp instruction; instead every CTA stages its 128 rows of each
// operand through shared memory with fully coalesced global loads/stores (odd row widths are
// bank-conflict free, the 4/6-wide ones 2-4 way: irrelevant next to the HBM time).
#include "lie.cuh"

namespace dpvo {
namespace {

using namespace lie;

enum LieOp {
  OP_EXP, OP_EXP_BWD, OP_LOG, OP_LOG_BWD, OP_INV, OP_INV_BWD, OP_MUL, OP_MUL_BWD, OP_ADJ, OP_ADJ_BWD,
  OP_ADJT, OP_ADJT_BWD, OP_ACT, OP_ACT_BWD, OP_ACT4, OP_ACT4_BWD, OP_MATRIX, OP_PROJ, OP_JINV
};

struct Widths { int i0, i1, i2, o0, o1; };

template <class G, int OP> __host__ __device__ constexpr Widths widths() {
  constexpr int K = G::K, N = G::N;
  switch (OP) {
    case OP_EXP:      return {K, 0, 0, N, 0};      // a -> X
    case OP_EXP_BWD:  return {N, K, 0, K, 0};      // grad[N], a -> da[K]
    case OP_LOG:      return {N, 0, 0, K, 0};      // X -> a
    case OP_LOG_BWD:  return {K, N, 0, N, 0};      // grad[K], X -> dX[N]
    case OP_INV:      return {N, 0, 0, N, 0};
    case OP_INV_BWD:  return {N, N, 0, N, 0};      // grad[N], X -> dX
    case OP_MUL:      return {N, N, 0, N, 0};
    case OP_MUL_BWD:  return {N, N, N, N, N};      // grad, X, Y -> dX, dY
    case OP_ADJ:      return {N, K, 0, K, 0};      // X, a -> b
    case OP_ADJ_BWD:  return {K, N, K, N, K};      // grad[K], X, a -> dX, da
    case OP_ADJT:     return {N, K, 0, K, 0};
    case OP_ADJT_BWD: return {K, N, K, N, K};
    case OP_ACT:      return {N, 3, 0, 3, 0};      // X, p -> q
    case OP_ACT_BWD:  return {3, N, 3, N, 3};      // grad[3], X, p -> dX, dp
    case OP_ACT4:     return {N, 4, 0, 4, 0};
    case OP_ACT4_BWD: return {4, N, 4, N, 4};
    case OP_MATRIX:   return {N, 0, 0, 16, 0};
    case OP_PROJ:     return {N, 0, 0, N * N, 0};
    case OP_JINV:     return {N, K, 0, K, 0};      // X, a -> b
  }
  return {0, 0, 0, 0, 0};
}

// row vector (K) times KxK matrix
template <typename T, int K> LIE_HD void rowvec(const T* g, const T (&A)[K][K], T* out) {
#pragma unroll
  for (int j = 0; j < K; ++j) {
    T s = T(0);
#pragma unroll
    for (int i = 0; i < K; ++i) s += g[i] * A[i][j];
    out[j] = s;
  }
}
template <typename T, int K> LIE_HD void matvec(const T (&A)[K][K], const T* a, T* out) {
#pragma unroll
  for (int i = 0; i < K; ++i) {
    T s = T(0);
#pragma unroll
    for (int j = 0; j < K; ++j) s += A[i][j] * a[j];
    out[i] = s;
  }
}
template <typename T, int K> LIE_HD void matTvec(const T (&A)[K][K], const T* a, T* out) {
  rowvec<T, K>(a, A, out);
}
template <typename T, int K, int N> LIE_HD void pad_store(const T* v, T* row, T sign = T(1)) {
#pragma unroll
  for (int i = 0; i < N; ++i) row[i] = (i < K) ? sign * v[i] : T(0);
}

template <class G, typename T, int OP>
LIE_HD void apply(const T* r0, const T* r1, const T* r2, T* w0, T* w1) {
  constexpr int K = G::K, N = G::N;
  if constexpr (OP == OP_EXP) {
    G::Exp(r0).store(w0);
  } else if constexpr (OP == OP_EXP_BWD) {
    T J[K][K], d[K];
    G::left_jacobian(r1, J);
    rowvec<T, K>(r0, J, d);
#pragma unroll
    for (int i = 0; i < K; ++i) w0[i] = d[i];
  } else if constexpr (OP == OP_LOG) {
    T a[K];
    G::load(r0).Log(a);
#pragma unroll
    for (int i = 0; i < K; ++i) w0[i] = a[i];
  } else if constexpr (OP == OP_LOG_BWD) {
    T a[K], J[K][K], d[K];
    G::load(r1).Log(a);
    G::left_jacobian_inverse(a, J);
    rowvec<T, K>(r0, J, d);
    pad_store<T, K, N>(d, w0);
  } else if constexpr (OP == OP_INV) {
    G::load(r0).inv().store(w0);
  } else if constexpr (OP == OP_INV_BWD) {
    T A[K][K], d[K];
    G::load(r1).inv().Adj(A);
    rowvec<T, K>(r0, A, d);
    pad_store<T, K, N>(d, w0, T(-1));
  } else if constexpr (OP == OP_MUL) {
    G::load(r0).mul(G::load(r1)).store(w0);
  } else if constexpr (OP == OP_MUL_BWD) {
    T A[K][K], d[K];
    pad_store<T, K, N>(r0, w0);
    G::load(r1).Adj(A);
    rowvec<T, K>(r0, A, d);
    pad_store<T, K, N>(d, w1);
  } else if constexpr (OP == OP_ADJ) {
    T A[K][K], b[K];
    G::load(r0).Adj(A);
    matvec<T, K>(A, r1, b);
#pragma unroll
    for (int i = 0; i < K; ++i) w0[i] = b[i];
  } else if constexpr (OP == OP_ADJ_BWD) {
    T A[K][K], b[K], B[K][K], d[K];
    G::load(r1).Adj(A);
    matvec<T, K>(A, r2, b);
    rowvec<T, K>(r0, A, d);
#pragma unroll
    for (int i = 0; i < K; ++i) w1[i] = d[i];
    G::ad(b, B);
    rowvec<T, K>(r0, B, d);
    pad_store<T, K, N>(d, w0, T(-1));
  } else if constexpr (OP == OP_ADJT) {
    T A[K][K], b[K];
    G::load(r0).Adj(A);
    matTvec<T, K>(A, r1, b);
#pragma unroll
    for (int i = 0; i < K; ++i) w0[i] = b[i];
  } else if constexpr (OP == OP_ADJT_BWD) {
    T A[K][K], db[K], B[K][K], d[K];
    G::load(r1).Adj(A);
    matvec<T, K>(A, r0, db);
#pragma unroll
    for (int i = 0; i < K; ++i) w1[i] = db[i];
    G::ad(db, B);
    rowvec<T, K>(r2, B, d);
    pad_store<T, K, N>(d, w0, T(-1));
  } else if constexpr (OP == OP_ACT) {
    G::load(r0).act(r1, w0);
  } else if constexpr (OP == OP_ACT_BWD) {
    const G X = G::load(r1);
    T q[3], J[3][K], d[K], M[4][4];
    X.act(r2, q);
    G::act_jacobian(q, J);
#pragma unroll
    for (int j = 0; j < K; ++j) d[j] = r0[0] * J[0][j] + r0[1] * J[1][j] + r0[2] * J[2][j];
    pad_store<T, K, N>(d, w0);
    X.matrix4(M);
#pragma unroll
    for (int j = 0; j < 3; ++j) w1[j] = r0[0] * M[0][j] + r0[1] * M[1][j] + r0[2] * M[2][j];
  } else if constexpr (OP == OP_ACT4) {
    G::load(r0).act4(r1, w0);
  } else if constexpr (OP == OP_ACT4_BWD) {
    const G X = G::load(r1);
    T q[4], J[4][K], d[K], M[4][4];
    X.act4(r2, q);
    G::act4_jacobian(q, J);
#pragma unroll
    for (int j = 0; j < K; ++j) d[j] = r0[0] * J[0][j] + r0[1] * J[1][j] + r0[2] * J[2][j] + r0[3] * J[3][j];
    pad_store<T, K, N>(d, w0);
    X.matrix4(M);
#pragma unroll
    for (int j = 0; j < 4; ++j) w1[j] = r0[0] * M[0][j] + r0[1] * M[1][j] + r0[2] * M[2][j] + r0[3] * M[3][j];
  } else if constexpr (OP == OP_MATRIX) {
    T M[4][4];
    G::load(r0).matrix4(M);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) w0[i * 4 + j] = M[i][j];
  } else if constexpr (OP == OP_PROJ) {
    T P[N][N];
    G::load(r0).projector(P);
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
      for (int j = 0; j < N; ++j) w0[i * N + j] = P[i][j];
  } else if constexpr (OP == OP_JINV) {
    T a[K], J[K][K], b[K];
    G::load(r0).Log(a);
    G::left_jacobian_inverse(a, J);
    matvec<T, K>(J, r1, b);
#pragma unroll
    for (int i = 0; i < K; ++i) w0[i] = b[i];
  }
}

template <int W, int ROWS, typename T>
__device__ __forceinline__ void stage_in(T* s, const T* g, int64_t row0, int rows) {
  if constexpr (W > 0) {
    const T* src = g + row0 * W;
    for (int i = threadIdx.x; i < rows * W; i += ROWS) s[i] = src[i];
  }
}
template <int W, int ROWS, typename T>
__device__ __forceinline__ void stage_out(const T* s, T* g, int64_t row0, int rows) {
  if constexpr (W > 0) {
    T* dst = g + row0 * W;
    for (int i = threadIdx.x; i < rows * W; i += ROWS) dst[i] = s[i];
  }
}

template <class G, typename T, int OP, int ROWS>
__global__ void __launch_bounds__(ROWS) lie_kernel(const T* __restrict__ in0, const T* __restrict__ in1,
                                                   const T* __restrict__ in2, T* __restrict__ out0,
                                                   T* __restrict__ out1, int64_t n) {
  constexpr Widths w = widths<G, OP>();
  constexpr int TOTAL = w.i0 + w.i1 + w.i2 + w.o0 + w.o1;
  __shared__ T smem[ROWS * TOTAL];
  T* s_i0 = smem;
  T* s_i1 = s_i0 + ROWS * w.i0;
  T* s_i2 = s_i1 + ROWS * w.i1;
  T* s_o0 = s_i2 + ROWS * w.i2;
  T* s_o1 = s_o0 + ROWS * w.o0;
  const int64_t tiles = (n + ROWS - 1) / ROWS;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int64_t row0 = tile * ROWS;
    const int rows = (int)min((int64_t)ROWS, n - row0);
    stage_in<w.i0, ROWS>(s_i0, in0, row0, rows);
    stage_in<w.i1, ROWS>(s_i1, in1, row0, rows);
    stage_in<w.i2, ROWS>(s_i2, in2, row0, rows);
    __syncthreads();
    if ((int)threadIdx.x < rows) {
      const int r = threadIdx.x;
      T a0[w.i0 > 0 ? w.i0 : 1], a1[w.i1 > 0 ? w.i1 : 1], a2[w.i2 > 0 ? w.i2 : 1];
#pragma unroll
      for (int k = 0; k < w.i0; ++k) a0[k] = s_i0[r * w.i0 + k];
#pragma unroll
      for (int k = 0; k < w.i1; ++k) a1[k] = s_i1[r * w.i1 + k];
#pragma unroll
      for (int k = 0; k < w.i2; ++k) a2[k] = s_i2[r * w.i2 + k];
      apply<G, T, OP>(a0, a1, a2, s_o0 + r * w.o0, s_o1 + r * w.o1);
    }
    __syncthreads();
    stage_out<w.o0, ROWS>(s_o0, out0, row0, rows);
    stage_out<w.o1, ROWS>(s_o1, out1, row0, rows);
    __syncthreads();
  }
}

template <class G, typename T, int OP>
int launch_t(const void* i0, const void* i1, const void* i2, void* o0, void* o1, int64_t n, cudaStream_t st) {
  constexpr int ROWS = (OP == OP_PROJ) ? 32 : 128;
  const int64_t tiles = (n + ROWS - 1) / ROWS;
  const int64_t cap = (int64_t)kNumSMs * 16;
  const int grid = (int)(tiles < cap ? tiles : cap);
  lie_kernel<G, T, OP, ROWS><<<grid, ROWS, 0, st>>>(static_cast<const T*>(i0), static_cast<const T*>(i1),
                                                   static_cast<const T*>(i2), static_cast<T*>(o0),
                                                   static_cast<T*>(o1), n);
  note_launches(1);
  return check_launch("lietorch");
}

template <int OP>
int launch(const char* name, int group, dpvo_dtype dt, const void* i0, const void* i1, const void* i2, void* o0,
           void* o1, int64_t n, void* stream) {
  DPVO_REQUIRE(n >= 0, "%s: negative batch size", name);
  if (n == 0) return 0;
  DPVO_REQUIRE(i0 && o0, "%s: null pointer", name);
  DPVO_REQUIRE(dt == DPVO_F32 || dt == DPVO_F64, "%s: dtype must be float32 or float64", name);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (group == DPVO_LIE_SE3) {
    return dt == DPVO_F32 ? launch_t<SE3<float>, float, OP>(i0, i1, i2, o0, o1, n, st)
                          : launch_t<SE3<double>, double, OP>(i0, i1, i2, o0, o1, n, st);
  }
  if (group == DPVO_LIE_SO3) {
    return dt == DPVO_F32 ? launch_t<SO3<float>, float, OP>(i0, i1, i2, o0, o1, n, st)
                          : launch_t<SO3<double>, double, OP>(i0, i1, i2, o0, o1, n, st);
  }
  if (group == DPVO_LIE_RXSO3) {
    return dt == DPVO_F32 ? launch_t<RxSO3<float>, float, OP>(i0, i1, i2, o0, o1, n, st)
                          : launch_t<RxSO3<double>, double, OP>(i0, i1, i2, o0, o1, n, st);
  }
  if (group == DPVO_LIE_SIM3) {
    return dt == DPVO_F32 ? launch_t<Sim3<float>, float, OP>(i0, i1, i2, o0, o1, n, st)
                          : launch_t<Sim3<double>, double, OP>(i0, i1, i2, o0, o1, n, st);
  }
  set_error("%s: unknown group %d (SO3=1, RxSO3=2, SE3=3, Sim3=4)", name, group);
  return 1;
}

// ------------------------------------------------------------------------------------ host path
template <class G, typename T, int OP>
int host_t(const void* i0, const void* i1, const void* i2, void* o0, void* o1, int64_t n) {
  constexpr Widths w = widths<G, OP>();
  const T* a0 = static_cast<const T*>(i0);
  const T* a1 = static_cast<const T*>(i1);
  const T* a2 = static_cast<const T*>(i2);
  T* b0 = static_cast<T*>(o0);
  T* b1 = static_cast<T*>(o1);
  T z0[1] = {T(0)};
  for (int64_t r = 0; r < n; ++r)
    apply<G, T, OP>(w.i0 ? a0 + r * w.i0 : z0, w.i1 ? a1 + r * w.i1 : z0, w.i2 ? a2 + r * w.i2 : z0,
                    b0 + r * w.o0, w.o1 ? b1 + r * w.o1 : b0);
  return 0;
}

template <int OP>
int host_op(int group, dpvo_dtype dt, const void* i0, const void* i1, const void* i2, void* o0, void* o1, int64_t n) {
  constexpr bool two_in = (OP != OP_EXP && OP != OP_LOG && OP != OP_INV && OP != OP_MATRIX && OP != OP_PROJ);
  constexpr bool three_in = (OP == OP_MUL_BWD || OP == OP_ADJ_BWD || OP == OP_ADJT_BWD || OP == OP_ACT_BWD || OP == OP_ACT4_BWD);
  DPVO_REQUIRE(i0 && o0 && (!two_in || i1) && (!three_in || (i2 && o1)), "lie_host: null pointer");
  DPVO_REQUIRE(dt == DPVO_F32 || dt == DPVO_F64, "lie_host: dtype must be float32 or float64");
  const bool f = dt == DPVO_F32;
  switch (group) {
    case DPVO_LIE_SO3:   return f ? host_t<SO3<float>, float, OP>(i0, i1, i2, o0, o1, n) : host_t<SO3<double>, double, OP>(i0, i1, i2, o0, o1, n);
    case DPVO_LIE_RXSO3: return f ? host_t<RxSO3<float>, float, OP>(i0, i1, i2, o0, o1, n) : host_t<RxSO3<double>, double, OP>(i0, i1, i2, o0, o1, n);
    case DPVO_LIE_SE3:   return f ? host_t<SE3<float>, float, OP>(i0, i1, i2, o0, o1, n) : host_t<SE3<double>, double, OP>(i0, i1, i2, o0, o1, n);
    case DPVO_LIE_SIM3:  return f ? host_t<Sim3<float>, float, OP>(i0, i1, i2, o0, o1, n) : host_t<Sim3<double>, double, OP>(i0, i1, i2, o0, o1, n);
  }
  set_error("lie_host: unknown group %d (SO3=1, RxSO3=2, SE3=3, Sim3=4)", group);
  return 1;
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

#define LIE_UNARY(NAME, OP)                                                                                   \
  extern "C" int NAME(int32_t group, dpvo_dtype dt, const void* in, void* out, int64_t n, void* stream) {    \
    return launch<OP>(#NAME, group, dt, in, nullptr, nullptr, out, nullptr, n, stream);                      \
  }
#define LIE_UNARY_BWD(NAME, OP)                                                                               \
  extern "C" int NAME(int32_t group, dpvo_dtype dt, const void* grad, const void* in, void* din, int64_t n,  \
                      void* stream) {                                                                         \
    DPVO_REQUIRE(n == 0 || in, #NAME ": null pointer");                                                      \
    return launch<OP>(#NAME, group, dt, grad, in, nullptr, din, nullptr, n, stream);                         \
  }
#define LIE_BINARY(NAME, OP)                                                                                  \
  extern "C" int NAME(int32_t group, dpvo_dtype dt, const void* X, const void* b, void* out, int64_t n,      \
                      void* stream) {                                                                         \
    DPVO_REQUIRE(n == 0 || b, #NAME ": null pointer");                                                       \
    return launch<OP>(#NAME, group, dt, X, b, nullptr, out, nullptr, n, stream);                             \
  }
#define LIE_BINARY_BWD(NAME, OP)                                                                              \
  extern "C" int NAME(int32_t group, dpvo_dtype dt, const void* grad, const void* X, const void* b,          \
                      void* dX, void* db, int64_t n, void* stream) {                                          \
    DPVO_REQUIRE(n == 0 || (X && b && db), #NAME ": null pointer");                                          \
    return launch<OP>(#NAME, group, dt, grad, X, b, dX, db, n, stream);                                      \
  }

LIE_UNARY(dpvo_lie_expm, OP_EXP)
LIE_UNARY_BWD(dpvo_lie_expm_backward, OP_EXP_BWD)
LIE_UNARY(dpvo_lie_logm, OP_LOG)
LIE_UNARY_BWD(dpvo_lie_logm_backward, OP_LOG_BWD)
LIE_UNARY(dpvo_lie_inv, OP_INV)
LIE_UNARY_BWD(dpvo_lie_inv_backward, OP_INV_BWD)
LIE_BINARY(dpvo_lie_mul, OP_MUL)
LIE_BINARY_BWD(dpvo_lie_mul_backward, OP_MUL_BWD)
LIE_BINARY(dpvo_lie_adj, OP_ADJ)
LIE_BINARY_BWD(dpvo_lie_adj_backward, OP_ADJ_BWD)
LIE_BINARY(dpvo_lie_adjT, OP_ADJT)
LIE_BINARY_BWD(dpvo_lie_adjT_backward, OP_ADJT_BWD)
LIE_BINARY(dpvo_lie_act, OP_ACT)
LIE_BINARY_BWD(dpvo_lie_act_backward, OP_ACT_BWD)
LIE_BINARY(dpvo_lie_act4, OP_ACT4)
LIE_BINARY_BWD(dpvo_lie_act4_backward, OP_ACT4_BWD)
LIE_UNARY(dpvo_lie_as_matrix, OP_MATRIX)
LIE_UNARY(dpvo_lie_projector, OP_PROJ)
LIE_BINARY(dpvo_lie_jinv, OP_JINV)

// Host (CPU-tensor) form of every operator above: `op` is a dpvo_lie_op code, pointers are HOST memory, rows as in the
// device entry points (in0 / in1 / in2 and out0 / out1 in the order the device function of that op takes them).
extern "C" int dpvo_lie_host(int32_t op, int32_t group, dpvo_dtype dt, const void* in0, const void* in1, const void* in2,
                             void* out0, void* out1, int64_t n) {
  DPVO_REQUIRE(n >= 0, "lie_host: negative batch size");
  if (n == 0) return 0;
  switch (op) {
#define LIE_HOST_CASE(CODE, OP) case CODE: return host_op<OP>(group, dt, in0, in1, in2, out0, out1, n);
    LIE_HOST_CASE(DPVO_LIE_OP_EXP, OP_EXP) LIE_HOST_CASE(DPVO_LIE_OP_EXP_BWD, OP_EXP_BWD)
    LIE_HOST_CASE(DPVO_LIE_OP_LOG, OP_LOG) LIE_HOST_CASE(DPVO_LIE_OP_LOG_BWD, OP_LOG_BWD)
    LIE_HOST_CASE(DPVO_LIE_OP_INV, OP_INV) LIE_HOST_CASE(DPVO_LIE_OP_INV_BWD, OP_INV_BWD)
    LIE_HOST_CASE(DPVO_LIE_OP_MUL, OP_MUL) LIE_HOST_CASE(DPVO_LIE_OP_MUL_BWD, OP_MUL_BWD)
    LIE_HOST_CASE(DPVO_LIE_OP_ADJ, OP_ADJ) LIE_HOST_CASE(DPVO_LIE_OP_ADJ_BWD, OP_ADJ_BWD)
    LIE_HOST_CASE(DPVO_LIE_OP_ADJT, OP_ADJT) LIE_HOST_CASE(DPVO_LIE_OP_ADJT_BWD, OP_ADJT_BWD)
    LIE_HOST_CASE(DPVO_LIE_OP_ACT, OP_ACT) LIE_HOST_CASE(DPVO_LIE_OP_ACT_BWD, OP_ACT_BWD)
    LIE_HOST_CASE(DPVO_LIE_OP_ACT4, OP_ACT4) LIE_HOST_CASE(DPVO_LIE_OP_ACT4_BWD, OP_ACT4_BWD)
    LIE_HOST_CASE(DPVO_LIE_OP_MATRIX, OP_MATRIX) LIE_HOST_CASE(DPVO_LIE_OP_PROJECTOR, OP_PROJ)
    LIE_HOST_CASE(DPVO_LIE_OP_JINV, OP_JINV)
#undef LIE_HOST_CASE
  }
  set_error("lie_host: unknown op %d", op);
  return 1;
}
