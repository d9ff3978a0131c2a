// Fused projective transform of patches between frames (inference path of projective_ops.py).
//
// Replaces the operator chain of /root/reference/methods/dpvo/dpvo/projective_ops.py:53-130
// (`transform`, `flow_mag`, `point_cloud`): index_select of patches / intrinsics / poses, `iproj`,
// lietorch Inv + Mul + Act4 (three launches, with Gij materialised 9x by broadcasting.py:26-29),
// `proj`, and for jacobian=True the Ja / Jp / adjT / matrix() products — about twenty small torch
// kernels per call — by ONE launch, one thread per edge.
//
// Semantics kept from the reference: per-frame intrinsics (frame ii for the back-projection, frame jj
// for the projection), quaternions normalised at every group-element load (so3.h:35-37), projection
// depth clamp d = 1 / max(Z, 0.1) (projective_ops.py:43), validity Z > 0.2, Jacobians at the centre
// pixel with d = 1/Z only where |Z| > 0.2 (:76-77), Ji = -Adj(Gij)^T Jj (:103), Jz = Jp Gij[:, 3].
// Output layout [E, P, P, 2|3] like the reference (DPVO.reproject permutes it afterwards).
#include <cooperative_groups.h>

#include "lie.cuh"

namespace dpvo {
namespace {

using lie::SE3;
using lie::SO3;
using lie::V3;

struct TransformArgs {
  const float* poses; const float* patches; const float* intr;
  const int64_t* ii; const int64_t* jj; const int64_t* kk;
  int64_t E, nframes, npatches;
  int P, depth, tonly;
  float* coords;   // [E,P,P,2+depth]
  float* valid;    // [E,P,P] or null
  float* Ji; float* Jj; float* Jz; float* jvalid;   // [E,2,6], [E,2,6], [E,2], [E] or null
};

// Gij = G_j * G_i^-1 with lietorch's load / normalise points (Inv, then Mul, then the consumer's load)
__device__ __forceinline__ SE3<float> relative_pose(const float* pi, const float* pj, bool tonly) {
  SE3<float> G = SE3<float>::load(pj).mul(SE3<float>::load(pi).inv());
  G.q = SO3<float>::from_quat(G.q.x, G.q.y, G.q.z, G.q.w);
  if (tonly) G.q = SO3<float>{0.f, 0.f, 0.f, 1.f};
  return G;
}

__global__ void __launch_bounds__(128)
transform_kernel(const TransformArgs a) {
  const int PP = a.P * a.P;
  const int nc = 2 + a.depth;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < a.E; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t fi = a.ii[e], fj = a.jj[e], kf = a.kk[e];
    float* out = a.coords + e * PP * nc;
    if (fi < 0 || fi >= a.nframes || fj < 0 || fj >= a.nframes || kf < 0 || kf >= a.npatches) {
      for (int i = 0; i < PP * nc; ++i) out[i] = 0.f;     // out-of-range indices: defined output, nothing read
      if (a.valid) for (int i = 0; i < PP; ++i) a.valid[e * PP + i] = 0.f;
      if (a.Ji) {
        for (int i = 0; i < 12; ++i) { a.Ji[e * 12 + i] = 0.f; a.Jj[e * 12 + i] = 0.f; }
        a.Jz[e * 2] = 0.f; a.Jz[e * 2 + 1] = 0.f; a.jvalid[e] = 0.f;
      }
      continue;
    }
    const SE3<float> G = relative_pose(a.poses + fi * 7, a.poses + fj * 7, a.tonly != 0);
    const float* Ki = a.intr + fi * 4;
    const float* Kj = a.intr + fj * 4;
    const float fxi = Ki[0], fyi = Ki[1], cxi = Ki[2], cyi = Ki[3];
    const float fxj = Kj[0], fyj = Kj[1], cxj = Kj[2], cyj = Kj[3];
    const float* pt = a.patches + kf * 3 * PP;
    for (int i = 0; i < PP; ++i) {
      const float X0[4] = {(pt[i] - cxi) / fxi, (pt[PP + i] - cyi) / fyi, 1.0f, pt[2 * PP + i]};
      float X1[4];
      G.act4(X0, X1);
      const float d = 1.0f / fmaxf(X1[2], 0.1f);
      out[i * nc + 0] = fxj * (d * X1[0]) + cxj;
      out[i * nc + 1] = fyj * (d * X1[1]) + cyj;
      if (a.depth) out[i * nc + 2] = d;
      if (a.valid) a.valid[e * PP + i] = (X1[2] > 0.2f) ? 1.0f : 0.0f;
      if (a.Ji && i == (a.P / 2) * a.P + a.P / 2) {
        const float X = X1[0], Y = X1[1], Z = X1[2], H = X1[3];
        const float dj = (fabsf(Z) > 0.2f) ? 1.0f / Z : 0.0f;
        // Jp (2x4) * Ja (4x6)
        const float p00 = fxj * dj, p02 = -fxj * X * dj * dj;
        const float p11 = fyj * dj, p12 = -fyj * Y * dj * dj;
        float J[2][6];
        J[0][0] = p00 * H; J[0][1] = 0.f;     J[0][2] = p02 * H; J[0][3] = p02 * Y;            J[0][4] = p00 * Z - p02 * X; J[0][5] = -p00 * Y;
        J[1][0] = 0.f;     J[1][1] = p11 * H; J[1][2] = p12 * H; J[1][3] = -p11 * Z + p12 * Y; J[1][4] = -p12 * X;          J[1][5] = p11 * X;
        float Ad[6][6];
        G.Adj(Ad);
#pragma unroll
        for (int r = 0; r < 2; ++r) {
#pragma unroll
          for (int c = 0; c < 6; ++c) {
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < 6; ++k) s += Ad[k][c] * J[r][k];       // (Adj^T J_r)[c]
            a.Ji[e * 12 + r * 6 + c] = -s;
            a.Jj[e * 12 + r * 6 + c] = J[r][c];
          }
        }
        a.Jz[e * 2 + 0] = p00 * G.t.x + p02 * G.t.z;
        a.Jz[e * 2 + 1] = p11 * G.t.y + p12 * G.t.z;
        a.jvalid[e] = (Z > 0.2f) ? 1.0f : 0.0f;
      }
    }
  }
}

// flow_mag (projective_ops.py:116-126): beta * |x_ij - x_ii| + (1 - beta) * |x_ij(translation only) - x_ii|
__global__ void __launch_bounds__(128)
flow_mag_kernel(const TransformArgs a, float beta, float* __restrict__ flow, unsigned char* __restrict__ val) {
  const int PP = a.P * a.P;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < a.E; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t fi = a.ii[e], fj = a.jj[e], kf = a.kk[e];
    if (fi < 0 || fi >= a.nframes || fj < 0 || fj >= a.nframes || kf < 0 || kf >= a.npatches) {
      for (int i = 0; i < PP; ++i) { flow[e * PP + i] = 0.f; val[e * PP + i] = 0; }
      continue;
    }
    const SE3<float> G0 = relative_pose(a.poses + fi * 7, a.poses + fi * 7, false);
    const SE3<float> G1 = relative_pose(a.poses + fi * 7, a.poses + fj * 7, false);
    const SE3<float> G2 = relative_pose(a.poses + fi * 7, a.poses + fj * 7, true);
    const float* Ki = a.intr + fi * 4;
    const float* Kj = a.intr + fj * 4;
    const float* pt = a.patches + kf * 3 * PP;
    for (int i = 0; i < PP; ++i) {
      const float X0[4] = {(pt[i] - Ki[2]) / Ki[0], (pt[PP + i] - Ki[3]) / Ki[1], 1.0f, pt[2 * PP + i]};
      float A0[4], A1[4], A2[4];
      G0.act4(X0, A0); G1.act4(X0, A1); G2.act4(X0, A2);
      const float d0 = 1.0f / fmaxf(A0[2], 0.1f), d1 = 1.0f / fmaxf(A1[2], 0.1f), d2 = 1.0f / fmaxf(A2[2], 0.1f);
      const float x0 = Ki[0] * (d0 * A0[0]) + Ki[2], y0 = Ki[1] * (d0 * A0[1]) + Ki[3];
      const float x1 = Kj[0] * (d1 * A1[0]) + Kj[2], y1 = Kj[1] * (d1 * A1[1]) + Kj[3];
      const float x2 = Kj[0] * (d2 * A2[0]) + Kj[2], y2 = Kj[1] * (d2 * A2[1]) + Kj[3];
      const float f1 = sqrtf((x1 - x0) * (x1 - x0) + (y1 - y0) * (y1 - y0));
      const float f2 = sqrtf((x2 - x0) * (x2 - x0) + (y2 - y0) * (y2 - y0));
      flow[e * PP + i] = beta * f1 + (1.0f - beta) * f2;
      val[e * PP + i] = (A1[2] > 0.2f) ? 1 : 0;
    }
  }
}

// motionmag(i, j) + motionmag(j, i) (dpvo.py:257-264, 266-272): the keyframe test reads the mean flow magnitude of the
// edges i -> j and of the edges j -> i.  ONE cluster of 8 CTAs scans the live edge list (a single CTA took 34 us for
// 45k edges, right in front of the frame's host read), evaluates flow_mag only on the matching edges (about two frames'
// worth of patches) and reduces in a fixed order — warp shuffles, then shared memory, then CTA 0 reads the other CTAs'
// partial sums through distributed shared memory: no atomics, no scratch, the decision repeats bit for bit.  An empty
// selection gives 0 / 0 = NaN like torch's mean over an empty tensor (=> the comparison is false, nothing is removed).
//
// With a KeyframePlan the same pass also COUNTS what keyframe() will do to the edge list under both outcomes of that
// test (dpvo.py:272-310), so the host learns the decision and the edge counts that follow from it in ONE read and issues
// the edits without reading their result back:
//   outcome A (no removal):   edges moved to the inactive list by the window rule at n            -> out[1], min ii -> out[2]
//   outcome B (frame k goes): edges touching k dropped -> out[3]; of the rest, renumbered (ii, jj > k shift down, kk by
//                             M), those moved by the window rule at n - 1 -> out[4], min ii of the survivors -> out[5]
// (the rules are edges_mask_keyframe / edges_renumber / edges_mask_window of csrc/edges.cu, evaluated without editing).
constexpr int kMotionCluster = 8;
constexpr int kPlanNone = 0x7fffffff;      // "no live edge" for the two minima
struct KeyframePlan {
  const int64_t* ix; int64_t n_ix;        // patch -> source frame table
  int64_t k, n, M, removal_window, optimization_window;
  int loop_closure;
  const int64_t* live_count; const int32_t* info;    // optional: the edge store's device count / error flag, echoed
};

__device__ __forceinline__ float edge_flow_mean(const TransformArgs& a, int64_t fi, int64_t fj, int64_t kf, float beta) {
  const int PP = a.P * a.P;
  if (!(fi >= 0 && fi < a.nframes && fj >= 0 && fj < a.nframes && kf >= 0 && kf < a.npatches)) return 0.f;
  const SE3<float> G0 = relative_pose(a.poses + fi * 7, a.poses + fi * 7, false);
  const SE3<float> G1 = relative_pose(a.poses + fi * 7, a.poses + fj * 7, false);
  const SE3<float> G2 = relative_pose(a.poses + fi * 7, a.poses + fj * 7, true);
  const float* Ki = a.intr + fi * 4;
  const float* Kj = a.intr + fj * 4;
  const float* pt = a.patches + kf * 3 * PP;
  float m = 0.f;
  for (int i = 0; i < PP; ++i) {
    const float X0[4] = {(pt[i] - Ki[2]) / Ki[0], (pt[PP + i] - Ki[3]) / Ki[1], 1.0f, pt[2 * PP + i]};
    float A0[4], A1[4], A2[4];
    G0.act4(X0, A0); G1.act4(X0, A1); G2.act4(X0, A2);
    const float d0 = 1.0f / fmaxf(A0[2], 0.1f), d1 = 1.0f / fmaxf(A1[2], 0.1f), d2 = 1.0f / fmaxf(A2[2], 0.1f);
    const float x0 = Ki[0] * (d0 * A0[0]) + Ki[2], y0 = Ki[1] * (d0 * A0[1]) + Ki[3];
    const float x1 = Kj[0] * (d1 * A1[0]) + Kj[2], y1 = Kj[1] * (d1 * A1[1]) + Kj[3];
    const float x2 = Kj[0] * (d2 * A2[0]) + Kj[2], y2 = Kj[1] * (d2 * A2[1]) + Kj[3];
    const float f1 = sqrtf((x1 - x0) * (x1 - x0) + (y1 - y0) * (y1 - y0));
    const float f2 = sqrtf((x2 - x0) * (x2 - x0) + (y2 - y0) * (y2 - y0));
    m += beta * f1 + (1.0f - beta) * f2;
  }
  return m / (float)PP;
}

// window rule of keyframe() (dpvo.py:304-309) for an edge (i, j) whose patch belongs to frame src, at frame count n
__device__ __forceinline__ bool window_rule(const KeyframePlan& q, int64_t src, int64_t i, int64_t j, int64_t n) {
  bool rem = src < n - q.removal_window;
  if (q.loop_closure) rem = rem && !((j - i > 30) && (j > n - q.optimization_window));
  return rem;
}

template <bool PLAN>
__global__ void __cluster_dims__(kMotionCluster, 1, 1) __launch_bounds__(1024)
motionmag_pair_kernel(const TransformArgs a, float beta, int64_t fa, int64_t fb, const KeyframePlan q,
                      float* __restrict__ out) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  float s[2] = {0.f, 0.f}, c[2] = {0.f, 0.f};
  int cnt[3] = {0, 0, 0};                  // moved (A), dropped (B), moved (B)
  int mn[2] = {kPlanNone, kPlanNone};      // min ii of the survivors (A), (B)
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < a.E; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t fi = a.ii[e], fj = a.jj[e];
    const int dir = (fi == fa && fj == fb) ? 0 : ((fi == fb && fj == fa) ? 1 : -1);
    if (!PLAN && dir < 0) continue;
    const int64_t kf = a.kk[e];
    if (dir >= 0) {
      s[dir] += edge_flow_mean(a, fi, fj, kf, beta);
      c[dir] += 1.f;
    }
    if (PLAN) {
      const int64_t src = (kf >= 0 && kf < q.n_ix) ? q.ix[kf] : 0;
      if (window_rule(q, src, fi, fj, q.n)) ++cnt[0];
      else mn[0] = min(mn[0], (int)fi);
      if (fi == q.k || fj == q.k) {
        ++cnt[1];
      } else {
        const int64_t i2 = fi - (fi > q.k ? 1 : 0), j2 = fj - (fj > q.k ? 1 : 0);
        const int64_t k2 = kf - (fi > q.k ? q.M : 0);
        const int64_t src2 = (k2 >= 0 && k2 < q.n_ix) ? q.ix[k2] : 0;
        if (window_rule(q, src2, i2, j2, q.n - 1)) ++cnt[2];
        else mn[1] = min(mn[1], (int)i2);
      }
    }
  }
  __shared__ float red[32][4];
  __shared__ int redi[32][5];
  __shared__ float part[4];
  __shared__ int parti[5];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float v[4] = {s[0], c[0], s[1], c[1]};
  int w[5] = {cnt[0], cnt[1], cnt[2], mn[0], mn[1]};
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int t = 0; t < 4; ++t) v[t] += __shfl_xor_sync(0xffffffffu, v[t], o);
    if (PLAN) {
#pragma unroll
      for (int t = 0; t < 3; ++t) w[t] += __shfl_xor_sync(0xffffffffu, w[t], o);
#pragma unroll
      for (int t = 3; t < 5; ++t) w[t] = min(w[t], __shfl_xor_sync(0xffffffffu, w[t], o));
    }
  }
  if (lane == 0) {
    for (int t = 0; t < 4; ++t) red[warp][t] = v[t];
    for (int t = 0; t < 5; ++t) redi[warp][t] = w[t];
  }
  __syncthreads();
  if (warp == 0) {
    const int nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int t = 0; t < 4; ++t) v[t] = lane < nw ? red[lane][t] : 0.f;
#pragma unroll
    for (int t = 0; t < 5; ++t) w[t] = lane < nw ? redi[lane][t] : (t < 3 ? 0 : kPlanNone);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int t = 0; t < 4; ++t) v[t] += __shfl_xor_sync(0xffffffffu, v[t], o);
#pragma unroll
      for (int t = 0; t < 3; ++t) w[t] += __shfl_xor_sync(0xffffffffu, w[t], o);
#pragma unroll
      for (int t = 3; t < 5; ++t) w[t] = min(w[t], __shfl_xor_sync(0xffffffffu, w[t], o));
    }
    if (lane == 0) {
      for (int t = 0; t < 4; ++t) part[t] = v[t];
      for (int t = 0; t < 5; ++t) parti[t] = w[t];
    }
  }
  cluster.sync();
  if (cluster.block_rank() == 0 && threadIdx.x == 0) {
    float t4[4] = {0.f, 0.f, 0.f, 0.f};
    int t5[5] = {0, 0, 0, kPlanNone, kPlanNone};
    for (int r = 0; r < kMotionCluster; ++r) {
      const float* pr = cluster.map_shared_rank(part, r);
      const int* pi = cluster.map_shared_rank(parti, r);
      for (int t = 0; t < 4; ++t) t4[t] += pr[t];
      for (int t = 0; t < 3; ++t) t5[t] += pi[t];
      for (int t = 3; t < 5; ++t) t5[t] = min(t5[t], pi[t]);
    }
    out[0] = t4[0] / t4[1] + t4[2] / t4[3];
    if (PLAN) {
      int* oi = reinterpret_cast<int*>(out);
      oi[1] = t5[0]; oi[2] = t5[3]; oi[3] = t5[1]; oi[4] = t5[2]; oi[5] = t5[4];
      oi[6] = q.live_count ? (int)min(*q.live_count, (int64_t)kPlanNone) : 0;
      oi[7] = q.info ? *q.info : 0;
    }
  }
  cluster.sync();                       // the other CTAs' shared memory stays alive until CTA 0 has read it
}

// point_cloud (projective_ops.py:111-113): G_ix^-1 * iproj(patch n, intrinsics[ix[n]])
__global__ void __launch_bounds__(128)
point_cloud_kernel(const float* __restrict__ poses, const float* __restrict__ patches, const float* __restrict__ intr,
                   const int64_t* __restrict__ ix, int64_t n, int64_t nframes, int P, float* __restrict__ out) {
  const int PP = P * P;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t f = ix[e];
    float* o = out + e * PP * 4;
    if (f < 0 || f >= nframes) { for (int i = 0; i < PP * 4; ++i) o[i] = 0.f; continue; }
    SE3<float> G = SE3<float>::load(poses + f * 7).inv();
    G.q = SO3<float>::from_quat(G.q.x, G.q.y, G.q.z, G.q.w);
    const float* K = intr + f * 4;
    const float* pt = patches + e * 3 * PP;
    for (int i = 0; i < PP; ++i) {
      const float X0[4] = {(pt[i] - K[2]) / K[0], (pt[PP + i] - K[3]) / K[1], 1.0f, pt[2 * PP + i]};
      G.act4(X0, o + 4 * i);
    }
  }
}

inline int grid_for(int64_t n) {
  const int64_t g = (n + 127) / 128;
  return (int)(g < (int64_t)kNumSMs * 16 ? (g > 0 ? g : 1) : (int64_t)kNumSMs * 16);
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

extern "C" int dpvo_transform(const float* poses, const float* patches, const float* intrinsics,
                              const int64_t* ii, const int64_t* jj, const int64_t* kk,
                              int64_t n_edges, int64_t n_frames, int64_t n_patches, int32_t P,
                              int32_t depth, int32_t tonly, float* coords, float* valid,
                              float* Ji, float* Jj, float* Jz, float* jvalid, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && n_frames > 0 && n_patches > 0 && P > 0 && P <= 8, "transform: bad sizes");
  if (n_edges == 0) return 0;
  DPVO_REQUIRE(poses && patches && intrinsics && ii && jj && kk && coords, "transform: null pointer");
  DPVO_REQUIRE((Ji == nullptr) == (Jj == nullptr) && (Ji == nullptr) == (Jz == nullptr) && (Ji == nullptr) == (jvalid == nullptr),
               "transform: Ji, Jj, Jz, jvalid must be given together");
  TransformArgs a{poses, patches, intrinsics, ii, jj, kk, n_edges, n_frames, n_patches, P, depth ? 1 : 0, tonly ? 1 : 0,
                  coords, valid, Ji, Jj, Jz, jvalid};
  transform_kernel<<<grid_for(n_edges), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
  note_launches(1);
  return check_launch("transform");
}

extern "C" int dpvo_flow_mag(const float* poses, const float* patches, const float* intrinsics,
                             const int64_t* ii, const int64_t* jj, const int64_t* kk,
                             int64_t n_edges, int64_t n_frames, int64_t n_patches, int32_t P, float beta,
                             float* flow, uint8_t* valid, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && n_frames > 0 && n_patches > 0 && P > 0 && P <= 8, "flow_mag: bad sizes");
  if (n_edges == 0) return 0;
  DPVO_REQUIRE(poses && patches && intrinsics && ii && jj && kk && flow && valid, "flow_mag: null pointer");
  TransformArgs a{poses, patches, intrinsics, ii, jj, kk, n_edges, n_frames, n_patches, P, 0, 0,
                  nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  flow_mag_kernel<<<grid_for(n_edges), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, beta, flow, valid);
  note_launches(1);
  return check_launch("flow_mag");
}

extern "C" int dpvo_motionmag_pair(const float* poses, const float* patches, const float* intrinsics,
                                  const int64_t* ii, const int64_t* jj, const int64_t* kk,
                                  int64_t n_edges, int64_t n_frames, int64_t n_patches, int32_t P, float beta,
                                  int64_t frame_a, int64_t frame_b, float* out, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && n_frames > 0 && n_patches > 0 && P > 0 && P <= 8, "motionmag_pair: bad sizes");
  DPVO_REQUIRE(poses && patches && intrinsics && out && (n_edges == 0 || (ii && jj && kk)), "motionmag_pair: null pointer");
  TransformArgs a{poses, patches, intrinsics, ii, jj, kk, n_edges, n_frames, n_patches, P, 0, 0,
                  nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  motionmag_pair_kernel<false><<<kMotionCluster, 1024, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, beta, frame_a, frame_b,
                                                                                                    KeyframePlan{}, out);
  note_launches(1);
  return check_launch("motionmag_pair");
}

extern "C" int dpvo_keyframe_plan(const float* poses, const float* patches, const float* intrinsics,
                                  const int64_t* ii, const int64_t* jj, const int64_t* kk, const int64_t* ix,
                                  int64_t n_edges, int64_t n_frames, int64_t n_patches, int32_t P, float beta,
                                  int64_t frame_a, int64_t frame_b, int64_t k, int64_t n, int64_t patches_per_frame,
                                  int64_t removal_window, int32_t loop_closure, int64_t optimization_window,
                                  const int64_t* live_count, const int32_t* info, int32_t* out, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && n_edges < (int64_t(1) << 31) && n_frames > 0 && n_patches > 0 && P > 0 && P <= 8 &&
               patches_per_frame > 0, "keyframe_plan: bad sizes");
  DPVO_REQUIRE(poses && patches && intrinsics && ix && out && (n_edges == 0 || (ii && jj && kk)), "keyframe_plan: null pointer");
  TransformArgs a{poses, patches, intrinsics, ii, jj, kk, n_edges, n_frames, n_patches, P, 0, 0,
                  nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  const KeyframePlan q{ix, n_patches, k, n, patches_per_frame, removal_window, optimization_window, loop_closure ? 1 : 0,
                       live_count, info};
  motionmag_pair_kernel<true><<<kMotionCluster, 1024, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      a, beta, frame_a, frame_b, q, reinterpret_cast<float*>(out));
  note_launches(1);
  return check_launch("keyframe_plan");
}

extern "C" int dpvo_point_cloud(const float* poses, const float* patches, const float* intrinsics, const int64_t* ix,
                                int64_t n, int64_t n_frames, int32_t P, float* points, void* stream) {
  DPVO_REQUIRE(n >= 0 && n_frames > 0 && P > 0 && P <= 8, "point_cloud: bad sizes");
  if (n == 0) return 0;
  DPVO_REQUIRE(poses && patches && intrinsics && ix && points, "point_cloud: null pointer");
  point_cloud_kernel<<<grid_for(n), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(poses, patches, intrinsics, ix, n,
                                                                                     n_frames, P, points);
  note_launches(1);
  return check_launch("point_cloud");
}
