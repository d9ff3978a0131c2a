// Sparse patch-graph bundle adjustment (fastba.BA) for sm_100a — dense-E path and shared pieces.
//
// Replaces /root/reference/methods/dpvo/dpvo/fastba/ba_cuda.cu:232-376 (per-edge
// reprojection residual + analytic Jacobians + accumulation) and :433-582 (normal
// equations, Schur complement over patch depths, Cholesky solve, retraction).
//
// Pipeline per call (no host synchronisation, no allocation, all on `stream`):
//   mark + prefix     patch compaction: kx = sorted unique(kk), ku = rank (bitmap + popcount
//                     prefix; never materialises ku — it is recomputed from two L1-resident
//                     tables where needed)                                (ba_cuda.cu:447-449)
//   per iteration:
//     zero            accumulators, bounded by the device-side patch count
//     assemble        one thread per edge: Gij, reprojection, r, Ji/Jj/Jz, validity mask;
//                     both residual rows are combined in registers, pose blocks are reduced
//                     in shared memory per CTA (upper-triangular, halved diagonal so that
//                     B = U + U^T) and flushed once; E^T rows / C / u use global reductions
//     schur           G = E Q E^T and g = E Q u as a split-K tiled fp32 contraction
//     solve           S = B - G (+ damping), in-CTA Cholesky + triangular solves, pose retraction
//     backsub         dZ = Q (u - E^T dX), depth retraction (ba_cuda.cu:209-229)
// Deviation from the reference, on purpose: a non-positive Cholesky pivot (which the
// reference ignores, producing NaN poses) leaves poses/depths of that iteration untouched.
#include "se3.cuh"
#include "ba_internal.cuh"

namespace dpvo {

// ------------------------------------------------------------------ compaction
__global__ void __launch_bounds__(256)
ba_mark_kernel(const int64_t* __restrict__ kk, int64_t E, uint32_t* __restrict__ bitmap, int64_t npatches) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = kk[e];
    if (k >= 0 && k < npatches) atomicOr(&bitmap[k >> 5], 1u << (k & 31));
  }
}

// One CTA: exclusive popcount prefix over the bitmap words, list of set bits (kx), total M.
__global__ void __launch_bounds__(1024)
ba_prefix_kernel(const uint32_t* __restrict__ bitmap, int64_t W, int* __restrict__ prefix,
                 int64_t* __restrict__ kx, int* __restrict__ meta, int64_t Mmax) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < W; base += 1024) {
    const int64_t w = base + tid;
    const uint32_t bits = (w < W) ? bitmap[w] : 0u;
    const int c = __popc(bits);
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    if (wid == 0) {
      int x = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += t;
      }
      s_warp[lane] = x;
    }
    __syncthreads();
    const int excl = s_carry + (wid ? s_warp[wid - 1] : 0) + inc - c;
    if (w < W) {
      prefix[w] = excl;
      uint32_t b = bits;
      int r = excl;
      while (b) {
        const int bit = __ffs(b) - 1;
        b &= b - 1;
        if (r < Mmax) kx[r] = w * 32 + bit;
        ++r;
      }
    }
    __syncthreads();
    if (tid == 0) s_carry += s_warp[31];
    __syncthreads();
  }
  if (tid == 0) meta[0] = (int)(s_carry < Mmax ? s_carry : Mmax);
}

// --------------------------------------------------------------------- zeroing
__global__ void __launch_bounds__(256)
ba_zero_kernel(const __grid_constant__ BADev p) {
  const int M = p.meta[0];
  const int64_t n = p.n;
  const int64_t nfix = 2 * n * n + 2 * n;  // U, v, G, g are contiguous
  const int64_t total = nfix + 2 * (int64_t)M + (int64_t)M * n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    if (i < nfix) p.U[i] = 0.f;
    else if (i < nfix + M) p.C[i - nfix] = 0.f;
    else if (i < nfix + 2 * (int64_t)M) p.u[i - nfix - M] = 0.f;
    else p.Et[i - nfix - 2 * (int64_t)M] = 0.f;
  }
}

// -------------------------------------------------------------------- assembly
template <bool SMEM>
__global__ void __launch_bounds__(256)
ba_assemble_kernel(const __grid_constant__ BADev p) {
  extern __shared__ float sm[];
  const int n = p.n, N = p.N;
  float* Us = SMEM ? sm : p.U;
  float* vs = SMEM ? sm + n * n : p.v;
  if (SMEM) {
    for (int i = threadIdx.x; i < n * n + n; i += blockDim.x) sm[i] = 0.f;
    __syncthreads();
  }
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];

  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < p.E; e += (int64_t)gridDim.x * blockDim.x) {
    EdgeTerms T;
    if (!edge_terms(p, e, fx, fy, cx, cy, T)) continue;
    const int k = patch_rank(p, T.kf);
    const bool iv = (T.ix >= 0) && (T.ix < N);
    const bool jv = (T.jx >= 0) && (T.jx < N);

    float cc = 0.f, uu = 0.f;
    float vi[6] = {0, 0, 0, 0, 0, 0}, vj[6] = {0, 0, 0, 0, 0, 0};
    float ei[6] = {0, 0, 0, 0, 0, 0}, ej[6] = {0, 0, 0, 0, 0, 0};
    float Hii[21], Hjj[21], Hij[36];
#pragma unroll
    for (int i = 0; i < 21; ++i) { Hii[i] = 0.f; Hjj[i] = 0.f; }
#pragma unroll
    for (int i = 0; i < 36; ++i) Hij[i] = 0.f;
#pragma unroll
    for (int row = 0; row < 2; ++row) {
      const float w = T.w[row], r = T.r[row], Jz = T.Jz[row];
      const float(&Ji)[6] = T.Ji[row];
      const float(&Jj)[6] = T.Jj[row];
      int idx = 0;
#pragma unroll
      for (int a = 0; a < 6; ++a) {
#pragma unroll
        for (int b = a; b < 6; ++b) {
          Hii[idx] += w * Ji[a] * Ji[b];
          Hjj[idx] += w * Jj[a] * Jj[b];
          ++idx;
        }
#pragma unroll
        for (int b = 0; b < 6; ++b) Hij[a * 6 + b] += w * Ji[a] * Jj[b];
        vi[a] += -w * r * Ji[a];
        vj[a] += w * r * Jj[a];
        ei[a] += -w * Jz * Ji[a];
        ej[a] += w * Jz * Jj[a];
      }
      cc += w * Jz * Jz;
      uu += w * r * Jz;
    }

    if (iv) {
      const int o = 6 * T.ix;
      int idx = 0;
#pragma unroll
      for (int a = 0; a < 6; ++a) {
#pragma unroll
        for (int b = a; b < 6; ++b) {
          atomicAdd(&Us[(o + a) * n + o + b], (a == b ? 0.5f : 1.0f) * Hii[idx]);
          ++idx;
        }
        atomicAdd(&vs[o + a], vi[a]);
        atomicAdd(&p.Et[(int64_t)k * n + o + a], ei[a]);
      }
    }
    if (jv) {
      const int o = 6 * T.jx;
      int idx = 0;
#pragma unroll
      for (int a = 0; a < 6; ++a) {
#pragma unroll
        for (int b = a; b < 6; ++b) {
          atomicAdd(&Us[(o + a) * n + o + b], (a == b ? 0.5f : 1.0f) * Hjj[idx]);
          ++idx;
        }
        atomicAdd(&vs[o + a], vj[a]);
        atomicAdd(&p.Et[(int64_t)k * n + o + a], ej[a]);
      }
    }
    if (iv && jv) {
      const int oi = 6 * T.ix, oj = 6 * T.jx;
#pragma unroll
      for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int b = 0; b < 6; ++b) atomicAdd(&Us[(oi + a) * n + oj + b], -Hij[a * 6 + b]);
    }
    atomicAdd(&p.C[k], cc);
    atomicAdd(&p.u[k], uu);
  }

  if (SMEM) {
    __syncthreads();
    for (int i = threadIdx.x; i < n * n; i += blockDim.x) {
      const float x = sm[i];
      if (x != 0.f) atomicAdd(&p.U[i], x);
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float x = sm[n * n + i];
      if (x != 0.f) atomicAdd(&p.v[i], x);
    }
  }
}

// ----------------------------------------------------------------------- Schur
// G[a][b] += sum_k Et[k][a] Q[k] Et[k][b];  g[a] += sum_k Et[k][a] Q[k] u[k].
// grid.x = split over patch chunks, grid.y = 64x64 output tiles.
constexpr int kSchurKC = 32;

__global__ void __launch_bounds__(256)
ba_schur_kernel(const __grid_constant__ BADev p) {
  __shared__ float As[kSchurKC][64 + 1];
  __shared__ float Bs[kSchurKC][64 + 1];
  __shared__ float us[kSchurKC];
  const int n = p.n;
  const int M = p.meta[0];
  const int tiles = (n + 63) / 64;
  const int ta = blockIdx.y / tiles, tb = blockIdx.y % tiles;
  const int a0 = ta * 64, b0 = tb * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const float lm = p.lmbda[0];
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float gacc = 0.f;

  for (int k0 = blockIdx.x * kSchurKC; k0 < M; k0 += gridDim.x * kSchurKC) {
    for (int i = threadIdx.x; i < kSchurKC * 64; i += blockDim.x) {
      const int kk_ = i >> 6, c = i & 63;
      const int k = k0 + kk_;
      float a = 0.f, b = 0.f;
      if (k < M) {
        const float q = 1.0f / (p.C[k] + lm);
        if (a0 + c < n) a = p.Et[(int64_t)k * n + a0 + c] * q;
        if (b0 + c < n) b = p.Et[(int64_t)k * n + b0 + c];
      }
      As[kk_][c] = a;
      Bs[kk_][c] = b;
    }
    if (threadIdx.x < kSchurKC) {
      const int k = k0 + threadIdx.x;
      us[threadIdx.x] = (k < M) ? p.u[k] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int kk_ = 0; kk_ < kSchurKC; ++kk_) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { av[i] = As[kk_][ty * 4 + i]; bv[i] = Bs[kk_][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] += av[i] * bv[j];
    }
    if (tb == 0 && threadIdx.x < 64) {
      for (int kk_ = 0; kk_ < kSchurKC; ++kk_) gacc += As[kk_][threadIdx.x] * us[kk_];
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int a = a0 + ty * 4 + i, b = b0 + tx * 4 + j;
      if (a < n && b < n && acc[i][j] != 0.f) atomicAdd(&p.G[(int64_t)a * n + b], acc[i][j]);
    }
  if (tb == 0 && threadIdx.x < 64 && a0 + threadIdx.x < n && gacc != 0.f) atomicAdd(&p.g[a0 + threadIdx.x], gacc);
}

// ----------------------------------------------------------------------- solve
// One CTA.  S = (U + U^T) - G, y = v - g, S_dd += 1e-4 S_dd + 1, Cholesky, solves,
// dX -> global, pose retraction.  S lives in shared memory when it fits.
__global__ void __launch_bounds__(512)
ba_solve_kernel(const __grid_constant__ BADev p, int s_in_smem) {
  extern __shared__ float sm[];
  const int n = p.n, N = p.N;
  const int T = blockDim.x, tid = threadIdx.x;
  float* y = sm;                       // n
  float* S = s_in_smem ? sm + n : p.Sg;  // n*n (lower triangle used)
  __shared__ int s_fail;
  if (tid == 0) s_fail = 0;
  for (int i = tid; i < n * n; i += T) {
    const int a = i / n, b = i - a * n;
    if (b <= a) {
      float s = p.U[(int64_t)a * n + b] + p.U[(int64_t)b * n + a] - 0.5f * (p.G[(int64_t)a * n + b] + p.G[(int64_t)b * n + a]);
      if (a == b) s += 1e-4f * s + 1.0f;
      S[(int64_t)a * n + b] = s;
    }
  }
  for (int i = tid; i < n; i += T) y[i] = p.v[i] - p.g[i];
  __syncthreads();

  for (int j = 0; j < n; ++j) {
    if (tid == 0) {
      const float d = S[(int64_t)j * n + j];
      if (!(d > 0.f)) s_fail = 1;
      S[(int64_t)j * n + j] = sqrtf(d);
    }
    __syncthreads();
    if (s_fail) break;
    const float inv = 1.0f / S[(int64_t)j * n + j];
    for (int i = j + 1 + tid; i < n; i += T) S[(int64_t)i * n + j] *= inv;
    __syncthreads();
    const int m = n - j - 1;
    for (int idx = tid; idx < m * m; idx += T) {
      const int r = idx / m, c = idx - r * m;
      if (c <= r) {
        const int i = j + 1 + r, k = j + 1 + c;
        S[(int64_t)i * n + k] -= S[(int64_t)i * n + j] * S[(int64_t)k * n + j];
      }
    }
    __syncthreads();
  }
  if (s_fail) {
    for (int i = tid; i < n; i += T) p.dX[i] = 0.f;
    if (tid == 0) p.meta[1] = 1;  // tells backsub to skip the depth update too
    return;
  }
  // forward: L z = y
  for (int j = 0; j < n; ++j) {
    if (tid == 0) y[j] /= S[(int64_t)j * n + j];
    __syncthreads();
    const float yj = y[j];
    for (int i = j + 1 + tid; i < n; i += T) y[i] -= S[(int64_t)i * n + j] * yj;
    __syncthreads();
  }
  // backward: L^T x = z
  for (int j = n - 1; j >= 0; --j) {
    if (tid == 0) y[j] /= S[(int64_t)j * n + j];
    __syncthreads();
    const float yj = y[j];
    for (int i = tid; i < j; i += T) y[i] -= S[(int64_t)j * n + i] * yj;
    __syncthreads();
  }
  for (int i = tid; i < n; i += T) p.dX[i] = y[i];
  if (tid == 0) p.meta[1] = 0;
  // pose retraction (ba_cuda.cu:178-206)
  for (int i = tid; i < N; i += T) {
    float* ps = p.poses + (int64_t)(p.t0 + i) * 7;
    const Vec3 t = {ps[0], ps[1], ps[2]};
    const Quat q = {ps[3], ps[4], ps[5], ps[6]};
    const float xi[6] = {y[6 * i + 0], y[6 * i + 1], y[6 * i + 2], y[6 * i + 3], y[6 * i + 4], y[6 * i + 5]};
    Vec3 t1;
    Quat q1;
    retr_se3(xi, t, q, t1, q1);
    ps[0] = t1.x; ps[1] = t1.y; ps[2] = t1.z;
    ps[3] = q1.x; ps[4] = q1.y; ps[5] = q1.z; ps[6] = q1.w;
  }
}

// --------------------------------------------------------------------- backsub
__global__ void __launch_bounds__(256)
ba_backsub_kernel(const __grid_constant__ BADev p) {
  extern __shared__ float dXs[];
  const int n = p.n;
  const int M = p.meta[0];
  if (n > 0 && p.meta[1]) return;  // failed factorisation: leave the depths alone
  for (int i = threadIdx.x; i < n; i += blockDim.x) dXs[i] = p.dX[i];
  __syncthreads();
  const float lm = p.lmbda[0];
  const int PP = p.P * p.P;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < M; k += gridDim.x * blockDim.x) {
    float dot = 0.f;
    const float* er = p.Et + (int64_t)k * n;
    for (int a = 0; a < n; ++a) dot += er[a] * dXs[a];
    const float q = 1.0f / (p.C[k] + lm);
    const float dz = q * (p.u[k] - dot);
    depth_retract(p.patches, p.kx[k], PP, dz);
  }
}

// ------------------------------------------------------------------- reproject
__global__ void __launch_bounds__(256)
reproject_kernel(const float* __restrict__ poses, const float* __restrict__ patches,
                 const float* __restrict__ intr, const int64_t* __restrict__ ii,
                 const int64_t* __restrict__ jj, const int64_t* __restrict__ kk,
                 int64_t E, int P, float* __restrict__ coords) {
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int PP = P * P;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
    const float* pi = poses + ii[e] * 7;
    const float* pj = poses + jj[e] * 7;
    const Vec3 ti = {pi[0], pi[1], pi[2]};
    const Quat qi = {pi[3], pi[4], pi[5], pi[6]};
    const Vec3 tj = {pj[0], pj[1], pj[2]};
    const Quat qj = {pj[3], pj[4], pj[5], pj[6]};
    Vec3 tij;
    Quat qij;
    rel_se3(ti, qi, tj, qj, tij, qij);
    const float* pt = patches + kk[e] * 3 * PP;
    float* out = coords + e * 2 * PP;
    for (int i = 0; i < PP; ++i) {
      const Vec3 X = {(pt[i] - cx) / fx, (pt[PP + i] - cy) / fy, 1.0f};
      const float d = pt[2 * PP + i];
      Vec3 Y = rot(qij, X);
      Y.x += d * tij.x; Y.y += d * tij.y; Y.z += d * tij.z;
      out[i] = fx * (Y.x / Y.z) + cx;
      out[PP + i] = fy * (Y.y / Y.z) + cy;
    }
  }
}

static int64_t align256(int64_t x) { return (x + 255) & ~(int64_t)255; }

// Carve the workspace; returns total bytes.  With base==nullptr only sizes are computed.
int64_t ba_carve(BADev& p, char* base, int64_t E, int64_t npatches, int N, int eff, int ppf, int64_t nframes) {
  const int64_t n = 6 * (int64_t)N;
  const int64_t W = (npatches + 31) / 32;
  const int64_t Mmax = E < npatches ? E : npatches;
  int64_t off = 0;
  auto take = [&](int64_t bytes) { int64_t o = off; off += align256(bytes); return base ? base + o : (char*)nullptr; };
  p.bitmap = (uint32_t*)take(W * 4);
  p.prefix = (int*)take(W * 4);
  p.meta = (int*)take(64);
  p.kx = (int64_t*)take(Mmax * 8);
  // U, v, G, g contiguous (zeroed as one range)
  char* fixed = take((2 * n * n + 2 * n) * 4);
  p.U = (float*)fixed;
  p.v = p.U + n * n;
  p.G = p.v + n;
  p.g = p.G + n * n;
  p.C = (float*)take(Mmax * 4);
  p.u = (float*)take(Mmax * 4);
  p.dX = (float*)take((n + 8) * 4);
  p.Sg = (float*)take(n * n * 4);
  p.Mmax = Mmax;
  p.W = W;
  if (!eff) {
    p.Et = (float*)take(Mmax * n * 4);
  } else {
    p.Et = nullptr;
    off += ba_eff_carve(p, base ? base + off : nullptr, E, npatches, N, ppf, nframes);
  }
  return off;
}

}  // namespace dpvo

using namespace dpvo;

extern "C" int64_t dpvo_ba_workspace_bytes(int64_t n_edges, int64_t n_patches_total, int32_t n_poses_opt,
                                           int32_t eff_impl, int32_t ppf, int64_t n_frames_total) {
  BADev p;
  memset(&p, 0, sizeof(p));
  if (n_poses_opt < 0) n_poses_opt = 0;
  return ba_carve(p, nullptr, n_edges, n_patches_total, n_poses_opt, eff_impl, ppf, n_frames_total) + 256;
}

extern "C" int dpvo_ba_forward(float* poses, float* patches, const float* intrinsics,
                               const float* target, const float* weight, const float* lmbda,
                               const int64_t* ii, const int64_t* jj, const int64_t* kk,
                               int64_t n_edges, int64_t n_frames_total, int64_t n_patches_total, int32_t P,
                               int32_t ppf, int32_t t0, int32_t t1, int32_t iterations, int32_t eff_impl,
                               void* workspace, int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && n_frames_total > 0 && n_patches_total > 0 && P > 0, "ba_forward: bad sizes");
  DPVO_REQUIRE(t1 >= t0 && t0 >= 0 && t1 <= n_frames_total, "ba_forward: need 0 <= t0 <= t1 <= n_frames (t0=%d t1=%d)", t0, t1);
  DPVO_REQUIRE(P * P <= 64, "ba_forward: patch too large");
  if (n_edges == 0 || iterations <= 0) return 0;
  DPVO_REQUIRE(poses && patches && intrinsics && target && weight && lmbda && ii && jj && kk && workspace,
               "ba_forward: null pointer");
  if (eff_impl) DPVO_REQUIRE(ppf > 0, "ba_forward: eff_impl needs patches-per-frame > 0");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int N = t1 - t0;
  BADev p;
  memset(&p, 0, sizeof(p));
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  const int64_t need = ba_carve(p, base, n_edges, n_patches_total, N, eff_impl, ppf, n_frames_total);
  DPVO_REQUIRE(base + need <= reinterpret_cast<char*>(workspace) + workspace_bytes,
               "ba_forward: workspace too small (%lld needed)", (long long)(need + 256));
  p.poses = poses; p.patches = patches; p.intr = intrinsics; p.target = target; p.weight = weight; p.lmbda = lmbda;
  p.ii = ii; p.jj = jj; p.kk = kk; p.E = n_edges; p.nframes = n_frames_total; p.npatches = n_patches_total;
  p.P = P; p.t0 = t0; p.N = N; p.n = 6 * N; p.ppf = ppf;

  const int egrid = (int)(ceil_div64(n_edges, 256) < kNumSMs * 8 ? ceil_div64(n_edges, 256) : kNumSMs * 8);
  cudaMemsetAsync(p.bitmap, 0, p.W * 4, st);
  cudaMemsetAsync(p.meta, 0, 64, st);
  ba_mark_kernel<<<egrid, 256, 0, st>>>(kk, n_edges, p.bitmap, n_patches_total);
  ba_prefix_kernel<<<1, 1024, 0, st>>>(p.bitmap, p.W, p.prefix, p.kx, p.meta, p.Mmax);
  if (check_launch("ba_prefix")) return 1;

  if (eff_impl) return ba_eff_run(p, iterations, st);

  const int n = p.n;
  const size_t asm_smem = (size_t)(n * n + n) * sizeof(float);
  const bool asm_in_smem = N > 0 && asm_smem <= 200 * 1024;
  if (asm_in_smem && asm_smem > 48 * 1024)
    cudaFuncSetAttribute(ba_assemble_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)asm_smem);
  const size_t solve_full = (size_t)(n * (size_t)n + n) * sizeof(float);
  const int s_in_smem = solve_full <= 200 * 1024;
  const size_t solve_smem = s_in_smem ? solve_full : (size_t)n * sizeof(float);
  if (solve_smem > 48 * 1024)
    cudaFuncSetAttribute(ba_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)solve_smem);
  const int64_t zero_total = 2 * (int64_t)n * n + 2 * n + 2 * p.Mmax + p.Mmax * n;
  const int zgrid = (int)(ceil_div64(zero_total, 256 * 8) < kNumSMs * 4 ? ceil_div64(zero_total, 256 * 8) : kNumSMs * 4);
  const int agrid = asm_in_smem ? (int)(ceil_div64(n_edges, 256) < kNumSMs ? ceil_div64(n_edges, 256) : kNumSMs) : egrid;
  const int tiles = (n + 63) / 64;
  int ksplit = (int)ceil_div64(p.Mmax, kSchurKC);
  const int kcap = (2 * kNumSMs) / (tiles * tiles) > 1 ? (2 * kNumSMs) / (tiles * tiles) : 1;
  if (ksplit > kcap) ksplit = kcap;
  const int bgrid = (int)(ceil_div64(p.Mmax, 256) < kNumSMs * 4 ? ceil_div64(p.Mmax, 256) : kNumSMs * 4);

  for (int it = 0; it < iterations; ++it) {
    ba_zero_kernel<<<zgrid > 0 ? zgrid : 1, 256, 0, st>>>(p);
    if (asm_in_smem)
      ba_assemble_kernel<true><<<agrid, 256, asm_smem, st>>>(p);
    else
      ba_assemble_kernel<false><<<agrid, 256, 0, st>>>(p);
    if (N > 0) {
      ba_schur_kernel<<<dim3(ksplit, tiles * tiles), 256, 0, st>>>(p);
      ba_solve_kernel<<<1, 512, solve_smem, st>>>(p, s_in_smem);
    }
    ba_backsub_kernel<<<bgrid, 256, (size_t)(n + 1) * sizeof(float), st>>>(p);
  }
  note_launches(2 + iterations * (N > 0 ? 5 : 3));
  return check_launch("ba_forward");
}

extern "C" int dpvo_reproject(const float* poses, const float* patches, const float* intrinsics,
                              const int64_t* ii, const int64_t* jj, const int64_t* kk,
                              int64_t n_edges, int32_t P, float* coords, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && P > 0, "reproject: bad sizes");
  if (n_edges == 0) return 0;
  DPVO_REQUIRE(poses && patches && intrinsics && ii && jj && kk && coords, "reproject: null pointer");
  const int grid = (int)(ceil_div64(n_edges, 256) < kNumSMs * 8 ? ceil_div64(n_edges, 256) : kNumSMs * 8);
  note_launches(1);
  reproject_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(poses, patches, intrinsics, ii, jj, kk,
                                                                            n_edges, P, coords);
  return check_launch("reproject");
}
