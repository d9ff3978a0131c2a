// Row-indexed pieces of DPVO's Update operator around the tensor-core layers (update_gemm.cu):
//   gather_rows   mask * net[:, ix]                          net.py:76-80
//   softagg       scatter_softmax / scatter_sum over groups   blocks.py:42-50, torch_scatter.py:75-132
//   expand_add    net + h(y)[:, jx]  (+ the LayerNorm that follows the second aggregation)
//   heads         d = Linear(ReLU(net)), w = sigmoid(Linear(ReLU(net)))   net.py:57-66,88
// All HBM-bound row traffic: 16- / 8-byte vector accesses, one warp (or one small CTA) per row group.
#include "common.cuh"

namespace dpvo {
namespace {

constexpr int kDim = 384;

// ------------------------------------------------------------------ gather_rows
__global__ void __launch_bounds__(256)
gather_rows_kernel(const uint4* __restrict__ src, const int64_t* __restrict__ index, uint4* __restrict__ out,
                   int64_t total, int chunks) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = i / chunks;
    const int c = (int)(i - e * chunks);
    const int64_t j = index[e];
    out[i] = j >= 0 ? __ldg(src + j * chunks + c) : make_uint4(0, 0, 0, 0);
  }
}

// ------------------------------------------------------------ group bucket sort
__global__ void __launch_bounds__(256)
agg_hist_kernel(const int64_t* __restrict__ ids, int64_t E, int* __restrict__ count) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x)
    atomicAdd(count + ids[e], 1);
}

// exclusive scan of count[0..G) -> offs[0..G], cursor = offs (one CTA)
__global__ void __launch_bounds__(1024)
agg_scan_kernel(const int* __restrict__ count, int G, int* __restrict__ offs, int* __restrict__ cursor) {
  __shared__ int warp_sums[32];
  __shared__ int carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int base = 0; base < G; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < G ? count[i] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += t;
    }
    if (lane == 31) warp_sums[wid] = x;
    __syncthreads();
    if (wid == 0) {
      int w = warp_sums[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      warp_sums[lane] = w;
    }
    __syncthreads();
    const int carry = carry_s;
    const int excl = carry + (wid ? warp_sums[wid - 1] : 0) + x - v;
    if (i < G) {
      offs[i] = excl;
      cursor[i] = excl;
    }
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = carry + warp_sums[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) offs[G] = carry_s;
}

__global__ void __launch_bounds__(256)
agg_fill_kernel(const int64_t* __restrict__ ids, int64_t E, int* __restrict__ cursor, int* __restrict__ perm) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x)
    perm[atomicAdd(cursor + ids[e], 1)] = (int)e;
}

// --------------------------------------------------------------------- softagg
// One CTA per group, one thread per 4 channels.  Single pass with a running maximum:
//   y = sum_e f_e exp(g_e - max) / (sum_e exp(g_e - max) + 1e-12)
__device__ __forceinline__ void load4(const __half* p, float (&f)[4]) {
  const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
  const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}
__device__ __forceinline__ void store4(__half* p, const float (&f)[4]) {
  uint2 u;
  *reinterpret_cast<__half2*>(&u.x) = __floats2half2_rn(f[0], f[1]);
  *reinterpret_cast<__half2*>(&u.y) = __floats2half2_rn(f[2], f[3]);
  *reinterpret_cast<uint2*>(p) = u;
}

// kAggSlices: how many ways a group's edges are dealt out inside its CTA (more for few large groups)
template <int kAggSlices>
__global__ void __launch_bounds__(96 * kAggSlices)
softagg_kernel(const __half* __restrict__ F, const __half* __restrict__ Gm, const int* __restrict__ offs,
               const int* __restrict__ perm, __half* __restrict__ Y) {
  __shared__ float part[kAggSlices - 1][3][kDim];     // (max, numerator, denominator) of slices 1..
  const int grp = blockIdx.x;
  const int beg = offs[grp], end = offs[grp + 1];
  if (beg == end) return;
  const int slice = threadIdx.x / 96;                 // the group's edges are dealt round-robin to the slices
  const int ch = (threadIdx.x % 96) * 4;
  float m[4], num[4], den[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) { m[c] = -INFINITY; num[c] = 0.f; den[c] = 0.f; }
  for (int t = beg + slice; t < end; t += 2 * kAggSlices) {
    // two edges per trip: both rows are requested before either is folded in
    const bool two = t + kAggSlices < end;
    const int64_t e0 = perm[t];
    const int64_t e1 = two ? perm[t + kAggSlices] : e0;
    float fv[2][4], gv[2][4];
    load4(F + e0 * kDim + ch, fv[0]);
    load4(Gm + e0 * kDim + ch, gv[0]);
    load4(F + e1 * kDim + ch, fv[1]);
    load4(Gm + e1 * kDim + ch, gv[1]);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (k == 1 && !two) break;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float mn = fmaxf(m[c], gv[k][c]);
        const float sc = __expf(m[c] - mn);       // exp(-inf) = 0 on the first edge
        const float pe = __expf(gv[k][c] - mn);
        num[c] = num[c] * sc + fv[k][c] * pe;
        den[c] = den[c] * sc + pe;
        m[c] = mn;
      }
    }
  }
  if (slice > 0) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      part[slice - 1][0][ch + c] = m[c];
      part[slice - 1][1][ch + c] = num[c];
      part[slice - 1][2][ch + c] = den[c];
    }
  }
  __syncthreads();
  if (slice == 0) {                                   // slice 0 always holds the group's first edge: m is finite
    float y[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float mx = m[c];
#pragma unroll
      for (int s = 0; s < kAggSlices - 1; ++s) mx = fmaxf(mx, part[s][0][ch + c]);
      float nu = num[c] * __expf(m[c] - mx), de = den[c] * __expf(m[c] - mx);
#pragma unroll
      for (int s = 0; s < kAggSlices - 1; ++s) {
        const float sc = __expf(part[s][0][ch + c] - mx);   // empty slice: exp(-inf) = 0
        nu += part[s][1][ch + c] * sc;
        de += part[s][2][ch + c] * sc;
      }
      y[c] = nu / (de + 1e-12f);
    }
    store4(Y + (int64_t)grp * kDim + ch, y);
  }
}

// ------------------------------------------------------------------ expand_add
// out[e] = net[e] + hy[ids[e]], then LayerNorm when gamma != nullptr.  One warp per row, lane l owns
// channels 4l + 128c, c = 0..2.
__global__ void __launch_bounds__(256)
expand_add_kernel(const __half* __restrict__ net, const __half* __restrict__ hy, const int64_t* __restrict__ ids,
                  int64_t E, const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                  __half* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < E; e += warps) {
    const __half* a = net + e * kDim;
    const __half* b = hy + ids[e] * kDim;
    float x[3][4];
    float sum = 0.f, sq = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float u[4], v[4];
      load4(a + 128 * c + 4 * lane, u);
      load4(b + 128 * c + 4 * lane, v);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        x[c][j] = u[j] + v[j];
        sum += x[c][j];
        sq += x[c][j] * x[c][j];
      }
    }
    if (gamma) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        sq += __shfl_xor_sync(0xffffffffu, sq, o);
      }
      const float mean = sum * (1.0f / kDim);
      const float rstd = rsqrtf(fmaxf(sq * (1.0f / kDim) - mean * mean, 0.f) + eps);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float4 g4 = __ldg(reinterpret_cast<const float4*>(gamma + 128 * c + 4 * lane));
        const float4 b4 = __ldg(reinterpret_cast<const float4*>(beta + 128 * c + 4 * lane));
        x[c][0] = (x[c][0] - mean) * rstd * g4.x + b4.x;
        x[c][1] = (x[c][1] - mean) * rstd * g4.y + b4.y;
        x[c][2] = (x[c][2] - mean) * rstd * g4.z + b4.z;
        x[c][3] = (x[c][3] - mean) * rstd * g4.w + b4.w;
      }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) store4(out + e * kDim + 128 * c + 4 * lane, x[c]);
  }
}

// ----------------------------------------------------------------------- heads
// delta[e] = Wd relu(net[e]) + bd;  weight[e] = sigmoid(Ww relu(net[e]) + bw).  One warp per row.
__global__ void __launch_bounds__(256)
heads_kernel(const __half* __restrict__ net, const float* __restrict__ wd, const float* __restrict__ bd,
             const float* __restrict__ ww, const float* __restrict__ bw, int64_t E, float* __restrict__ delta,
             float* __restrict__ weight) {
  __shared__ __align__(16) float w_s[4][kDim];
  for (int i = threadIdx.x; i < 2 * kDim; i += blockDim.x) {
    w_s[i / kDim][i % kDim] = wd[i];
    w_s[2 + i / kDim][i % kDim] = ww[i];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < E; e += warps) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float u[4];
      load4(net + e * kDim + 128 * c + 4 * lane, u);
#pragma unroll
      for (int h = 0; h < 4; ++h) {                   // 16-byte shared loads: consecutive lanes, consecutive banks
        const float4 w4 = *reinterpret_cast<const float4*>(&w_s[h][128 * c + 4 * lane]);
        acc[h] += fmaxf(u[0], 0.f) * w4.x + fmaxf(u[1], 0.f) * w4.y + fmaxf(u[2], 0.f) * w4.z + fmaxf(u[3], 0.f) * w4.w;
      }
    }
#pragma unroll
    for (int h = 0; h < 4; ++h)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc[h] += __shfl_xor_sync(0xffffffffu, acc[h], o);
    if (lane == 0) {
      delta[e * 2] = acc[0] + bd[0];
      delta[e * 2 + 1] = acc[1] + bd[1];
      weight[e * 2] = 1.0f / (1.0f + __expf(-(acc[2] + bw[0])));
      weight[e * 2 + 1] = 1.0f / (1.0f + __expf(-(acc[3] + bw[1])));
    }
  }
}

// finish of the heads accumulated by the last layer's epilogue: acc [E,4] = (d0, d1, w0, w1) dot products
__global__ void __launch_bounds__(256)
heads_finish_kernel(const float4* __restrict__ acc, const float* __restrict__ bd, const float* __restrict__ bw,
                    int64_t E, float2* __restrict__ delta, float2* __restrict__ weight) {
  const float b0 = bd[0], b1 = bd[1], c0 = bw[0], c1 = bw[1];
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = acc[e];
    delta[e] = make_float2(a.x + b0, a.y + b1);
    weight[e] = make_float2(1.0f / (1.0f + __expf(-(a.z + c0))), 1.0f / (1.0f + __expf(-(a.w + c1))));
  }
}

inline int grid_for(int64_t items, int per_block, int cap_mult) {
  const int64_t b = ceil_div64(items, per_block);
  const int64_t cap = (int64_t)kNumSMs * cap_mult;
  return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

extern "C" int dpvo_gather_rows_f16(const void* src, const int64_t* index, int64_t n_rows, int32_t dim, void* out,
                                    void* stream) {
  DPVO_REQUIRE(n_rows >= 0 && dim > 0 && dim % 8 == 0, "gather_rows: dim must be a positive multiple of 8");
  if (n_rows == 0) return 0;
  DPVO_REQUIRE(src && index && out, "gather_rows: null pointer");
  const int chunks = dim / 8;
  const int64_t total = n_rows * chunks;
  note_launches(1);
  gather_rows_kernel<<<grid_for(total, 256, 16), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(src), index, static_cast<uint4*>(out), total, chunks);
  return check_launch("gather_rows");
}

extern "C" int64_t dpvo_softagg_workspace_bytes(int64_t n_edges, int64_t group_bound) {
  if (n_edges < 0 || group_bound < 0) return -1;
  // count[G] | offs[G+1] | cursor[G] | perm[E]
  return (3 * group_bound + 1 + n_edges) * (int64_t)sizeof(int) + 64;
}

namespace {
struct AggWs { int* count; int* offs; int* cursor; int* perm; };
AggWs carve_agg(void* workspace, int G) {
  AggWs w;
  w.count = static_cast<int*>(workspace);
  w.offs = w.count + G;
  w.cursor = w.offs + G + 1;
  w.perm = w.cursor + G;
  return w;
}
}  // namespace

extern "C" int dpvo_softagg_prepare(const int64_t* ids, int64_t n_edges, int64_t group_bound, void* workspace,
                                    int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && group_bound >= 0 && group_bound < (int64_t(1) << 30) && n_edges < (int64_t(1) << 31),
               "softagg_prepare: bad sizes");
  if (n_edges == 0 || group_bound == 0) return 0;
  DPVO_REQUIRE(ids && workspace, "softagg_prepare: null pointer");
  DPVO_REQUIRE(workspace_bytes >= dpvo_softagg_workspace_bytes(n_edges, group_bound),
               "softagg_prepare: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int G = (int)group_bound;
  const AggWs w = carve_agg(workspace, G);
  cudaMemsetAsync(w.count, 0, sizeof(int) * G, st);
  agg_hist_kernel<<<grid_for(n_edges, 256, 8), 256, 0, st>>>(ids, n_edges, w.count);
  agg_scan_kernel<<<1, 1024, 0, st>>>(w.count, G, w.offs, w.cursor);
  agg_fill_kernel<<<grid_for(n_edges, 256, 8), 256, 0, st>>>(ids, n_edges, w.cursor, w.perm);
  note_launches(3);
  return check_launch("softagg_prepare");
}

extern "C" int dpvo_softagg_f16(const void* f, const void* g, const int64_t* ids, int64_t n_edges, int64_t group_bound,
                                void* y, int32_t prepared, void* workspace, int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && group_bound >= 0 && group_bound < (int64_t(1) << 30) && n_edges < (int64_t(1) << 31),
               "softagg: bad sizes");
  if (n_edges == 0 || group_bound == 0) return 0;
  DPVO_REQUIRE(f && g && y && workspace, "softagg: null pointer");
  DPVO_REQUIRE(workspace_bytes >= dpvo_softagg_workspace_bytes(n_edges, group_bound), "softagg: workspace too small");
  if (!prepared && dpvo_softagg_prepare(ids, n_edges, group_bound, workspace, workspace_bytes, stream)) return 1;
  const int G = (int)group_bound;
  const AggWs w = carve_agg(workspace, G);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const __half* fp = static_cast<const __half*>(f);
  const __half* gp = static_cast<const __half*>(g);
  __half* yp = static_cast<__half*>(y);
  const int64_t avg = n_edges / group_bound;          // lower bound on the mean group size
  if (avg >= 24) softagg_kernel<8><<<G, 96 * 8, 0, st>>>(fp, gp, w.offs, w.perm, yp);
  else if (avg >= 6) softagg_kernel<4><<<G, 96 * 4, 0, st>>>(fp, gp, w.offs, w.perm, yp);
  else softagg_kernel<2><<<G, 96 * 2, 0, st>>>(fp, gp, w.offs, w.perm, yp);
  note_launches(1);
  return check_launch("softagg");
}

extern "C" int dpvo_expand_add_f16(const void* net, const void* hy, const int64_t* ids, int64_t n_edges,
                                   const float* ln_gamma, const float* ln_beta, float ln_eps, void* out, void* stream) {
  DPVO_REQUIRE(n_edges >= 0, "expand_add: bad sizes");
  if (n_edges == 0) return 0;
  DPVO_REQUIRE(net && hy && ids && out, "expand_add: null pointer");
  DPVO_REQUIRE((ln_gamma == nullptr) == (ln_beta == nullptr), "expand_add: LayerNorm needs both gamma and beta");
  note_launches(1);
  expand_add_kernel<<<grid_for(n_edges, 8, 16), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      static_cast<const __half*>(net), static_cast<const __half*>(hy), ids, n_edges, ln_gamma, ln_beta, ln_eps,
      static_cast<__half*>(out));
  return check_launch("expand_add");
}

extern "C" int dpvo_update_heads_f16(const void* net, const float* wd, const float* bd, const float* ww,
                                     const float* bw, int64_t n_edges, float* delta, float* weight, void* stream) {
  DPVO_REQUIRE(n_edges >= 0, "update_heads: bad sizes");
  if (n_edges == 0) return 0;
  DPVO_REQUIRE(net && wd && bd && ww && bw && delta && weight, "update_heads: null pointer");
  note_launches(1);
  heads_kernel<<<grid_for(n_edges, 8, 64), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      static_cast<const __half*>(net), wd, bd, ww, bw, n_edges, delta, weight);
  return check_launch("update_heads");
}

extern "C" int dpvo_update_heads_finish(const float* acc, const float* bd, const float* bw, int64_t n_edges,
                                        float* delta, float* weight, void* stream) {
  DPVO_REQUIRE(n_edges >= 0, "update_heads_finish: bad sizes");
  if (n_edges == 0) return 0;
  DPVO_REQUIRE(acc && bd && bw && delta && weight, "update_heads_finish: null pointer");
  note_launches(1);
  heads_finish_kernel<<<grid_for(n_edges, 256, 8), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(acc), bd, bw, n_edges, reinterpret_cast<float2*>(delta),
      reinterpret_cast<float2*>(weight));
  return check_launch("update_heads_finish");
}
