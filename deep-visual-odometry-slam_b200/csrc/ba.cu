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

// ------------------------------------------------- compaction + edge grouping
// Pose-pair key of an edge for the dense path: (ix+1)*(N+1) + (jx+1) with out-of-window
// (fixed) poses mapped to 0.  Edges with the same key add into the same Hessian blocks.
__device__ __forceinline__ int pair_key(const BADev& p, int64_t fi, int64_t fj) {
  const int64_t ix = fi - p.t0, jx = fj - p.t0;
  const int a = (ix >= 0 && ix < p.N) ? (int)ix + 1 : 0;
  const int b = (jx >= 0 && jx < p.N) ? (int)jx + 1 : 0;
  return a * (p.N + 1) + b;
}

// Marks the patches referenced by kk (bitmap) and histograms the pose-pair keys
// (one atomic per distinct key per warp).
constexpr int kPrepSmemKeys = 2048;

__global__ void __launch_bounds__(256)
ba_prep_kernel(const __grid_constant__ BADev p) {
  __shared__ int s_hist[kPrepSmemKeys];
  const bool use_smem = p.K <= kPrepSmemKeys;
  if (use_smem) {
    for (int i = threadIdx.x; i < p.K; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
  }
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  for (int64_t e0 = (int64_t)blockIdx.x * blockDim.x; e0 < p.E; e0 += stride) {
    const int64_t e = e0 + threadIdx.x;
    const bool act = e < p.E;
    int key = -1;
    int word = -1;
    uint32_t bit = 0;
    if (act) {
      const int64_t k = p.kk[e];
      if (k >= 0 && k < p.npatches) { word = (int)(k >> 5); bit = 1u << (k & 31); }
      key = pair_key(p, p.ii[e], p.jj[e]);
      p.ekey[e] = key;
    }
    // one atomicOr per distinct bitmap word per warp, and only if it would change the word
    // (bits only ever get set: a stale read can only cause a redundant atomic)
    const unsigned wpeers = __match_any_sync(0xffffffffu, word);
    const uint32_t bits = __reduce_or_sync(wpeers, bit);
    if (word >= 0 && lane == __ffs(wpeers) - 1 && (__ldcg(&p.bitmap[word]) & bits) != bits)
      atomicOr(&p.bitmap[word], bits);
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    if (act && lane == __ffs(peers) - 1) atomicAdd(use_smem ? &s_hist[key] : &p.hist[key], __popc(peers));
  }
  if (use_smem) {
    __syncthreads();
    for (int i = threadIdx.x; i < p.K; i += blockDim.x) {
      const int c = s_hist[i];
      if (c) atomicAdd(&p.hist[i], c);
    }
  }
}

constexpr int kKeyScanDirect = 8192;   // up to this many keys ba_prefix_kernel scans the histogram itself
constexpr int kKeyScanBlock = 4096;    // keys per CTA of the blocked scan (1024 threads x 4)

// One CTA: (1) exclusive popcount prefix over the bitmap words, list of set bits (kx), total M;
// (2) exclusive scans of the key histogram: edge offsets `off` and warp-task offsets `toff`
// (a task = up to 32 edges of one key).
__global__ void __launch_bounds__(1024)
ba_prefix_kernel(const __grid_constant__ BADev p) {
  __shared__ int s_warp[32];
  __shared__ int s_warp2[32];
  __shared__ int s_carry, s_carry2;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t* bitmap = p.bitmap;
  if (tid == 0) { s_carry = 0; s_carry2 = 0; }
  __syncthreads();
  for (int64_t base = 0; base < p.W; base += 1024) {
    const int64_t w = base + tid;
    const uint32_t bits = (w < p.W) ? bitmap[w] : 0u;
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
    if (w < p.W) {
      p.prefix[w] = excl;
      uint32_t b = bits;
      int r = excl;
      while (b) {
        const int bit = __ffs(b) - 1;
        b &= b - 1;
        if (r < p.Mmax) p.kx[r] = w * 32 + bit;
        ++r;
      }
    }
    __syncthreads();
    if (tid == 0) s_carry += s_warp[31];
    __syncthreads();
  }
  if (tid == 0) p.meta[0] = (int)(s_carry < p.Mmax ? s_carry : p.Mmax);

  // ---- key histogram scans.  Many keys (global BA: K = (N+1)^2 ~ 10^6): the histogram itself was reduced per
  // 4096-key block by ba_keysum_kernel; only the block sums are scanned here, ba_keyscan_kernel finishes the job.
  const bool blocked = p.K > kKeyScanDirect;
  const int nscan = blocked ? (p.K + kKeyScanBlock - 1) / kKeyScanBlock : p.K;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < nscan; base += 1024) {
    const int k = base + tid;
    const int c = (k < nscan) ? (blocked ? p.ksum[2 * k] : p.hist[k]) : 0;
    const int t = blocked ? ((k < nscan) ? p.ksum[2 * k + 1] : 0) : (c + 31) >> 5;
    int inc = c, inc2 = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int a = __shfl_up_sync(0xffffffffu, inc, o);
      const int b = __shfl_up_sync(0xffffffffu, inc2, o);
      if (lane >= o) { inc += a; inc2 += b; }
    }
    if (lane == 31) { s_warp[wid] = inc; s_warp2[wid] = inc2; }
    __syncthreads();
    if (wid == 0) {
      int x = s_warp[lane], y = s_warp2[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int a = __shfl_up_sync(0xffffffffu, x, o);
        const int b = __shfl_up_sync(0xffffffffu, y, o);
        if (lane >= o) { x += a; y += b; }
      }
      s_warp[lane] = x; s_warp2[lane] = y;
    }
    __syncthreads();
    if (k < nscan) {
      const int e0 = s_carry + (wid ? s_warp[wid - 1] : 0) + inc - c;
      const int e1 = s_carry2 + (wid ? s_warp2[wid - 1] : 0) + inc2 - t;
      if (blocked) { p.ksum[2 * k] = e0; p.ksum[2 * k + 1] = e1; }
      else {
        p.off[k] = e0; p.toff[k] = e1;
        for (int i = 0; i < t; ++i) p.tkey[e1 + i] = k;     // task -> key, so the assembly does not search toff
      }
    }
    __syncthreads();
    if (tid == 0) { s_carry += s_warp[31]; s_carry2 += s_warp2[31]; }
    __syncthreads();
  }
  if (tid == 0) {
    p.off[p.K] = s_carry;
    p.toff[p.K] = s_carry2;
    p.meta[2] = s_carry2;  // number of warp tasks
  }
}

// Blocked form of the key scans (K > kKeyScanDirect): sums of (edge count, warp-task count) per block of 4096 keys ..
__global__ void __launch_bounds__(1024)
ba_keysum_kernel(const __grid_constant__ BADev p) {
  __shared__ int s_a[32], s_b[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int k0 = blockIdx.x * kKeyScanBlock + tid * 4;
  int a = 0, b = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = (k0 + i < p.K) ? p.hist[k0 + i] : 0;
    a += c;
    b += (c + 31) >> 5;
  }
  a = __reduce_add_sync(0xffffffffu, a);
  b = __reduce_add_sync(0xffffffffu, b);
  if (lane == 0) { s_a[wid] = a; s_b[wid] = b; }
  __syncthreads();
  if (wid == 0) {
    a = __reduce_add_sync(0xffffffffu, s_a[lane]);
    b = __reduce_add_sync(0xffffffffu, s_b[lane]);
    if (lane == 0) { p.ksum[2 * blockIdx.x] = a; p.ksum[2 * blockIdx.x + 1] = b; }
  }
}

// .. and, after ba_prefix_kernel turned the block sums into block offsets, the scan inside every block.
__global__ void __launch_bounds__(1024)
ba_keyscan_kernel(const __grid_constant__ BADev p) {
  __shared__ int s_warp[32], s_warp2[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int k0 = blockIdx.x * kKeyScanBlock + tid * 4;
  int c[4], t[4], a = 0, b = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    c[i] = (k0 + i < p.K) ? p.hist[k0 + i] : 0;
    t[i] = (c[i] + 31) >> 5;
    a += c[i];
    b += t[i];
  }
  int inc = a, inc2 = b;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int x = __shfl_up_sync(0xffffffffu, inc, o);
    const int y = __shfl_up_sync(0xffffffffu, inc2, o);
    if (lane >= o) { inc += x; inc2 += y; }
  }
  if (lane == 31) { s_warp[wid] = inc; s_warp2[wid] = inc2; }
  __syncthreads();
  if (wid == 0) {
    int x = s_warp[lane], y = s_warp2[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int x1 = __shfl_up_sync(0xffffffffu, x, o);
      const int y1 = __shfl_up_sync(0xffffffffu, y, o);
      if (lane >= o) { x += x1; y += y1; }
    }
    s_warp[lane] = x; s_warp2[lane] = y;
  }
  __syncthreads();
  int e0 = p.ksum[2 * blockIdx.x] + (wid ? s_warp[wid - 1] : 0) + inc - a;
  int e1 = p.ksum[2 * blockIdx.x + 1] + (wid ? s_warp2[wid - 1] : 0) + inc2 - b;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (k0 + i < p.K) {
      p.off[k0 + i] = e0; p.toff[k0 + i] = e1;
      for (int q = 0; q < t[i]; ++q) p.tkey[e1 + q] = k0 + i;
    }
    e0 += c[i];
    e1 += t[i];
  }
}

// perm[off[key] + rank] = e  (rank from a per-key cursor; one atomic per distinct key per warp)
__global__ void __launch_bounds__(256)
ba_scatter_kernel(const __grid_constant__ BADev p) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  for (int64_t e0 = (int64_t)blockIdx.x * blockDim.x; e0 < p.E; e0 += stride) {
    const int64_t e = e0 + threadIdx.x;
    const bool act = e < p.E;
    const int key = act ? p.ekey[e] : -1;
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const int leader = __ffs(peers) - 1;
    int base = 0;
    if (act && lane == leader) base = atomicAdd(&p.cursor[key], __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (act) p.perm[p.off[key] + base + __popc(peers & ((1u << lane) - 1u))] = (int)e;
  }
}

// --------------------------------------------------------------------- zeroing
__global__ void __launch_bounds__(256)
ba_zero_kernel(const __grid_constant__ BADev p) {
  const int M = p.meta[0];
  const int64_t n = p.n;
  const int64_t nfix = 2 * n * n + 2 * n;  // U, v, G, g are contiguous
  const int64_t total = nfix + 2 * (int64_t)M + (int64_t)M * n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    if (i < nfix) p.U[i] = 0.0;
    else if (i < nfix + M) p.C[i - nfix] = 0.f;
    else if (i < nfix + 2 * (int64_t)M) p.u[i - nfix - M] = 0.f;
    else p.Et[i - nfix - 2 * (int64_t)M] = 0.f;
  }
}

// -------------------------------------------------------------------- assembly
// One warp per task = up to 32 edges that share a pose-pair key.  Every lane evaluates one
// edge (residual, Ji/Jj/Jz, validity), forms its contribution to the pair's Hessian blocks
// (Hii, Hjj upper triangles, Hij, vi, vj: 90 values, both residual rows combined), and the
// warp reduces them with a shuffle reduce-scatter (93 shuffles for 96 values instead of 5 per
// value); each lane then adds 3 fully reduced values to U / v.  Patch-keyed terms (E^T rows,
// C, u) go to memory directly: their addresses are spread over M patches.
__constant__ unsigned char c_tri_a[21] = {0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 4, 4, 5};
__constant__ unsigned char c_tri_b[21] = {0, 1, 2, 3, 4, 5, 1, 2, 3, 4, 5, 2, 3, 4, 5, 3, 4, 5, 4, 5, 5};

template <int HALF>
__device__ __forceinline__ void reduce_scatter_step(float (&v)[96], bool hi, int off) {
#pragma unroll
  for (int i = 0; i < HALF; ++i) {
    const float keep = hi ? v[i + HALF] : v[i];
    const float send = hi ? v[i] : v[i + HALF];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
  }
}

constexpr int kAsmWarps = 4;

// The task count lives on the device (meta[2]); its host-side bound E / 32 + K is ~10^6 for a global BA (K = (N+1)^2
// pose-pair keys, nearly all of them empty), so the grid is capped and the warps stride over the tasks: 250,000 CTAs
// that only read the count and exit cost more than the assembly itself.
__device__ __forceinline__ void ba_assemble_task(const BADev& p, int task, int lane);

__global__ void __launch_bounds__(kAsmWarps * 32)
ba_assemble_kernel(const __grid_constant__ BADev p) {
  const int lane = threadIdx.x & 31;
  const int ntasks = p.meta[2];
  for (int task = blockIdx.x * kAsmWarps + (threadIdx.x >> 5); task < ntasks; task += gridDim.x * kAsmWarps)
    ba_assemble_task(p, task, lane);
}

__device__ __forceinline__ void ba_assemble_task(const BADev& p, int task, int lane) {
  // the task's pose-pair key: written next to the scan of the task offsets (a binary search over toff here was ~20
  // dependent loads per warp at the 10^6 keys of a global BA)
  const int key = p.tkey[task];
  const int chunk = task - p.toff[key];
  const int pos = p.off[key] + chunk * 32 + lane;
  const bool act = pos < p.off[key + 1];
  const int n = p.n;
  const int ix = key / (p.N + 1) - 1;   // -1: fixed pose
  const int jx = key % (p.N + 1) - 1;
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];

  float v[96];
#pragma unroll
  for (int i = 0; i < 96; ++i) v[i] = 0.f;

  EdgeTerms T;
  if (act && edge_terms(p, p.perm[pos], fx, fy, cx, cy, T)) {
    const int k = patch_rank(p, T.kf);
    float cc = 0.f, uu = 0.f;
    float ei[6] = {0, 0, 0, 0, 0, 0}, ej[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int row = 0; row < 2; ++row) {
      const float w = T.w[row], r = T.r[row], Jz = T.Jz[row];
      const float(&Ji)[6] = T.Ji[row];
      const float(&Jj)[6] = T.Jj[row];
      int idx = 0;
#pragma unroll
      for (int a = 0; a < 6; ++a) {
        const float wi = w * Ji[a], wj = w * Jj[a];
#pragma unroll
        for (int b = a; b < 6; ++b) {
          v[idx] += wi * Ji[b];
          v[21 + idx] += wj * Jj[b];
          ++idx;
        }
#pragma unroll
        for (int b = 0; b < 6; ++b) v[42 + a * 6 + b] += wi * Jj[b];
        v[78 + a] += -r * wi;
        v[84 + a] += r * wj;
        ei[a] += -Jz * wi;
        ej[a] += Jz * wj;
      }
      cc += w * Jz * Jz;
      uu += w * r * Jz;
    }
    if (p.Et) {
      float* et = p.Et + (int64_t)k * n;
      if (ix >= 0) {
#pragma unroll
        for (int a = 0; a < 6; ++a) atomicAdd(&et[6 * ix + a], ei[a]);
      }
      if (jx >= 0) {
#pragma unroll
        for (int a = 0; a < 6; ++a) atomicAdd(&et[6 * jx + a], ej[a]);
      }
    } else {
      // block-E path: the edge's two E entries (w.r.t. its source and target pose) stay per edge, stored in the
      // (source frame, slot) order the Schur kernel walks them in
      float4* ee = reinterpret_cast<float4*>(p.Ee + (int64_t)p.finv[p.perm[pos]] * 12);
      ee[0] = make_float4(ei[0], ei[1], ei[2], ei[3]);
      ee[1] = make_float4(ei[4], ei[5], ej[0], ej[1]);
      ee[2] = make_float4(ej[2], ej[3], ej[4], ej[5]);
    }
    atomicAdd(&p.C[k], cc);
    atomicAdd(&p.u[k], uu);
  } else if (act && !p.Et) {
    float4* ee = reinterpret_cast<float4*>(p.Ee + (int64_t)p.finv[p.perm[pos]] * 12);
    ee[0] = ee[1] = ee[2] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (ix < 0 && jx < 0) return;  // both poses fixed: no pose blocks (warp-uniform)

  reduce_scatter_step<48>(v, lane & 16, 16);
  reduce_scatter_step<24>(v, lane & 8, 8);
  reduce_scatter_step<12>(v, lane & 4, 4);
  reduce_scatter_step<6>(v, lane & 2, 2);
  reduce_scatter_step<3>(v, lane & 1, 1);
  const int base = ((lane >> 4) & 1) * 48 + ((lane >> 3) & 1) * 24 + ((lane >> 2) & 1) * 12 +
                   ((lane >> 1) & 1) * 6 + (lane & 1) * 3;
  const int oi = 6 * ix, oj = 6 * jx;
#pragma unroll
  for (int t = 0; t < 3; ++t) {
    const int idx = base + t;
    const float val = v[t];
    if (val == 0.f) continue;
    if (idx < 21) {
      if (ix >= 0) {
        const int a = c_tri_a[idx], b = c_tri_b[idx];
        atomicAdd(&p.U[(oi + a) * n + oi + b], (double)((a == b ? 0.5f : 1.0f) * val));
      }
    } else if (idx < 42) {
      if (jx >= 0) {
        const int a = c_tri_a[idx - 21], b = c_tri_b[idx - 21];
        atomicAdd(&p.U[(oj + a) * n + oj + b], (double)((a == b ? 0.5f : 1.0f) * val));
      }
    } else if (idx < 78) {
      if (ix >= 0 && jx >= 0) {
        const int a = (idx - 42) / 6, b = (idx - 42) % 6;
        atomicAdd(&p.U[(oi + a) * n + oj + b], (double)-val);
      }
    } else if (idx < 84) {
      if (ix >= 0) atomicAdd(&p.v[oi + idx - 78], (double)val);
    } else if (idx < 90) {
      if (jx >= 0) atomicAdd(&p.v[oj + idx - 84], (double)val);
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
  double acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
  double gacc = 0.0;

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
    // fp32 inside a chunk of kSchurKC patches, fp64 across chunks (and across CTAs, in the atomics)
    float part[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) part[i][j] = 0.f;
#pragma unroll 8
    for (int kk_ = 0; kk_ < kSchurKC; ++kk_) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { av[i] = As[kk_][ty * 4 + i]; bv[i] = Bs[kk_][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) part[i][j] += av[i] * bv[j];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] += (double)part[i][j];
    if (tb == 0 && threadIdx.x < 64) {
      float gp = 0.f;
      for (int kk_ = 0; kk_ < kSchurKC; ++kk_) gp += As[kk_][threadIdx.x] * us[kk_];
      gacc += (double)gp;
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int a = a0 + ty * 4 + i, b = b0 + tx * 4 + j;
      if (a < n && b < n && acc[i][j] != 0.0) atomicAdd(&p.G[(int64_t)a * n + b], acc[i][j]);
    }
  if (tb == 0 && threadIdx.x < 64 && a0 + threadIdx.x < n && gacc != 0.0) atomicAdd(&p.g[a0 + threadIdx.x], gacc);
}

// ------------------------------------------------------------- solve (registers)
// n + 1 <= 16*TS.  256 threads as a 16 x 16 grid; thread (ti, tk) keeps elements (i, k),
// i = ti + 16 r, k = tk + 16 c, of the augmented system [S ; y^T] in registers (TS x TS tile,
// cyclic distribution keeps the shrinking trailing matrix balanced).  Per column j its owners
// publish the unscaled column through a double-buffered shared-memory vector — one barrier per
// column — and every thread applies the rank-1 update to its tile.  The scaled columns
// (L, with z = L^-1 y as row n) go to shared memory for the back substitution L^T x = z,
// done column-oriented by one warp.  Then dX -> global, pose retraction (ba_cuda.cu:178-206);
// U, v, G, g are cleared for the next iteration.
template <int TS>
__global__ void __launch_bounds__(256)
ba_solve_reg_kernel(const __grid_constant__ BADev p) {
  extern __shared__ float sm[];
  const int n = p.n, N = p.N;
  const int ldL = (n + 2) | 1;
  float* Ls = sm;                    // n * ldL : Ls[j*ldL + i] = L[i][j], i in (j, n]
  float* col = Ls + n * ldL;         // 2 * 16*TS
  float* dinv = col + 2 * 16 * TS;   // n
  float* x = dinv + n;               // n
  __shared__ int s_fail;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int ti = tid >> 4, tk = tid & 15;
  if (tid == 0) s_fail = 0;

  float a[TS][TS];
#pragma unroll
  for (int r = 0; r < TS; ++r)
#pragma unroll
    for (int c = 0; c < TS; ++c) {
      const int i = ti + 16 * r, k = tk + 16 * c;
      float s = 0.f;
      if (i < n && k <= i) {
        s = (float)(p.U[i * n + k] + p.U[k * n + i] - 0.5 * (p.G[i * n + k] + p.G[k * n + i]));
        if (i == k) s += 1e-4f * s + 1.0f;
      } else if (i == n && k < n) {
        s = (float)(p.v[k] - p.g[k]);
      }
      a[r][c] = s;
    }
  __syncthreads();   // every element has been read: U, v, G, g (contiguous) can be cleared
  for (int i = tid; i < 2 * n * n + 2 * n; i += 256) p.U[i] = 0.0;

  bool fail = false;
#pragma unroll
  for (int jb = 0; jb < TS; ++jb) {
    for (int jj = 0; jj < 16; ++jj) {
      const int j = jb * 16 + jj;
      if (j >= n || fail) break;
      float* cb = col + (j & 1) * 16 * TS;
      if (tk == jj) {
#pragma unroll
        for (int r = 0; r < TS; ++r) cb[ti + 16 * r] = a[r][jb];
      }
      __syncthreads();
      const float d = cb[j];
      if (!(d > 0.f)) { fail = true; break; }   // uniform
      // 1/sqrt(d) by MUFU.RSQ + one Newton step (<= 1 ulp) instead of an IEEE sqrt and two divisions:
      // this chain sits on the critical path of every column
      float rs = rsqrtf(d);
      rs = rs * (1.5f - 0.5f * d * rs * rs);
      const float invd = rs * rs;
      float li[TS];
#pragma unroll
      for (int r = 0; r < TS; ++r) li[r] = cb[ti + 16 * r];
#pragma unroll
      for (int c = jb; c < TS; ++c) {
        const int k = tk + 16 * c;
        const float lk = (k > j) ? cb[k < 16 * TS ? k : 0] * invd : 0.f;
#pragma unroll
        for (int r = 0; r < TS; ++r) a[r][c] -= li[r] * lk;
      }
      if (tk == jj) {
#pragma unroll
        for (int r = 0; r < TS; ++r) {
          const int i = ti + 16 * r;
          if (i > j && i <= n) Ls[j * ldL + i] = li[r] * rs;
        }
        if (ti == 0) dinv[j] = rs;
      }
    }
  }
  if (fail && tid == 0) s_fail = 1;
  __syncthreads();
  if (s_fail) {
    for (int i = tid; i < n; i += 256) p.dX[i] = 0.f;
    if (tid == 0) p.meta[1] = 1;  // tells backsub to skip the depth update too
    return;
  }
  if (TS == 4) {
    // n < 64: L^T x = z by one warp with x in registers (lane l holds x[l] and x[l + 32]); the only chain per column
    // is shuffle -> multiply -> fused multiply-add, the L entries are independent loads the scheduler hoists
    __syncthreads();
    if (wid == 0) {
      const int i0 = lane, i1 = lane + 32;
      float x0 = i0 < n ? Ls[i0 * ldL + n] : 0.f, x1 = i1 < n ? Ls[i1 * ldL + n] : 0.f;   // z
      const float d0 = i0 < n ? dinv[i0] : 0.f, d1 = i1 < n ? dinv[i1] : 0.f;
      for (int j = n - 1; j >= 32; --j) {
        const float xj = __shfl_sync(0xffffffffu, x1 * d1, j - 32);
        if (i1 == j) x1 = xj;
        x0 -= Ls[i0 * ldL + j] * xj;                       // i0 < 32 <= j
        if (i1 < j) x1 -= Ls[i1 * ldL + j] * xj;
      }
      for (int j = (n < 32 ? n : 32) - 1; j >= 0; --j) {
        const float xj = __shfl_sync(0xffffffffu, x0 * d0, j);
        if (i0 == j) x0 = xj;
        if (i0 < j) x0 -= Ls[i0 * ldL + j] * xj;
      }
      if (i0 < n) x[i0] = x0;
      if (i1 < n) x[i1] = x1;
    }
    __syncthreads();
  } else {
    for (int i = tid; i < n; i += 256) x[i] = Ls[i * ldL + n];   // z
    __syncthreads();
    if (wid == 0) {
      for (int j = n - 1; j >= 0; --j) {
        const float xj = x[j] * dinv[j];
        __syncwarp();
        if (lane == 0) x[j] = xj;
        for (int i = lane; i < j; i += 32) x[i] -= Ls[i * ldL + j] * xj;
        __syncwarp();
      }
    }
    __syncthreads();
  }
  for (int i = tid; i < n; i += 256) p.dX[i] = x[i];
  if (tid == 0) p.meta[1] = 0;
  for (int i = tid; i < N; i += 256) {
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

// --------------------------------------------------------------------- backsub
__global__ void __launch_bounds__(256)
ba_backsub_kernel(const __grid_constant__ BADev p) {
  extern __shared__ float dXs[];
  const int n = p.n;
  const int M = p.meta[0];
  const bool failed = n > 0 && p.meta[1];  // failed factorisation: leave the depths alone, still clear
  for (int i = threadIdx.x; i < n; i += blockDim.x) dXs[i] = p.dX[i];
  __syncthreads();
  const float lm = p.lmbda[0];
  const int PP = p.P * p.P;
  const int lane = threadIdx.x & 31;
  const int nwarp = (gridDim.x * blockDim.x) >> 5;
  for (int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; k < M; k += nwarp) {
    float* er = p.Et + (int64_t)k * n;
    // everything the retraction needs is requested before the dot product, so the row of E^T, the
    // diagonal terms and the patch depth are in flight together instead of one after the other
    const float ck = p.C[k], uk = p.u[k];
    float* dp = p.patches + p.kx[k] * 3 * PP + 2 * PP;
    const float d0 = dp[0];
    float dot = 0.f;
    for (int a = lane; a < n; a += 32) {
      dot += er[a] * dXs[a];
      er[a] = 0.f;                       // clear this patch's accumulators for the next iteration
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0) {
      p.C[k] = 0.f;
      p.u[k] = 0.f;
    }
    if (!failed) {
      // d <- d + dz; d > 20 -> 1; d <- max(d, 1e-4); broadcast over the patch (depth_retract, one pixel per lane)
      float d = d0 + (uk - dot) * (1.0f / (ck + lm));
      d = (d > 20.f) ? 1.0f : d;
      d = fmaxf(d, 1e-4f);
      for (int i = lane; i < PP; i += 32) dp[i] = d;
    }
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

// S (lower triangle) and y row for the blocked solver; clears nothing (a memset follows).
__global__ void __launch_bounds__(256)
ba_build_s_kernel(const __grid_constant__ BADev p, float* __restrict__ A, int lda) {
  const int n = p.n;
  const int64_t total = (int64_t)(n + 1) * n;
  if (blockIdx.x == 0 && threadIdx.x == 0) p.meta[1] = 0;   // failure flag of THIS iteration's factorisation
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int a = (int)(idx / n), b = (int)(idx - (int64_t)a * n);
    if (a == n) { A[(int64_t)n * lda + b] = (float)(p.v[b] - p.g[b]); continue; }
    if (b > a) continue;
    float s = (float)(p.U[(int64_t)a * n + b] + p.U[(int64_t)b * n + a] - 0.5 * (p.G[(int64_t)a * n + b] + p.G[(int64_t)b * n + a]));
    if (a == b) s += 1e-4f * s + 1.0f;
    A[(int64_t)a * lda + b] = s;
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
  // bitmap, hist, cursor, meta are contiguous: cleared by one memset per call
  const int64_t K = ((int64_t)N + 1) * (N + 1);
  p.K = (int)K;
  p.bitmap = (uint32_t*)take(W * 4);
  p.hist = (int*)take(K * 4);
  p.cursor = (int*)take(K * 4);
  p.meta = (int*)take(64);
  p.prefix = (int*)take(W * 4);
  p.off = (int*)take((K + 1) * 4);
  p.toff = (int*)take((K + 1) * 4);
  p.tkey = (int*)take((E / 32 + K + 2) * 4);
  p.ksum = (int*)take((2 * ((K + kKeyScanBlock - 1) / kKeyScanBlock) + 2) * 4);
  p.perm = (int*)take(E * 4);
  p.ekey = (int*)take(E * 4);
  p.kx = (int64_t*)take(Mmax * 8);
  // U, v, G, g contiguous (zeroed as one range)
  char* fixed = take((2 * n * n + 2 * n) * 8);
  p.U = (double*)fixed;
  p.v = p.U + n * n;
  p.G = p.v + n;
  p.g = p.G + n * n;
  p.C = (float*)take(Mmax * 4);
  p.u = (float*)take(Mmax * 4);
  p.dX = (float*)take((n + 8) * 4);
  p.Sg = (float*)take(((n + 1) * (int64_t)chol_lda((int)n) + chol_scratch_floats((int)n) + 128) * 4);
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


void ba_launch_assemble(const BADev& p, cudaStream_t st) {
  const int64_t max_tasks = p.E / 32 + p.K + 1;
  const int64_t cap = (int64_t)kNumSMs * 32;       // 148 x 32 CTAs of 4 warps: several waves of resident CTAs
  const int agrid = (int)(ceil_div64(max_tasks, kAsmWarps) < cap ? ceil_div64(max_tasks, kAsmWarps) : cap);
  ba_assemble_kernel<<<agrid, kAsmWarps * 32, 0, st>>>(p);
  note_launches(1);
}

int ba_launch_solve(const BADev& p, cudaStream_t st) {
  const int n = p.n;
  const int ld = (n + 2) | 1;
  const size_t reg_smem = ((size_t)n * ld + 2 * 16 * 8 + 2 * (size_t)n) * sizeof(float);
  if (n + 1 <= 64) {
    ba_solve_reg_kernel<4><<<1, 256, reg_smem, st>>>(p);
    note_launches(1);
  } else if (n + 1 <= 128) {
    if (reg_smem > 48 * 1024)
      cudaFuncSetAttribute(ba_solve_reg_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)reg_smem);
    ba_solve_reg_kernel<8><<<1, 256, reg_smem, st>>>(p);
    note_launches(1);
  } else {
    // blocked multi-CTA Cholesky on a global copy of S; U, v, G, g are cleared for the next iteration
    const int lda = chol_lda(n);
    const int64_t tot = (int64_t)(n + 1) * n;
    const int sgrid = (int)(ceil_div64(tot, 256) < kNumSMs * 8 ? ceil_div64(tot, 256) : kNumSMs * 8);
    ba_build_s_kernel<<<sgrid, 256, 0, st>>>(p, p.Sg, lda);
    cudaMemsetAsync(p.U, 0, (size_t)(2 * (int64_t)n * n + 2 * n) * sizeof(double), st);
    note_launches(1);
    if (ba_chol_solve(p, p.Sg, lda, p.Sg + (int64_t)(n + 1) * lda, st)) return 1;
  }
  return check_launch("ba_solve");
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

// mode 0: everything; 1: only the part that depends on the edge list and the window (patch bitmap,
// compaction, per-patch edge lists, cleared accumulators); 2: the iterations, on a workspace mode 1 filled
static int ba_forward_impl(int mode, float* poses, float* patches, const float* intrinsics,
                           const float* target, const float* weight, const float* lmbda,
                           const int64_t* ii, const int64_t* jj, const int64_t* kk,
                           int64_t n_edges, int64_t n_frames_total, int64_t n_patches_total, int32_t P,
                           int32_t ppf, int32_t t0, int32_t t1, int32_t iterations, int32_t eff_impl,
                           void* workspace, int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(n_edges >= 0 && n_frames_total > 0 && n_patches_total > 0 && P > 0, "ba_forward: bad sizes");
  DPVO_REQUIRE(t1 >= t0 && t0 >= 0 && t1 <= n_frames_total, "ba_forward: need 0 <= t0 <= t1 <= n_frames (t0=%d t1=%d)", t0, t1);
  DPVO_REQUIRE(P * P <= 64, "ba_forward: patch too large");
  if (n_edges == 0 || iterations <= 0) return 0;
  DPVO_REQUIRE(ii && jj && kk && workspace, "ba_forward: null pointer");
  DPVO_REQUIRE(mode == 1 || (poses && patches && intrinsics && target && weight && lmbda), "ba_forward: null pointer");
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

  DPVO_REQUIRE(eff_impl || N <= 2048, "ba_forward: dense path supports at most 2048 free poses (use eff_impl)");
  // the dense float64 U and G accumulators are (6N)^2 each: 9.7 GB at the reference's BUFFER_SIZE = 4096 (config.py:6)
  DPVO_REQUIRE(N <= 4096, "ba_forward: at most 4096 free poses (the dense Schur accumulators are (6N)^2 float64)");
  const int egrid = (int)(ceil_div64(n_edges, 256) < kNumSMs * 8 ? ceil_div64(n_edges, 256) : kNumSMs * 8);
  if (mode != 2) {
    cudaMemsetAsync(p.bitmap, 0, reinterpret_cast<char*>(p.meta) + 64 - reinterpret_cast<char*>(p.bitmap), st);
    ba_prep_kernel<<<egrid, 256, 0, st>>>(p);
    const int kblocks = p.K > kKeyScanDirect ? (p.K + kKeyScanBlock - 1) / kKeyScanBlock : 0;
    if (kblocks) ba_keysum_kernel<<<kblocks, 1024, 0, st>>>(p);
    ba_prefix_kernel<<<1, 1024, 0, st>>>(p);
    if (kblocks) ba_keyscan_kernel<<<kblocks, 1024, 0, st>>>(p);
    if (check_launch("ba_prefix")) return 1;
    ba_scatter_kernel<<<egrid, 256, 0, st>>>(p);
    note_launches(kblocks ? 5 : 3);
  }
  if (eff_impl) {
    if (mode == 1) return check_launch("ba_prepare");
    return ba_eff_run(p, iterations, st);
  }
  const int n = p.n;
  const int64_t zero_total = 2 * (int64_t)n * n + 2 * n + 2 * p.Mmax + p.Mmax * n;
  const int zgrid = (int)(ceil_div64(zero_total, 256 * 8) < kNumSMs * 4 ? ceil_div64(zero_total, 256 * 8) : kNumSMs * 4);
  const int tiles = n > 0 ? (n + 63) / 64 : 1;
  int ksplit = (int)ceil_div64(p.Mmax, kSchurKC);
  const int kcap = (2 * kNumSMs) / (tiles * tiles) > 1 ? (2 * kNumSMs) / (tiles * tiles) : 1;
  if (ksplit > kcap) ksplit = kcap;
  const int bgrid = (int)(ceil_div64(p.Mmax, 8) < kNumSMs * 8 ? ceil_div64(p.Mmax, 8) : kNumSMs * 8);

  // accumulators are cleared once here; solve / backsub clear what they consumed for the next iteration
  if (mode != 2) {
    ba_zero_kernel<<<zgrid > 0 ? zgrid : 1, 256, 0, st>>>(p);
    note_launches(1);
  }
  if (mode == 1) return check_launch("ba_prepare");
  for (int it = 0; it < iterations; ++it) {
    ba_launch_assemble(p, st);
    if (N > 0) {
      ba_schur_kernel<<<dim3(ksplit, tiles * tiles), 256, 0, st>>>(p);
      note_launches(1);
      if (ba_launch_solve(p, st)) return 1;
    }
    ba_backsub_kernel<<<bgrid, 256, (size_t)(n + 1) * sizeof(float), st>>>(p);
  }
  note_launches(iterations);
  return check_launch("ba_forward");
}

extern "C" int dpvo_ba_forward(float* poses, float* patches, const float* intrinsics,
                               const float* target, const float* weight, const float* lmbda,
                               const int64_t* ii, const int64_t* jj, const int64_t* kk,
                               int64_t n_edges, int64_t n_frames_total, int64_t n_patches_total, int32_t P,
                               int32_t ppf, int32_t t0, int32_t t1, int32_t iterations, int32_t eff_impl,
                               void* workspace, int64_t workspace_bytes, void* stream) {
  return ba_forward_impl(0, poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, n_edges, n_frames_total,
                         n_patches_total, P, ppf, t0, t1, iterations, eff_impl, workspace, workspace_bytes, stream);
}

extern "C" int dpvo_ba_prepare(const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t n_edges,
                               int64_t n_frames_total, int64_t n_patches_total, int32_t P, int32_t ppf, int32_t t0,
                               int32_t t1, int32_t eff_impl, void* workspace, int64_t workspace_bytes, void* stream) {
  return ba_forward_impl(1, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, ii, jj, kk, n_edges, n_frames_total,
                         n_patches_total, P, ppf, t0, t1, 1, eff_impl, workspace, workspace_bytes, stream);
}

extern "C" int dpvo_ba_forward_prepared(float* poses, float* patches, const float* intrinsics,
                                        const float* target, const float* weight, const float* lmbda,
                                        const int64_t* ii, const int64_t* jj, const int64_t* kk,
                                        int64_t n_edges, int64_t n_frames_total, int64_t n_patches_total, int32_t P,
                                        int32_t ppf, int32_t t0, int32_t t1, int32_t iterations, int32_t eff_impl,
                                        void* workspace, int64_t workspace_bytes, void* stream) {
  return ba_forward_impl(2, poses, patches, intrinsics, target, weight, lmbda, ii, jj, kk, n_edges, n_frames_total,
                         n_patches_total, P, ppf, t0, t1, iterations, eff_impl, workspace, workspace_bytes, stream);
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
