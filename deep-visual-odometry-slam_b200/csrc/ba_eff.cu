// Block-E ("efficient E") Schur path for global bundle adjustment (eff_impl=True).
//
// Replaces /root/reference/methods/dpvo/dpvo/fastba/block_e.cu:38-145 (EfficentE constructor: host
// round trips, CPU loops, unordered_set, index tensors rebuilt and uploaded on every call) and
// :147-300 (EEt / Ev / Etv kernels: one thread per (frame-pair-pair, patch) with 36 global atomics
// each, ~2.3e9 atomics at 1,000 keyframes), used from ba_cuda.cu:475-478,537-551.
//
// Same mathematics as the dense path (S = B - E Q E^T, y = v - E Q u, dZ = Q (u - E^T dX)); only
// the representation of E differs.  Here E never exists as a matrix or as a per-frame-pair table:
//   * the assembly kernel leaves every edge's two 6-vector entries (w.r.t. its source and its
//     target pose) in Ee[e][12] — no atomics, no zeroing, sized by E alone, so no host
//     synchronisation is needed to size anything;
//   * edges are counting-sorted by source frame once per call (device side);
//   * Schur product: one CTA per source frame i.  All patches of frame i connect to the same few
//     target frames conn(i), so G restricted to conn(i) x conn(i) is a small GEMM
//     A_i diag(q) A_i^T with A_i = [6*|conn(i)|, ppf] gathered into shared memory from Ee; each
//     6x6 block is reduced over the frame's patches in registers and leaves the CTA as 36 atomics
//     (~100x fewer global atomics than the reference).  conn(i) is tiled so that frames with
//     hundreds of loop-closure partners still fit;
//   * back substitution: per-edge dot products with dX, one atomic per edge.
// Like the reference (block_e.cu:57-62, ba_cuda.cu:352-353) the slot of a patch inside its frame is
// kk % ppf; unlike it, the patch's Q / u / C entry is always the one of kk itself.
#include "se3.cuh"
#include "ba_internal.cuh"
#include "scan.cuh"

namespace dpvo {

static int64_t align256e(int64_t x) { return (x + 255) & ~(int64_t)255; }

// Edges are counting-sorted by (source frame, slot of the patch in its frame) once per call; key nframes * ppf
// collects the edges the Schur product ignores (frame or patch id out of range).
int64_t ba_eff_carve(BADev& p, char* base, int64_t E, int64_t npatches, int N, int ppf, int64_t nframes) {
  (void)N;
  const int64_t Mmax = E < npatches ? E : npatches;
  const int64_t nkeys = nframes * ppf + 1;
  int64_t off = 0;
  auto take = [&](int64_t bytes) { int64_t o = off; off += align256e(bytes); return base ? base + o : (char*)nullptr; };
  // fhist and fcursor are contiguous (one memset)
  p.fhist = (int*)take(nkeys * 4);
  p.fcursor = (int*)take(nkeys * 4);
  p.foff = (int*)take((nkeys + 1) * 4);
  p.fscan = (int*)take(scan_workspace_ints(nkeys) * 4);
  p.fperm = (int*)take(E * 4);
  p.finv = (int*)take(E * 4);
  p.ejx = (int*)take(E * 4);
  p.Ee = (float*)take(E * 12 * 4);
  p.w = (float*)take(Mmax * 4);
  return off;
}

namespace {

// --------------------------------------------------- counting sort by (source frame, slot)
__device__ __forceinline__ int eff_key(const BADev& p, int64_t e) {
  const int64_t fi = p.ii[e], kf = p.kk[e];
  if (fi < 0 || fi >= p.nframes || kf < 0 || kf >= p.npatches) return (int)(p.nframes * p.ppf);
  return (int)(fi * p.ppf + kf % p.ppf);
}

__global__ void __launch_bounds__(256)
eff_fhist_kernel(const __grid_constant__ BADev p) {
  const int lane = threadIdx.x & 31;
  for (int64_t e0 = (int64_t)blockIdx.x * blockDim.x; e0 < p.E; e0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = e0 + threadIdx.x;
    const int key = (e < p.E) ? eff_key(p, e) : -1;
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    if (key >= 0 && lane == __ffs(peers) - 1) atomicAdd(&p.fhist[key], __popc(peers));
  }
}

// sorted position of every edge (fperm / finv) and, in sorted order, its target pose relative to t0 (-1: fixed)
__global__ void __launch_bounds__(256)
eff_fscatter_kernel(const __grid_constant__ BADev p) {
  const int lane = threadIdx.x & 31;
  for (int64_t e0 = (int64_t)blockIdx.x * blockDim.x; e0 < p.E; e0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = e0 + threadIdx.x;
    const int key = (e < p.E) ? eff_key(p, e) : -1;
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const int leader = __ffs(peers) - 1;
    int base = 0;
    if (key >= 0 && lane == leader) base = atomicAdd(&p.fcursor[key], __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (key >= 0) {
      const int dst = p.foff[key] + base + __popc(peers & ((1u << lane) - 1u));
      p.fperm[dst] = (int)e;
      p.finv[e] = dst;
      const int64_t jx = p.jj[e] - p.t0;
      p.ejx[dst] = (jx >= 0 && jx < p.N) ? (int)jx : -1;
    }
  }
}

// --------------------------------------------------------------- Schur product
// One CTA per source frame.  Shared memory: cidx[N] (pose -> compact id in conn(i)), conn[N],
// q[ppf], uq[ppf], poff[ppf + 1] (the frame's per-slot edge ranges), two tiles A1/A2 [ppf][TC*6] of E entries
// (target-pose compact id x 6).
struct EffSmem {
  unsigned short* cidx;
  unsigned short* conn;
  float* q;
  float* uq;
  int* poff;
  float* A1;
  float* A2;
};

__host__ __device__ __forceinline__ int eff_pad4(int x) { return (x + 3) & ~3; }
__host__ __device__ __forceinline__ size_t eff_fixed_smem(int N, int ppf) {
  const int Npad = (N + 7) & ~7;
  return (size_t)Npad * 2 * sizeof(unsigned short) + 2 * (size_t)eff_pad4(ppf) * sizeof(float) +
         (size_t)eff_pad4(ppf + 1) * sizeof(int);
}

__device__ __forceinline__ EffSmem eff_carve_smem(char* sm, int N, int ppf, int TC) {
  EffSmem s;
  const int Npad = (N + 7) & ~7;
  s.cidx = reinterpret_cast<unsigned short*>(sm);
  s.conn = s.cidx + Npad;
  s.q = reinterpret_cast<float*>(s.conn + Npad);
  s.uq = s.q + eff_pad4(ppf);
  s.poff = reinterpret_cast<int*>(s.uq + eff_pad4(ppf));
  s.A1 = reinterpret_cast<float*>(s.poff + eff_pad4(ppf + 1));   // 16-byte aligned: the products read the tiles as float4
  s.A2 = s.A1 + (size_t)ppf * TC * 6;
  return s;
}

// Gather the E entries of the frame's edges whose pose falls into conn tile `t` into A[slot][c*6+a].  No atomics: a
// warp owns a patch slot (slots w, w + 8, ..), the slot's edges are contiguous in the sorted order, so the lanes read
// ejx / Ee coalesced with no dependent index chain, every (slot, target pose) entry has one writer — two edges with
// the same target in one 32-edge batch (duplicate edges are legal; the reference sums them) are found with match.any
// and take turns — and the self-pose column (every edge of the slot contributes) is a shuffle reduction.  (The first
// version accumulated into shared memory with atomicAdd(float): a compare-and-swap loop per value, 56-fold contended
// on the self column — 45 % of the kernel in ncu.)
__device__ __forceinline__ void eff_fill_tile(const BADev& p, const EffSmem& s, float* A, int t, int TC, int ix_self) {
  constexpr unsigned FULL = 0xffffffffu;
  const int ppf = p.ppf;
  const int row = TC * 6;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int lo = t * TC, hi = lo + TC;
  const int ci = (ix_self >= 0) ? (int)s.cidx[ix_self] : 0xFFFF;
  const bool self_here = ci >= lo && ci < hi;
  for (int slot = warp; slot < ppf; slot += nwarps) {
    float* arow = A + slot * row;
    for (int i = lane; i < row / 4; i += 32) reinterpret_cast<float4*>(arow)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    const int b = s.poff[slot], e = s.poff[slot + 1];
    float se[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // software pipeline: the next batch's loads are in flight while this one is scattered
    int pos = b + lane;
    int jx = -1;
    float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0, v2 = v0;
    if (pos < e) {
      jx = p.ejx[pos];
      const float4* ee = reinterpret_cast<const float4*>(p.Ee + (int64_t)pos * 12);
      v0 = ee[0]; v1 = ee[1]; v2 = ee[2];
    }
    for (int base = b; base < e; base += 32) {
      const bool act = base + lane < e;
      const int cjx = jx;
      const float4 c0 = v0, c1 = v1, c2 = v2;
      pos = base + 32 + lane;
      if (pos < e) {
        jx = p.ejx[pos];
        const float4* ee = reinterpret_cast<const float4*>(p.Ee + (int64_t)pos * 12);
        v0 = ee[0]; v1 = ee[1]; v2 = ee[2];
      }
      if (act && self_here) {
        se[0] += c0.x; se[1] += c0.y; se[2] += c0.z; se[3] += c0.w; se[4] += c1.x; se[5] += c1.y;
      }
      const int cj = (act && cjx >= 0) ? (int)s.cidx[cjx] : 0xFFFF;
      const bool here = cj >= lo && cj < hi;
      const unsigned peers = __match_any_sync(FULL, here ? cj : -1 - lane);
      const int rank = __popc(peers & ((1u << lane) - 1u));
      const int rounds = __reduce_max_sync(FULL, here ? rank : 0);
      float* dst = arow + (here ? (cj - lo) * 6 : 0);
      for (int r = 0; r <= rounds; ++r) {
        if (here && rank == r) {
          dst[0] += c1.z; dst[1] += c1.w; dst[2] += c2.x; dst[3] += c2.y; dst[4] += c2.z; dst[5] += c2.w;
        }
        __syncwarp();
      }
    }
    if (self_here) {
#pragma unroll
      for (int a = 0; a < 6; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) se[a] += __shfl_xor_sync(FULL, se[a], o);
      }
      if (lane < 6) {
        float v = se[0];
#pragma unroll
        for (int a = 1; a < 6; ++a) v = (lane == a) ? se[a] : v;
        arow[(ci - lo) * 6 + lane] += v;          // "+=": an edge whose target is its own source frame lands here too
      }
      __syncwarp();
    }
  }
  __syncthreads();
}

// G is consumed as 0.5 (G[a][b] + G[b][a]) (ba_build_s_kernel, ba_solve_reg_kernel): a 6x6 block between two
// DIFFERENT poses is therefore written once, doubled, at its upper position (row pose < column pose) and its mirror
// is left at zero — half the float64 reductions of writing both; diagonal blocks are written in full.
__global__ void __launch_bounds__(256, 1)
eff_schur_kernel(const __grid_constant__ BADev p, int TC) {
  extern __shared__ __align__(16) char eff_sm[];
  __shared__ int s_cnt[256];
  __shared__ int s_nconn;
  const int fi = blockIdx.x;
  const int N = p.N, n = p.n, ppf = p.ppf, tid = threadIdx.x;
  const int e_begin = p.foff[fi * ppf], e_end = p.foff[(fi + 1) * ppf];
  if (e_begin == e_end) return;
  const EffSmem s = eff_carve_smem(eff_sm, N, ppf, TC);
  const int ix_self = (fi - p.t0 >= 0 && fi - p.t0 < N) ? fi - p.t0 : -1;
  const float lm = p.lmbda[0];

  // ---- conn(i): free poses this frame's patches have E entries for; per slot: edge range, Q and Q u of its patch
  for (int j = tid; j < N; j += 256) s.cidx[j] = 0;
  for (int k = tid; k <= ppf; k += 256) {
    const int b = p.foff[fi * ppf + k];
    s.poff[k] = b;
    if (k < ppf) {
      float q = 0.f, uq = 0.f;
      if (b < p.foff[fi * ppf + k + 1]) {
        const int r = patch_rank(p, p.kk[p.fperm[b]]);
        q = 1.0f / (p.C[r] + lm);
        uq = p.u[r] * q;
      }
      s.q[k] = q;
      s.uq[k] = uq;
    }
  }
  __syncthreads();
  if (tid == 0 && ix_self >= 0) s.cidx[ix_self] = 1;
  for (int pos = e_begin + tid; pos < e_end; pos += 256) {
    const int jx = p.ejx[pos];
    if (jx >= 0) s.cidx[jx] = 1;
  }
  __syncthreads();
  // compact ids in ascending pose order (deterministic): chunked count + block scan
  const int chunk = (N + 255) / 256;
  const int j0 = tid * chunk, j1 = min(N, j0 + chunk);
  int cnt = 0;
  for (int j = j0; j < j1; ++j) cnt += s.cidx[j];
  s_cnt[tid] = cnt;
  __syncthreads();
  if (tid < 32) {
    int run = 0;
    for (int b = 0; b < 8; ++b) {          // 8 x 32 = 256 counters
      const int v = s_cnt[b * 32 + tid];
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (tid >= o) inc += t;
      }
      s_cnt[b * 32 + tid] = run + inc - v;
      run += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (tid == 0) s_nconn = run;
  }
  __syncthreads();
  {
    int id = s_cnt[tid];
    for (int j = j0; j < j1; ++j) {
      if (s.cidx[j]) { s.cidx[j] = (unsigned short)id; s.conn[id] = (unsigned short)j; ++id; }
      else s.cidx[j] = 0xFFFF;
    }
  }
  __syncthreads();
  const int nconn = s_nconn;
  if (nconn == 0) return;
  const int ntiles = (nconn + TC - 1) / TC;
  const int row = TC * 6;
  const int H = TC >> 1;                   // pose PAIRS per tile: a thread owns a 2 x 2 group of 6x6 blocks

  for (int t1 = 0; t1 < ntiles; ++t1) {
    eff_fill_tile(p, s, s.A1, t1, TC, ix_self);
    const int n1 = min(TC, nconn - t1 * TC);
    // g = E Q u restricted to this tile
    if (tid < n1 * 6) {
      float acc = 0.f;
      for (int k = 0; k < ppf; ++k) acc += s.A1[k * row + tid] * s.uq[k];
      if (acc != 0.f) atomicAdd(&p.g[6 * (int)s.conn[t1 * TC + tid / 6] + tid % 6], (double)acc);
    }
    for (int t2 = t1; t2 < ntiles; ++t2) {
      const float* A2 = s.A1;
      if (t2 != t1) { eff_fill_tile(p, s, s.A2, t2, TC, ix_self); A2 = s.A2; }
      const int n2 = min(TC, nconn - t2 * TC);
      const bool diag = t2 == t1;
      // 12 x 12 outputs per thread: per patch 6 LDS.128 feed 144 FMAs (one 6x6 block per thread and scalar reads made
      // this loop shared-memory bound: 12 LDS for 36 FMAs)
      for (int pr = tid; pr < H * H; pr += 256) {
        const int i2 = pr / H, j2 = pr - i2 * H;
        if (2 * i2 >= n1 || 2 * j2 >= n2 || (diag && j2 < i2)) continue;
        float acc[12][12];
#pragma unroll
        for (int a = 0; a < 12; ++a)
#pragma unroll
          for (int b = 0; b < 12; ++b) acc[a][b] = 0.f;
        const float4* r1 = reinterpret_cast<const float4*>(s.A1 + i2 * 12);
        const float4* r2 = reinterpret_cast<const float4*>(A2 + j2 * 12);
        const int row4 = row >> 2;
#pragma unroll 2
        for (int k = 0; k < ppf; ++k) {
          const float qk = s.q[k];
          const float4 x0 = r1[k * row4], x1 = r1[k * row4 + 1], x2 = r1[k * row4 + 2];
          const float4 y0 = r2[k * row4], y1 = r2[k * row4 + 1], y2 = r2[k * row4 + 2];
          const float x[12] = {x0.x * qk, x0.y * qk, x0.z * qk, x0.w * qk, x1.x * qk, x1.y * qk,
                               x1.z * qk, x1.w * qk, x2.x * qk, x2.y * qk, x2.z * qk, x2.w * qk};
          const float y[12] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w, y2.x, y2.y, y2.z, y2.w};
#pragma unroll
          for (int a = 0; a < 12; ++a)
#pragma unroll
            for (int b = 0; b < 12; ++b) acc[a][b] = fmaf(x[a], y[b], acc[a][b]);
        }
#pragma unroll
        for (int h1 = 0; h1 < 2; ++h1)
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            const int b1 = 2 * i2 + h1, b2 = 2 * j2 + h2;
            if (b1 >= n1 || b2 >= n2 || (diag && b2 < b1)) continue;
            const int ja = 6 * (int)s.conn[t1 * TC + b1], jb = 6 * (int)s.conn[t2 * TC + b2];
            const double scale = (diag && b1 == b2) ? 1.0 : 2.0;     // conn is ascending: ja < jb off the diagonal
#pragma unroll
            for (int a = 0; a < 6; ++a)
#pragma unroll
              for (int b = 0; b < 6; ++b) {
                const float v = acc[6 * h1 + a][6 * h2 + b];
                if (v != 0.f) atomicAdd(&p.G[(int64_t)(ja + a) * n + jb + b], scale * (double)v);
              }
          }
      }
      __syncthreads();   // A2 (or A1's readers) done before the next fill
    }
  }
}

// ------------------------------------------------------------ back substitution
// w[k] += ei . dX[ix] + ej . dX[jx] per edge  (E^T dX, ba_cuda.cu:549 / block_e.cu:255-281)
__global__ void __launch_bounds__(256)
eff_etdx_kernel(const __grid_constant__ BADev p) {
  if (p.meta[1]) return;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < p.E; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t fi = p.ii[e], fj = p.jj[e], kf = p.kk[e];
    if (fi < 0 || fi >= p.nframes || fj < 0 || fj >= p.nframes || kf < 0 || kf >= p.npatches) continue;
    const float4* ee = reinterpret_cast<const float4*>(p.Ee + (int64_t)p.finv[e] * 12);   // Ee is in sorted order
    const float4 a = ee[0], b = ee[1], c = ee[2];
    const int64_t ix = fi - p.t0, jx = fj - p.t0;
    float dot = 0.f;
    if (ix >= 0 && ix < p.N) {
      const float* d = p.dX + 6 * ix;
      dot += a.x * d[0] + a.y * d[1] + a.z * d[2] + a.w * d[3] + b.x * d[4] + b.y * d[5];
    }
    if (jx >= 0 && jx < p.N) {
      const float* d = p.dX + 6 * jx;
      dot += b.z * d[0] + b.w * d[1] + c.x * d[2] + c.y * d[3] + c.z * d[4] + c.w * d[5];
    }
    if (dot != 0.f) atomicAdd(&p.w[patch_rank(p, kf)], dot);
  }
}

// dZ = Q (u - E^T dX), depth retraction (ba_cuda.cu:209-229); clears C, u, w for the next iteration
__global__ void __launch_bounds__(256)
eff_depth_kernel(const __grid_constant__ BADev p) {
  const int M = p.meta[0];
  const bool failed = p.n > 0 && p.meta[1];
  const float lm = p.lmbda[0];
  const int PP = p.P * p.P;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < M; k += gridDim.x * blockDim.x) {
    const float q = 1.0f / (p.C[k] + lm);
    const float dz = q * (p.u[k] - p.w[k]);
    p.C[k] = 0.f; p.u[k] = 0.f; p.w[k] = 0.f;
    if (!failed) depth_retract(p.patches, p.kx[k], PP, dz);
  }
}

}  // namespace

int ba_eff_run(BADev& p, int iterations, cudaStream_t st) {
  const int N = p.N, n = p.n, ppf = p.ppf;
  const int64_t E = p.E;
  const int egrid = (int)(ceil_div64(E, 256) < kNumSMs * 8 ? ceil_div64(E, 256) : kNumSMs * 8);
  // edges grouped by source frame (once per call)
  const int64_t nkeys = p.nframes * ppf + 1;
  DPVO_REQUIRE(nkeys < ((int64_t)1 << 30), "ba_forward(eff_impl): frames x patches-per-frame too large");
  cudaMemsetAsync(p.fhist, 0, reinterpret_cast<char*>(p.foff) - reinterpret_cast<char*>(p.fhist), st);
  eff_fhist_kernel<<<egrid, 256, 0, st>>>(p);
  if (exclusive_scan_int(p.fhist, p.foff, nkeys, p.fscan, p.foff + nkeys, st)) return 1;
  eff_fscatter_kernel<<<egrid, 256, 0, st>>>(p);
  note_launches(5);

  // Schur kernel shared memory: tables + two conn tiles of TC poses
  const size_t fixed = eff_fixed_smem(N, ppf);
  const size_t budget = 200 * 1024;
  int TC = 32;                               // even: a thread of the Schur product owns pairs of poses
  while (TC > 2 && fixed + 2 * (size_t)ppf * TC * 6 * sizeof(float) > budget) TC >>= 1;
  const size_t schur_smem = fixed + 2 * (size_t)ppf * TC * 6 * sizeof(float);
  DPVO_REQUIRE(schur_smem <= budget, "ba_forward(eff_impl): patches-per-frame %d too large for the Schur tiles", ppf);
  if (N > 0) cudaFuncSetAttribute(eff_schur_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)schur_smem);

  // accumulators: cleared once; solve / depth kernels clear what they consumed
  if (n > 0) cudaMemsetAsync(p.U, 0, (size_t)(2 * (int64_t)n * n + 2 * n) * sizeof(double), st);
  cudaMemsetAsync(p.C, 0, (size_t)p.Mmax * sizeof(float), st);
  cudaMemsetAsync(p.u, 0, (size_t)p.Mmax * sizeof(float), st);
  cudaMemsetAsync(p.w, 0, (size_t)p.Mmax * sizeof(float), st);
  const int mgrid = (int)(ceil_div64(p.Mmax, 256) < kNumSMs * 8 ? ceil_div64(p.Mmax, 256) : kNumSMs * 8);
  for (int it = 0; it < iterations; ++it) {
    ba_launch_assemble(p, st);
    if (N > 0) {
      eff_schur_kernel<<<(unsigned)p.nframes, 256, schur_smem, st>>>(p, TC);
      note_launches(1);
      if (ba_launch_solve(p, st)) return 1;
      eff_etdx_kernel<<<egrid, 256, 0, st>>>(p);
      note_launches(1);
    }
    eff_depth_kernel<<<mgrid > 0 ? mgrid : 1, 256, 0, st>>>(p);
    note_launches(1);
  }
  return check_launch("ba_forward(eff_impl)");
}

}  // namespace dpvo
