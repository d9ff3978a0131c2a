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

namespace dpvo {

static int64_t align256e(int64_t x) { return (x + 255) & ~(int64_t)255; }

int64_t ba_eff_carve(BADev& p, char* base, int64_t E, int64_t npatches, int N, int ppf, int64_t nframes) {
  (void)N; (void)ppf;
  const int64_t Mmax = E < npatches ? E : npatches;
  int64_t off = 0;
  auto take = [&](int64_t bytes) { int64_t o = off; off += align256e(bytes); return base ? base + o : (char*)nullptr; };
  // fhist and fcursor are contiguous (one memset)
  p.fhist = (int*)take((nframes + 1) * 4);
  p.fcursor = (int*)take((nframes + 1) * 4);
  p.foff = (int*)take((nframes + 2) * 4);
  p.fperm = (int*)take(E * 4);
  p.Ee = (float*)take(E * 12 * 4);
  p.w = (float*)take(Mmax * 4);
  return off;
}

namespace {

// --------------------------------------------------- counting sort by source frame
__global__ void __launch_bounds__(256)
eff_fhist_kernel(const __grid_constant__ BADev p) {
  const int lane = threadIdx.x & 31;
  for (int64_t e0 = (int64_t)blockIdx.x * blockDim.x; e0 < p.E; e0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = e0 + threadIdx.x;
    int key = -1;
    if (e < p.E) {
      const int64_t fi = p.ii[e];
      key = (fi >= 0 && fi < p.nframes) ? (int)fi : (int)p.nframes;   // last bucket: ignored edges
    }
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    if (key >= 0 && lane == __ffs(peers) - 1) atomicAdd(&p.fhist[key], __popc(peers));
  }
}

// exclusive scan of fhist[0..nframes] -> foff[0..nframes+1]   (one CTA)
__global__ void __launch_bounds__(1024)
eff_fscan_kernel(const __grid_constant__ BADev p) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int K = (int)p.nframes + 1;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < K; base += 1024) {
    const int k = base + tid;
    const int c = (k < K) ? p.fhist[k] : 0;
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
    if (k < K) p.foff[k] = s_carry + (wid ? s_warp[wid - 1] : 0) + inc - c;
    __syncthreads();
    if (tid == 0) s_carry += s_warp[31];
    __syncthreads();
  }
  if (tid == 0) p.foff[K] = s_carry;
}

__global__ void __launch_bounds__(256)
eff_fscatter_kernel(const __grid_constant__ BADev p) {
  const int lane = threadIdx.x & 31;
  for (int64_t e0 = (int64_t)blockIdx.x * blockDim.x; e0 < p.E; e0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = e0 + threadIdx.x;
    int key = -1;
    if (e < p.E) {
      const int64_t fi = p.ii[e];
      key = (fi >= 0 && fi < p.nframes) ? (int)fi : (int)p.nframes;
    }
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const int leader = __ffs(peers) - 1;
    int base = 0;
    if (key >= 0 && lane == leader) base = atomicAdd(&p.fcursor[key], __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (key >= 0) p.fperm[p.foff[key] + base + __popc(peers & ((1u << lane) - 1u))] = (int)e;
  }
}

// --------------------------------------------------------------- Schur product
// One CTA per source frame.  Shared memory: cidx[N] (pose -> compact id in conn(i)), conn[N],
// q[ppf], uq[ppf], two tiles A1/A2 [ppf][TC*6] of E entries (target-pose compact id x 6).
struct EffSmem {
  unsigned short* cidx;
  unsigned short* conn;
  float* q;
  float* uq;
  float* A1;
  float* A2;
};

__device__ __forceinline__ EffSmem eff_carve_smem(char* sm, int N, int ppf, int TC) {
  EffSmem s;
  const int Npad = (N + 7) & ~7;
  s.cidx = reinterpret_cast<unsigned short*>(sm);
  s.conn = s.cidx + Npad;
  s.q = reinterpret_cast<float*>(s.conn + Npad);
  s.uq = s.q + ppf;
  s.A1 = s.uq + ppf;
  s.A2 = s.A1 + (size_t)ppf * TC * 6;
  return s;
}

// gather the E entries of the frame's edges whose pose falls into conn tile `t` into A[slot][c*6+a]
__device__ __forceinline__ void eff_fill_tile(const BADev& p, const EffSmem& s, float* A, int t, int TC,
                                              int e_begin, int e_end, int ix_self) {
  const int ppf = p.ppf;
  const int row = TC * 6;
  for (int i = threadIdx.x; i < ppf * row; i += blockDim.x) A[i] = 0.f;
  __syncthreads();
  const int lo = t * TC, hi = lo + TC;
  const int ci = (ix_self >= 0) ? (int)s.cidx[ix_self] : 0xFFFF;
  for (int pos = e_begin + threadIdx.x; pos < e_end; pos += blockDim.x) {
    const int e = p.fperm[pos];
    const int64_t kf = p.kk[e];
    if (kf < 0 || kf >= p.npatches) continue;
    const int slot = (int)(kf % ppf);
    const float* ee = p.Ee + (int64_t)e * 12;
    if (ci >= lo && ci < hi) {
      float* dst = A + slot * row + (ci - lo) * 6;
#pragma unroll
      for (int a = 0; a < 6; ++a) { const float v = ee[a]; if (v != 0.f) atomicAdd(&dst[a], v); }
    }
    const int64_t jx = p.jj[e] - p.t0;
    if (jx >= 0 && jx < p.N) {
      const int cj = s.cidx[jx];
      if (cj >= lo && cj < hi) {
        float* dst = A + slot * row + (cj - lo) * 6;
#pragma unroll
        for (int a = 0; a < 6; ++a) { const float v = ee[6 + a]; if (v != 0.f) atomicAdd(&dst[a], v); }
      }
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256)
eff_schur_kernel(const __grid_constant__ BADev p, int TC) {
  extern __shared__ __align__(16) char eff_sm[];
  __shared__ int s_cnt[256];
  __shared__ int s_nconn;
  const int fi = blockIdx.x;
  const int e_begin = p.foff[fi], e_end = p.foff[fi + 1];
  if (e_begin == e_end) return;
  const int N = p.N, n = p.n, ppf = p.ppf, tid = threadIdx.x;
  const EffSmem s = eff_carve_smem(eff_sm, N, ppf, TC);
  const int ix_self = (fi - p.t0 >= 0 && fi - p.t0 < N) ? fi - p.t0 : -1;
  const float lm = p.lmbda[0];

  // ---- conn(i): free poses this frame's patches have E entries for
  for (int j = tid; j < N; j += 256) s.cidx[j] = 0;
  for (int k = tid; k < ppf; k += 256) { s.q[k] = 0.f; s.uq[k] = 0.f; }
  __syncthreads();
  if (tid == 0 && ix_self >= 0) s.cidx[ix_self] = 1;
  for (int pos = e_begin + tid; pos < e_end; pos += 256) {
    const int e = p.fperm[pos];
    const int64_t jx = p.jj[e] - p.t0;
    if (jx >= 0 && jx < N) s.cidx[jx] = 1;
    const int64_t kf = p.kk[e];
    if (kf >= 0 && kf < p.npatches) {
      const int k = patch_rank(p, kf);
      const float q = 1.0f / (p.C[k] + lm);
      const int slot = (int)(kf % ppf);
      s.q[slot] = q;
      s.uq[slot] = p.u[k] * q;
    }
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

  for (int t1 = 0; t1 < ntiles; ++t1) {
    eff_fill_tile(p, s, s.A1, t1, TC, e_begin, e_end, ix_self);
    const int n1 = min(TC, nconn - t1 * TC);
    // g = E Q u restricted to this tile
    if (tid < n1 * 6) {
      float acc = 0.f;
      for (int k = 0; k < ppf; ++k) acc += s.A1[k * row + tid] * s.uq[k];
      if (acc != 0.f) atomicAdd(&p.g[6 * (int)s.conn[t1 * TC + tid / 6] + tid % 6], (double)acc);
    }
    for (int t2 = t1; t2 < ntiles; ++t2) {
      const float* A2 = s.A1;
      if (t2 != t1) { eff_fill_tile(p, s, s.A2, t2, TC, e_begin, e_end, ix_self); A2 = s.A2; }
      const int n2 = min(TC, nconn - t2 * TC);
      for (int pr = tid; pr < n1 * n2; pr += 256) {
        const int b1 = pr / n2, b2 = pr - b1 * n2;
        float acc[6][6];
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
          for (int b = 0; b < 6; ++b) acc[a][b] = 0.f;
        const float* r1 = s.A1 + b1 * 6;
        const float* r2 = A2 + b2 * 6;
        for (int k = 0; k < ppf; ++k) {
          const float qk = s.q[k];
          float x[6], y[6];
#pragma unroll
          for (int a = 0; a < 6; ++a) { x[a] = r1[k * row + a] * qk; y[a] = r2[k * row + a]; }
#pragma unroll
          for (int a = 0; a < 6; ++a)
#pragma unroll
            for (int b = 0; b < 6; ++b) acc[a][b] += x[a] * y[b];
        }
        const int ja = 6 * (int)s.conn[t1 * TC + b1], jb = 6 * (int)s.conn[t2 * TC + b2];
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
          for (int b = 0; b < 6; ++b) {
            const float v = acc[a][b];
            if (v == 0.f) continue;
            atomicAdd(&p.G[(int64_t)(ja + a) * n + jb + b], (double)v);
            if (t2 != t1) atomicAdd(&p.G[(int64_t)(jb + b) * n + ja + a], (double)v);
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
    const float4* ee = reinterpret_cast<const float4*>(p.Ee + e * 12);
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
  cudaMemsetAsync(p.fhist, 0, reinterpret_cast<char*>(p.foff) - reinterpret_cast<char*>(p.fhist), st);
  eff_fhist_kernel<<<egrid, 256, 0, st>>>(p);
  eff_fscan_kernel<<<1, 1024, 0, st>>>(p);
  eff_fscatter_kernel<<<egrid, 256, 0, st>>>(p);
  note_launches(3);

  // Schur kernel shared memory: tables + two conn tiles of TC poses
  const int Npad = (N + 7) & ~7;
  const size_t fixed = (size_t)Npad * 2 * sizeof(unsigned short) + 2 * (size_t)ppf * sizeof(float);
  const size_t budget = 200 * 1024;
  int TC = 32;
  while (TC > 1 && fixed + 2 * (size_t)ppf * TC * 6 * sizeof(float) > budget) TC >>= 1;
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
