// Edge-MLP layer of DPVO's Update operator (net.py:23-88, blocks.py:15-48): Y = epilogue(A W^T + b)
// for A [E, K] fp16 (K = 384 or 882) and W [384, K] fp16 (nn.Linear layout), fp32 accumulation in
// tensor memory.
//
//   * one persistent CTA per SM walks 128-edge row tiles; the full 384-wide output row of a tile
//     lives in TMEM (384 of 512 columns), so LayerNorm / gating / residuals are row-local;
//   * warp 0 streams K-chunks of 64 (A: 128x64, W: 384x64 as two 192-row boxes) through a 2-stage
//     shared-memory ring with TMA (128-byte swizzle), warp 1's lane 0 issues tcgen05.mma
//     (M128 x N192 x K16, two per K step);
//   * warps 2-13 are the epilogue: TMEM lanes are only reachable from warp (id % 4), so each lane
//     quarter gets three warps that split the 384 columns; one thread = one row x 128 columns.
//     LayerNorm statistics of the three thirds meet in spare TMEM columns, the pre-normalisation
//     row is parked in TMEM between the two passes;
//   * the fp16 output tile is assembled in shared memory in the swizzled 128x64 box layout and
//     written with six TMA stores (coalesced, row tails clipped by the tensor map);
//   * K tails (882 = 13.78 chunks) and row tails are zero-filled by TMA's out-of-bounds handling.
#include "common.cuh"
#include "sm100.cuh"
#include <stdlib.h>

namespace dpvo {
namespace {
using namespace sm100;

constexpr int kDim = 384;
constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kStages = 2;
constexpr int kABytes = BM * BK * 2;             // 16 KB
constexpr int kWBytes = kDim * BK * 2;           // 48 KB
constexpr int kStageBytes = kABytes + kWBytes;   // 64 KB
constexpr int kOutBytes = BM * kDim * 2;         // 96 KB: six 128x64 boxes of 16 KB
constexpr int kEpiWarps = 12;
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreads = 64 + kEpiThreads;       // 448
constexpr uint32_t kTmemCols = 512;
constexpr int kSmemBytes = kStages * kStageBytes + kOutBytes + 1024;

struct LinearDev {
  const float* bias;
  const __half* gate;
  const __half* res1;
  const __half* res2;
  const int64_t* res2_index;
  const float* ln_gamma;
  const float* ln_beta;
  float ln_eps;
  int act;
  int64_t M;
  int kblocks;
};

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __half22float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

// x[j] = (acc[j] + bias) * gate + res1 + res2 for 32 consecutive columns starting at n0 of one row
__device__ __forceinline__ void apply_pre(float (&x)[32], const float* bias, int n0, const __half* gate,
                                          const __half* res1, const __half* res2) {
  if (bias) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float4 b = *(reinterpret_cast<const float4*>(bias + n0) + q);
      x[4 * q] += b.x; x[4 * q + 1] += b.y; x[4 * q + 2] += b.z; x[4 * q + 3] += b.w;
    }
  }
  if (gate) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float f[8];
      unpack8(ldg128(gate + n0 + 8 * q), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) x[8 * q + j] *= f[j];
    }
  }
  if (res1) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float f[8];
      unpack8(ldg128(res1 + n0 + 8 * q), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) x[8 * q + j] += f[j];
    }
  }
  if (res2) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float f[8];
      unpack8(ldg128(res2 + n0 + 8 * q), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) x[8 * q + j] += f[j];
    }
  }
}

// CL = CTAs per cluster.  The CL CTAs of a cluster work on CL consecutive row tiles in lock step and
// share every W chunk: each loads 384/CL of its rows and multicasts them into all CL shared memories,
// so W crosses the L2 -> SM fabric once per cluster instead of once per tile.
template <int CL>
__global__ void __launch_bounds__(kThreads, 1)
linear384_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmY, const LinearDev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* out_tile = smem + kStages * kStageBytes;
  __shared__ uint64_t full[kStages], empty[kStages], tmem_full, tmem_empty;
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t tiles = (p.M + BM - 1) / BM;
  const int64_t groups = (tiles + CL - 1) / CL;       // groups of CL consecutive row tiles, one per cluster pass
  const uint32_t crank = CL > 1 ? cluster_ctarank() : 0u;
  const uint32_t cid = CL > 1 ? cluster_id_x() : blockIdx.x;
  const uint32_t ncl = CL > 1 ? cluster_count_x() : gridDim.x;
  constexpr uint16_t kMask = static_cast<uint16_t>((1u << CL) - 1u);
  const bool ln = p.ln_gamma != nullptr;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmY);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kStages; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], CL);                     // one arrival per CTA of the cluster
      }
      mbar_init(&tmem_full, 1);
      mbar_init(&tmem_empty, kEpiThreads);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(&tmem_base_s, kTmemCols);
  }
  tc_fence_before();
  if (CL > 1) cluster_sync_all(); else __syncthreads();   // barriers initialised cluster-wide before any multicast
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t grp = cid; grp < groups; grp += ncl) {
        const int64_t tile = grp * CL + crank;        // may lie past the last tile: loads are zero-filled
        for (int kb = 0; kb < p.kblocks; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1;
          mbar_wait(&empty[s], ph ^ 1);               // every CTA of the cluster has released stage s
          mbar_arrive_expect_tx(&full[s], kStageBytes);
          uint8_t* st = smem + s * kStageBytes;
          tma_load_2d(st, &tmA, &full[s], kb * BK, static_cast<int32_t>(tile * BM));
          if (CL == 1) {
            tma_load_2d(st + kABytes, &tmW, &full[s], kb * BK, 0);
            tma_load_2d(st + kABytes + kWBytes / 2, &tmW, &full[s], kb * BK, kDim / 2);
          } else {
            tma_load_2d_mc(st + kABytes + crank * (kWBytes / CL), &tmW, &full[s], kb * BK, crank * (kDim / CL), kMask);
          }
        }
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = instr_desc_f16(BM, kDim / 2);
      uint32_t it = 0, tcount = 0;
      for (int64_t grp = cid; grp < groups; grp += ncl, ++tcount) {
        mbar_wait(&tmem_empty, (tcount & 1) ^ 1);     // epilogue has drained the previous tile
        tc_fence_after();
        for (int kb = 0; kb < p.kblocks; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t a0 = smem_u32(smem + s * kStageBytes);
          const uint32_t w0 = a0 + kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t ad = smem_desc_sw128(a0 + k * 32);
            const uint64_t b0 = smem_desc_sw128(w0 + k * 32);
            const uint64_t b1 = smem_desc_sw128(w0 + kWBytes / 2 + k * 32);
            const uint32_t acc = (kb | k) != 0 ? 1u : 0u;
            mma_f16_ss(tmem_base, ad, b0, idesc, acc);
            mma_f16_ss(tmem_base + kDim / 2, ad, b1, idesc, acc);
          }
          // stage s is free once these MMAs have read it — tell every producer of the cluster
          if (CL == 1) mma_commit(&empty[s]); else mma_commit_mc(&empty[s], kMask);
        }
        mma_commit(&tmem_full);                       // accumulator complete
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue
    const int q = warp & 3;                           // TMEM lane quarter this warp may access
    const int t = (warp - 2) >> 2;                    // column third: [128 t, 128 t + 128)
    const int r_tile = q * 32 + lane;                 // row within the tile
    const bool issuer = (warp == 2 && lane == 0);
    uint32_t tcount = 0;
    for (int64_t grp = cid; grp < groups; grp += ncl, ++tcount) {
      const int64_t tile = grp * CL + crank;
      mbar_wait(&tmem_full, tcount & 1);
      tc_fence_after();
      const int64_t row = tile * BM + r_tile;
      const bool ok = row < p.M;
      const int64_t r = ok ? row : 0;
      const uint32_t trow = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + t * 128;
      const __half* gate = p.gate ? p.gate + r * kDim : nullptr;
      const __half* res1 = p.res1 ? p.res1 + r * kDim : nullptr;
      const __half* res2 = p.res2 ? p.res2 + (p.res2_index ? p.res2_index[r] : r) * kDim : nullptr;
      float mean = 0.f, rstd = 1.f;
      if (ln) {
        float sum = 0.f, sq = 0.f;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          float x[32];
          tmem_ld32(trow + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]);
          apply_pre(x, p.bias, t * 128 + c * 32, gate, res1, res2);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            sum += x[j];
            sq += x[j] * x[j];
            v[j] = __float_as_uint(x[j]);
          }
          tmem_st32(trow + c * 32, v);                // park the pre-normalisation row in TMEM
        }
        // partial statistics of the three column thirds meet in spare TMEM columns 384..389 of the row
        tmem_st2(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + kDim + 2 * t, __float_as_uint(sum),
                 __float_as_uint(sq));
        tmem_st_wait();
        tc_fence_before();
        named_barrier(2, kEpiThreads);
        tc_fence_after();
        uint32_t ps[8];
        tmem_ld8(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + kDim, ps);
        tmem_ld_wait();
        sum = __uint_as_float(ps[0]) + __uint_as_float(ps[2]) + __uint_as_float(ps[4]);
        sq = __uint_as_float(ps[1]) + __uint_as_float(ps[3]) + __uint_as_float(ps[5]);
        mean = sum * (1.0f / kDim);
        const float var = fmaxf(sq * (1.0f / kDim) - mean * mean, 0.f);
        rstd = rsqrtf(var + p.ln_eps);
      }
      named_barrier(1, kEpiThreads);                  // previous tile's TMA stores have read out_tile (issuer waited)
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        float x[32];
        tmem_ld32(trow + c * 32, v);
        tmem_ld_wait();
        const int n0 = t * 128 + c * 32;
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]);
        if (ln) {
#pragma unroll
          for (int qd = 0; qd < 8; ++qd) {
            const float4 g4 = __ldg(reinterpret_cast<const float4*>(p.ln_gamma + n0) + qd);
            const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.ln_beta + n0) + qd);
            x[4 * qd] = (x[4 * qd] - mean) * rstd * g4.x + b4.x;
            x[4 * qd + 1] = (x[4 * qd + 1] - mean) * rstd * g4.y + b4.y;
            x[4 * qd + 2] = (x[4 * qd + 2] - mean) * rstd * g4.z + b4.z;
            x[4 * qd + 3] = (x[4 * qd + 3] - mean) * rstd * g4.w + b4.w;
          }
        } else {
          apply_pre(x, p.bias, n0, gate, res1, res2);
        }
        if (p.act == 1) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
        } else if (p.act == 2) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = __fdividef(1.0f, 1.0f + __expf(-x[j]));
        }
        // 32 columns = four 16-byte units of this row in box n0 / 64, XOR-swizzled like the TMA layout
        uint8_t* box = out_tile + (n0 >> 6) * (BM * 128) + r_tile * 128;
        const int u0 = (n0 & 63) >> 3;
#pragma unroll
        for (int qd = 0; qd < 4; ++qd) {
          uint4 o;
          __half2* h = reinterpret_cast<__half2*>(&o);
#pragma unroll
          for (int j = 0; j < 4; ++j) h[j] = __floats2half2_rn(x[8 * qd + 2 * j], x[8 * qd + 2 * j + 1]);
          *reinterpret_cast<uint4*>(box + (((u0 + qd) ^ (r_tile & 7)) << 4)) = o;
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty);                       // TMEM drained: the next tile's MMAs may start
      fence_proxy_async_smem();
      named_barrier(3, kEpiThreads);                  // whole tile is in out_tile
      if (issuer && tile < tiles) {
#pragma unroll
        for (int b = 0; b < kDim / 64; ++b)
          tma_store_2d(&tmY, out_tile + b * (BM * 128), b * 64, static_cast<int32_t>(tile * BM));
        tma_store_commit();
        tma_store_wait_read();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();                     // no CTA leaves while peers may still signal its barriers
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// Row operands (gate / residuals) for a 32-row x 32-column chunk, fetched the coalesced way round: four
// lanes read one row's 64 bytes (a warp load covers 8 whole row segments instead of 32 scattered
// 16-byte pieces); the pieces are parked in the warp's 2 KB buffer and every thread picks up its own
// row.  (Issuing these loads a chunk ahead was tried: no gain — the kernel is bound by L2 throughput,
// not by load latency — and the extra registers spilled.)
// `myrow`: this lane's row (already offset to the CTA's first column); rows may be gathered.
__device__ __forceinline__ void rows_issue(const __half* myrow, int n0, int lane, uint4 (&pc)[4]) {
  const unsigned long long mine = reinterpret_cast<unsigned long long>(myrow + n0);
#pragma unroll
  for (int ps = 0; ps < 4; ++ps) {
    const unsigned long long rp = __shfl_sync(0xffffffffu, mine, ps * 8 + (lane >> 2));
    pc[ps] = ldg128(reinterpret_cast<const __half*>(rp) + (lane & 3) * 8);
  }
}
template <bool MUL>
__device__ __forceinline__ void rows_apply(uint8_t* sw, const uint4 (&pc)[4], int lane, float (&x)[32]) {
  __syncwarp();                                       // the buffer's previous contents have been consumed
#pragma unroll
  for (int ps = 0; ps < 4; ++ps) {
    const int rr = ps * 8 + (lane >> 2), u = lane & 3;
    *reinterpret_cast<uint4*>(sw + rr * 64 + ((u ^ ((rr >> 1) & 3)) << 4)) = pc[ps];
  }
  __syncwarp();
#pragma unroll
  for (int qd = 0; qd < 4; ++qd) {
    const uint4 o = *reinterpret_cast<const uint4*>(sw + lane * 64 + ((qd ^ ((lane >> 1) & 3)) << 4));
    float f[8];
    unpack8(o, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[8 * qd + j] = MUL ? x[8 * qd + j] * f[j] : x[8 * qd + j] + f[j];
  }
}
template <bool MUL>
__device__ __forceinline__ void rows_fetch_apply(uint8_t* sw, const __half* myrow, int n0, int lane, float (&x)[32]) {
  uint4 pc[4];
  rows_issue(myrow, n0, lane, pc);
  rows_apply<MUL>(sw, pc, lane, x);
}
__device__ __forceinline__ void add_bias32(float (&x)[32], const float* bias) {
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float4 b = *(reinterpret_cast<const float4*>(bias) + q);
    x[4 * q] += b.x; x[4 * q + 1] += b.y; x[4 * q + 2] += b.z; x[4 * q + 3] += b.w;
  }
}

// ------------------------------------------------------ resident-W variant (K = 384, no LayerNorm)
// The streaming kernel above is bound by what one SM can pull in from L2 / HBM (measured ~30 bytes per
// clock per SM, whatever the ring depth or multicast):
// every 128-row tile re-reads all of W (288 KB) next to its 96 KB of A.  Here the output columns are
// split between the two CTAs of a pair: each keeps ITS half of W (192 x 384 fp16 = 144 KB) in shared
// memory for the whole kernel and streams only A tiles through a 3-stage ring, so a tile costs 96 KB
// of ingest per SM instead of 384 KB.  The half-width accumulator (192 TMEM columns) is double
// buffered: the epilogue of tile i overlaps the MMAs of tile i+1.  Output chunks are transposed through
// a 2 KB per-warp buffer so that every global store instruction writes whole 64-byte row segments.
constexpr int kHalf = kDim / 2;                      // 192 output columns per CTA
constexpr int kRwStages = 3;
constexpr int kRwWBytes = kHalf * kDim * 2;          // 144 KB: six 192x64 chunks of 24 KB
constexpr int kRwWChunk = kHalf * BK * 2;            // 24 KB
constexpr int kRwStageOut = kEpiWarps * 2048;       // 2 KB transposition buffer per epilogue warp
constexpr int kRwSmemBytes = kRwWBytes + kRwStages * kABytes + kRwStageOut + 1024;

// Per-role wait-time counters of CTA 0 (DPVO_LINEAR_PROFILE=1; read with dpvo_linear_profile_read):
// [0] producer: cycles waiting for a free ring stage   [1] producer: total
// [2] MMA: waiting for A stages   [3] MMA: waiting for a drained accumulator   [4] MMA: waiting for W   [5] MMA: total
// [6] epilogue: waiting for an accumulator   [7] epilogue: total   [8] tiles of this CTA
__device__ long long g_linear_profile[16];

struct LinearRwDev {
  int profile;
  const float* bias;
  const __half* gate;
  const __half* res1;
  const __half* res2;
  const int64_t* res2_index;
  const float* ln_gamma;
  const float* ln_beta;
  float ln_eps;
  int act;
  __half* Y;
  int64_t M;
  const float* head_w;   // [4][384]: rows of the two 384->2 heads, applied to ReLU(output)
  float* head_acc;       // [M][4], accumulated with atomics (both column halves contribute)
};

// LN: LayerNorm over the full 384-wide row.  The two CTAs of a pair then run as a cluster and swap the
// per-row (sum, sum of squares) of their halves through each other's shared memory once per tile.
// AMC: the pair runs as a cluster and shares every A chunk: each CTA loads 64 of its 128 rows and
// multicasts them into both shared memories.
template <bool LN, bool HEADS, bool AMC>
__global__ void __launch_bounds__(kThreads, 1)
linear384_rw_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                    const LinearRwDev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* w_res = smem;                              // [6][192 rows][128 B]
  uint8_t* ring = smem + kRwWBytes;                   // [stages][128 rows][128 B]
  uint8_t* stage_out = ring + kRwStages * kABytes;    // [epilogue warp][32 rows][64 B]
  __shared__ uint64_t wfull, full[kRwStages], empty[kRwStages], tmem_full[2], tmem_empty[2];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_bias[kHalf];       // this CTA's half of the bias (L1 is tiny next to 220 KB of smem)
  __shared__ __align__(16) float s_gamma[LN ? kHalf : 4], s_beta[LN ? kHalf : 4];
  __shared__ __align__(8) float2 peer_part[2][LN ? BM : 1];   // written by the other CTA of the pair
  __shared__ uint64_t ln_bar[2];
  __shared__ __align__(16) float s_head[HEADS ? 4 : 1][HEADS ? kHalf : 4];   // this half's slice of the head weights

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t tiles = (p.M + BM - 1) / BM;
  const int h = blockIdx.x & 1;                       // which half of the output columns
  const int64_t first = blockIdx.x >> 1, step = gridDim.x >> 1;
  constexpr int kKb = kDim / BK;                      // 6 K chunks

  for (int i = threadIdx.x; i < kHalf; i += kThreads) {
    s_bias[i] = p.bias ? p.bias[(blockIdx.x & 1) * kHalf + i] : 0.f;
    if (HEADS) {
#pragma unroll
      for (int hd = 0; hd < 4; ++hd) s_head[HEADS ? hd : 0][i] = p.head_w[hd * kDim + (blockIdx.x & 1) * kHalf + i];
    }
    if (LN) {
      s_gamma[i] = p.ln_gamma[(blockIdx.x & 1) * kHalf + i];
      s_beta[i] = p.ln_beta[(blockIdx.x & 1) * kHalf + i];
    }
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(&wfull, 1);
      for (int s = 0; s < kRwStages; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], AMC ? 2 : 1);
      }
      for (int b = 0; b < 2; ++b) {
        mbar_init(&tmem_full[b], 1);
        mbar_init(&tmem_empty[b], kEpiThreads);
        mbar_init(&ln_bar[b], BM);                    // one arrival per row from the peer CTA
      }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(&tmem_base_s, kTmemCols);
  }
  tc_fence_before();
  if (LN || AMC) cluster_sync_all(); else __syncthreads();   // the peer's barriers must exist before the first remote arrive
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      const bool prof = p.profile && blockIdx.x == 0;
      long long t_wait = 0;
      const long long t_begin = clock64();
      mbar_arrive_expect_tx(&wfull, kRwWBytes);
      for (int kb = 0; kb < kKb; ++kb) tma_load_2d(w_res + kb * kRwWChunk, &tmW, &wfull, kb * BK, h * kHalf);
      uint32_t it = 0;
      for (int64_t tile = first; tile < tiles; tile += step) {
        for (int kb = 0; kb < kKb; ++kb, ++it) {
          const int s = it % kRwStages;
          const uint32_t ph = (it / kRwStages) & 1;
          const long long t0 = clock64();
          mbar_wait(&empty[s], ph ^ 1);
          t_wait += clock64() - t0;
          mbar_arrive_expect_tx(&full[s], kABytes);
          if (AMC)
            tma_load_2d_mc(ring + s * kABytes + h * (kABytes / 2), &tmA, &full[s], kb * BK,
                           static_cast<int32_t>(tile * BM + h * (BM / 2)), 3);
          else
            tma_load_2d(ring + s * kABytes, &tmA, &full[s], kb * BK, static_cast<int32_t>(tile * BM));
        }
      }
      if (prof) {
        g_linear_profile[0] = t_wait;
        g_linear_profile[1] = clock64() - t_begin;
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = instr_desc_f16(BM, kHalf);
      const bool prof = p.profile && blockIdx.x == 0;
      long long t_full = 0, t_acc = 0;
      const long long t_begin = clock64();
      mbar_wait(&wfull, 0);
      const long long t_w = clock64() - t_begin;
      tc_fence_after();
      const uint32_t w0 = smem_u32(w_res);
      uint32_t it = 0, tcount = 0;
      for (int64_t tile = first; tile < tiles; tile += step, ++tcount) {
        const uint32_t buf = tcount & 1;
        long long t0 = clock64();
        mbar_wait(&tmem_empty[buf], ((tcount >> 1) & 1) ^ 1);   // epilogue has drained this accumulator
        t_acc += clock64() - t0;
        tc_fence_after();
        const uint32_t d = tmem_base + buf * kHalf;
        for (int kb = 0; kb < kKb; ++kb, ++it) {
          const int s = it % kRwStages;
          const uint32_t ph = (it / kRwStages) & 1;
          t0 = clock64();
          mbar_wait(&full[s], ph);
          t_full += clock64() - t0;
          tc_fence_after();
          const uint32_t a0 = smem_u32(ring + s * kABytes);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            mma_f16_ss(d, smem_desc_sw128(a0 + k * 32), smem_desc_sw128(w0 + kb * kRwWChunk + k * 32), idesc,
                       (kb | k) != 0 ? 1u : 0u);
          if (AMC) mma_commit_mc(&empty[s], 3); else mma_commit(&empty[s]);
        }
        mma_commit(&tmem_full[buf]);
      }
      if (prof) {
        g_linear_profile[2] = t_full;
        g_linear_profile[3] = t_acc;
        g_linear_profile[4] = t_w;
        g_linear_profile[5] = clock64() - t_begin;
        g_linear_profile[8] = tcount;
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue
    const int q = warp & 3;                           // TMEM lane quarter
    const int t = (warp - 2) >> 2;                    // column third of the half: [64 t, 64 t + 64)
    const bool prof = p.profile && blockIdx.x == 0 && threadIdx.x == 64;
    long long t_accw = 0;
    const long long t_begin = clock64();
    const int r_tile = q * 32 + lane;
    uint32_t tcount = 0;
    for (int64_t tile = first; tile < tiles; tile += step, ++tcount) {
      const uint32_t buf = tcount & 1;
      const int64_t row = tile * BM + r_tile;
      const bool ok = row < p.M;
      const int64_t r = ok ? row : 0;
      const uint32_t trow = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * kHalf + t * 64;
      const __half* gate = p.gate ? p.gate + r * kDim + h * kHalf : nullptr;
      const __half* res1 = p.res1 ? p.res1 + r * kDim + h * kHalf : nullptr;
      const __half* res2 = p.res2 ? p.res2 + (p.res2_index ? p.res2_index[r] : r) * kDim + h * kHalf : nullptr;
      uint8_t* sw = stage_out + (warp - 2) * 2048;
      const long long t0 = clock64();
      mbar_wait(&tmem_full[buf], (tcount >> 1) & 1);
      t_accw += clock64() - t0;
      tc_fence_after();
      float mean = 0.f, rstd = 1.f;
      if (LN) {
        float sum = 0.f, sq = 0.f;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          uint32_t v[32];
          float x[32];
          tmem_ld32(trow + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]);
          add_bias32(x, s_bias + t * 64 + c * 32);
          if (gate) rows_fetch_apply<true>(sw, gate, t * 64 + c * 32, lane, x);
          if (res1) rows_fetch_apply<false>(sw, res1, t * 64 + c * 32, lane, x);
          if (res2) rows_fetch_apply<false>(sw, res2, t * 64 + c * 32, lane, x);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            sum += x[j];
            sq += x[j] * x[j];
            v[j] = __float_as_uint(x[j]);
          }
          tmem_st32(trow + c * 32, v);                // park the pre-normalisation row in TMEM
        }
        // the three column thirds of this half meet in spare TMEM columns 384 + 8 buf ..
        const uint32_t tstat = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + kDim + 8 * buf;
        tmem_st2(tstat + 2 * t, __float_as_uint(sum), __float_as_uint(sq));
        tmem_st_wait();
        tc_fence_before();
        named_barrier(2, kEpiThreads);
        tc_fence_after();
        uint32_t ps[8];
        tmem_ld8(tstat, ps);
        tmem_ld_wait();
        sum = __uint_as_float(ps[0]) + __uint_as_float(ps[2]) + __uint_as_float(ps[4]);
        sq = __uint_as_float(ps[1]) + __uint_as_float(ps[3]) + __uint_as_float(ps[5]);
        // .. and the two halves meet through the peer CTA's shared memory
        const uint32_t par = (tcount >> 1) & 1;
        if (t == 0) {
          st_remote_f32x2(map_to_cta(&peer_part[buf][r_tile], h ^ 1), sum, sq);
          mbar_arrive_remote(map_to_cta(&ln_bar[buf], h ^ 1));
        }
        mbar_wait_cluster(&ln_bar[buf], par);
        const float2 other = peer_part[buf][r_tile];
        sum += other.x;
        sq += other.y;
        mean = sum * (1.0f / kDim);
        rstd = rsqrtf(fmaxf(sq * (1.0f / kDim) - mean * mean, 0.f) + p.ln_eps);
      }
      float hacc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        float x[32];
        tmem_ld32(trow + c * 32, v);
        tmem_ld_wait();
        const int n0 = t * 64 + c * 32;               // column within this CTA's half
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]);
        if (LN) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = (x[j] - mean) * rstd * s_gamma[n0 + j] + s_beta[n0 + j];
        } else {
          add_bias32(x, s_bias + n0);
          if (gate) rows_fetch_apply<true>(sw, gate, t * 64 + c * 32, lane, x);
          if (res1) rows_fetch_apply<false>(sw, res1, t * 64 + c * 32, lane, x);
          if (res2) rows_fetch_apply<false>(sw, res2, t * 64 + c * 32, lane, x);
        }
        if (p.act == 1) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
        } else if (p.act == 2) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = __fdividef(1.0f, 1.0f + __expf(-x[j]));
        }
        if (HEADS) {                                  // the two 384->2 heads ride on the last layer (net.py:88)
#pragma unroll
          for (int hd = 0; hd < 4; ++hd) {
#pragma unroll
            for (int qd = 0; qd < 8; ++qd) {
              const float4 w4 = *reinterpret_cast<const float4*>(&s_head[HEADS ? hd : 0][n0 + 4 * qd]);
              hacc[hd] += fmaxf(x[4 * qd], 0.f) * w4.x + fmaxf(x[4 * qd + 1], 0.f) * w4.y +
                          fmaxf(x[4 * qd + 2], 0.f) * w4.z + fmaxf(x[4 * qd + 3], 0.f) * w4.w;
            }
          }
        }
        // transpose through shared memory so that global stores are whole 64-byte row segments:
        // thread r parks its four 16-byte units (XOR-swizzled by row pair: conflict-free both ways) ...
        __syncwarp();                                 // previous chunk has been read out
#pragma unroll
        for (int qd = 0; qd < 4; ++qd) {
          uint4 o;
          __half2* hh = reinterpret_cast<__half2*>(&o);
#pragma unroll
          for (int j = 0; j < 4; ++j) hh[j] = __floats2half2_rn(x[8 * qd + 2 * j], x[8 * qd + 2 * j + 1]);
          *reinterpret_cast<uint4*>(sw + lane * 64 + ((qd ^ ((lane >> 1) & 3)) << 4)) = o;
        }
        __syncwarp();
        // ... then four lanes write one row: a warp store covers 8 rows x 64 contiguous bytes
        if (p.Y) {
          const int64_t row0 = tile * BM + q * 32;
#pragma unroll
          for (int ps = 0; ps < 4; ++ps) {
            const int rr = ps * 8 + (lane >> 2);
            const int u = lane & 3;
            const uint4 o = *reinterpret_cast<const uint4*>(sw + rr * 64 + ((u ^ ((rr >> 1) & 3)) << 4));
            if (row0 + rr < p.M)
              *reinterpret_cast<uint4*>(p.Y + (row0 + rr) * kDim + h * kHalf + n0 + u * 8) = o;
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty[buf]);
      if (HEADS && ok) {
#pragma unroll
        for (int hd = 0; hd < 4; ++hd) atomicAdd(p.head_acc + row * 4 + hd, hacc[hd]);
      }
    }
    if (prof) {
      g_linear_profile[6] = t_accw;
      g_linear_profile[7] = clock64() - t_begin;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (LN || AMC) cluster_sync_all();                  // the peer may still be signalling this CTA's barriers
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}

// fp16 [rows, K] row-major with row stride ld (elements); tiles of 64 x box_rows, 128-byte swizzle
int make_map(CUtensorMap* m, const void* base, int64_t K, int64_t rows, int64_t ld, int box_rows) {
  EncodeTiledFn enc = encode_fn();
  DPVO_REQUIRE(enc != nullptr, "linear384: cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t gdim[2] = {static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(rows)};
  const cuuint64_t gstride[1] = {static_cast<cuuint64_t>(ld) * 2};
  const cuuint32_t box[2] = {BK, static_cast<cuuint32_t>(box_rows)};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DPVO_REQUIRE(r == CUDA_SUCCESS, "linear384: cuTensorMapEncodeTiled failed (%d)", static_cast<int>(r));
  return 0;
}

// DPVO_LINEAR_RW=0 routes every layer through the streaming kernel (A/B measurements)
bool use_resident_w() {
  static int v = -1;
  if (v < 0) {
    const char* e = debug_env("DPVO_LINEAR_RW");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

// CTAs per cluster: DPVO_LINEAR_CLUSTER = 1 | 2 | 4 | 8 overrides the default
int cluster_size() {
  static int cl = 0;
  if (!cl) {
    const char* e = debug_env("DPVO_LINEAR_CLUSTER");
    const int v = e ? atoi(e) : 0;
    cl = (v == 1 || v == 2 || v == 4 || v == 8) ? v : 1;
  }
  return cl;
}

template <int CL>
int launch_linear(const CUtensorMap& tmA, const CUtensorMap& tmW, const CUtensorMap& tmY, const LinearDev& p,
                  cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(linear384_kernel<CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  DPVO_REQUIRE(e == cudaSuccess, "linear384: cannot reserve %d bytes of shared memory: %s", kSmemBytes,
               cudaGetErrorString(e));
  const int64_t tiles = (p.M + BM - 1) / BM;
  const int64_t groups = (tiles + CL - 1) / CL;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int max_clusters = kNumSMs / CL;
  if (CL > 1) {
    static int cached[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (!cached[CL]) {
      cfg.gridDim = dim3(kNumSMs / CL * CL);
      int n = 0;
      e = cudaOccupancyMaxActiveClusters(&n, linear384_kernel<CL>, &cfg);
      DPVO_REQUIRE(e == cudaSuccess && n > 0, "linear384: no cluster of %d CTAs fits on this device: %s", CL,
                   cudaGetErrorString(e));
      cached[CL] = n;
    }
    if (cached[CL] < max_clusters) max_clusters = cached[CL];
  }
  const int64_t clusters = groups < max_clusters ? groups : max_clusters;
  cfg.gridDim = dim3(static_cast<unsigned>(clusters * CL));
  note_launches(1);
  e = cudaLaunchKernelEx(&cfg, linear384_kernel<CL>, tmA, tmW, tmY, p);
  if (e != cudaSuccess) {
    set_error("linear384: launch failed: %s", cudaGetErrorString(e));
    return 1;
  }
  return check_launch("linear384");
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

extern "C" int dpvo_linear384_f16(const dpvo_linear_params* lp, void* stream) {
  DPVO_REQUIRE(lp != nullptr, "linear384: null params");
  DPVO_REQUIRE(lp->M >= 0 && lp->K > 0, "linear384: bad sizes");
  if (lp->M == 0) return 0;
  DPVO_REQUIRE(lp->A && lp->W && lp->Y, "linear384: null pointer");
  DPVO_REQUIRE(lp->lda >= lp->K && lp->ldw >= lp->K && lp->lda % 8 == 0 && lp->ldw % 8 == 0,
               "linear384: row strides must be >= K and multiples of 8 elements (16 bytes), got lda=%lld ldw=%lld",
               (long long)lp->lda, (long long)lp->ldw);
  DPVO_REQUIRE((reinterpret_cast<uintptr_t>(lp->A) & 15) == 0 && (reinterpret_cast<uintptr_t>(lp->W) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(lp->Y) & 15) == 0,
               "linear384: A, W and Y must be 16-byte aligned");
  DPVO_REQUIRE(lp->act >= 0 && lp->act <= 2, "linear384: act must be 0 (none), 1 (relu) or 2 (sigmoid)");
  DPVO_REQUIRE((lp->head_w == nullptr) == (lp->head_acc == nullptr), "linear384: head_w and head_acc go together");
  DPVO_REQUIRE(lp->head_w == nullptr || (lp->K == kDim && lp->ln_gamma == nullptr && use_resident_w()),
               "linear384: fused heads need K = 384 and no LayerNorm");
  DPVO_REQUIRE(lp->M < (int64_t(1) << 31) - BM, "linear384: too many rows");
  CUtensorMap tmA, tmW, tmY;
  if (make_map(&tmA, lp->A, lp->K, lp->M, lp->lda, BM)) return 1;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (lp->K == kDim && use_resident_w()) {
    if (make_map(&tmW, lp->W, lp->K, kDim, lp->ldw, kHalf)) return 1;
    LinearRwDev q;
    q.profile = debug_env("DPVO_LINEAR_PROFILE") ? 1 : 0;
    q.bias = lp->bias;
    q.gate = static_cast<const __half*>(lp->gate);
    q.res1 = static_cast<const __half*>(lp->res1);
    q.res2 = static_cast<const __half*>(lp->res2);
    q.res2_index = lp->res2_index;
    q.ln_gamma = lp->ln_gamma;
    q.ln_beta = lp->ln_beta;
    q.ln_eps = lp->ln_eps;
    q.act = lp->act;
    q.Y = debug_env("DPVO_LINEAR_DEBUG_NOSTORE") ? nullptr : static_cast<__half*>(lp->Y);
    q.M = lp->M;
    q.head_w = lp->head_w;
    q.head_acc = lp->head_acc;
    const int64_t tiles = (lp->M + BM - 1) / BM;
    const int pairs = static_cast<int>(tiles < kNumSMs / 2 ? tiles : kNumSMs / 2);
    const bool ln = lp->ln_gamma != nullptr;
    if (ln) DPVO_REQUIRE(lp->ln_beta != nullptr, "linear384: LayerNorm needs both gamma and beta");
    const bool heads = lp->head_w != nullptr;
    static const bool amc = debug_env("DPVO_LINEAR_AMC") != nullptr && debug_env("DPVO_LINEAR_AMC")[0] == '1';
    if (amc && make_map(&tmA, lp->A, lp->K, lp->M, lp->lda, BM / 2)) return 1;
    const void* fn =
        amc ? (ln ? reinterpret_cast<const void*>(linear384_rw_kernel<true, false, true>)
                  : heads ? reinterpret_cast<const void*>(linear384_rw_kernel<false, true, true>)
                          : reinterpret_cast<const void*>(linear384_rw_kernel<false, false, true>))
            : (ln ? reinterpret_cast<const void*>(linear384_rw_kernel<true, false, false>)
                  : heads ? reinterpret_cast<const void*>(linear384_rw_kernel<false, true, false>)
                          : reinterpret_cast<const void*>(linear384_rw_kernel<false, false, false>));
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kRwSmemBytes);
    DPVO_REQUIRE(e == cudaSuccess, "linear384: cannot reserve %d bytes of shared memory: %s", kRwSmemBytes,
                 cudaGetErrorString(e));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = kRwSmemBytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (ln || amc) ? 2 : 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    note_launches(1);
    void* args[3] = {&tmA, &tmW, &q};
    e = cudaLaunchKernelExC(&cfg, fn, args);
    if (e != cudaSuccess) {
      set_error("linear384: launch failed: %s", cudaGetErrorString(e));
      return 1;
    }
    return check_launch("linear384");
  }
  const int cl = cluster_size();
  if (make_map(&tmW, lp->W, lp->K, kDim, lp->ldw, cl == 1 ? kDim / 2 : kDim / cl)) return 1;
  if (make_map(&tmY, lp->Y, kDim, lp->M, kDim, BM)) return 1;
  LinearDev p;
  p.bias = lp->bias;
  p.gate = static_cast<const __half*>(lp->gate);
  p.res1 = static_cast<const __half*>(lp->res1);
  p.res2 = static_cast<const __half*>(lp->res2);
  p.res2_index = lp->res2_index;
  p.ln_gamma = lp->ln_gamma;
  p.ln_beta = lp->ln_beta;
  p.ln_eps = lp->ln_eps;
  p.act = lp->act;
  p.M = lp->M;
  p.kblocks = (lp->K + BK - 1) / BK;
  if (cl == 8) return launch_linear<8>(tmA, tmW, tmY, p, st);
  if (cl == 4) return launch_linear<4>(tmA, tmW, tmY, p, st);
  if (cl == 2) return launch_linear<2>(tmA, tmW, tmY, p, st);
  return launch_linear<1>(tmA, tmW, tmY, p, st);
}

/* Diagnostics: per-role cycle counters of CTA 0 of the last resident-W launch made with
 * DPVO_LINEAR_PROFILE=1 (see g_linear_profile in this file).  Synchronises the device. */
extern "C" int dpvo_linear_profile_read(long long* out16) {
  DPVO_REQUIRE(out16 != nullptr, "linear_profile_read: null pointer");
  cudaError_t e = cudaMemcpyFromSymbol(out16, g_linear_profile, sizeof(long long) * 16);
  DPVO_REQUIRE(e == cudaSuccess, "linear_profile_read: %s", cudaGetErrorString(e));
  return 0;
}
