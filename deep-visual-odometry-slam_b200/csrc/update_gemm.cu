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
      const float4 b = __ldg(reinterpret_cast<const float4*>(bias + n0) + q);
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
        mbar_init(&empty[s], 1);
      }
      mbar_init(&tmem_full, 1);
      mbar_init(&tmem_empty, kEpiThreads);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc(&tmem_base_s, kTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        for (int kb = 0; kb < p.kblocks; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&full[s], kStageBytes);
          uint8_t* st = smem + s * kStageBytes;
          tma_load_2d(st, &tmA, &full[s], kb * BK, static_cast<int32_t>(tile * BM));
          tma_load_2d(st + kABytes, &tmW, &full[s], kb * BK, 0);
          tma_load_2d(st + kABytes + kWBytes / 2, &tmW, &full[s], kb * BK, kDim / 2);
        }
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = instr_desc_f16(BM, kDim / 2);
      uint32_t it = 0, tcount = 0;
      for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++tcount) {
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
          mma_commit(&empty[s]);                      // stage s is free once these MMAs have read it
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
    for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++tcount) {
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
      if (issuer) {
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
  DPVO_REQUIRE(lp->M < (int64_t(1) << 31) - BM, "linear384: too many rows");
  CUtensorMap tmA, tmW, tmY;
  if (make_map(&tmA, lp->A, lp->K, lp->M, lp->lda, BM)) return 1;
  if (make_map(&tmW, lp->W, lp->K, kDim, lp->ldw, kDim / 2)) return 1;
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
  {
    cudaError_t e = cudaFuncSetAttribute(linear384_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    DPVO_REQUIRE(e == cudaSuccess, "linear384: cannot reserve %d bytes of shared memory: %s", kSmemBytes,
                 cudaGetErrorString(e));
  }
  const int64_t tiles = (lp->M + BM - 1) / BM;
  const int grid = static_cast<int>(tiles < kNumSMs ? tiles : kNumSMs);
  note_launches(1);
  linear384_kernel<<<grid, kThreads, kSmemBytes, reinterpret_cast<cudaStream_t>(stream)>>>(tmA, tmW, tmY, p);
  return check_launch("linear384");
}
