// Per-frame feature encoders: BasicEncoder4 (methods/dpvo/dpvo/extractor.py:200-264) as used
// by Patchifier (net.py:95-96, 108-109): fnet (instance norm, 128 outputs) and inet (no norm, 384 outputs).
//
//   conv1 7x7 s2 3->32 | layer1: 2 x ResidualBlock(32) | layer2: ResidualBlock(32->64, s2) + ResidualBlock(64) | conv2 1x1
//
// Activations are channels-last fp16 ([img, H, W, C]) so a pixel's channels are one contiguous K-slice of the implicit
// GEMM and the final map can be written straight into the NHWC feature ring.  Convolutions accumulate in fp32 on the
// tensor cores (mma.sync m16n8k16: N is 32-64 wide and K = 9*C <= 576, far too small for a tcgen05 128xN tile to pay
// for its TMEM round trip) and emit per-(image, channel) sum / sum of squares (float64 atomics) for the instance norm
// that follows; the norm, ReLU and residual add are one elementwise pass over the fp16 map.
#include "common.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace dpvo {
namespace {

constexpr int kStatReplicas = DPVO_ENCODER_STATS_REPLICAS;   // atomics of different CTAs land on different L2 lines

__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool pred) {
  const unsigned dst = (unsigned)__cvta_generic_to_shared(smem);
  const int bytes = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(gmem), "r"(bytes) : "memory");
}

// mean / rstd of one (image, channel) plane from the replicated float64 sums (biased variance, like InstanceNorm2d)
__device__ __forceinline__ void plane_stats(const double* st, int64_t n, int C, int c, double inv_n, float eps, float& mean,
                                            float& rstd) {
  double s = 0, q = 0;
#pragma unroll 4
  for (int r = 0; r < kStatReplicas; ++r) {
    const double* p = st + (((n * kStatReplicas + r) * C) + c) * 2;
    s += p[0]; q += p[1];
  }
  const double m = s * inv_n;
  double var = q * inv_n - m * m;
  var = var < 0 ? 0 : var;
  mean = (float)m;
  rstd = (float)(1.0 / sqrt(var + (double)eps));
}

// ------------------------------------------------------------------------------------------------ conv1: 7x7 s2, 3 -> 32
// fp32 image [img, 3, H, W] (NCHW, as dpvo.py:389 builds it) -> fp16 [img, OH, OW, 32] on the tensor cores.
// The image patch is staged in shared memory as fp16 with the 3 channels interleaved, so for a fixed kernel row ky
// the 7 x 3 = 21 inputs of an output pixel are CONTIGUOUS (starting 6 halves after its left neighbour's): that row is
// one K-slice of 32 (21 live + 11 columns whose weights are zero), two k16 steps.  K = 7 x 32 = 224 in all.
// One CTA = 8 x 16 output pixels, 4 warps x 2 rows, N = 32.  Weights: fp16 [ky][n][32] (k = kx * 3 + c, zero padded).
constexpr int kC1Out = 32, kC1TH = 8, kC1TW = 16;
constexpr int kC1Rows = (kC1TH - 1) * 2 + 7;          // 21 input rows
constexpr int kC1Cols = (kC1TW - 1) * 2 + 7;          // 37 input pixels
constexpr int kC1RowHalves = 136;                     // >= 6 * 15 + 32 = 122, 68 words: rows land on different banks
constexpr int kC1WP = 40;                             // padded K stride of the weight rows

__global__ void __launch_bounds__(128)
conv7x7s2_kernel(const float* __restrict__ img, const __half* __restrict__ wgt /*[7][32][32]*/,
                 const float* __restrict__ bias, int H, int W, int OH, int OW, __half* __restrict__ out,
                 double* __restrict__ stats, int relu) {
  __shared__ __align__(16) __half s_img[kC1Rows][kC1RowHalves];
  __shared__ __align__(16) __half s_w[7][kC1Out][kC1WP];
  __shared__ float s_red[4 * kC1Out * 2];
  const int n = blockIdx.z;
  const int ox0 = blockIdx.x * kC1TW, oy0 = blockIdx.y * kC1TH;
  const int ix0 = ox0 * 2 - 3, iy0 = oy0 * 2 - 3;
  for (int i = threadIdx.x; i < kC1Rows * kC1RowHalves; i += 128) {
    const int r = i / kC1RowHalves, e = i % kC1RowHalves;
    const int x = e / 3, c = e % 3;
    const int iy = iy0 + r, ix = ix0 + x;
    float v = 0.f;
    if (x < kC1Cols && iy >= 0 && iy < H && ix >= 0 && ix < W) v = img[((int64_t)(n * 3 + c) * H + iy) * W + ix];
    s_img[r][e] = __float2half_rn(v);
  }
  for (int i = threadIdx.x; i < 7 * kC1Out * 4; i += 128) {          // 32 halves = 4 x 16 bytes per (ky, n)
    const int v = i % 4, o = (i / 4) % kC1Out, ky = i / (4 * kC1Out);
    *reinterpret_cast<uint4*>(&s_w[ky][o][v * 8]) = *reinterpret_cast<const uint4*>(wgt + ((ky * kC1Out + o) * 32 + v * 8));
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  float acc[2][4][4];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[mt][nt][c] = 0.f;
#pragma unroll
  for (int ky = 0; ky < 7; ++ky)
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      uint32_t a[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const __half* r0 = &s_img[(wid * 2 + mt) * 2 + ky][6 * g + 16 * ks + 2 * t];
        const __half* r1 = r0 + 6 * 8;
        a[mt][0] = *reinterpret_cast<const uint32_t*>(r0);
        a[mt][1] = *reinterpret_cast<const uint32_t*>(r1);
        a[mt][2] = *reinterpret_cast<const uint32_t*>(r0 + 8);
        a[mt][3] = *reinterpret_cast<const uint32_t*>(r1 + 8);
      }
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        const __half* br = &s_w[ky][nt * 8 + g][16 * ks + 2 * t];
        const uint32_t b0 = *reinterpret_cast<const uint32_t*>(br);
        const uint32_t b1 = *reinterpret_cast<const uint32_t*>(br + 8);
        mma16816(acc[0][nt], a[0], b0, b1);
        mma16816(acc[1][nt], a[1], b0, b1);
      }
    }
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) {
    const int ch = nt * 8 + 2 * t;
    const float b0 = bias[ch], b1 = bias[ch + 1];
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      const int oy = oy0 + wid * 2 + mt;
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const int ox = ox0 + g + 8 * hf;
        if (oy < OH && ox < OW) {
          float v0 = acc[mt][nt][2 * hf] + b0, v1 = acc[mt][nt][2 * hf + 1] + b1;
          if (relu) { v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); }
          const __half2 h = __floats2half2_rn(v0, v1);
          *reinterpret_cast<__half2*>(out + (((int64_t)n * OH + oy) * OW + ox) * kC1Out + ch) = h;
          const float2 f = __half22float2(h);
          s0 += f.x; q0 += f.x * f.x; s1 += f.y; q1 += f.y * f.y;
        }
      }
    }
    if (stats) {
#pragma unroll
      for (int d = 4; d < 32; d <<= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, d); q0 += __shfl_xor_sync(0xffffffffu, q0, d);
        s1 += __shfl_xor_sync(0xffffffffu, s1, d); q1 += __shfl_xor_sync(0xffffffffu, q1, d);
      }
      if (g == 0) {
        float* r = s_red + (wid * kC1Out + ch) * 2;
        r[0] = s0; r[1] = q0; r[2] = s1; r[3] = q1;
      }
    }
  }
  if (stats) {
    __syncthreads();
    if (threadIdx.x < kC1Out * 2) {
      const double tot = (double)s_red[threadIdx.x] + (double)s_red[kC1Out * 2 + threadIdx.x] +
                         (double)s_red[kC1Out * 4 + threadIdx.x] + (double)s_red[kC1Out * 6 + threadIdx.x];
      const int rep = (blockIdx.x + blockIdx.y * gridDim.x) % kStatReplicas;
      atomicAdd(stats + ((int64_t)n * kStatReplicas + rep) * kC1Out * 2 + threadIdx.x, tot);
    }
  }
}

// ------------------------------------------------------------------------- KSxKS conv (3x3 / 1x1), stride 1 or 2, mma
// One CTA: 8 x 16 output pixels x NC output channels (NC <= 64; wider layers split over blockIdx.z).  4 warps, each 2
// output rows (two m16 tiles) x NC/8 n8 tiles.  Shared memory: the input patch [rows][cols][CIN + 8] and the weights
// [tap][NC][CIN + 8] (the +8 halves make both fragment load patterns bank-conflict free).
struct ConvParams {
  const __half* in;     // [img, H, W, CIN]
  const __half* wgt;    // [COUT, KS*KS, CIN]
  const float* bias;    // [COUT]
  __half* out;          // [img, OH, OW, COUT]
  double* stats;        // [img, replicas, COUT, 2] or null
  const double* in_stats;  // statistics of `in`: instance norm + ReLU applied while the patch is staged, or null
  const __half* skip;   // [img, OH, OW, COUT] or null: out = relu(skip + relu(conv)) (ResidualBlock without norm)
  int relu;             // out = relu(conv) (norm_fn='none' only: with a norm the ReLU follows the statistics)
  float eps;
  int H, W, OH, OW, COUT, nchunks;
  float scale;          // applied after the bias (conv2's "/ 4")
};

template <int KS, int S, int CIN, int NC, int MT>
__global__ void __launch_bounds__(128)
conv_mma_kernel(const ConvParams p) {
  constexpr int TH = 4 * MT, TW = 16, IH = (TH - 1) * S + KS, IW = (TW - 1) * S + KS, CP = CIN + 8, NT = NC / 8;
  constexpr int PAD = KS / 2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __half* s_in = reinterpret_cast<__half*>(smem_raw);                 // [IH][IW][CP]
  __half* s_w = s_in + IH * IW * CP;                                  // [KS*KS][NC][CP]
  __shared__ float s_red[4 * NC * 2];
  __shared__ float s_norm[CIN][2];
  const int n = blockIdx.z / p.nchunks, chunk = blockIdx.z % p.nchunks;
  if (p.in_stats) {
    if (threadIdx.x < CIN)
      plane_stats(p.in_stats, n, CIN, threadIdx.x, 1.0 / ((double)p.H * p.W), p.eps, s_norm[threadIdx.x][0], s_norm[threadIdx.x][1]);
    __syncthreads();
  }
  const int ox0 = blockIdx.x * TW, oy0 = blockIdx.y * TH;
  const int ix0 = ox0 * S - PAD, iy0 = oy0 * S - PAD;
  constexpr int CV = CIN / 8;                                         // 16-byte vectors per pixel
  // staging with cp.async: every 16-byte piece of the patch and of the weights is in flight at once (with plain
  // loads the kernel sat on the long scoreboard: ~50 dependent L2 round trips per thread, 4 warps per SM)
  for (int i = threadIdx.x; i < IH * IW * CV; i += 128) {
    const int v = i % CV, x = (i / CV) % IW, r = i / (CV * IW);
    const int iy = iy0 + r, ix = ix0 + x;
    const bool live = iy >= 0 && iy < p.H && ix >= 0 && ix < p.W;
    const __half* src = live ? p.in + (((int64_t)n * p.H + iy) * p.W + ix) * CIN + v * 8 : p.in;
    cp_async16(s_in + (r * IW + x) * CP + v * 8, src, live);          // out of bounds: zero fill
  }
  for (int i = threadIdx.x; i < KS * KS * NC * CV; i += 128) {
    const int v = i % CV, o = (i / CV) % NC, tap = i / (CV * NC);
    cp_async16(s_w + (tap * NC + o) * CP + v * 8, p.wgt + ((int64_t)(chunk * NC + o) * KS * KS + tap) * CIN + v * 8, true);
  }
  asm volatile("cp.async.commit_group;\n" ::: "memory");
  asm volatile("cp.async.wait_group 0;\n" ::: "memory");
  __syncthreads();
  if (p.in_stats) {
    // relu(InstanceNorm(in)) in place on the live pixels (the zero padding applies to the NORMALISED map)
    for (int i = threadIdx.x; i < IH * IW * CV; i += 128) {
      const int v = i % CV, x = (i / CV) % IW, r = i / (CV * IW);
      const int iy = iy0 + r, ix = ix0 + x;
      if (iy >= 0 && iy < p.H && ix >= 0 && ix < p.W) {
        uint4* q = reinterpret_cast<uint4*>(s_in + (r * IW + x) * CP + v * 8);
        uint4 val = *q;
        __half* h = reinterpret_cast<__half*>(&val);
#pragma unroll
        for (int c = 0; c < 8; ++c)
          h[c] = __float2half_rn(fmaxf((__half2float(h[c]) - s_norm[v * 8 + c][0]) * s_norm[v * 8 + c][1], 0.f));
        *q = val;
      }
    }
    __syncthreads();
  }

  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  float acc[MT][NT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[mt][nt][c] = 0.f;

  for (int tap = 0; tap < KS * KS; ++tap) {
    const int ky = tap / KS, kx = tap % KS;
#pragma unroll
    for (int ks = 0; ks < CIN / 16; ++ks) {
      uint32_t a[MT][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        const int py = wid * MT + mt;
        const __half* r0 = s_in + ((py * S + ky) * IW + (g * S + kx)) * CP + ks * 16 + 2 * t;
        const __half* r1 = r0 + 8 * S * CP;
        a[mt][0] = *reinterpret_cast<const uint32_t*>(r0);
        a[mt][1] = *reinterpret_cast<const uint32_t*>(r1);
        a[mt][2] = *reinterpret_cast<const uint32_t*>(r0 + 8);
        a[mt][3] = *reinterpret_cast<const uint32_t*>(r1 + 8);
      }
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const __half* br = s_w + (tap * NC + nt * 8 + g) * CP + ks * 16 + 2 * t;
        const uint32_t b0 = *reinterpret_cast<const uint32_t*>(br);
        const uint32_t b1 = *reinterpret_cast<const uint32_t*>(br + 8);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) mma16816(acc[mt][nt], a[mt], b0, b1);
      }
    }
  }

  // epilogue: bias, scale, fp16 store, per-channel statistics of the ROUNDED values over the valid pixels
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    const int ch = chunk * NC + nt * 8 + 2 * t;
    const float b0 = p.bias[ch], b1 = p.bias[ch + 1];
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      const int oy = oy0 + wid * MT + mt;
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const int ox = ox0 + g + 8 * hf;
        if (oy < p.OH && ox < p.OW) {
          float v0 = (acc[mt][nt][2 * hf] + b0) * p.scale, v1 = (acc[mt][nt][2 * hf + 1] + b1) * p.scale;
          const int64_t o = (((int64_t)n * p.OH + oy) * p.OW + ox) * p.COUT + ch;
          if (p.relu) { v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); }
          if (p.skip) {
            const float2 sk = __half22float2(*reinterpret_cast<const __half2*>(p.skip + o));
            v0 = fmaxf(sk.x + v0, 0.f); v1 = fmaxf(sk.y + v1, 0.f);
          }
          const __half2 h = __floats2half2_rn(v0, v1);
          *reinterpret_cast<__half2*>(p.out + o) = h;
          const float2 f = __half22float2(h);
          s0 += f.x; q0 += f.x * f.x; s1 += f.y; q1 += f.y * f.y;
        }
      }
    }
    if (p.stats) {
#pragma unroll
      for (int d = 4; d < 32; d <<= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, d); q0 += __shfl_xor_sync(0xffffffffu, q0, d);
        s1 += __shfl_xor_sync(0xffffffffu, s1, d); q1 += __shfl_xor_sync(0xffffffffu, q1, d);
      }
      if (g == 0) {
        float* r = s_red + (wid * NC + nt * 8 + 2 * t) * 2;
        r[0] = s0; r[1] = q0; r[2] = s1; r[3] = q1;
      }
    }
  }
  if (p.stats) {
    // the four warps' partial sums -> one float64 atomic per (channel, moment) per CTA, on this CTA's replica
    __syncthreads();
    if (threadIdx.x < NC * 2) {
      const double tot = (double)s_red[threadIdx.x] + (double)s_red[NC * 2 + threadIdx.x] +
                         (double)s_red[NC * 4 + threadIdx.x] + (double)s_red[NC * 6 + threadIdx.x];
      const int rep = (blockIdx.x + blockIdx.y * gridDim.x) % kStatReplicas;
      atomicAdd(p.stats + (((int64_t)n * kStatReplicas + rep) * p.COUT + chunk * NC) * 2 + threadIdx.x, tot);
    }
  }
}

template <int KS, int S, int CIN, int NC, int MT>
int launch_conv_mt(const ConvParams& p, int n_img, cudaStream_t st) {
  constexpr int IH = (4 * MT - 1) * S + KS, IW = 15 * S + KS, CP = CIN + 8;
  constexpr int smem = (IH * IW * CP + KS * KS * NC * CP) * 2;
  static bool once = false;
  if (!once) {
    cudaFuncSetAttribute(conv_mma_kernel<KS, S, CIN, NC, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    once = true;
  }
  dim3 grid((p.OW + 15) / 16, (p.OH + 4 * MT - 1) / (4 * MT), n_img * p.nchunks);
  conv_mma_kernel<KS, S, CIN, NC, MT><<<grid, 128, smem, st>>>(p);
  note_launches(1);
  return check_launch("encoder_conv");
}

// rows per warp: 2 (8 x 16 pixel tiles).  4 x 16 tiles (DPVO_ENC_MT=1: twice the CTAs, two resident per SM) were measured
// slower on every layer of a 480x640 frame (fnet 0.216 vs 0.191 ms): each CTA re-stages the layer's weights.
template <int KS, int S, int CIN, int NC>
int launch_conv(const ConvParams& p, int n_img, cudaStream_t st) {
  static const int forced = debug_env("DPVO_ENC_MT") ? atoi(debug_env("DPVO_ENC_MT")) : 0;
  const int mt = forced == 1 ? 1 : 2;
  return mt == 2 ? launch_conv_mt<KS, S, CIN, NC, 2>(p, n_img, st) : launch_conv_mt<KS, S, CIN, NC, 1>(p, n_img, st);
}

// ------------------------------------------------------------------- instance norm + ReLU (+ residual add + ReLU)
// out = relu(norm(raw))                                   (skip == null)
// out = relu(skipv + relu(norm(raw))), skipv = skip or norm(skip) when skip_stats is given (the down-sampled branch)
// norm(x) = (x - mean) * rsqrt(var + eps) with the biased variance of the (image, channel) plane; identity when
// stats == null (norm_fn='none').  fp32 arithmetic, one rounding to fp16.  8 channels (16 bytes) per thread.
constexpr int kNormMaxC = 128;

__global__ void __launch_bounds__(256)
norm_act_kernel(const __half* __restrict__ raw, const double* __restrict__ stats, const __half* __restrict__ skip,
                const double* __restrict__ skip_stats, int64_t hw, int C, float eps, __half* __restrict__ out,
                int64_t total_vec) {
  // mean / rstd of the image this block starts in, once per block (a block that straddles two images recomputes
  // the second image's values per thread)
  __shared__ float s_tab[2][kNormMaxC][2];
  const int cv = C / 8;
  const double inv_n = 1.0 / (double)hw;
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x;
  const int64_t n0 = i0 / ((int64_t)cv * hw);
  if (stats || skip_stats) {
    for (int x = threadIdx.x; x < 2 * C; x += blockDim.x) {
      const int which = x / C, c = x % C;
      const double* st = which ? skip_stats : stats;
      if (st) plane_stats(st, n0, C, c, inv_n, eps, s_tab[which][c][0], s_tab[which][c][1]);
    }
    __syncthreads();
  }
  const int64_t i = i0 + threadIdx.x;
  if (i >= total_vec) return;
  const int c0 = (int)(i % cv) * 8;
  const int64_t n = i / ((int64_t)cv * hw);
  const uint4 rv = reinterpret_cast<const uint4*>(raw)[i];
  uint4 sv = make_uint4(0, 0, 0, 0);
  if (skip) sv = reinterpret_cast<const uint4*>(skip)[i];
  const __half* rh = reinterpret_cast<const __half*>(&rv);
  const __half* sh = reinterpret_cast<const __half*>(&sv);
  __half oh[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    float v = __half2float(rh[c]);
    if (stats) {
      float m, r;
      if (n == n0) { m = s_tab[0][c0 + c][0]; r = s_tab[0][c0 + c][1]; }
      else plane_stats(stats, n, C, c0 + c, inv_n, eps, m, r);
      v = (v - m) * r;
    }
    v = fmaxf(v, 0.f);
    if (skip) {
      float s = __half2float(sh[c]);
      if (skip_stats) {
        float m, r;
        if (n == n0) { m = s_tab[1][c0 + c][0]; r = s_tab[1][c0 + c][1]; }
        else plane_stats(skip_stats, n, C, c0 + c, inv_n, eps, m, r);
        s = (s - m) * r;
      }
      v = fmaxf(s + v, 0.f);
    }
    oh[c] = __float2half_rn(v);
  }
  reinterpret_cast<uint4*>(out)[i] = *reinterpret_cast<const uint4*>(oh);
}

// F.avg_pool2d(x, 4, 4) on a channels-last fp16 map (dpvo.py:438 builds pyramid level 2 this way): fp32 sum, one rounding
__global__ void __launch_bounds__(256)
avgpool4_kernel(const __half* __restrict__ in, int H, int W, int C, int OH, int OW, __half* __restrict__ out, int64_t total_vec) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total_vec) return;
  const int cv = C / 8;
  const int v = (int)(i % cv);
  const int ox = (int)((i / cv) % OW), oy = (int)((i / ((int64_t)cv * OW)) % OH);
  const int64_t n = i / ((int64_t)cv * OW * OH);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int dy = 0; dy < 4; ++dy)
    for (int dx = 0; dx < 4; ++dx) {
      const uint4 q = *reinterpret_cast<const uint4*>(in + ((n * H + oy * 4 + dy) * W + ox * 4 + dx) * C + v * 8);
      const __half* h = reinterpret_cast<const __half*>(&q);
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[c] += __half2float(h[c]);
    }
  __half oh[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) oh[c] = __float2half_rn(acc[c] * (1.f / 16.f));
  reinterpret_cast<uint4*>(out)[i] = *reinterpret_cast<const uint4*>(oh);
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

extern "C" int dpvo_encoder_conv7x7s2(const float* images, const void* weight, const float* bias, int32_t n_img, int32_t H,
                                      int32_t W, int32_t relu, void* out, double* stats, void* stream) {
  DPVO_REQUIRE(n_img >= 0 && H >= 0 && W >= 0, "encoder_conv7x7s2: bad sizes");
  DPVO_REQUIRE(!(relu && stats), "encoder_conv7x7s2: relu and stats are exclusive (the ReLU follows the norm)");
  if (n_img == 0 || H == 0 || W == 0) return 0;
  DPVO_REQUIRE(images && weight && bias && out, "encoder_conv7x7s2: null pointer");
  DPVO_REQUIRE(((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(weight)) & 15) == 0,
               "encoder_conv7x7s2: weight / output must be 16-byte aligned");
  const int OH = (H + 6 - 7) / 2 + 1, OW = (W + 6 - 7) / 2 + 1;
  dim3 grid((OW + kC1TW - 1) / kC1TW, (OH + kC1TH - 1) / kC1TH, n_img);
  conv7x7s2_kernel<<<grid, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      images, reinterpret_cast<const __half*>(weight), bias, H, W, OH, OW, reinterpret_cast<__half*>(out), stats, relu);
  note_launches(1);
  return check_launch("encoder_conv7x7s2");
}

extern "C" int dpvo_encoder_conv(const void* in, const void* weight, const float* bias, int32_t n_img, int32_t H, int32_t W,
                                 int32_t cin, int32_t cout, int32_t ksize, int32_t stride, float scale,
                                 const double* in_stats, float eps, int32_t relu, const void* skip, void* out,
                                 double* stats, void* stream) {
  DPVO_REQUIRE(n_img >= 0 && H >= 0 && W >= 0, "encoder_conv: bad sizes");
  DPVO_REQUIRE(!((relu || skip) && stats), "encoder_conv: relu / skip and stats are exclusive (the ReLU follows the norm)");
  DPVO_REQUIRE((reinterpret_cast<uintptr_t>(skip) & 3) == 0, "encoder_conv: skip must be 4-byte aligned");
  if (n_img == 0 || H == 0 || W == 0) return 0;
  DPVO_REQUIRE(in && weight && bias && out, "encoder_conv: null pointer");
  DPVO_REQUIRE(((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(weight) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
               "encoder_conv: pointers must be 16-byte aligned");
  ConvParams p;
  p.in = reinterpret_cast<const __half*>(in); p.wgt = reinterpret_cast<const __half*>(weight); p.bias = bias;
  p.out = reinterpret_cast<__half*>(out); p.stats = stats;
  p.in_stats = in_stats; p.eps = eps; p.relu = relu; p.skip = reinterpret_cast<const __half*>(skip);
  const int pad = ksize / 2;
  p.H = H; p.W = W; p.OH = (H + 2 * pad - ksize) / stride + 1; p.OW = (W + 2 * pad - ksize) / stride + 1;
  p.COUT = cout; p.scale = scale;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int key = ksize * 100000 + stride * 10000 + cin * 100 + (cout == 32 ? 32 : 64);
  DPVO_REQUIRE(cout == 32 || cout % 64 == 0, "encoder_conv: cout must be 32 or a multiple of 64");
  p.nchunks = cout == 32 ? 1 : cout / 64;
  switch (key) {
    case 310000 + 3232: return launch_conv<3, 1, 32, 32>(p, n_img, st);
    case 320000 + 3264: return launch_conv<3, 2, 32, 64>(p, n_img, st);
    case 310000 + 6464: return launch_conv<3, 1, 64, 64>(p, n_img, st);
    case 120000 + 3264: return launch_conv<1, 2, 32, 64>(p, n_img, st);
    case 110000 + 6464: return launch_conv<1, 1, 64, 64>(p, n_img, st);
  }
  set_error("encoder_conv: unsupported layer (k=%d s=%d cin=%d cout=%d); BasicEncoder4 uses 3x3 s1 32->32, 3x3 s2 32->64, "
            "3x3 s1 64->64, 1x1 s2 32->64, 1x1 s1 64->64k", ksize, stride, cin, cout);
  return 1;
}

extern "C" int dpvo_encoder_norm_act(const void* raw, const double* stats, const void* skip, const double* skip_stats,
                                     int32_t n_img, int64_t hw, int32_t channels, float eps, void* out, void* stream) {
  DPVO_REQUIRE(n_img >= 0 && hw >= 0 && channels > 0 && channels % 8 == 0 && channels <= kNormMaxC,
               "encoder_norm_act: bad sizes");
  DPVO_REQUIRE(!skip_stats || skip, "encoder_norm_act: skip_stats without skip");
  const int64_t total = (int64_t)n_img * hw * (channels / 8);
  if (total == 0) return 0;
  DPVO_REQUIRE(raw && out, "encoder_norm_act: null pointer");
  norm_act_kernel<<<(unsigned)ceil_div64(total, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __half*>(raw), stats, reinterpret_cast<const __half*>(skip), skip_stats, hw, channels, eps,
      reinterpret_cast<__half*>(out), total);
  note_launches(1);
  return check_launch("encoder_norm_act");
}

extern "C" int dpvo_avgpool4_nhwc_f16(const void* in, int32_t n_img, int32_t H, int32_t W, int32_t channels, void* out,
                                      void* stream) {
  DPVO_REQUIRE(n_img >= 0 && H >= 0 && W >= 0 && channels > 0 && channels % 8 == 0, "avgpool4: bad sizes");
  const int OH = H / 4, OW = W / 4;
  const int64_t total = (int64_t)n_img * OH * OW * (channels / 8);
  if (total == 0) return 0;
  DPVO_REQUIRE(in && out, "avgpool4: null pointer");
  avgpool4_kernel<<<(unsigned)ceil_div64(total, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __half*>(in), H, W, channels, OH, OW, reinterpret_cast<__half*>(out), total);
  note_launches(1);
  return check_launch("avgpool4");
}
