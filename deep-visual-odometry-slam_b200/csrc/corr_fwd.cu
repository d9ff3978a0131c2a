// Correlation lookup forward (altcorr.corr) for sm_100a.
//
// Replaces /root/reference/methods/dpvo/dpvo/altcorr/correlation_kernel.cu:83-137
// (raw (2R+2)^2 window of C-channel dot products per patch pixel, fp32 accumulate,
// rounded to the feature dtype) fused with :222-233 (4-term bilinear blend done in
// the feature dtype, one rounding per elementwise op) and the final permute.
//
// Fast path (fp16, C=128, 3x3 patches, channels-last feature ring):
//   one warp per patch-graph edge, both pyramid levels in the same pass.
//   * the edge's gmap patch (9 x 128 halves, 2304 B) is staged once in shared
//     memory and turned into mma.m16n8k16 A fragments (rows = patch pixels);
//   * the union bounding box of the 9 patch pixels' windows (10x10 pixels
//     typical) is walked in tiles of 8 pixels; each quad of lanes loads one
//     pixel's 128 channels as 4 x 16-byte coalesced loads (64 contiguous bytes
//     per quad per load) straight into B fragments — the channel order inside
//     the contraction is permuted identically for A and B so no shuffle or
//     shared-memory staging of the fmap tile is needed;
//   * fp32 accumulators are rounded to fp16 (as the reference stores them) into
//     a per-warp shared-memory tile, then blended with the reference's rounding
//     points and written in the consumer's order.
// Out-of-bounds window pixels contribute 0 (never dereferenced).
//
// Generic path: any dtype (f16/f32/f64), any strides, any C/P/R; one CTA per edge.
#include "common.cuh"
#include <limits.h>

namespace dpvo {

struct CorrLevelDev {
  const void* fmap;
  int64_t sb, sn, sc, sy, sx;
  int N2, H2, W2;
  float scale;
  void* out;
  int64_t oes, oel;
};

struct CorrParams {
  const void* fmap1;
  int64_t g_sb, g_sn, g_sc, g_sy, g_sx;
  int N1, C, PH, PW;
  const float* coords;
  const int64_t* us;
  const int64_t* vs;
  int B, M, R, nlev;
  int pair;  // nlev==2 and the two outputs interleave element-wise (half2 stores)
  const int32_t* perm;  // optional processing order (csrc/corr_order.cu); results still land at the edge's own index
  CorrLevelDev lev[2];
};

// ---------------------------------------------------------------- fp16 helpers
// ATen evaluates every elementwise half op in fp32 and rounds once; mirror that.
__device__ __forceinline__ __half h_mul(__half a, __half b) {
  return __float2half_rn(__half2float(a) * __half2float(b));
}
__device__ __forceinline__ __half h_add(__half a, __half b) {
  return __float2half_rn(__half2float(a) + __half2float(b));
}
__device__ __forceinline__ __half h_one_minus(__half a) {
  return __float2half_rn(1.0f - __half2float(a));
}

__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// ------------------------------------------------------------------ fast path
constexpr int kWarps = 4;     // warps (edges in flight) per CTA
constexpr int kUMax = 160;    // max pixels in a union bounding box
constexpr int kRawStride = 168;  // halves; 84 words -> conflict-free fragment stores

// Native fp16 multiply / add are bit-identical to ATen's "compute in fp32, round once"
// for these two operations (the fp32 product of two halves is exact; for the sum, 24 bits
// >= 2*11 + 2 makes the double rounding innocuous), so the blend can use packed HMUL2 / HADD2.
// The _rn forms forbid contraction into HFMA2, which would drop a rounding.
__device__ __forceinline__ __half2 h2_mul(__half2 a, __half2 b) { return __hmul2_rn(a, b); }
__device__ __forceinline__ __half2 h2_add(__half2 a, __half2 b) { return __hadd2_rn(a, b); }
__device__ __forceinline__ __half2 as_h2(uint32_t v) { return *reinterpret_cast<__half2*>(&v); }
__device__ __forceinline__ uint32_t as_u32(__half2 v) { return *reinterpret_cast<uint32_t*>(&v); }

// Warp-uniform description of one pyramid level of one edge.
struct LevelBox {
  int x0, y0;      // top-left fmap pixel of the union bounding box (window origin included)
  int bw, nU;      // box width, number of box pixels
  bool uni;        // the 9 windows fit one box of <= kUMax pixels
  bool interior;   // ... and the box lies inside the feature map
};

// 32-byte read-only global load (sm_100: one LDG.256 per lane; four lanes cover a pixel's 128-byte line half)
__device__ __forceinline__ void ldg256(const void* p, uint32_t* w) {
  asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
               : "l"(p));
}

// WIDE: feature maps 32-byte aligned -> each lane reads 16 consecutive channels per load (two LDG.256 per
// tile, every 128-byte L1 wavefront fully used) instead of 8 (four LDG.128, half-used wavefronts).
template <int NLEV, bool WIDE>
__global__ void __launch_bounds__(kWarps * 32, 4)
corr_fwd_mma_f16_kernel(const __grid_constant__ CorrParams prm) {
  constexpr int R = 3, P = 3;
  constexpr int PP = P * P;
  constexpr int D = 2 * R + 2;
  constexpr int DO = D - 1;
  constexpr int C = 128;
  static_assert(PP <= 16 && D * D <= kUMax, "fast path limits");

  __shared__ __align__(16) __half s_patch[kWarps][PP * C];
  __shared__ __align__(16) __half s_raw[kWarps][NLEV][PP][kRawStride];
  __shared__ int s_bofs[kWarps][NLEV][PP];
  __shared__ int s_fxy[kWarps][NLEV][PP][2];
  __shared__ __align__(8) __half s_w[kWarps][NLEV][PP][4];

  const int lane = threadIdx.x & 31;
  const int wid = threadIdx.x >> 5;
  const int g = lane >> 2;
  const int t = lane & 3;
  const int64_t nedge = (int64_t)prm.B * prm.M;
  const __half* fmap1 = reinterpret_cast<const __half*>(prm.fmap1);

  for (int64_t es = (int64_t)blockIdx.x * kWarps + wid; es < nedge; es += (int64_t)gridDim.x * kWarps) {
    const int64_t e = prm.perm ? (int64_t)prm.perm[es] : es;
    const int b = (int)(e / prm.M);
    const int m = (int)(e - (int64_t)b * prm.M);
    const int64_t u = prm.us[m];
    const int64_t v = prm.vs[m];

    // ---- stage the gmap patch (dense blob of PP*C halves)
    {
      const bool u_ok = (u >= 0) && (u < prm.N1);
      const __half* gp = fmap1 + (int64_t)b * prm.g_sb + (u_ok ? u : 0) * prm.g_sn;
      uint4* dst = reinterpret_cast<uint4*>(&s_patch[wid][0]);
#pragma unroll
      for (int i = 0; i < (PP * C / 8 + 31) / 32; ++i)
        if (lane + 32 * i < PP * C / 8)
          dst[lane + 32 * i] = u_ok ? ldg128(gp + 8 * (lane + 32 * i)) : make_uint4(0, 0, 0, 0);
    }

    // ---- per level: window origins, blend weights, union bounding box (overlaps the patch load)
    LevelBox box[NLEV];
#pragma unroll
    for (int l = 0; l < NLEV; ++l) {
      const CorrLevelDev& L = prm.lev[l];
      float x = 0.f, y = 0.f;
      if (lane < PP) {
        const float* cp = prm.coords + e * (2 * PP) + lane;
        x = cp[0] * L.scale;
        y = cp[PP] * L.scale;
      }
      const int fx = floor_to_int(x);
      const int fy = floor_to_int(y);
      const int mnx = __reduce_min_sync(0xffffffffu, lane < PP ? fx : INT_MAX);
      const int mny = __reduce_min_sync(0xffffffffu, lane < PP ? fy : INT_MAX);
      const int mxx = __reduce_max_sync(0xffffffffu, lane < PP ? fx : INT_MIN);
      const int mxy = __reduce_max_sync(0xffffffffu, lane < PP ? fy : INT_MIN);
      const int ubw = mxx - mnx + D;
      const int ubh = mxy - mny + D;
      const bool uni = (ubw <= kUMax) && (ubh <= kUMax) && (ubw * ubh <= kUMax);
      if (lane < PP) {
        const __half hdx = __float2half_rn(x - floorf(x));
        const __half hdy = __float2half_rn(y - floorf(y));
        const __half mdx = h_one_minus(hdx), mdy = h_one_minus(hdy);
        s_w[wid][l][lane][0] = h_mul(mdx, mdy);
        s_w[wid][l][lane][1] = h_mul(hdx, mdy);
        s_w[wid][l][lane][2] = h_mul(mdx, hdy);
        s_w[wid][l][lane][3] = h_mul(hdx, hdy);
        s_bofs[wid][l][lane] = uni ? (fy - mny) * ubw + (fx - mnx) : 0;
        s_fxy[wid][l][lane][0] = fx;
        s_fxy[wid][l][lane][1] = fy;
      }
      box[l].x0 = mnx - R;
      box[l].y0 = mny - R;
      box[l].bw = uni ? ubw : D;
      box[l].nU = ubw * ubh;
      box[l].uni = uni;
      box[l].interior = uni && (mnx - R >= 0) && (mny - R >= 0) && (mnx - R + ubw <= L.W2) && (mny - R + ubh <= L.H2);
    }
    __syncwarp();

    // ---- A fragments: rows = patch pixels, contraction order permuted identically to the B loads
    uint32_t a[8][4];
    {
      const unsigned short* sp = reinterpret_cast<const unsigned short*>(&s_patch[wid][0]);
      const int p_lo = g, p_hi = g + 8;
      const int off_lo = (p_lo < PP) ? (int)((p_lo / P) * prm.g_sy + (p_lo % P) * prm.g_sx) : -1;
      const int off_hi = (p_hi < PP) ? (int)((p_hi / P) * prm.g_sy + (p_hi % P) * prm.g_sx) : -1;
      const int sc = (int)prm.g_sc;
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const int c0 = WIDE ? 64 * (s >> 2) + 16 * t + 4 * (s & 3) : 32 * (s >> 1) + 8 * t + 4 * (s & 1);
        uint32_t lo01 = 0, lo23 = 0, hi01 = 0, hi23 = 0;
        if (off_lo >= 0) {
          lo01 = (uint32_t)sp[off_lo + (c0 + 0) * sc] | ((uint32_t)sp[off_lo + (c0 + 1) * sc] << 16);
          lo23 = (uint32_t)sp[off_lo + (c0 + 2) * sc] | ((uint32_t)sp[off_lo + (c0 + 3) * sc] << 16);
        }
        if (off_hi >= 0) {
          hi01 = (uint32_t)sp[off_hi + (c0 + 0) * sc] | ((uint32_t)sp[off_hi + (c0 + 1) * sc] << 16);
          hi23 = (uint32_t)sp[off_hi + (c0 + 2) * sc] | ((uint32_t)sp[off_hi + (c0 + 3) * sc] << 16);
        }
        a[s][0] = lo01; a[s][1] = hi01; a[s][2] = lo23; a[s][3] = hi23;
      }
    }

    // ---- contraction over the window pixels, 8 pixels per mma tile, loads two tiles ahead
#pragma unroll
    for (int l = 0; l < NLEV; ++l) {
      const CorrLevelDev& L = prm.lev[l];
      const __half* fm = reinterpret_cast<const __half*>(L.fmap);
      const bool v_ok = (v >= 0) && (v < L.N2) && (u >= 0) && (u < prm.N1);
      const __half* fmv = fm + (int64_t)b * L.sb + (v_ok ? v : 0) * L.sn;
      if (box[l].uni) {
        const int x0 = box[l].x0, y0 = box[l].y0, bw = box[l].bw, nU = box[l].nU;
        const bool interior = box[l].interior;
        const int ntile = (nU + 7) >> 3;
        const int sx = (int)L.sx, sy = (int)L.sy;
        // this lane's pixel walks the box row-major in steps of 8 (bw >= 8: at most one wrap per step)
        int uidx = g, ux = g, py = y0;
        const __half* ptr = fmv + (int64_t)y0 * sy + (int64_t)(x0 + g) * sx + (WIDE ? 16 : 8) * t;
        const int step = 8 * sx, wrap = sy - bw * sx;

        auto load_tile = [&](uint32_t (&q)[16]) {
          bool ok = v_ok && (uidx < nU);
          if (!interior) {
            const int px = x0 + ux;
            ok = ok && (py >= 0) && (py < L.H2) && (px >= 0) && (px < L.W2);
          }
          if (ok) {
            if (WIDE) {
              ldg256(ptr, q);
              ldg256(ptr + 64, q + 8);
            } else {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const uint4 v4 = ldg128(ptr + 32 * j);
                q[4 * j] = v4.x; q[4 * j + 1] = v4.y; q[4 * j + 2] = v4.z; q[4 * j + 3] = v4.w;
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) q[j] = 0u;
          }
          uidx += 8;
          ux += 8;
          ptr += step;
          if (ux >= bw) { ux -= bw; ++py; ptr += wrap; }
        };
        auto compute_tile = [&](const uint32_t (&q)[16], int tile) {
          float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 8; ++j) mma_16816(acc, a[j], q[2 * j], q[2 * j + 1]);
          const int col = tile * 8 + 2 * t;
          *reinterpret_cast<__half2*>(&s_raw[wid][l][g][col]) = __floats2half2_rn(acc[0], acc[1]);   // g < 8 < PP
          if (g + 8 < PP)
            *reinterpret_cast<__half2*>(&s_raw[wid][l][g + 8][col]) = __floats2half2_rn(acc[2], acc[3]);
        };

        uint32_t q0[16], q1[16], q2[16];
        load_tile(q0);
        load_tile(q1);
        for (int tile = 0; tile < ntile; tile += 3) {
          load_tile(q2);
          compute_tile(q0, tile);
          if (tile + 1 < ntile) { load_tile(q0); compute_tile(q1, tile + 1); }
          if (tile + 2 < ntile) { load_tile(q1); compute_tile(q2, tile + 2); }
        }
      } else {
        // Patch pixels too far apart for one box (large scale change): plain per-window dot
        // products.  Rare; correctness path.
        const unsigned short* sp = reinterpret_cast<const unsigned short*>(&s_patch[wid][0]);
        const int sc = (int)prm.g_sc;
        for (int idx = lane; idx < PP * D * D; idx += 32) {
          const int p = idx / (D * D);
          const int w = idx - p * D * D;
          const int py = s_fxy[wid][l][p][1] - R + w / D;
          const int px = s_fxy[wid][l][p][0] - R + w % D;
          float s = 0.f;
          if (v_ok && py >= 0 && py < L.H2 && px >= 0 && px < L.W2) {
            const __half* fp = fmv + (int64_t)py * L.sy + (int64_t)px * L.sx;
            const int off = (int)((p / P) * prm.g_sy + (p % P) * prm.g_sx);
            for (int c8 = 0; c8 < C / 8; ++c8) {
              const uint4 q = ldg128(fp + 8 * c8);
              const __half2* qh = reinterpret_cast<const __half2*>(&q);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float2 f = __half22float2(qh[i]);
                const __half_raw r0 = {sp[off + (8 * c8 + 2 * i) * sc]};
                const __half_raw r1 = {sp[off + (8 * c8 + 2 * i + 1) * sc]};
                s = fmaf(__half2float(__half(r0)), f.x, s);
                s = fmaf(__half2float(__half(r1)), f.y, s);
              }
            }
          }
          s_raw[wid][l][p][w] = __float2half_rn(s);
        }
      }
    }
    __syncwarp();

    // ---- bilinear blend (reference rounding points) + write in consumer order.
    // lane <-> (window row yo, patch pixel p): 63 pairs in two passes; each lane produces the 7
    // window columns of its row for all levels, as packed (level0, level1) half2 values.  For a
    // fixed column the 63 lanes write 63 consecutive outputs -> coalesced stores.
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
      const int idx = pass * 32 + lane;
      if (idx < DO * PP) {
        const int yo = idx / PP;
        const int p = idx - yo * PP;
        uint32_t hv[NLEV][2][4];   // [level][row yo / yo+1][4 words = 8 window columns]
        uint32_t wv[NLEV][2];      // blend weights (w00,w01),(w10,w11)
#pragma unroll
        for (int l = 0; l < NLEV; ++l) {
          const uint32_t* rw = reinterpret_cast<const uint32_t*>(&s_raw[wid][l][0][0]);
          const int bw = box[l].bw;
          const int h0 = p * kRawStride + s_bofs[wid][l][p] + yo * bw;
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            const int h = h0 + r * bw;
            const uint32_t* src = rw + (h >> 1);
            uint32_t w5[5];
#pragma unroll
            for (int i = 0; i < 5; ++i) w5[i] = src[i];
            const uint32_t sel = (h & 1) ? 0x5432u : 0x3210u;
#pragma unroll
            for (int i = 0; i < 4; ++i) hv[l][r][i] = __byte_perm(w5[i], w5[i + 1], sel);
          }
          const uint2 wq = *reinterpret_cast<const uint2*>(&s_w[wid][l][p][0]);
          wv[l][0] = wq.x;
          wv[l][1] = wq.y;
        }
        constexpr int L1 = NLEV - 1;
        const __half2 W00 = as_h2(__byte_perm(wv[0][0], wv[L1][0], 0x5410));
        const __half2 W01 = as_h2(__byte_perm(wv[0][0], wv[L1][0], 0x7632));
        const __half2 W10 = as_h2(__byte_perm(wv[0][1], wv[L1][1], 0x5410));
        const __half2 W11 = as_h2(__byte_perm(wv[0][1], wv[L1][1], 0x7632));
        __half2 c0[D], c1[D];      // (level0, level1) raw values of rows yo and yo+1
#pragma unroll
        for (int x = 0; x < D; ++x) {
          const uint32_t sel = (x & 1) ? 0x7632u : 0x5410u;
          c0[x] = as_h2(__byte_perm(hv[0][0][x >> 1], hv[L1][0][x >> 1], sel));
          c1[x] = as_h2(__byte_perm(hv[0][1][x >> 1], hv[L1][1][x >> 1], sel));
        }
#pragma unroll
        for (int xo = 0; xo < DO; ++xo) {
          __half2 acc = h2_mul(W00, c0[xo]);
          acc = h2_add(acc, h2_mul(W01, c0[xo + 1]));
          acc = h2_add(acc, h2_mul(W10, c1[xo]));
          acc = h2_add(acc, h2_mul(W11, c1[xo + 1]));
          const int64_t o = (xo * DO + yo) * PP + p;
          if (NLEV == 2 && prm.pair) {
            __half* op = reinterpret_cast<__half*>(prm.lev[0].out) + e * prm.lev[0].oes + 2 * o;
            *reinterpret_cast<__half2*>(op) = acc;
          } else {
            __half* op0 = reinterpret_cast<__half*>(prm.lev[0].out) + e * prm.lev[0].oes + o * prm.lev[0].oel;
            *op0 = __low2half(acc);
            if (NLEV == 2) {
              __half* op1 = reinterpret_cast<__half*>(prm.lev[L1].out) + e * prm.lev[L1].oes + o * prm.lev[L1].oel;
              *op1 = __high2half(acc);
            }
          }
        }
      }
    }
    __syncwarp();
  }
}

// --------------------------------------------------------------- generic path
template <typename T> struct Ar;
template <> struct Ar<float> {
  static __device__ __forceinline__ float from(float v) { return v; }
  static __device__ __forceinline__ float to_f(float v) { return v; }
  static __device__ __forceinline__ float mul(float a, float b) { return a * b; }
  static __device__ __forceinline__ float add(float a, float b) { return a + b; }
  static __device__ __forceinline__ float one_minus(float a) { return 1.0f - a; }
};
template <> struct Ar<double> {
  static __device__ __forceinline__ double from(float v) { return (double)v; }
  static __device__ __forceinline__ float to_f(double v) { return (float)v; }
  static __device__ __forceinline__ double mul(double a, double b) { return a * b; }
  static __device__ __forceinline__ double add(double a, double b) { return a + b; }
  static __device__ __forceinline__ double one_minus(double a) { return 1.0 - a; }
};
template <> struct Ar<__half> {
  static __device__ __forceinline__ __half from(float v) { return __float2half_rn(v); }
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half mul(__half a, __half b) { return h_mul(a, b); }
  static __device__ __forceinline__ __half add(__half a, __half b) { return h_add(a, b); }
  static __device__ __forceinline__ __half one_minus(__half a) { return h_one_minus(a); }
};

// One CTA per edge.  Dynamic smem: T raw[PP*D*D], T w[PP*4], int fxy[PP*2].
template <typename T>
__global__ void __launch_bounds__(128)
corr_fwd_generic_kernel(const __grid_constant__ CorrParams prm) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int PP = prm.PH * prm.PW;
  const int R = prm.R, D = 2 * R + 2, DO = D - 1;
  T* s_raw = reinterpret_cast<T*>(smem_raw);
  T* s_w = s_raw + PP * D * D;
  int* s_f = reinterpret_cast<int*>(s_w + PP * 4);
  const T* fmap1 = reinterpret_cast<const T*>(prm.fmap1);
  const int64_t nedge = (int64_t)prm.B * prm.M;

  for (int64_t e = blockIdx.x; e < nedge; e += gridDim.x) {
    const int b = (int)(e / prm.M);
    const int m = (int)(e - (int64_t)b * prm.M);
    const int64_t u = prm.us[m];
    const int64_t v = prm.vs[m];
    const bool u_ok = (u >= 0) && (u < prm.N1);
    for (int l = 0; l < prm.nlev; ++l) {
      const CorrLevelDev& L = prm.lev[l];
      const T* fm = reinterpret_cast<const T*>(L.fmap);
      const bool v_ok = u_ok && (v >= 0) && (v < L.N2);
      __syncthreads();
      if (threadIdx.x < PP) {
        const int p = threadIdx.x;
        const float x = prm.coords[e * (2 * PP) + p] * L.scale;
        const float y = prm.coords[e * (2 * PP) + PP + p] * L.scale;
        s_f[2 * p] = floor_to_int(x);
        s_f[2 * p + 1] = floor_to_int(y);
        const T dx = Ar<T>::from(x - floorf(x));
        const T dy = Ar<T>::from(y - floorf(y));
        const T mdx = Ar<T>::one_minus(dx), mdy = Ar<T>::one_minus(dy);
        s_w[4 * p + 0] = Ar<T>::mul(mdx, mdy);
        s_w[4 * p + 1] = Ar<T>::mul(dx, mdy);
        s_w[4 * p + 2] = Ar<T>::mul(mdx, dy);
        s_w[4 * p + 3] = Ar<T>::mul(dx, dy);
      }
      __syncthreads();
      for (int idx = threadIdx.x; idx < PP * D * D; idx += blockDim.x) {
        const int p = idx / (D * D);
        const int w = idx - p * D * D;
        const int wy = w / D, wx = w - wy * D;
        const int i1 = s_f[2 * p + 1] + wy - R;
        const int j1 = s_f[2 * p] + wx - R;
        float s = 0.f;
        if (v_ok && i1 >= 0 && i1 < L.H2 && j1 >= 0 && j1 < L.W2) {
          const T* p1 = fmap1 + (int64_t)b * prm.g_sb + u * prm.g_sn + (p / prm.PW) * prm.g_sy + (p % prm.PW) * prm.g_sx;
          const T* p2 = fm + (int64_t)b * L.sb + v * L.sn + (int64_t)i1 * L.sy + (int64_t)j1 * L.sx;
          for (int c = 0; c < prm.C; ++c)
            s += Ar<T>::to_f(p1[c * prm.g_sc]) * Ar<T>::to_f(p2[c * L.sc]);
        }
        s_raw[idx] = Ar<T>::from(s);
      }
      __syncthreads();
      T* out = reinterpret_cast<T*>(L.out) + e * L.oes;
      for (int o = threadIdx.x; o < DO * DO * PP; o += blockDim.x) {
        const int p = o % PP;
        const int r = o / PP;
        const int yo = r % DO, xo = r / DO;
        const T* rw = s_raw + p * D * D + yo * D + xo;
        T acc = Ar<T>::mul(s_w[4 * p + 0], rw[0]);
        acc = Ar<T>::add(acc, Ar<T>::mul(s_w[4 * p + 1], rw[1]));
        acc = Ar<T>::add(acc, Ar<T>::mul(s_w[4 * p + 2], rw[D]));
        acc = Ar<T>::add(acc, Ar<T>::mul(s_w[4 * p + 3], rw[D + 1]));
        out[(int64_t)o * L.oel] = acc;
      }
    }
  }
}

template <typename T>
static int launch_generic(const CorrParams& prm, cudaStream_t st) {
  const int PP = prm.PH * prm.PW, D = 2 * prm.R + 2;
  size_t smem = (size_t)PP * D * D * sizeof(T) + (size_t)PP * 4 * sizeof(T) + (size_t)PP * 2 * sizeof(int) + 16;
  if (smem > 200 * 1024) { set_error("corr_forward: window too large for shared memory"); return 1; }
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(corr_fwd_generic_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int64_t nedge = (int64_t)prm.B * prm.M;
  int grid = (int)(nedge < (int64_t)kNumSMs * 64 ? nedge : (int64_t)kNumSMs * 64);
  note_launches(1);
  corr_fwd_generic_kernel<T><<<grid, 128, smem, st>>>(prm);
  return check_launch("corr_fwd_generic");
}

static bool is_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace dpvo

using namespace dpvo;

static const char* g_last_corr_kernel = "";
extern "C" const char* dpvo_corr_last_kernel(void) { return g_last_corr_kernel; }

static int corr_forward_impl(dpvo_dtype dtype,
                             const void* fmap1, int64_t g_sb, int64_t g_sn, int64_t g_sc, int64_t g_sy, int64_t g_sx,
                             int32_t N1, int32_t C, int32_t PH, int32_t PW,
                             const float* coords, const int64_t* us, const int64_t* vs,
                             int32_t B, int32_t M, int32_t radius,
                             const dpvo_corr_level* levels, int32_t nlev, const int32_t* perm, void* stream) {
  DPVO_REQUIRE(nlev == 1 || nlev == 2, "corr_forward: nlev must be 1 or 2 (got %d)", nlev);
  DPVO_REQUIRE(radius >= 0 && PH > 0 && PW > 0 && C > 0, "corr_forward: bad sizes");
  DPVO_REQUIRE(B >= 0 && M >= 0, "corr_forward: negative batch/edge count");
  if ((int64_t)B * M == 0) return 0;
  DPVO_REQUIRE(fmap1 && coords && us && vs && levels, "corr_forward: null pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);

  CorrParams prm;
  memset(&prm, 0, sizeof(prm));
  prm.fmap1 = fmap1;
  prm.g_sb = g_sb; prm.g_sn = g_sn; prm.g_sc = g_sc; prm.g_sy = g_sy; prm.g_sx = g_sx;
  prm.N1 = N1; prm.C = C; prm.PH = PH; prm.PW = PW;
  prm.coords = coords; prm.us = us; prm.vs = vs;
  prm.B = B; prm.M = M; prm.R = radius; prm.nlev = nlev;
  prm.perm = perm;
  const int D = 2 * radius + 2;
  for (int l = 0; l < nlev; ++l) {
    DPVO_REQUIRE(levels[l].fmap && levels[l].out, "corr_forward: null level pointer");
    prm.lev[l].fmap = levels[l].fmap;
    prm.lev[l].sb = levels[l].sb; prm.lev[l].sn = levels[l].sn; prm.lev[l].sc = levels[l].sc;
    prm.lev[l].sy = levels[l].sy; prm.lev[l].sx = levels[l].sx;
    prm.lev[l].N2 = levels[l].N2; prm.lev[l].H2 = levels[l].H2; prm.lev[l].W2 = levels[l].W2;
    prm.lev[l].scale = levels[l].coord_scale;
    prm.lev[l].out = levels[l].out;
    prm.lev[l].oes = levels[l].out_edge_stride; prm.lev[l].oel = levels[l].out_elem_stride;
  }

  // Fast path eligibility: fp16, C=128, 3x3 patch stored as one dense 16B-aligned
  // blob per patch, radius 3, channels-last 16B-aligned feature maps.
  bool fast = (dtype == DPVO_F16) && C == 128 && PH == 3 && PW == 3 && radius == 3;
  if (fast) {
    const int64_t span = (C - 1) * g_sc + (PH - 1) * g_sy + (PW - 1) * g_sx + 1;
    fast = g_sc > 0 && g_sy > 0 && g_sx > 0 && span == (int64_t)C * PH * PW &&
           is_aligned16(fmap1) && (g_sn % 8 == 0) && (g_sb % 8 == 0);
    for (int l = 0; l < nlev && fast; ++l) {
      const CorrLevelDev& L = prm.lev[l];
      fast = L.sc == 1 && is_aligned16(L.fmap) && (L.sx % 8 == 0) && (L.sy % 8 == 0) &&
             (L.sn % 8 == 0) && (L.sb % 8 == 0);
    }
  }
  if (fast) {
    prm.pair = 0;
    if (nlev == 2) {
      const char* o0 = reinterpret_cast<const char*>(prm.lev[0].out);
      const char* o1 = reinterpret_cast<const char*>(prm.lev[1].out);
      prm.pair = (o1 == o0 + 2) && prm.lev[0].oel == 2 && prm.lev[1].oel == 2 &&
                 prm.lev[0].oes == prm.lev[1].oes && (prm.lev[0].oes % 2 == 0) &&
                 ((reinterpret_cast<uintptr_t>(o0) & 3) == 0);
    }
    const int64_t nedge = (int64_t)B * M;
    int64_t want = ceil_div64(nedge, kWarps);
    int grid = (int)(want < (int64_t)kNumSMs * 64 ? want : (int64_t)kNumSMs * 64);
    static const bool narrow_env = debug_env("DPVO_CORR_NARROW") != nullptr;   // read once (debug switch)
    bool wide = !narrow_env;
    for (int l = 0; l < nlev && wide; ++l) {
      const CorrLevelDev& L = prm.lev[l];
      wide = ((reinterpret_cast<uintptr_t>(L.fmap) & 31) == 0) && (L.sx % 16 == 0) && (L.sy % 16 == 0) &&
             (L.sn % 16 == 0) && (L.sb % 16 == 0);
    }
    if (nlev == 1) {
      if (wide) corr_fwd_mma_f16_kernel<1, true><<<grid, kWarps * 32, 0, st>>>(prm);
      else corr_fwd_mma_f16_kernel<1, false><<<grid, kWarps * 32, 0, st>>>(prm);
    } else {
      if (wide) corr_fwd_mma_f16_kernel<2, true><<<grid, kWarps * 32, 0, st>>>(prm);
      else corr_fwd_mma_f16_kernel<2, false><<<grid, kWarps * 32, 0, st>>>(prm);
    }
    g_last_corr_kernel = nlev == 1 ? (wide ? "corr_fwd_mma_f16_kernel<1, 1>" : "corr_fwd_mma_f16_kernel<1, 0>")
                                   : (wide ? "corr_fwd_mma_f16_kernel<2, 1>" : "corr_fwd_mma_f16_kernel<2, 0>");
    note_launches(1);
    return check_launch("corr_fwd_mma_f16");
  }
  (void)D;
  g_last_corr_kernel = dtype == DPVO_F16 ? "corr_fwd_generic_kernel<__half>"
                       : dtype == DPVO_F32 ? "corr_fwd_generic_kernel<float>" : "corr_fwd_generic_kernel<double>";
  switch (dtype) {
    case DPVO_F16: return launch_generic<__half>(prm, st);
    case DPVO_F32: return launch_generic<float>(prm, st);
    case DPVO_F64: return launch_generic<double>(prm, st);
  }
  set_error("corr_forward: unknown dtype %d", (int)dtype);
  return 1;
}

extern "C" int dpvo_corr_forward(dpvo_dtype dtype,
                                 const void* fmap1, int64_t g_sb, int64_t g_sn, int64_t g_sc, int64_t g_sy, int64_t g_sx,
                                 int32_t N1, int32_t C, int32_t PH, int32_t PW,
                                 const float* coords, const int64_t* us, const int64_t* vs,
                                 int32_t B, int32_t M, int32_t radius,
                                 const dpvo_corr_level* levels, int32_t nlev, void* stream) {
  return corr_forward_impl(dtype, fmap1, g_sb, g_sn, g_sc, g_sy, g_sx, N1, C, PH, PW, coords, us, vs, B, M, radius, levels,
                           nlev, nullptr, stream);
}

// Same lookup walked in the order `perm` (a permutation of [0, B*M) from dpvo_corr_order, or any other); every edge's
// result is written at its own index, so the output is bit-identical to dpvo_corr_forward.  The generic (non fp16 /
// non channels-last) path ignores `perm`.
extern "C" int dpvo_corr_forward_ordered(dpvo_dtype dtype,
                                         const void* fmap1, int64_t g_sb, int64_t g_sn, int64_t g_sc, int64_t g_sy,
                                         int64_t g_sx, int32_t N1, int32_t C, int32_t PH, int32_t PW,
                                         const float* coords, const int64_t* us, const int64_t* vs,
                                         int32_t B, int32_t M, int32_t radius,
                                         const dpvo_corr_level* levels, int32_t nlev, const int32_t* perm, void* stream) {
  return corr_forward_impl(dtype, fmap1, g_sb, g_sn, g_sc, g_sy, g_sx, N1, C, PH, PW, coords, us, vs, B, M, radius, levels,
                           nlev, perm, stream);
}
