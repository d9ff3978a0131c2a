// Locality order for the correlation lookup.
//
// DPVO's edge list is in append order (dpvo.py:215-221,362-375: "patches of the last 13 frames -> the new frame" is
// target-frame-major, "patches of the new frame -> the last 13 frames" is patch-major), so consecutive edges of the
// second kind hit 13 different frames of the feature ring.  Every frame is the target of ~2,000 edges whose 8x8
// windows cover its level-2 map ~135x and its level-1 map ~10x over; if the warps resident on an SM work on edges of
// the SAME target frame and the SAME spatial tile at the same time, most of those re-reads hit L1 (level 2: one
// 8x8-pixel tile + halo is 83 KB) instead of going to L2.  This file produces that order: a counting sort of the
// edges by (target-frame slot, 32x32-pixel tile of the patch centre at level 1) — three small launches, no host
// synchronisation; the lookup kernel then walks `perm` and still writes every edge's result at its own index, so
// results are bit-identical to the unordered launch.
#include "common.cuh"

namespace dpvo {
namespace {

constexpr int kOrdThreads = 256;

__device__ __forceinline__ int order_key(const float* __restrict__ coords, const int64_t* __restrict__ vs, int64_t e,
                                         int N2, int H1, int W1, int tile, int tiles_x, int tiles_y) {
  // centre pixel (i0 = j0 = 1) of the 3x3 patch: coords[e][0][1][1] (x), coords[e][1][1][1] (y)
  const float x = coords[e * 18 + 4], y = coords[e * 18 + 9 + 4];
  int tx = floor_to_int(x) / tile, ty = floor_to_int(y) / tile;
  tx = max(0, min(tiles_x - 1, tx));
  ty = max(0, min(tiles_y - 1, ty));
  int64_t v = vs[e];
  if (v < 0 || v >= N2) v = 0;                       // invalid edges produce zeros wherever they are scheduled
  (void)H1; (void)W1;
  return ((int)v * tiles_y + ty) * tiles_x + tx;
}

__global__ void __launch_bounds__(kOrdThreads)
order_hist_kernel(const float* __restrict__ coords, const int64_t* __restrict__ vs, int64_t M, int N2, int H1, int W1,
                  int tile, int tiles_x, int tiles_y, int* __restrict__ hist, int* __restrict__ keys) {
  for (int64_t e = (int64_t)blockIdx.x * kOrdThreads + threadIdx.x; e < M; e += (int64_t)gridDim.x * kOrdThreads) {
    const int k = order_key(coords, vs, e, N2, H1, W1, tile, tiles_x, tiles_y);
    keys[e] = k;
    atomicAdd(&hist[k], 1);
  }
}

// single CTA: exclusive scan of hist[nb] -> offs[nb] (nb is a few hundred .. a few thousand)
__global__ void __launch_bounds__(1024)
order_scan_kernel(const int* __restrict__ hist, int nb, int* __restrict__ offs) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += 1024) {
    const int k = base + tid;
    const int c = (k < nb) ? hist[k] : 0;
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
    if (k < nb) offs[k] = s_carry + (wid ? s_warp[wid - 1] : 0) + inc - c;
    __syncthreads();
    if (tid == 0) s_carry += s_warp[31];
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kOrdThreads)
order_scatter_kernel(const int* __restrict__ keys, int64_t M, int* __restrict__ cursor /* = offs, consumed */,
                     int32_t* __restrict__ perm) {
  for (int64_t e = (int64_t)blockIdx.x * kOrdThreads + threadIdx.x; e < M; e += (int64_t)gridDim.x * kOrdThreads) {
    const int pos = atomicAdd(&cursor[keys[e]], 1);
    perm[pos] = (int32_t)e;
  }
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

static inline int64_t align256o(int64_t x) { return (x + 255) & ~(int64_t)255; }

extern "C" int64_t dpvo_corr_order_workspace_bytes(int64_t M, int32_t N2, int32_t H1, int32_t W1, int32_t tile) {
  if (M < 0 || N2 <= 0 || H1 <= 0 || W1 <= 0 || tile <= 0) return -1;
  const int64_t nb = (int64_t)N2 * ((H1 + tile - 1) / tile) * ((W1 + tile - 1) / tile);
  return align256o(nb * 4) * 2 + align256o(M * 4);
}

extern "C" int dpvo_corr_order(const float* coords, const int64_t* vs, int64_t M, int32_t N2, int32_t H1, int32_t W1,
                               int32_t tile, int32_t* perm, void* ws, int64_t ws_bytes, void* stream) {
  DPVO_REQUIRE(M >= 0 && N2 > 0 && H1 > 0 && W1 > 0 && tile > 0, "corr_order: bad sizes");
  if (M == 0) return 0;
  DPVO_REQUIRE(coords && vs && perm && ws, "corr_order: null pointer");
  DPVO_REQUIRE(M < (int64_t(1) << 31), "corr_order: too many edges");
  const int tiles_x = (W1 + tile - 1) / tile, tiles_y = (H1 + tile - 1) / tile;
  const int64_t nb = (int64_t)N2 * tiles_x * tiles_y;
  DPVO_REQUIRE(nb < (int64_t(1) << 24), "corr_order: too many bins");
  DPVO_REQUIRE(ws_bytes >= dpvo_corr_order_workspace_bytes(M, N2, H1, W1, tile), "corr_order: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  char* base = reinterpret_cast<char*>(ws);
  int* hist = reinterpret_cast<int*>(base);
  int* offs = reinterpret_cast<int*>(base + align256o(nb * 4));
  int* keys = reinterpret_cast<int*>(base + 2 * align256o(nb * 4));
  if (cudaMemsetAsync(hist, 0, (size_t)nb * 4, st) != cudaSuccess) { set_error("corr_order: memset failed"); return 1; }
  const int grid = (int)(ceil_div64(M, kOrdThreads) < kNumSMs * 4 ? ceil_div64(M, kOrdThreads) : kNumSMs * 4);
  note_launches(3);
  order_hist_kernel<<<grid, kOrdThreads, 0, st>>>(coords, vs, M, N2, H1, W1, tile, tiles_x, tiles_y, hist, keys);
  order_scan_kernel<<<1, 1024, 0, st>>>(hist, (int)nb, offs);
  order_scatter_kernel<<<grid, kOrdThreads, 0, st>>>(keys, M, offs, perm);
  return check_launch("corr_order");
}
