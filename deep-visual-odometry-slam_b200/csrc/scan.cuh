// Device-wide exclusive prefix sum over int32 (three small kernels, no host sync).
// Used for CSR construction in `neighbors` and the block-E index build.
#pragma once
#include "common.cuh"

namespace dpvo {

constexpr int kScanThreads = 1024;
constexpr int kScanItems = 4;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int block_exclusive_scan(int val, int* s_warp, int& total) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = val;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    int x = (lane < (int)(blockDim.x >> 5)) ? s_warp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += t;
    }
    s_warp[lane] = x;
  }
  __syncthreads();
  total = s_warp[31];
  const int res = (wid ? s_warp[wid - 1] : 0) + inc - val;
  __syncthreads();
  return res;
}

static __global__ void __launch_bounds__(kScanThreads)
scan_reduce_kernel(const int* __restrict__ in, int64_t n, int* __restrict__ block_sums) {
  __shared__ int s_warp[32];
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i)
    if (base + i < n) s += in[base + i];
  int total;
  block_exclusive_scan(s, s_warp, total);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// single CTA: in-place exclusive scan of block_sums[nb]; writes grand total to *total_out (may be null)
static __global__ void __launch_bounds__(kScanThreads)
scan_sums_kernel(int* __restrict__ block_sums, int nb, int* __restrict__ total_out) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += kScanThreads) {
    const int i = base + threadIdx.x;
    const int v = (i < nb) ? block_sums[i] : 0;
    int total;
    const int ex = block_exclusive_scan(v, s_warp, total);
    if (i < nb) block_sums[i] = s_carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) s_carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0 && total_out) *total_out = s_carry;
}

static __global__ void __launch_bounds__(kScanThreads)
scan_apply_kernel(const int* __restrict__ in, int* __restrict__ out, int64_t n, const int* __restrict__ block_sums) {
  __shared__ int s_warp[32];
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int v[kScanItems];
  int s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = (base + i < n) ? in[base + i] : 0;
    s += v[i];
  }
  int total;
  int ex = block_exclusive_scan(s, s_warp, total) + block_sums[blockIdx.x];
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) out[base + i] = ex;
    ex += v[i];
  }
}

inline int64_t scan_workspace_ints(int64_t n) { return ceil_div64(n, kScanTile) + 8; }

// out may alias in.  block_sums: scan_workspace_ints(n) ints.
inline int exclusive_scan_int(const int* in, int* out, int64_t n, int* block_sums, int* total_out, cudaStream_t st) {
  if (n <= 0) return 0;
  const int64_t nb = ceil_div64(n, kScanTile);
  if (nb > (int64_t)1 << 30) { set_error("scan: input too large"); return 1; }
  scan_reduce_kernel<<<(unsigned)nb, kScanThreads, 0, st>>>(in, n, block_sums);
  scan_sums_kernel<<<1, kScanThreads, 0, st>>>(block_sums, (int)nb, total_out);
  scan_apply_kernel<<<(unsigned)nb, kScanThreads, 0, st>>>(in, out, n, block_sums);
  return check_launch("exclusive_scan");
}

}  // namespace dpvo
