// Feature-ring layout change: [N, C, HW] (reference NCHW ring, dpvo.py:72-73) ->
// [N, HW, C] (channels-last; a pixel's C channels are contiguous so the
// correlation gather reads whole 16-byte vectors).  Tiled through shared memory
// so both the read (along HW) and the write (along C) are coalesced.
#include "common.cuh"

namespace dpvo {

template <typename T>
__global__ void __launch_bounds__(256)
nchw_to_nhwc_kernel(const T* __restrict__ src, T* __restrict__ dst, int C, int64_t HW, int64_t tiles_hw, int tiles_c) {
  // tile: 64 pixels x 64 channels
  __shared__ T tile[64][64 + 2];
  const int64_t blk = blockIdx.x;
  const int64_t per_img = tiles_hw * tiles_c;
  const int64_t n = blk / per_img;
  const int64_t r = blk - n * per_img;
  const int tc = (int)(r / tiles_hw);
  const int64_t th = r - (int64_t)tc * tiles_hw;
  const int64_t p0 = th * 64;
  const int c0 = tc * 64;
  const T* s = src + n * (int64_t)C * HW;
  T* d = dst + n * (int64_t)C * HW;
  const int tx = threadIdx.x & 63;
  const int ty = threadIdx.x >> 6;  // 0..3
  for (int cc = ty; cc < 64; cc += 4) {
    const int c = c0 + cc;
    const int64_t p = p0 + tx;
    if (c < C && p < HW) tile[cc][tx] = s[(int64_t)c * HW + p];
  }
  __syncthreads();
  for (int pp = ty; pp < 64; pp += 4) {
    const int64_t p = p0 + pp;
    const int c = c0 + tx;
    if (c < C && p < HW) d[p * C + c] = tile[tx][pp];
  }
}

// 16-bit elements, even HW: 64 pixels x 128 channels per CTA through shared memory as packed channel
// pairs.  Reads: one warp instruction = 64 consecutive pixels of one channel (128 B); writes: one warp
// instruction = 64 consecutive channels of one pixel (128 B).  4-byte accesses on both sides.
__global__ void __launch_bounds__(256)
nchw_to_nhwc_u16_kernel(const unsigned short* __restrict__ src, unsigned short* __restrict__ dst, int C, int64_t HW,
                        int64_t tiles_hw, int tiles_c) {
  __shared__ uint32_t tile[64][64 + 1];          // [pixel][channel pair]
  const int64_t blk = blockIdx.x;
  const int64_t per_img = tiles_hw * tiles_c;
  const int64_t n = blk / per_img;
  const int64_t r = blk - n * per_img;
  const int tc = (int)(r / tiles_hw);
  const int64_t p0 = (r - (int64_t)tc * tiles_hw) * 64;
  const int c0 = tc * 128;
  const unsigned short* s = src + n * (int64_t)C * HW;
  unsigned short* d = dst + n * (int64_t)C * HW;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  // load: warp handles channel pair cp = wid, wid+8, ...; lane handles pixels 2*lane, 2*lane+1
  const int64_t p = p0 + 2 * lane;
  for (int cp = wid; cp < 64; cp += 8) {
    const int c = c0 + 2 * cp;
    uint32_t a = 0, b = 0;                        // a: channel c (pixels p, p+1), b: channel c+1
    if (p < HW) {                                 // HW even and p even: p+1 < HW too
      if (c < C) a = *reinterpret_cast<const uint32_t*>(s + (int64_t)c * HW + p);
      if (c + 1 < C) b = *reinterpret_cast<const uint32_t*>(s + (int64_t)(c + 1) * HW + p);
    }
    tile[2 * lane][cp] = __byte_perm(a, b, 0x5410);        // pixel p:   (c, c+1)
    tile[2 * lane + 1][cp] = __byte_perm(a, b, 0x7632);    // pixel p+1: (c, c+1)
  }
  __syncthreads();
  // store: warp handles pixel pp = wid, wid+8, ...; lane handles channel pairs lane, lane+32
  for (int pp = wid; pp < 64; pp += 8) {
    const int64_t q = p0 + pp;
    if (q >= HW) break;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int cp = lane + 32 * h;
      const int c = c0 + 2 * cp;
      const uint32_t v = tile[pp][cp];
      if (c + 1 < C) *reinterpret_cast<uint32_t*>(d + q * C + c) = v;
      else if (c < C) d[q * C + c] = (unsigned short)(v & 0xffffu);
    }
  }
}

template <typename T>
static int launch(const void* src, void* dst, int64_t N, int C, int64_t HW, cudaStream_t st) {
  const int64_t tiles_hw = ceil_div64(HW, 64);
  const int tiles_c = (C + 63) / 64;
  const int64_t blocks = N * tiles_hw * tiles_c;
  if (blocks == 0) return 0;
  if (blocks > 0x7fffffffLL) { set_error("nchw_to_nhwc: too many tiles"); return 1; }
  if (sizeof(T) == 2 && (HW % 2) == 0 && (C % 2) == 0) {
    const int tc2 = (C + 127) / 128;
    const int64_t blocks2 = N * tiles_hw * tc2;
    note_launches(1);
    nchw_to_nhwc_u16_kernel<<<(unsigned)blocks2, 256, 0, st>>>(
        reinterpret_cast<const unsigned short*>(src), reinterpret_cast<unsigned short*>(dst), C, HW, tiles_hw, tc2);
    return check_launch("nchw_to_nhwc");
  }
  note_launches(1);
  nchw_to_nhwc_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(
      reinterpret_cast<const T*>(src), reinterpret_cast<T*>(dst), C, HW, tiles_hw, tiles_c);
  return check_launch("nchw_to_nhwc");
}

}  // namespace dpvo

extern "C" int dpvo_nchw_to_nhwc(dpvo_dtype dtype, const void* src, void* dst,
                                 int64_t N, int32_t C, int64_t HW, void* stream) {
  DPVO_REQUIRE(src && dst, "nchw_to_nhwc: null pointer");
  DPVO_REQUIRE(N >= 0 && C > 0 && HW >= 0, "nchw_to_nhwc: bad sizes");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
    case DPVO_F16: return dpvo::launch<unsigned short>(src, dst, N, C, HW, st);
    case DPVO_F32: return dpvo::launch<float>(src, dst, N, C, HW, st);
    case DPVO_F64: return dpvo::launch<double>(src, dst, N, C, HW, st);
  }
  dpvo::set_error("nchw_to_nhwc: unknown dtype");
  return 1;
}
