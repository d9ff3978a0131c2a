// Feature-ring layout change: [N, C, HW] (reference NCHW ring, dpvo.py:72-73) ->
// [N, HW, C] (channels-last; a pixel's C channels are contiguous so the
// correlation gather reads whole 16-byte vectors).  Tiled through shared memory
// so both the read (along HW) and the write (along C) are coalesced.
#include "common.cuh"
#include <algorithm>

namespace dpvo {

template <typename T>
__global__ void __launch_bounds__(256)
nchw_to_nhwc_kernel(const T* __restrict__ src, T* __restrict__ dst, int C, int64_t HW, int64_t tiles_hw, int tiles_c,
                    const unsigned long long* __restrict__ hnew, const unsigned long long* __restrict__ hold) {
  // tile: 64 pixels x 64 channels
  __shared__ T tile[64][64 + 2];
  const int64_t blk = blockIdx.x;
  const int64_t per_img = tiles_hw * tiles_c;
  const int64_t n = blk / per_img;
  if (hnew != nullptr && hnew[n] == hold[n]) return;   // slot unchanged since the last sync
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
                        int64_t tiles_hw, int tiles_c,
                        const unsigned long long* __restrict__ hnew, const unsigned long long* __restrict__ hold) {
  __shared__ uint32_t tile[64][64 + 1];          // [pixel][channel pair]
  const int64_t blk = blockIdx.x;
  const int64_t per_img = tiles_hw * tiles_c;
  const int64_t n = blk / per_img;
  if (hnew != nullptr && hnew[n] == hold[n]) return;   // slot unchanged since the last sync
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

// Per-slot 64-bit content hash of the source ring: sum over 16-byte words of a position-keyed
// splitmix64 finaliser (order-independent accumulation, so CTAs add with one atomic each).
// One pass over the source (the cheapest possible "which slots changed" test: the slot writes of
// the unmodified reference, dpvo.py:435-438,296-299, are plain torch copies this library never sees).
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256)
slot_hash_kernel(const uint4* __restrict__ src, int64_t words_per_slot, int chunks, unsigned long long* __restrict__ hash) {
  const int64_t n = blockIdx.x / chunks;
  const int chunk = blockIdx.x - (int)(n * chunks);
  const int64_t per = ceil_div64(words_per_slot, chunks);
  const int64_t w0 = (int64_t)chunk * per;
  const int64_t w1 = min(words_per_slot, w0 + per);
  const uint4* s = src + n * words_per_slot;
  unsigned long long acc = 0;
  for (int64_t w = w0 + threadIdx.x; w < w1; w += 256) {
    const uint4 v = __ldg(s + w);
    const unsigned long long a = ((unsigned long long)v.y << 32) | v.x;
    const unsigned long long b = ((unsigned long long)v.w << 32) | v.z;
    const unsigned long long key = (unsigned long long)w * 0x9e3779b97f4a7c15ULL;
    acc += mix64(a + key) + mix64(b ^ (key + 0x632be59bd9b4e019ULL));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ unsigned long long part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += part[i];
    atomicAdd(hash + n, t + 1ULL);   // +1 per chunk: an all-zero slot does not hash to the memset value
  }
}

template <typename T>
static int launch(const void* src, void* dst, int64_t N, int C, int64_t HW, cudaStream_t st,
                  const unsigned long long* hnew = nullptr, const unsigned long long* hold = nullptr) {
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
        reinterpret_cast<const unsigned short*>(src), reinterpret_cast<unsigned short*>(dst), C, HW, tiles_hw, tc2,
        hnew, hold);
    return check_launch("nchw_to_nhwc");
  }
  note_launches(1);
  nchw_to_nhwc_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(
      reinterpret_cast<const T*>(src), reinterpret_cast<T*>(dst), C, HW, tiles_hw, tiles_c, hnew, hold);
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

extern "C" int dpvo_ring_sync(dpvo_dtype dtype, const void* src, void* dst, int64_t N, int32_t C, int64_t HW,
                              void* hashes, int32_t parity, void* stream) {
  DPVO_REQUIRE(src && dst && hashes, "ring_sync: null pointer");
  DPVO_REQUIRE(N >= 0 && C > 0 && HW >= 0, "ring_sync: bad sizes");
  DPVO_REQUIRE(parity == 0 || parity == 1, "ring_sync: parity must be 0 or 1");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t esz = dtype == DPVO_F16 ? 2 : dtype == DPVO_F32 ? 4 : 8;
  const int64_t slot_bytes = (int64_t)C * HW * esz;
  DPVO_REQUIRE(slot_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0,
               "ring_sync: slots must be 16-byte multiples and aligned (use dpvo_nchw_to_nhwc)");
  if (N == 0 || slot_bytes == 0) return 0;
  unsigned long long* hnew = reinterpret_cast<unsigned long long*>(hashes) + (int64_t)parity * N;
  const unsigned long long* hold = reinterpret_cast<unsigned long long*>(hashes) + (int64_t)(1 - parity) * N;
  if (cudaMemsetAsync(hnew, 0, sizeof(unsigned long long) * N, st) != cudaSuccess) {
    dpvo::set_error("ring_sync: memset failed");
    return 1;
  }
  const int64_t words = slot_bytes / 16;
  int chunks = (int)dpvo::ceil_div64(dpvo::kNumSMs * 8, N);
  chunks = (int)std::max<int64_t>(1, std::min<int64_t>(chunks, dpvo::ceil_div64(words, 1024)));
  dpvo::note_launches(1);
  dpvo::slot_hash_kernel<<<(unsigned)(N * chunks), 256, 0, st>>>(reinterpret_cast<const uint4*>(src), words, chunks, hnew);
  if (dpvo::check_launch("ring_sync: hash")) return 1;
  switch (dtype) {
    case DPVO_F16: return dpvo::launch<unsigned short>(src, dst, N, C, HW, st, hnew, hold);
    case DPVO_F32: return dpvo::launch<float>(src, dst, N, C, HW, st, hnew, hold);
    case DPVO_F64: return dpvo::launch<double>(src, dst, N, C, HW, st, hnew, hold);
  }
  dpvo::set_error("ring_sync: unknown dtype");
  return 1;
}
