// fastba.neighbors on the device, plus sorted-unique/inverse of bounded keys.
//
// Replaces /root/reference/methods/dpvo/dpvo/fastba/ba.cpp:59-97, which copies the edge
// lists to the host, buckets them per key, std::stable_sort's each bucket by jj and copies
// the result back — every update iteration.  Here: open-addressing hash table groups the
// edges by key (no bound on the key range needed), a prefix sum turns the group sizes into
// a CSR, and every edge scans its own (small) group for its predecessor / successor in
// (jj, edge index) order, which is exactly the stable-sort order.  Integer work, bit-exact.
#include "scan.cuh"

namespace dpvo {

constexpr unsigned long long kEmptyKey = 0x8080808080808080ull;  // memset(0x80) pattern

__device__ __forceinline__ uint32_t hash64(unsigned long long k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return (uint32_t)k;
}

__global__ void __launch_bounds__(256)
nb_insert_kernel(const int64_t* __restrict__ keys_in, int64_t n, unsigned long long* __restrict__ table,
                 uint32_t mask, int* __restrict__ slot_of, int* __restrict__ count) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long key = (unsigned long long)keys_in[e];
    uint32_t s = hash64(key) & mask;
    while (true) {
      const unsigned long long prev = atomicCAS(&table[s], kEmptyKey, key);
      if (prev == kEmptyKey || prev == key) break;
      s = (s + 1) & mask;
    }
    slot_of[e] = (int)s;
    atomicAdd(&count[s], 1);
  }
}

__global__ void __launch_bounds__(256)
nb_fill_kernel(int64_t n, const int* __restrict__ slot_of, const int* __restrict__ offset,
               int* __restrict__ cursor, int* __restrict__ list) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
    const int s = slot_of[e];
    const int pos = offset[s] + atomicAdd(&cursor[s], 1);
    list[pos] = (int)e;
  }
}

__global__ void __launch_bounds__(256)
nb_predsucc_kernel(const int64_t* __restrict__ jj, int64_t n, const int* __restrict__ slot_of,
                   const int* __restrict__ offset, const int* __restrict__ count,
                   const int* __restrict__ list, int64_t* __restrict__ ix, int64_t* __restrict__ jx) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
    const int s = slot_of[e];
    const int beg = offset[s], cnt = count[s];
    const int64_t myj = jj[e];
    int64_t pj = 0, sj = 0;
    int pe = -1, se = -1;
    for (int t = 0; t < cnt; ++t) {
      const int f = list[beg + t];
      if (f == (int)e) continue;
      const int64_t fj = jj[f];
      const bool less = (fj < myj) || (fj == myj && f < (int)e);
      if (less) {
        if (pe < 0 || fj > pj || (fj == pj && f > pe)) { pe = f; pj = fj; }
      } else {
        if (se < 0 || fj < sj || (fj == sj && f < se)) { se = f; sj = fj; }
      }
    }
    ix[e] = pe;
    jx[e] = se;
  }
}

static uint32_t table_size(int64_t n) {
  uint32_t t = 1024;
  while ((int64_t)t < 2 * n) t <<= 1;
  return t;
}

static int64_t a256(int64_t x) { return (x + 255) & ~(int64_t)255; }

// ------------------------------------------------------------ unique / inverse
// Edge lists are runs of equal / neighbouring keys, so tens of lanes of a warp (and hundreds of edges in
// total) hit the same bitmap word: lanes that share a word combine their bits, one lane per word issues
// the atomic, and only when the word does not already hold those bits.
__global__ void __launch_bounds__(256)
uq_mark_kernel(const int64_t* __restrict__ keys, int64_t n, uint32_t* __restrict__ bitmap, int64_t bound) {
  const int lane = threadIdx.x & 31;
  for (int64_t base = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) - lane; base < n;
       base += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = base + lane;
    const int64_t k = e < n ? keys[e] : -1;
    const bool valid = k >= 0 && k < bound;
    const int w = valid ? (int)(k >> 5) : -1;
    const unsigned peers = __match_any_sync(0xffffffffu, w);
    const unsigned bits = __reduce_or_sync(peers, valid ? (1u << (k & 31)) : 0u);
    if (valid && lane == __ffs(peers) - 1 && (__ldcg(bitmap + w) & bits) != bits) atomicOr(bitmap + w, bits);
  }
}
__global__ void __launch_bounds__(256)
uq_popc_kernel(const uint32_t* __restrict__ bitmap, int64_t W, int* __restrict__ cnt) {
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < W; w += (int64_t)gridDim.x * blockDim.x)
    cnt[w] = __popc(bitmap[w]);
}
__global__ void __launch_bounds__(256)
uq_emit_kernel(const uint32_t* __restrict__ bitmap, int64_t W, const int* __restrict__ prefix,
               int64_t* __restrict__ uniq) {
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < W; w += (int64_t)gridDim.x * blockDim.x) {
    uint32_t b = bitmap[w];
    int r = prefix[w];
    while (b) {
      const int bit = __ffs(b) - 1;
      b &= b - 1;
      uniq[r++] = w * 32 + bit;
    }
  }
}
__global__ void __launch_bounds__(256)
uq_inverse_kernel(const int64_t* __restrict__ keys, int64_t n, const uint32_t* __restrict__ bitmap,
                  const int* __restrict__ prefix, int64_t bound, int64_t* __restrict__ inverse) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = keys[e];
    int64_t r = -1;
    if (k >= 0 && k < bound) r = prefix[k >> 5] + __popc(bitmap[k >> 5] & ((1u << (k & 31)) - 1u));
    inverse[e] = r;
  }
}

}  // namespace dpvo

using namespace dpvo;

extern "C" int64_t dpvo_neighbors_workspace_bytes(int64_t n) {
  if (n < 0) n = 0;
  const int64_t T = table_size(n);
  return a256(T * 8) + 3 * a256(T * 4) + 2 * a256(n * 4 + 4) + a256(scan_workspace_ints(T) * 4) + 512;
}

extern "C" int dpvo_neighbors(const int64_t* ii, const int64_t* jj, int64_t n,
                              int64_t* ix, int64_t* jx, void* workspace, int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(n >= 0, "neighbors: negative length");
  if (n == 0) return 0;
  DPVO_REQUIRE(n < ((int64_t)1 << 30), "neighbors: too many edges");
  DPVO_REQUIRE(ii && jj && ix && jx && workspace, "neighbors: null pointer");
  DPVO_REQUIRE(workspace_bytes >= dpvo_neighbors_workspace_bytes(n), "neighbors: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t T = table_size(n);
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  unsigned long long* table = (unsigned long long*)base; base += a256(T * 8);
  int* count = (int*)base; base += a256(T * 4);
  int* cursor = (int*)base; base += a256(T * 4);   // count and cursor are adjacent: one memset
  int* offset = (int*)base; base += a256(T * 4);
  int* slot_of = (int*)base; base += a256(n * 4 + 4);
  int* list = (int*)base; base += a256(n * 4 + 4);
  int* sums = (int*)base;
  cudaMemsetAsync(table, 0x80, T * 8, st);
  cudaMemsetAsync(count, 0, 2 * a256(T * 4), st);
  const int grid = (int)(ceil_div64(n, 256) < kNumSMs * 8 ? ceil_div64(n, 256) : kNumSMs * 8);
  nb_insert_kernel<<<grid, 256, 0, st>>>(ii, n, table, (uint32_t)(T - 1), slot_of, count);
  if (exclusive_scan_int(count, offset, T, sums, nullptr, st)) return 1;
  nb_fill_kernel<<<grid, 256, 0, st>>>(n, slot_of, offset, cursor, list);
  nb_predsucc_kernel<<<grid, 256, 0, st>>>(jj, n, slot_of, offset, count, list, ix, jx);
  note_launches(6);
  return check_launch("neighbors");
}

extern "C" int64_t dpvo_unique_workspace_bytes(int64_t key_bound) {
  if (key_bound < 0) key_bound = 0;
  const int64_t W = (key_bound + 31) / 32;
  return a256(W * 4) + 2 * a256(W * 4) + a256(scan_workspace_ints(W) * 4) + 512;
}

extern "C" int dpvo_unique_inverse(const int64_t* keys, int64_t n, int64_t key_bound,
                                   int64_t* uniq, int64_t* inverse, int32_t* n_uniq,
                                   void* workspace, int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(n >= 0 && key_bound >= 0, "unique_inverse: bad sizes");
  DPVO_REQUIRE(uniq && inverse && n_uniq && workspace, "unique_inverse: null pointer");
  DPVO_REQUIRE(workspace_bytes >= dpvo_unique_workspace_bytes(key_bound), "unique_inverse: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t W = (key_bound + 31) / 32;
  if (n == 0 || W == 0) { cudaMemsetAsync(n_uniq, 0, 4, st); return 0; }
  DPVO_REQUIRE(keys, "unique_inverse: null keys");
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  uint32_t* bitmap = (uint32_t*)base; base += a256(W * 4);
  int* cnt = (int*)base; base += a256(W * 4);
  int* prefix = (int*)base; base += a256(W * 4);
  int* sums = (int*)base;
  cudaMemsetAsync(bitmap, 0, W * 4, st);
  const int grid = (int)(ceil_div64(n, 256) < kNumSMs * 8 ? ceil_div64(n, 256) : kNumSMs * 8);
  const int wgrid = (int)(ceil_div64(W, 256) < kNumSMs * 8 ? ceil_div64(W, 256) : kNumSMs * 8);
  uq_mark_kernel<<<grid, 256, 0, st>>>(keys, n, bitmap, key_bound);
  uq_popc_kernel<<<wgrid, 256, 0, st>>>(bitmap, W, cnt);
  if (exclusive_scan_int(cnt, prefix, W, sums, n_uniq, st)) return 1;
  uq_emit_kernel<<<wgrid, 256, 0, st>>>(bitmap, W, prefix, uniq);
  uq_inverse_kernel<<<grid, 256, 0, st>>>(keys, n, bitmap, prefix, key_bound, inverse);
  note_launches(7);
  return check_launch("unique_inverse");
}
