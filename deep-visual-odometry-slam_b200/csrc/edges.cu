// Patch-graph edge bookkeeping on the device.
//
// The reference keeps the edge list (ii, jj, kk, net, target, weight) as torch tensors that are
// re-allocated by torch.cat / boolean-mask indexing on every frame (dpvo.py:215-239, 262-310); each
// mask index is a device->host sync.  Here the list lives in fixed-capacity device arrays with the
// live count in device memory, and the same edits are kernels:
//
//   append_factors (dpvo.py:215-221)            -> dpvo_edges_append / dpvo_edges_append_grid
//   __edges_forw / __edges_back (:362-375)      -> dpvo_edges_append_grid (flatmeshgrid, 'ij' order)
//   remove_factors(m, store) (:223-239)         -> dpvo_edges_remove (stable compaction + inactive list)
//   keyframe()'s masks and renumbering (:280-286, 304-310) -> dpvo_edges_mask_keyframe / _mask_window / _renumber
//   keyframe()'s slot-by-slot ring shift (:288-297)        -> dpvo_frames_shift
//   reduce_edges (loop_closure/optim_utils.py:27-64)       -> dpvo_edges_reduce
//
// Everything is integer / byte movement: results are bit-identical to the reference's.
#include "common.cuh"

namespace dpvo {
namespace {

constexpr int kChunk = 1024;   // edges per CTA in the compaction kernels

__device__ __forceinline__ int64_t live_count(const dpvo_edge_store& s) {
  const int64_t c = *s.count;
  return c < 0 ? 0 : (c > s.capacity ? s.capacity : c);
}

__device__ __forceinline__ void zero_row(const dpvo_edge_store& s, int64_t p, int lane_in_row, int lanes) {
  if (s.net) {
    uint32_t* row = reinterpret_cast<uint32_t*>(reinterpret_cast<uint16_t*>(s.net) + p * s.net_ld);
    for (int c = lane_in_row; c < (s.net_dim >> 1); c += lanes) row[c] = 0u;   // net_dim, net_ld even
  }
  if (lane_in_row == 0) {
    if (s.target) { s.target[2 * p] = 0.f; s.target[2 * p + 1] = 0.f; }
    if (s.weight) { s.weight[2 * p] = 0.f; s.weight[2 * p + 1] = 0.f; }
  }
}

// one 32-lane group per new edge.  grid mode: t -> (kk, jj) = (a0 + t / nb, b0 + t % nb)
__global__ void __launch_bounds__(256)
edges_append_kernel(const dpvo_edge_store s, const int64_t* __restrict__ ix, int64_t n_ix,
                    const int64_t* __restrict__ kk_new, const int64_t* __restrict__ jj_new, int64_t a0, int64_t b0,
                    int64_t nb, int64_t n_new, int32_t* __restrict__ info) {
  const int lane = threadIdx.x & 31;
  const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (t >= n_new) return;
  const int64_t base = live_count(s);
  const int64_t p = base + t;
  if (p >= s.capacity) {
    if (lane == 0) atomicExch(info, 1);
    return;
  }
  const int64_t k = kk_new ? kk_new[t] : a0 + t / nb;
  const int64_t j = jj_new ? jj_new[t] : b0 + t % nb;
  if (lane == 0) {
    s.kk[p] = k;
    s.jj[p] = j;
    if (k < 0 || k >= n_ix) { atomicExch(info, 2); s.ii[p] = 0; }
    else s.ii[p] = ix[k];
  }
  zero_row(s, p, lane, 32);
}

__global__ void edges_count_add_kernel(const dpvo_edge_store s, int64_t n_new) {
  const int64_t c = live_count(s) + n_new;
  *s.count = c > s.capacity ? s.capacity : c;
}

// ---- remove_factors: kept edges compacted in order, removed ones appended (in order) to the inactive list
__global__ void __launch_bounds__(256)
edges_block_count_kernel(const dpvo_edge_store s, const uint8_t* __restrict__ mask, int64_t* __restrict__ block_kept) {
  const int64_t n = live_count(s);
  const int64_t lo = (int64_t)blockIdx.x * kChunk;
  int kept = 0;
  for (int t = threadIdx.x; t < kChunk; t += blockDim.x) kept += (lo + t < n && mask[lo + t] == 0) ? 1 : 0;
  __shared__ int red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) kept += __shfl_xor_sync(0xffffffffu, kept, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = kept;
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
    for (int w = 0; w < 8; ++w) tot += red[w];
    block_kept[blockIdx.x] = tot;
  }
}

// exclusive scan of the per-chunk counts (one CTA; a few thousand chunks at most); totals[0] = kept,
// totals[1] = removed, totals[2] = the inactive list's count before this call
__global__ void __launch_bounds__(1024)
edges_block_scan_kernel(const dpvo_edge_store s, const dpvo_edge_store inac, int has_inac, int64_t* __restrict__ block_kept,
                        int nblocks, int64_t* __restrict__ totals, int32_t* __restrict__ info) {
  __shared__ int64_t warp_tot[32];
  __shared__ int64_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int base = 0; base < nblocks; base += 1024) {
    const int i = base + threadIdx.x;
    const int64_t v = i < nblocks ? block_kept[i] : 0;
    int64_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int64_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_tot[wid] = x;
    __syncthreads();
    if (wid == 0) {
      int64_t w = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int64_t y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      warp_tot[lane] = w;
    }
    __syncthreads();
    const int64_t before = carry + (wid ? warp_tot[wid - 1] : 0) + x - v;
    if (i < nblocks) block_kept[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const int64_t n = live_count(s);
    totals[0] = carry;
    totals[1] = n - carry;
    int64_t ic = 0;
    if (has_inac) {
      ic = *inac.count;
      ic = ic < 0 ? 0 : ic;
      if (ic + (n - carry) > inac.capacity) atomicExch(info, 1);
    }
    totals[2] = ic;
  }
}

struct EdgeScratch {
  int64_t* ii;
  int64_t* jj;
  int64_t* kk;
  uint16_t* net;
  float* target;
  float* weight;
  int* dest;
};

// kept rows -> scratch at their new position; removed rows -> the inactive list.  256 threads, 1024 edges.
__global__ void __launch_bounds__(256)
edges_scatter_kernel(const dpvo_edge_store s, const dpvo_edge_store inac, int has_inac, const uint8_t* __restrict__ mask,
                     const int64_t* __restrict__ block_before, const int64_t* __restrict__ totals, EdgeScratch w) {
  __shared__ int pos[kChunk];       // rank among this chunk's kept edges, or -1-rank among its removed ones
  __shared__ int warp_cnt[8];
  const int64_t n = live_count(s);
  const int64_t lo = (int64_t)blockIdx.x * kChunk;
  if (lo >= n) return;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int run = 0;                      // kept edges of this chunk before the current 256-edge slab
  for (int slab = 0; slab < kChunk; slab += 256) {
    const int t = slab + threadIdx.x;
    const bool live = lo + t < n;
    const bool keep = live && mask[lo + t] == 0;
    const unsigned b = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_cnt[wid] = __popc(b);
    __syncthreads();
    int before = run;
    for (int x = 0; x < wid; ++x) before += warp_cnt[x];
    int slab_tot = 0;
    for (int x = 0; x < 8; ++x) slab_tot += warp_cnt[x];
    const int r = before + __popc(b & ((1u << lane) - 1u));
    pos[t] = !live ? INT_MIN : (keep ? r : -1 - (t - r));
    run += slab_tot;
    __syncthreads();
  }
  const int64_t kept_before = block_before[blockIdx.x];
  const int64_t removed_before = lo - kept_before;
  const int64_t inac_base = totals[2];
  // index / target / weight rows: one thread per edge
  for (int t = threadIdx.x; t < kChunk; t += 256) {
    const int p = pos[t];
    if (p == INT_MIN) continue;
    const int64_t e = lo + t;
    if (p >= 0) {
      const int64_t d = kept_before + p;
      w.ii[d] = s.ii[e]; w.jj[d] = s.jj[e]; w.kk[d] = s.kk[e];
      if (s.target) { w.target[2 * d] = s.target[2 * e]; w.target[2 * d + 1] = s.target[2 * e + 1]; }
      if (s.weight) { w.weight[2 * d] = s.weight[2 * e]; w.weight[2 * d + 1] = s.weight[2 * e + 1]; }
    } else if (has_inac) {
      const int64_t d = inac_base + removed_before + (-1 - p);
      if (d < inac.capacity) {
        inac.ii[d] = s.ii[e]; inac.jj[d] = s.jj[e]; inac.kk[d] = s.kk[e];
        if (inac.target && s.target) { inac.target[2 * d] = s.target[2 * e]; inac.target[2 * d + 1] = s.target[2 * e + 1]; }
        if (inac.weight && s.weight) { inac.weight[2 * d] = s.weight[2 * e]; inac.weight[2 * d + 1] = s.weight[2 * e + 1]; }
      }
    }
  }
  // destination of every edge for the wide row kernels: new position, or -1 for removed edges
  for (int t = threadIdx.x; t < kChunk; t += 256) {
    const int p = pos[t];
    if (p != INT_MIN) w.dest[lo + t] = p >= 0 ? (int)(kept_before + p) : -1;
  }
}

// hidden-state rows, one warp per row, 16 bytes per lane when the rows allow it.  TO_SCRATCH: row e -> scratch
// row dest[e] (rows that stay where they are are skipped: they form a prefix); else scratch row d -> row d.
template <bool TO_SCRATCH>
__global__ void __launch_bounds__(256)
edges_rows_kernel(const dpvo_edge_store s, const int64_t* __restrict__ totals, EdgeScratch w) {
  const int lane = threadIdx.x & 31;
  const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n = TO_SCRATCH ? live_count(s) : totals[0];
  if (e >= n) return;
  const int d = w.dest[e];
  if (d == (int)e || (TO_SCRATCH && d < 0)) return;
  const int64_t row_bytes = (int64_t)s.net_dim * 2;
  const char* src;
  char* dst;
  if (TO_SCRATCH) {
    src = reinterpret_cast<const char*>(s.net) + e * s.net_ld * 2;
    dst = reinterpret_cast<char*>(w.net) + (int64_t)d * row_bytes;
  } else {
    src = reinterpret_cast<const char*>(w.net) + e * row_bytes;
    dst = reinterpret_cast<char*>(s.net) + e * s.net_ld * 2;
  }
  if (((row_bytes | (s.net_ld * 2) | (int64_t)reinterpret_cast<uintptr_t>(s.net)) & 15) == 0) {
    for (int64_t c = lane * 16; c < row_bytes; c += 512)
      *reinterpret_cast<uint4*>(dst + c) = *reinterpret_cast<const uint4*>(src + c);
  } else {
    for (int64_t c = lane * 4; c < row_bytes; c += 128)
      *reinterpret_cast<uint32_t*>(dst + c) = *reinterpret_cast<const uint32_t*>(src + c);
  }
}

__global__ void __launch_bounds__(256)
edges_copy_back_kernel(const dpvo_edge_store s, const int64_t* __restrict__ totals, EdgeScratch w) {
  const int64_t kept = totals[0];
  const int64_t lo = (int64_t)blockIdx.x * kChunk;
  if (lo >= kept) return;
  for (int t = threadIdx.x; t < kChunk; t += 256) {
    const int64_t e = lo + t;
    if (e >= kept) break;
    s.ii[e] = w.ii[e]; s.jj[e] = w.jj[e]; s.kk[e] = w.kk[e];
    if (s.target) { s.target[2 * e] = w.target[2 * e]; s.target[2 * e + 1] = w.target[2 * e + 1]; }
    if (s.weight) { s.weight[2 * e] = w.weight[2 * e]; s.weight[2 * e + 1] = w.weight[2 * e + 1]; }
  }
}

__global__ void edges_commit_kernel(const dpvo_edge_store s, const dpvo_edge_store inac, int has_inac,
                                    const int64_t* __restrict__ totals) {
  *s.count = totals[0];
  if (has_inac) {
    const int64_t c = totals[2] + totals[1];
    *inac.count = c > inac.capacity ? inac.capacity : c;
  }
}

// ---- masks and renumbering of keyframe()
__global__ void __launch_bounds__(256)
edges_mask_keyframe_kernel(const dpvo_edge_store s, int64_t k, uint8_t* __restrict__ mask, int64_t bound) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= bound) return;
  mask[e] = (e < live_count(s) && (s.ii[e] == k || s.jj[e] == k)) ? 1 : 0;
}

__global__ void __launch_bounds__(256)
edges_mask_window_kernel(const dpvo_edge_store s, const int64_t* __restrict__ ix, int64_t n_ix, int64_t n,
                         int64_t removal_window, int loop_closure, int64_t optimization_window,
                         uint8_t* __restrict__ mask, int64_t bound) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= bound) return;
  uint8_t m = 0;
  if (e < live_count(s)) {
    const int64_t k = s.kk[e];
    const int64_t src = (k >= 0 && k < n_ix) ? ix[k] : 0;
    bool rem = src < n - removal_window;
    if (loop_closure) {
      const bool lc = (s.jj[e] - s.ii[e] > 30) && (s.jj[e] > n - optimization_window);
      rem = rem && !lc;
    }
    m = rem ? 1 : 0;
  }
  mask[e] = m;
}

__global__ void __launch_bounds__(256)
edges_renumber_kernel(const dpvo_edge_store s, int64_t k, int64_t M, int64_t bound) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= bound || e >= live_count(s)) return;
  const int64_t i = s.ii[e], j = s.jj[e];
  if (i > k) { s.kk[e] -= M; s.ii[e] = i - 1; }
  if (j > k) s.jj[e] = j - 1;
}

// ---- keyframe()'s "for i in range(k, n-1): buf[i % mod] = buf[(i+1) % mod]".  Every thread owns one
// 16-byte (or 1-byte) column of the slot and walks the frames in order, so the sequential semantics hold.
template <typename T>
__global__ void __launch_bounds__(256)
frames_shift_kernel(T* __restrict__ buf, int64_t slot_elems, int mod, int k, int n) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= slot_elems) return;
  T next = buf[(int64_t)((k + 1) % mod) * slot_elems + c];
  for (int i = k; i < n - 1; ++i) {
    const T cur = next;
    if (i + 2 < n) next = buf[(int64_t)((i + 2) % mod) * slot_elems + c];
    buf[(int64_t)(i % mod) * slot_elems + c] = cur;
  }
}

// ---- reduce_edges (optim_utils.py:27-64): candidates in ascending flow magnitude (ties: lower index
// first), greedy pick with suppression of the +-nms neighbouring source frames at the same target.
// One CTA: bitonic sort of (magnitude, index) keys in global scratch, then one thread walks the list.
__global__ void __launch_bounds__(1024)
edges_reduce_kernel(const float* __restrict__ flow_mag, const int64_t* __restrict__ ii, const int64_t* __restrict__ jj,
                    int n, int npow2, int max_num_edges, int nms, uint64_t* __restrict__ keys,
                    uint8_t* __restrict__ ignore, int64_t Ni, int64_t Nj, int64_t* __restrict__ out,
                    int64_t* __restrict__ out_count) {
  for (int t = threadIdx.x; t < npow2; t += blockDim.x) {
    uint64_t key = ~0ull;
    if (t < n) {
      // order-preserving map of the float (NaN sorts last among finite/inf like numpy's argsort)
      uint32_t u = __float_as_uint(flow_mag[t]);
      if (flow_mag[t] != flow_mag[t]) u = 0xffffffffu;
      else u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
      key = ((uint64_t)u << 32) | (uint32_t)t;
    }
    keys[t] = key;
  }
  for (int64_t t = threadIdx.x; t < Ni * Nj; t += blockDim.x) ignore[t] = 0;
  __syncthreads();
  for (int size = 2; size <= npow2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < npow2; t += blockDim.x) {
        const int p = t ^ stride;
        if (p > t) {
          const bool up = (t & size) == 0;
          const uint64_t a = keys[t], b = keys[p];
          if ((a > b) == up) { keys[t] = b; keys[p] = a; }
        }
      }
      __syncthreads();
    }
  }
  if (threadIdx.x == 0) {
    int64_t cnt = 0;
    for (int s = 0; s < n; ++s) {
      if (cnt + 1 > max_num_edges) break;            // the reference's list starts with a dummy entry
      const int idx = (int)(keys[s] & 0xffffffffu);
      const int64_t i = ii[idx], j = jj[idx];
      const float mag = flow_mag[idx];
      if (j - i < 30) continue;
      if (mag >= 1000.f) continue;
      if (mag != mag) continue;
      if (ignore[i * Nj + j]) continue;
      out[2 * cnt] = i;
      out[2 * cnt + 1] = j;
      ++cnt;
      for (int di = -nms; di <= nms; ++di) {
        const int64_t i1 = i + di;
        if (i1 >= 0 && i1 < Ni) ignore[i1 * Nj + j] = 1;
      }
    }
    *out_count = cnt;
  }
}

int64_t a256(int64_t x) { return (x + 255) & ~(int64_t)255; }

bool store_ok(const dpvo_edge_store* s) {
  return s && s->ii && s->jj && s->kk && s->count && s->capacity >= 0 &&
         (!s->net || (s->net_dim > 0 && (s->net_dim % 2) == 0 && s->net_ld >= s->net_dim && (s->net_ld % 2) == 0 &&
                      (reinterpret_cast<uintptr_t>(s->net) & 3) == 0));
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

extern "C" int dpvo_edges_append(const dpvo_edge_store* store, const int64_t* ix, int64_t n_ix, const int64_t* kk_new,
                                 const int64_t* jj_new, int64_t n_new, int32_t* info, void* stream) {
  DPVO_REQUIRE(store_ok(store), "edges_append: bad edge store");
  DPVO_REQUIRE(n_new >= 0 && info, "edges_append: bad arguments");
  if (n_new == 0) return 0;
  DPVO_REQUIRE(ix && kk_new && jj_new, "edges_append: null pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  edges_append_kernel<<<(unsigned)ceil_div64(n_new * 32, 256), 256, 0, st>>>(*store, ix, n_ix, kk_new, jj_new, 0, 0, 1,
                                                                              n_new, info);
  edges_count_add_kernel<<<1, 1, 0, st>>>(*store, n_new);
  note_launches(2);
  return check_launch("edges_append");
}

extern "C" int dpvo_edges_append_grid(const dpvo_edge_store* store, const int64_t* ix, int64_t n_ix, int64_t kk0,
                                      int64_t kk1, int64_t jj0, int64_t jj1, int32_t* info, void* stream) {
  DPVO_REQUIRE(store_ok(store), "edges_append_grid: bad edge store");
  DPVO_REQUIRE(info, "edges_append_grid: null info");
  const int64_t na = kk1 > kk0 ? kk1 - kk0 : 0, nb = jj1 > jj0 ? jj1 - jj0 : 0;
  const int64_t n_new = na * nb;
  if (n_new == 0) return 0;
  DPVO_REQUIRE(ix, "edges_append_grid: null pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  edges_append_kernel<<<(unsigned)ceil_div64(n_new * 32, 256), 256, 0, st>>>(*store, ix, n_ix, nullptr, nullptr, kk0,
                                                                              jj0, nb, n_new, info);
  edges_count_add_kernel<<<1, 1, 0, st>>>(*store, n_new);
  note_launches(2);
  return check_launch("edges_append_grid");
}

extern "C" int64_t dpvo_edges_workspace_bytes(int64_t capacity, int32_t net_dim) {
  if (capacity < 0 || net_dim < 0) return -1;
  const int64_t nblocks = (capacity + kChunk - 1) / kChunk + 1;
  // totals | per-chunk counts | ii jj kk | target weight | dest | net rows
  return 256 + a256(nblocks * 8) + 3 * a256(capacity * 8) + 2 * a256(capacity * 8) + a256(capacity * 4) +
         a256(capacity * net_dim * 2) + 256;
}

extern "C" int dpvo_edges_remove(const dpvo_edge_store* store, const uint8_t* mask, const dpvo_edge_store* inactive,
                                 int64_t bound, int32_t* info, void* workspace, int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(store_ok(store), "edges_remove: bad edge store");
  DPVO_REQUIRE(!inactive || store_ok(inactive), "edges_remove: bad inactive store");
  DPVO_REQUIRE(bound >= 0 && bound <= store->capacity && store->capacity < (int64_t(1) << 31) && info,
               "edges_remove: bad arguments");
  if (bound == 0) return 0;
  DPVO_REQUIRE(mask && workspace, "edges_remove: null pointer");
  DPVO_REQUIRE(workspace_bytes >= dpvo_edges_workspace_bytes(store->capacity, store->net ? store->net_dim : 0),
               "edges_remove: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t cap = store->capacity;
  const int nd = store->net ? store->net_dim : 0;
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  int64_t* totals = reinterpret_cast<int64_t*>(base); base += 256;
  const int64_t nblocks_cap = (cap + kChunk - 1) / kChunk + 1;
  int64_t* block_kept = reinterpret_cast<int64_t*>(base); base += a256(nblocks_cap * 8);
  EdgeScratch w;
  w.ii = reinterpret_cast<int64_t*>(base); base += a256(cap * 8);
  w.jj = reinterpret_cast<int64_t*>(base); base += a256(cap * 8);
  w.kk = reinterpret_cast<int64_t*>(base); base += a256(cap * 8);
  w.target = reinterpret_cast<float*>(base); base += a256(cap * 8);
  w.weight = reinterpret_cast<float*>(base); base += a256(cap * 8);
  w.dest = reinterpret_cast<int*>(base); base += a256(cap * 4);
  w.net = reinterpret_cast<uint16_t*>(base);
  (void)nd;
  const int nblocks = (int)ceil_div64(bound, kChunk);
  dpvo_edge_store none;
  memset(&none, 0, sizeof(none));
  const dpvo_edge_store& in = inactive ? *inactive : none;
  const int has = inactive ? 1 : 0;
  edges_block_count_kernel<<<nblocks, 256, 0, st>>>(*store, mask, block_kept);
  edges_block_scan_kernel<<<1, 1024, 0, st>>>(*store, in, has, block_kept, nblocks, totals, info);
  edges_scatter_kernel<<<nblocks, 256, 0, st>>>(*store, in, has, mask, block_kept, totals, w);
  const unsigned row_grid = (unsigned)ceil_div64(bound * 32, 256);
  if (store->net) edges_rows_kernel<true><<<row_grid, 256, 0, st>>>(*store, totals, w);
  edges_copy_back_kernel<<<nblocks, 256, 0, st>>>(*store, totals, w);
  if (store->net) edges_rows_kernel<false><<<row_grid, 256, 0, st>>>(*store, totals, w);
  edges_commit_kernel<<<1, 1, 0, st>>>(*store, in, has, totals);
  note_launches(store->net ? 7 : 5);
  return check_launch("edges_remove");
}

extern "C" int dpvo_edges_mask_keyframe(const dpvo_edge_store* store, int64_t k, uint8_t* mask, int64_t bound,
                                        void* stream) {
  DPVO_REQUIRE(store_ok(store) && bound >= 0 && bound <= store->capacity, "edges_mask_keyframe: bad arguments");
  if (bound == 0) return 0;
  DPVO_REQUIRE(mask, "edges_mask_keyframe: null mask");
  edges_mask_keyframe_kernel<<<(unsigned)ceil_div64(bound, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      *store, k, mask, bound);
  note_launches(1);
  return check_launch("edges_mask_keyframe");
}

extern "C" int dpvo_edges_mask_window(const dpvo_edge_store* store, const int64_t* ix, int64_t n_ix, int64_t n,
                                      int64_t removal_window, int32_t loop_closure, int64_t optimization_window,
                                      uint8_t* mask, int64_t bound, void* stream) {
  DPVO_REQUIRE(store_ok(store) && bound >= 0 && bound <= store->capacity, "edges_mask_window: bad arguments");
  if (bound == 0) return 0;
  DPVO_REQUIRE(mask && ix, "edges_mask_window: null pointer");
  edges_mask_window_kernel<<<(unsigned)ceil_div64(bound, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      *store, ix, n_ix, n, removal_window, loop_closure, optimization_window, mask, bound);
  note_launches(1);
  return check_launch("edges_mask_window");
}

extern "C" int dpvo_edges_renumber(const dpvo_edge_store* store, int64_t k, int64_t patches_per_frame, int64_t bound,
                                   void* stream) {
  DPVO_REQUIRE(store_ok(store) && bound >= 0 && bound <= store->capacity, "edges_renumber: bad arguments");
  if (bound == 0) return 0;
  edges_renumber_kernel<<<(unsigned)ceil_div64(bound, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      *store, k, patches_per_frame, bound);
  note_launches(1);
  return check_launch("edges_renumber");
}

extern "C" int dpvo_frames_shift(void* buf, int64_t slot_bytes, int32_t mod, int32_t k, int32_t n, void* stream) {
  DPVO_REQUIRE(slot_bytes >= 0 && mod > 0 && k >= 0, "frames_shift: bad arguments");
  if (n - 1 <= k || slot_bytes == 0) return 0;
  DPVO_REQUIRE(buf, "frames_shift: null buffer");
  DPVO_REQUIRE(n - k <= mod, "frames_shift: more frames to move than ring slots");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if ((slot_bytes % 16) == 0 && (reinterpret_cast<uintptr_t>(buf) & 15) == 0) {
    const int64_t el = slot_bytes / 16;
    frames_shift_kernel<uint4><<<(unsigned)ceil_div64(el, 256), 256, 0, st>>>(reinterpret_cast<uint4*>(buf), el, mod, k, n);
  } else {
    frames_shift_kernel<uint8_t><<<(unsigned)ceil_div64(slot_bytes, 256), 256, 0, st>>>(reinterpret_cast<uint8_t*>(buf),
                                                                                       slot_bytes, mod, k, n);
  }
  note_launches(1);
  return check_launch("frames_shift");
}

extern "C" int64_t dpvo_edges_reduce_workspace_bytes(int64_t n, int64_t ni, int64_t nj) {
  if (n < 0 || ni < 0 || nj < 0) return -1;
  int64_t p = 1;
  while (p < n) p <<= 1;
  return 256 + a256(p * 8) + a256(ni * nj) + 256;
}

extern "C" int dpvo_edges_reduce(const float* flow_mag, const int64_t* ii, const int64_t* jj, int64_t n, int64_t ni,
                                 int64_t nj, int32_t max_num_edges, int32_t nms, int64_t* out_edges,
                                 int64_t* out_count, void* workspace, int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(n >= 0 && n < (int64_t(1) << 24) && max_num_edges >= 0 && nms >= 0 && ni >= 0 && nj >= 0,
               "edges_reduce: bad sizes");
  DPVO_REQUIRE(out_count, "edges_reduce: null output");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  cudaMemsetAsync(out_count, 0, sizeof(int64_t), st);
  if (n == 0 || max_num_edges == 0) return check_launch("edges_reduce");
  DPVO_REQUIRE(flow_mag && ii && jj && out_edges && workspace, "edges_reduce: null pointer");
  DPVO_REQUIRE(workspace_bytes >= dpvo_edges_reduce_workspace_bytes(n, ni, nj), "edges_reduce: workspace too small");
  int npow2 = 1;
  while (npow2 < n) npow2 <<= 1;
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  uint64_t* keys = reinterpret_cast<uint64_t*>(base); base += a256((int64_t)npow2 * 8);
  uint8_t* ignore = reinterpret_cast<uint8_t*>(base);
  edges_reduce_kernel<<<1, 1024, 0, st>>>(flow_mag, ii, jj, (int)n, npow2, max_num_edges, nms, keys, ignore, ni, nj,
                                          out_edges, out_count);
  note_launches(1);
  return check_launch("edges_reduce");
}
