// The whole Update operator (net.py:69-88) behind ONE C-ABI call.
//
// update.py used to issue the operator's ~45 launches one ctypes call at a time: 1.46 ms of host time against
// 0.72 ms of GPU time at E = 45,312 (profiles/r02_time_update.json) — the operator was host-bound whenever it was not
// inside a CUDA graph, i.e. in every real tracking loop (the edge count changes from frame to frame, so a graph of the
// frame does not survive).  Here the same launches, in the same order, with the same operands, are issued from C++:
// the caller passes the weights once (dpvo_update_weights), the per-call tensors and one workspace; nothing is
// allocated and nothing synchronises.  The index-only work (temporal neighbours, the two SoftAgg groupings and their
// bucket sorts) is forked onto an internal side stream underneath the three layers of the correlation encoder, as
// before.  Results equal the launch-by-launch path to the operator's own run-to-run spread (its grouped sums and head
// dot products use float atomics; tests/test_update_gpu.py).
#include "common.cuh"

namespace dpvo {
namespace {

constexpr int kD = 384;

enum Lin { C1_0, C1_2, C2_0, C2_2, KK_F, KK_G, KK_H, IJ_F, IJ_G, IJ_H, G1_GATE, G1_RES0, G1_RES2, G3_GATE, G3_RES0,
           G3_RES2, CORR0, CORR2, CORR5, kNumLin };
enum Norm { N_NORM, N_GRU0, N_GRU2, N_CORR3, kNumNorm };

__global__ void __launch_bounds__(256)
pair_keys_kernel(const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int64_t n, int64_t n_frames,
                 int64_t* __restrict__ keys) {
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < n; e += (int64_t)gridDim.x * 256)
    keys[e] = ii[e] * n_frames + jj[e];
}

inline int64_t al(int64_t x) { return (x + 255) & ~(int64_t)255; }

struct Layout {
  int64_t act[4], y1, y2, acc, keys, ix, jx, uniq, inv_kk, inv_ij, cnt, uws, sws_kk, sws_ij, nws, total;
  int64_t Gkk, Gij, uws_bytes, sws_kk_bytes, sws_ij_bytes, nws_bytes;
};

Layout make_layout(int64_t E, int64_t n_frames, int64_t n_patches) {
  Layout L;
  const int64_t kb_kk = n_patches, kb_ij = n_frames * n_frames;
  L.Gkk = E < kb_kk ? E : kb_kk;
  L.Gij = E < kb_ij ? E : kb_ij;
  const int64_t Gmax = L.Gkk > L.Gij ? L.Gkk : L.Gij;
  L.uws_bytes = dpvo_unique_workspace_bytes(kb_kk > kb_ij ? kb_kk : kb_ij);
  L.sws_kk_bytes = dpvo_softagg_workspace_bytes(E, L.Gkk);
  L.sws_ij_bytes = dpvo_softagg_workspace_bytes(E, L.Gij);
  L.nws_bytes = dpvo_neighbors_workspace_bytes(E);
  int64_t off = 0;
  auto take = [&](int64_t bytes) { const int64_t o = off; off += al(bytes); return o; };
  for (int i = 0; i < 4; ++i) L.act[i] = take(E * kD * 2);
  L.y1 = take(Gmax * kD * 2);
  L.y2 = take(Gmax * kD * 2);
  L.acc = take(E * 4 * 4);
  L.keys = take(E * 8);
  L.ix = take(E * 8);
  L.jx = take(E * 8);
  L.uniq = take(Gmax * 8);
  L.inv_kk = take(E * 8);
  L.inv_ij = take(E * 8);
  L.cnt = take(256);
  L.uws = take(L.uws_bytes);
  L.sws_kk = take(L.sws_kk_bytes);
  L.sws_ij = take(L.sws_ij_bytes);
  L.nws = take(L.nws_bytes);
  L.total = off;
  return L;
}

struct Side {
  cudaStream_t stream = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
  bool ok = false;
};
Side* side_for_device() {
  static Side per_dev[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  Side& s = per_dev[dev];
  if (!s.ok) {
    if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    s.ok = true;
  }
  return &s;
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

extern "C" int64_t dpvo_update_workspace_bytes(int64_t n_edges, int64_t n_frames, int64_t n_patches) {
  if (n_edges < 0 || n_frames <= 0 || n_patches <= 0) return -1;
  return make_layout(n_edges, n_frames, n_patches).total + 256;
}

extern "C" int dpvo_update_forward(const dpvo_update_weights* w, const void* net, const void* table,
                                   const int64_t* table_index, const void* corr, int64_t ld_corr,
                                   const int64_t* ii, const int64_t* jj, const int64_t* kk, int64_t E,
                                   int64_t n_frames, int64_t n_patches, void* net_out, float* delta, float* weight,
                                   void* workspace, int64_t workspace_bytes, void* stream) {
  DPVO_REQUIRE(E >= 0 && n_frames > 0 && n_patches > 0, "update_forward: bad sizes");
  if (E == 0) return 0;
  DPVO_REQUIRE(w && net && table && corr && ii && jj && kk && net_out && delta && weight && workspace,
               "update_forward: null pointer");
  const Layout L = make_layout(E, n_frames, n_patches);
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  DPVO_REQUIRE(workspace_bytes >= L.total + (base - reinterpret_cast<char*>(workspace)), "update_forward: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  Side* sd = side_for_device();
  DPVO_REQUIRE(sd != nullptr, "update_forward: could not create the side stream");
  void* act[4] = {base + L.act[0], base + L.act[1], base + L.act[2], base + L.act[3]};
  void* y1 = base + L.y1;
  void* y2 = base + L.y2;
  float* acc = reinterpret_cast<float*>(base + L.acc);
  int64_t* keys = reinterpret_cast<int64_t*>(base + L.keys);
  int64_t* ix = reinterpret_cast<int64_t*>(base + L.ix);
  int64_t* jx = reinterpret_cast<int64_t*>(base + L.jx);
  int64_t* uniq = reinterpret_cast<int64_t*>(base + L.uniq);
  int64_t* inv_kk = reinterpret_cast<int64_t*>(base + L.inv_kk);
  int64_t* inv_ij = reinterpret_cast<int64_t*>(base + L.inv_ij);
  int32_t* cnt = reinterpret_cast<int32_t*>(base + L.cnt);

  auto lin = [&](int l, const void* A, int64_t lda, int64_t M, int K, void* Y, const void* gate, const void* res1,
                 const void* res2, const int64_t* idx, int norm, int actv, bool head) -> int {
    dpvo_linear_params p;
    memset(&p, 0, sizeof(p));
    p.A = A; p.lda = (M > 1) ? lda : (lda > K ? lda : K);
    p.W = w->W[l]; p.ldw = w->ldw[l];
    p.bias = w->bias[l];
    p.gate = gate; p.res1 = res1; p.res2 = res2; p.res2_index = idx;
    if (norm >= 0) { p.ln_gamma = w->ln_gamma[norm]; p.ln_beta = w->ln_beta[norm]; p.ln_eps = w->ln_eps; }
    p.act = actv;
    p.Y = Y; p.M = M; p.K = K;
    if (head) { p.head_w = w->head_w; p.head_acc = acc; }
    return dpvo_linear384_f16(&p, st);
  };

  // ---- index-only work on the side stream (net.py:74 neighbours, :82-83 groupings)
  cudaEventRecord(sd->fork, st);
  cudaStreamWaitEvent(sd->stream, sd->fork, 0);
  {
    void* ss = sd->stream;
    if (dpvo_neighbors(kk, jj, E, ix, jx, base + L.nws, L.nws_bytes, ss)) return 1;
    cudaMemsetAsync(cnt, 0, 8, sd->stream);
    if (dpvo_unique_inverse(kk, E, n_patches, uniq, inv_kk, cnt, base + L.uws, L.uws_bytes, ss)) return 1;
    if (dpvo_softagg_prepare(inv_kk, E, L.Gkk, base + L.sws_kk, L.sws_kk_bytes, ss)) return 1;
    const int grid = (int)(ceil_div64(E, 256) < kNumSMs * 4 ? ceil_div64(E, 256) : kNumSMs * 4);
    pair_keys_kernel<<<grid, 256, 0, sd->stream>>>(ii, jj, E, n_frames, keys);
    note_launches(1);
    if (dpvo_unique_inverse(keys, E, n_frames * n_frames, uniq, inv_ij, cnt + 1, base + L.uws, L.uws_bytes, ss)) return 1;
    if (dpvo_softagg_prepare(inv_ij, E, L.Gij, base + L.sws_ij, L.sws_ij_bytes, ss)) return 1;
    cudaEventRecord(sd->join, sd->stream);
  }
  // ---- correlation encoder + input injection + norm (net.py:71-72)
  if (lin(CORR0, corr, ld_corr, E, w->corr_features, act[0], nullptr, nullptr, nullptr, nullptr, -1, 1, false)) return 1;
  if (lin(CORR2, act[0], kD, E, kD, act[1], nullptr, nullptr, nullptr, nullptr, N_CORR3, 1, false)) return 1;
  if (lin(CORR5, act[1], kD, E, kD, act[0], nullptr, net, table, table_index, N_NORM, 0, false)) return 1;
  cudaStreamWaitEvent(st, sd->join, 0);
  // ---- temporal neighbours along each patch track (net.py:74-80)
  if (dpvo_gather_rows_f16(act[0], ix, E, kD, act[1], st)) return 1;
  if (lin(C1_0, act[1], kD, E, kD, act[2], nullptr, nullptr, nullptr, nullptr, -1, 1, false)) return 1;
  if (lin(C1_2, act[2], kD, E, kD, act[1], nullptr, act[0], nullptr, nullptr, -1, 0, false)) return 1;
  if (dpvo_gather_rows_f16(act[1], jx, E, kD, act[0], st)) return 1;
  if (lin(C2_0, act[0], kD, E, kD, act[2], nullptr, nullptr, nullptr, nullptr, -1, 1, false)) return 1;
  if (lin(C2_2, act[2], kD, E, kD, act[0], nullptr, act[1], nullptr, nullptr, -1, 0, false)) return 1;
  // ---- soft aggregation over patches, then over frame pairs with gru.0's LayerNorm (net.py:82-83, :42)
  if (lin(KK_F, act[0], kD, E, kD, act[1], nullptr, nullptr, nullptr, nullptr, -1, 0, false)) return 1;
  if (lin(KK_G, act[0], kD, E, kD, act[2], nullptr, nullptr, nullptr, nullptr, -1, 0, false)) return 1;
  cudaMemsetAsync(y1, 0, (size_t)L.Gkk * kD * 2, st);
  if (dpvo_softagg_f16(act[1], act[2], inv_kk, E, L.Gkk, y1, 1, base + L.sws_kk, L.sws_kk_bytes, st)) return 1;
  if (lin(KK_H, y1, kD, L.Gkk, kD, y2, nullptr, nullptr, nullptr, nullptr, -1, 0, false)) return 1;
  if (dpvo_expand_add_f16(act[0], y2, inv_kk, E, nullptr, nullptr, 0.f, act[1], st)) return 1;
  if (lin(IJ_F, act[1], kD, E, kD, act[0], nullptr, nullptr, nullptr, nullptr, -1, 0, false)) return 1;
  if (lin(IJ_G, act[1], kD, E, kD, act[2], nullptr, nullptr, nullptr, nullptr, -1, 0, false)) return 1;
  cudaMemsetAsync(y1, 0, (size_t)L.Gij * kD * 2, st);
  if (dpvo_softagg_f16(act[0], act[2], inv_ij, E, L.Gij, y1, 1, base + L.sws_ij, L.sws_ij_bytes, st)) return 1;
  if (lin(IJ_H, y1, kD, L.Gij, kD, y2, nullptr, nullptr, nullptr, nullptr, -1, 0, false)) return 1;
  if (dpvo_expand_add_f16(act[1], y2, inv_ij, E, w->ln_gamma[N_GRU0], w->ln_beta[N_GRU0], w->ln_eps, act[0], st)) return 1;
  // ---- gated residual blocks (net.py:41-46,85); the second accumulates the two 384->2 heads (net.py:57-66,88)
  if (lin(G1_GATE, act[0], kD, E, kD, act[1], nullptr, nullptr, nullptr, nullptr, -1, 2, false)) return 1;
  if (lin(G1_RES0, act[0], kD, E, kD, act[2], nullptr, nullptr, nullptr, nullptr, -1, 1, false)) return 1;
  if (lin(G1_RES2, act[2], kD, E, kD, act[3], act[1], act[0], nullptr, nullptr, N_GRU2, 0, false)) return 1;
  cudaMemsetAsync(acc, 0, (size_t)E * 4 * sizeof(float), st);
  if (lin(G3_GATE, act[3], kD, E, kD, act[0], nullptr, nullptr, nullptr, nullptr, -1, 2, false)) return 1;
  if (lin(G3_RES0, act[3], kD, E, kD, act[1], nullptr, nullptr, nullptr, nullptr, -1, 1, false)) return 1;
  if (lin(G3_RES2, act[1], kD, E, kD, net_out, act[0], act[3], nullptr, nullptr, -1, 0, true)) return 1;
  if (dpvo_update_heads_finish(acc, w->head_bias_d, w->head_bias_w, E, delta, weight, st)) return 1;
  return check_launch("update_forward");
}
