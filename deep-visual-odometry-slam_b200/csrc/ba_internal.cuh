// Shared device structures / per-edge math for the BA kernels.
#pragma once
#include "se3.cuh"

namespace dpvo {

struct BADev {
  // problem (caller memory)
  float* poses; float* patches;
  const float* intr; const float* target; const float* weight; const float* lmbda;
  const int64_t* ii; const int64_t* jj; const int64_t* kk;
  int64_t E, nframes, npatches;
  int P, t0, N, n, ppf;
  // compaction
  uint32_t* bitmap; int* prefix; int* meta; int64_t* kx;
  int64_t Mmax, W;
  // edge grouping by pose-pair key (counting sort): hist[K], cursor[K], off[K+1], toff[K+1], perm[E], ekey[E]
  int K;
  int *hist, *cursor, *off, *toff, *perm, *ekey;
  int* tkey;         // [E / 32 + K + 1] key of every warp task (a task = up to 32 edges of one key)
  int* ksum;         // [2 * ceil(K / 4096)] (edges, warp tasks) per block of keys: blocked scan of a large histogram
  // normal equations (dense path)
  // U, v, G, g are float64 accumulators: S = B - E Q E^T cancels most of B, so the sums over
  // ~10^3-10^4 edge / patch terms are kept exact to fp32 rounding and S, y are rounded once.
  double *U, *v, *G, *g;
  float *C, *u, *Et, *dX, *Sg;
  // block-E path (global BA): E is kept as two 6-vectors per edge (entries w.r.t. the edge's
  // source and target pose), edges are grouped by source frame for the Schur product
  float* Ee;         // [E][12]  (ei[6], ej[6]) in sorted order (below); zero for masked edges
  float* w;          // [Mmax]   E^T dX accumulator
  // counting sort of the edges by (source frame, slot of the patch in its frame): nframes * ppf keys + 1 for ignored
  // edges; fperm[pos] = edge, finv[edge] = pos, ejx[pos] = target pose - t0 or -1, fscan = scan scratch
  int *fhist, *foff, *fcursor, *fperm, *finv, *ejx, *fscan;
};

struct EdgeTerms {
  int ix, jx;        // frame indices relative to t0
  int fi, fj;        // absolute frame indices
  int64_t kf;        // patch id
  float w[2], r[2], Jz[2];
  float Ji[2][6], Jj[2][6];
};

// rank of patch id among the marked ones == torch::_unique's inverse index
__device__ __forceinline__ int patch_rank(const BADev& p, int64_t kf) {
  const int64_t w = kf >> 5;
  const uint32_t bits = p.bitmap[w];
  return p.prefix[w] + __popc(bits & ((1u << (kf & 31)) - 1u));
}

// Per-edge residual and Jacobians (ba_cuda.cu:265-333).  Returns false when the edge
// contributes nothing (invalid indices, or both row weights are zero after masking).
__device__ __forceinline__ bool edge_terms(const BADev& p, int64_t e, float fx, float fy, float cx, float cy,
                                           EdgeTerms& T) {
  const int64_t fi = p.ii[e], fj = p.jj[e], kf = p.kk[e];
  if (fi < 0 || fi >= p.nframes || fj < 0 || fj >= p.nframes || kf < 0 || kf >= p.npatches) return false;
  const float* pi = p.poses + fi * 7;
  const float* pj = p.poses + fj * 7;
  const Vec3 ti = {pi[0], pi[1], pi[2]};
  const Quat qi = {pi[3], pi[4], pi[5], pi[6]};
  const Vec3 tj = {pj[0], pj[1], pj[2]};
  const Quat qj = {pj[3], pj[4], pj[5], pj[6]};
  const int PP = p.P * p.P;
  const int ctr = (p.P / 2) * p.P + (p.P / 2);  // [1][1] for 3x3 (ba_cuda.cu:282-285)
  const float* pt = p.patches + kf * 3 * PP;
  const Vec3 Xi = {(pt[ctr] - cx) / fx, (pt[PP + ctr] - cy) / fy, 1.0f};
  const float Wd = pt[2 * PP + ctr];
  Vec3 tij;
  Quat qij;
  rel_se3(ti, qi, tj, qj, tij, qij);
  Vec3 Y3 = rot(qij, Xi);
  const float X = Y3.x + Wd * tij.x;
  const float Y = Y3.y + Wd * tij.y;
  const float Z = Y3.z + Wd * tij.z;
  const float d = (Z >= 0.2f) ? 1.0f / Z : 0.0f;
  const float d2 = d * d;
  const float x1 = fx * (X / Z) + cx;
  const float y1 = fy * (Y / Z) + cy;
  const float rx = p.target[2 * e + 0] - x1;
  const float ry = p.target[2 * e + 1] - y1;
  const bool inb = (sqrtf(rx * rx + ry * ry) < 128.f) && (Z > 0.2f) && (x1 > -64.f) && (y1 > -64.f) &&
                   (x1 < 2.f * cx + 64.f) && (y1 < 2.f * cy + 64.f);
  if (!inb) return false;
  T.w[0] = p.weight[2 * e + 0];
  T.w[1] = p.weight[2 * e + 1];
  if (T.w[0] == 0.f && T.w[1] == 0.f) return false;
  T.fi = (int)fi; T.fj = (int)fj; T.kf = kf;
  T.ix = (int)fi - p.t0;
  T.jx = (int)fj - p.t0;
  T.r[0] = rx; T.r[1] = ry;
  T.Jz[0] = fx * (tij.x * d - tij.z * (X * d2));
  T.Jz[1] = fy * (tij.y * d - tij.z * (Y * d2));
  T.Jj[0][0] = fx * Wd * d;  T.Jj[0][1] = 0.f;          T.Jj[0][2] = fx * -X * Wd * d2;
  T.Jj[0][3] = fx * -X * Y * d2; T.Jj[0][4] = fx * (1.f + X * X * d2); T.Jj[0][5] = fx * -Y * d;
  T.Jj[1][0] = 0.f;          T.Jj[1][1] = fy * Wd * d;  T.Jj[1][2] = fy * -Y * Wd * d2;
  T.Jj[1][3] = fy * (-1.f - Y * Y * d2); T.Jj[1][4] = fy * (X * Y * d2); T.Jj[1][5] = fy * X * d;
  adj_se3(tij, qij, T.Jj[0], T.Ji[0]);
  adj_se3(tij, qij, T.Jj[1], T.Ji[1]);
  return true;
}

// d <- d + dz; d > 20 -> 1; d <- max(d, 1e-4); broadcast over the patch (ba_cuda.cu:209-229)
__device__ __forceinline__ void depth_retract(float* patches, int64_t kx, int PP, float dz) {
  float* dp = patches + kx * 3 * PP + 2 * PP;
  float d = dp[0] + dz;
  d = (d > 20.f) ? 1.0f : d;
  d = fmaxf(d, 1e-4f);
  for (int i = 0; i < PP; ++i) dp[i] = d;
}

int64_t ba_eff_carve(BADev& p, char* base, int64_t E, int64_t npatches, int N, int ppf, int64_t nframes);
int ba_eff_run(BADev& p, int iterations, cudaStream_t st);
// shared with the block-E path (defined in ba.cu)
void ba_launch_assemble(const BADev& p, cudaStream_t st);
int ba_launch_solve(const BADev& p, cudaStream_t st);   // S = B - G (+damping), Cholesky, dX, pose retraction; clears U, v, G, g
int ba_chol_solve(const BADev& p, float* A, int lda, float* xbuf, cudaStream_t st);
inline int chol_lda(int n) { return (n + 63) & ~63; }
int64_t chol_scratch_floats(int n);

}  // namespace dpvo
