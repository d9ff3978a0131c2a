// DPViewer's two plain CUDA helpers, headless: results go to caller-owned device buffers,
// there is no window, GL buffer or interop (unavailable on this box).
//
//   pose_to_matrix_kernel (modules/DPViewer/dpviewer/viewer_cuda.cu:221-262)  -> dpvo_viewer_pose_to_matrix
//   iproj_kernel (viewer_cuda.cu:56-164)                                       -> dpvo_viewer_iproj
//
// Same arithmetic in the same order (float32; the reference's double literals only scale by exact powers of
// two or compare).  Differences in structure: the other frames' poses / intrinsics are staged in shared memory
// once per CTA instead of being re-read under a barrier inside divergent code, and border pixels, which the
// reference leaves at the zeros of its freshly allocated outputs, are written as zeros.
#include "common.cuh"

namespace dpvo {
namespace {

__device__ __forceinline__ void act_so3(const float* q, const float* X, float* Y) {
  float uv[3];
  uv[0] = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  uv[1] = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  uv[2] = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  Y[0] = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  Y[1] = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  Y[2] = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
}

__device__ __forceinline__ void act_se3(const float* t, const float* q, const float* X, float* Y) {
  act_so3(q, X, Y);
  Y[3] = X[3];
  Y[0] += X[3] * t[0];
  Y[1] += X[3] * t[1];
  Y[2] += X[3] * t[2];
}

__device__ __forceinline__ void inv_se3(const float* t, const float* q, float* tinv, float* qinv) {
  qinv[0] = -q[0]; qinv[1] = -q[1]; qinv[2] = -q[2]; qinv[3] = q[3];
  act_so3(qinv, t, tinv);
  tinv[0] = -tinv[0]; tinv[1] = -tinv[1]; tinv[2] = -tinv[2];
}

__global__ void __launch_bounds__(128)
pose_to_matrix_kernel(const float* __restrict__ poses, float* __restrict__ mat, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float t0[3], q0[4], t[3], q[4];
#pragma unroll
  for (int c = 0; c < 3; ++c) t0[c] = poses[i * 7 + c];
#pragma unroll
  for (int c = 0; c < 4; ++c) q0[c] = poses[i * 7 + 3 + c];
  inv_se3(t0, q0, t, q);
  float m[16];
  m[0] = 1 - 2 * q[1] * q[1] - 2 * q[2] * q[2];
  m[1] = 2 * q[0] * q[1] - 2 * q[3] * q[2];
  m[2] = 2 * q[0] * q[2] + 2 * q[3] * q[1];
  m[3] = t[0];
  m[4] = 2 * q[0] * q[1] + 2 * q[3] * q[2];
  m[5] = 1 - 2 * q[0] * q[0] - 2 * q[2] * q[2];
  m[6] = 2 * q[1] * q[2] - 2 * q[3] * q[0];
  m[7] = t[1];
  m[8] = 2 * q[0] * q[2] - 2 * q[3] * q[1];
  m[9] = 2 * q[1] * q[2] + 2 * q[3] * q[0];
  m[10] = 1 - 2 * q[0] * q[0] - 2 * q[1] * q[1];
  m[11] = t[2];
  m[12] = 0.f; m[13] = 0.f; m[14] = 0.f; m[15] = 1.f;
  float4* o = reinterpret_cast<float4*>(mat + i * 16);     // 64-byte rows of a cudaMalloc'd / torch buffer
#pragma unroll
  for (int r = 0; r < 4; ++r) o[r] = make_float4(m[4 * r], m[4 * r + 1], m[4 * r + 2], m[4 * r + 3]);
}

constexpr int kFrameTile = 64;   // other frames staged per pass: 64 x (7 + 4) floats of shared memory

__global__ void __launch_bounds__(128)
iproj_kernel(int index, int n_frames, float thresh, const uint8_t* __restrict__ images, const float* __restrict__ poses,
             const float* __restrict__ disps, const float* __restrict__ intrinsics, int ht, int wd,
             float* __restrict__ points, uint8_t* __restrict__ colors, float* __restrict__ count) {
  __shared__ float s_pose[kFrameTile][7];
  __shared__ float s_intr[kFrameTile][4];
  const int64_t hw = (int64_t)ht * wd;
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool in_image = k < hw;
  const int i = in_image ? (int)(k / wd) : 0;
  const int j = in_image ? (int)(k - (int64_t)i * wd) : 0;
  const bool inner = in_image && (i < ht - 1) && (j < wd - 1);

  float t[3], q[4], t1[3], q1[4], intr[4];
#pragma unroll
  for (int c = 0; c < 3; ++c) t[c] = poses[(int64_t)index * 7 + c];
#pragma unroll
  for (int c = 0; c < 4; ++c) q[c] = poses[(int64_t)index * 7 + 3 + c];
#pragma unroll
  for (int c = 0; c < 4; ++c) intr[c] = 8 * intrinsics[(int64_t)index * 4 + c];
  inv_se3(t, q, t1, q1);

  float cnt = 0.f, d = 0.f;
  float X0[4] = {0.f, 0.f, 0.f, 0.f}, X1[4] = {0.f, 0.f, 0.f, 0.f};
  if (inner) {
    const float* dp = disps + (int64_t)index * hw + (int64_t)i * wd + j;
    d = dp[0];
    const float dx = dp[1] - d;
    const float dy = dp[wd] - d;
    if ((double)sqrtf(dx * dx + dy * dy) > 0.01) cnt = -100.f;
    X0[0] = ((float)j - intr[2]) / intr[0];
    X0[1] = ((float)i - intr[3]) / intr[1];
    X0[2] = 1.f;
    X0[3] = d;
    act_se3(t1, q1, X0, X1);
    points[k * 3 + 0] = X0[0] / X0[3];
    points[k * 3 + 1] = X0[1] / X0[3];
    points[k * 3 + 2] = X0[2] / X0[3];
    const uint8_t* im = images + (int64_t)index * 3 * hw + (int64_t)i * wd + j;
    colors[k * 3 + 0] = im[2 * hw];
    colors[k * 3 + 1] = im[hw];
    colors[k * 3 + 2] = im[0];
  } else if (in_image) {
    points[k * 3 + 0] = 0.f; points[k * 3 + 1] = 0.f; points[k * 3 + 2] = 0.f;
    colors[k * 3 + 0] = 0; colors[k * 3 + 1] = 0; colors[k * 3 + 2] = 0;
  }

  for (int base = 0; base < n_frames; base += kFrameTile) {
    const int nt = min(kFrameTile, n_frames - base);
    __syncthreads();
    for (int x = threadIdx.x; x < nt * 7; x += blockDim.x) s_pose[x / 7][x % 7] = poses[(int64_t)base * 7 + x];
    for (int x = threadIdx.x; x < nt * 4; x += blockDim.x) s_intr[x / 4][x % 4] = 8 * intrinsics[(int64_t)base * 4 + x];
    __syncthreads();
    if (!inner) continue;
    for (int f = 0; f < nt; ++f) {
      const int jx = base + f;
      if (jx == index) continue;
      float X2[4];
      act_se3(&s_pose[f][0], &s_pose[f][3], X1, X2);
      const float x1 = s_intr[f][0] * (X2[0] / X2[2]) + s_intr[f][2];
      const float y1 = s_intr[f][1] * (X2[1] / X2[2]) + s_intr[f][3];
      const int i1 = (int)roundf(y1);
      const int j1 = (int)roundf(x1);
      if ((i1 >= 0) && (i1 < ht) && (j1 >= 0) && (j1 < wd) && ((double)d > 0.1)) {
        const float z1 = disps[(int64_t)jx * hw + (int64_t)i1 * wd + j1];
        const float z2 = X2[3] / X2[2];
        if (100 * (fmaxf(z1 / z2, z2 / z1) - 1) < thresh) cnt += 1.f;
      }
    }
  }
  if (in_image) count[k] = cnt;
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

extern "C" int dpvo_viewer_pose_to_matrix(const float* poses, float* mat4x4, int64_t n, void* stream) {
  DPVO_REQUIRE(n >= 0, "viewer_pose_to_matrix: negative count");
  if (n == 0) return 0;
  DPVO_REQUIRE(poses && mat4x4, "viewer_pose_to_matrix: null pointer");
  DPVO_REQUIRE((reinterpret_cast<uintptr_t>(mat4x4) & 15) == 0, "viewer_pose_to_matrix: output must be 16-byte aligned");
  pose_to_matrix_kernel<<<(unsigned)ceil_div64(n, 128), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(poses, mat4x4, n);
  note_launches(1);
  return check_launch("viewer_pose_to_matrix");
}

extern "C" int dpvo_viewer_iproj(int32_t index, int32_t n_frames, float thresh, const uint8_t* images, const float* poses,
                                 const float* disps, const float* intrinsics, int32_t ht, int32_t wd, float* points,
                                 uint8_t* colors, float* count, void* stream) {
  DPVO_REQUIRE(ht >= 0 && wd >= 0 && n_frames >= 0, "viewer_iproj: bad sizes");
  DPVO_REQUIRE(index >= 0 && index < n_frames, "viewer_iproj: index outside the frames");
  if ((int64_t)ht * wd == 0) return 0;
  DPVO_REQUIRE(images && poses && disps && intrinsics && points && colors && count, "viewer_iproj: null pointer");
  iproj_kernel<<<(unsigned)ceil_div64((int64_t)ht * wd, 128), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      index, n_frames, thresh, images, poses, disps, intrinsics, ht, wd, points, colors, count);
  note_launches(1);
  return check_launch("viewer_iproj");
}
