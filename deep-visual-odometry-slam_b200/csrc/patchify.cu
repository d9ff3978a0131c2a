// Patch gather / scatter (altcorr.patchify).
// Replaces /root/reference/methods/dpvo/dpvo/altcorr/correlation_kernel.cu:17-48
// (forward gather of a (2R+2)^2 window at floor(coords), zeros outside the map,
// output zero-initialised at :298) and :50-81 (backward atomic scatter-add).
// One thread per output element; the whole output tensor is written here so the
// caller does not need a separate zero fill.
#include "common.cuh"

namespace dpvo {

template <typename T>
__global__ void __launch_bounds__(256)
patchify_fwd_kernel(const T* __restrict__ net, int64_t sb, int64_t sc, int64_t sy, int64_t sx,
                    int C, int H, int W, const float* __restrict__ coords, int M, int R,
                    T* __restrict__ out, int64_t total) {
  const int D = 2 * R + 2;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    int64_t n = idx;
    const int jj = (int)(n % D); n /= D;
    const int ii = (int)(n % D); n /= D;
    const int c = (int)(n % C); n /= C;
    const int m = (int)(n % M); n /= M;
    const int b = (int)n;
    const float x = coords[((int64_t)b * M + m) * 2 + 0];
    const float y = coords[((int64_t)b * M + m) * 2 + 1];
    const int i = floor_to_int(y) + (ii - R);
    const int j = floor_to_int(x) + (jj - R);
    T v = T(0);
    if (i >= 0 && i < H && j >= 0 && j < W)
      v = net[(int64_t)b * sb + (int64_t)c * sc + (int64_t)i * sy + (int64_t)j * sx];
    out[idx] = v;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
patchify_bwd_kernel(const T* __restrict__ pg, const float* __restrict__ coords,
                    int M, int C, int H, int W, int R, T* __restrict__ grad, int64_t total) {
  const int D = 2 * R + 2;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    int64_t n = idx;
    const int jj = (int)(n % D); n /= D;
    const int ii = (int)(n % D); n /= D;
    const int c = (int)(n % C); n /= C;
    const int m = (int)(n % M); n /= M;
    const int b = (int)n;
    const float x = coords[((int64_t)b * M + m) * 2 + 0];
    const float y = coords[((int64_t)b * M + m) * 2 + 1];
    const int i = floor_to_int(y) + (ii - R);
    const int j = floor_to_int(x) + (jj - R);
    if (i >= 0 && i < H && j >= 0 && j < W)
      atomicAdd(&grad[(((int64_t)b * C + c) * H + i) * W + j], pg[idx]);
  }
}

static int grid_for(int64_t total) {
  int64_t g = ceil_div64(total, 256);
  const int64_t cap = (int64_t)kNumSMs * 32;
  return (int)(g < cap ? g : cap);
}

}  // namespace dpvo

using namespace dpvo;

extern "C" int dpvo_patchify_forward(dpvo_dtype dtype, const void* net,
                                     int64_t n_sb, int64_t n_sc, int64_t n_sy, int64_t n_sx,
                                     int32_t B, int32_t C, int32_t H, int32_t W,
                                     const float* coords, int32_t M, int32_t radius,
                                     void* patches, void* stream) {
  DPVO_REQUIRE(radius >= 0 && B >= 0 && M >= 0 && C >= 0, "patchify_forward: bad sizes");
  const int D = 2 * radius + 2;
  const int64_t total = (int64_t)B * M * C * D * D;
  if (total == 0) return 0;
  DPVO_REQUIRE(net && coords && patches, "patchify_forward: null pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = grid_for(total);
  switch (dtype) {
    case DPVO_F16:
      patchify_fwd_kernel<__half><<<grid, 256, 0, st>>>(reinterpret_cast<const __half*>(net), n_sb, n_sc, n_sy, n_sx,
                                                       C, H, W, coords, M, radius, reinterpret_cast<__half*>(patches), total);
      break;
    case DPVO_F32:
      patchify_fwd_kernel<float><<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(net), n_sb, n_sc, n_sy, n_sx,
                                                      C, H, W, coords, M, radius, reinterpret_cast<float*>(patches), total);
      break;
    case DPVO_F64:
      patchify_fwd_kernel<double><<<grid, 256, 0, st>>>(reinterpret_cast<const double*>(net), n_sb, n_sc, n_sy, n_sx,
                                                       C, H, W, coords, M, radius, reinterpret_cast<double*>(patches), total);
      break;
    default:
      set_error("patchify_forward: unknown dtype");
      return 1;
  }
  note_launches(1);
  return check_launch("patchify_fwd");
}

extern "C" int dpvo_patchify_backward(dpvo_dtype dtype, const void* patch_grad,
                                      const float* coords, int32_t B, int32_t M, int32_t C,
                                      int32_t H, int32_t W, int32_t radius,
                                      void* net_grad, void* stream) {
  DPVO_REQUIRE(radius >= 0 && B >= 0 && M >= 0 && C >= 0, "patchify_backward: bad sizes");
  const int D = 2 * radius + 2;
  const int64_t total = (int64_t)B * M * C * D * D;
  if (total == 0) return 0;
  DPVO_REQUIRE(patch_grad && coords && net_grad, "patchify_backward: null pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = grid_for(total);
  switch (dtype) {
    case DPVO_F16:
      patchify_bwd_kernel<__half><<<grid, 256, 0, st>>>(reinterpret_cast<const __half*>(patch_grad), coords, M, C, H, W,
                                                       radius, reinterpret_cast<__half*>(net_grad), total);
      break;
    case DPVO_F32:
      patchify_bwd_kernel<float><<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(patch_grad), coords, M, C, H, W,
                                                      radius, reinterpret_cast<float*>(net_grad), total);
      break;
    case DPVO_F64:
      patchify_bwd_kernel<double><<<grid, 256, 0, st>>>(reinterpret_cast<const double*>(patch_grad), coords, M, C, H, W,
                                                       radius, reinterpret_cast<double*>(net_grad), total);
      break;
    default:
      set_error("patchify_backward: unknown dtype");
      return 1;
  }
  note_launches(1);
  return check_launch("patchify_bwd");
}
