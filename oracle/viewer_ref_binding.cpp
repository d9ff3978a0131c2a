// TEST INFRASTRUCTURE.  pybind wrapper (this repo's code) around the reference's UNTOUCHED
// modules/DPViewer/dpviewer/viewer_cuda.cu, so that its two plain CUDA kernels (iproj_kernel via
// backproject_and_filter, pose_to_matrix_kernel via poseToMatrix) can be run headless on the GPU box as the
// parity oracle of csrc/viewer.cu.  The reference exposes them only through the Pangolin/OpenGL viewer class
// (viewer.cpp), which cannot be built here (no graphics stack).
#include <torch/extension.h>
#include "viewer_cuda.h"

static std::vector<torch::Tensor> backproject(int index, int nFrames, double thresh, bool fg, bool bg,
                                              torch::Tensor images, torch::Tensor poses, torch::Tensor disps,
                                              torch::Tensor masks, torch::Tensor intrinsics) {
  PointCloud pc = backproject_and_filter(index, nFrames, (float)thresh, fg, bg, images, poses, disps, masks, intrinsics);
  return {pc.points, pc.colors};
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("backproject_and_filter", &backproject);
  m.def("poseToMatrix", &poseToMatrix);
}
