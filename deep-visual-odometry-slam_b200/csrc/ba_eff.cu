// Block-E ("efficient E") Schur path for global bundle adjustment — placeholder until the
// kernels land; see ba.cu for the dense path.
#include "ba_internal.cuh"

namespace dpvo {

int64_t ba_eff_carve(BADev& p, char* base, int64_t E, int64_t npatches, int N, int ppf, int64_t nframes) {
  (void)p; (void)base; (void)E; (void)npatches; (void)N; (void)ppf; (void)nframes;
  return 0;
}

int ba_eff_run(BADev& p, int iterations, cudaStream_t st) {
  (void)p; (void)iterations; (void)st;
  set_error("ba_forward: eff_impl=True path is not built yet");
  return 1;
}

}  // namespace dpvo
