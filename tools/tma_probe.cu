// Diagnostics: how fast can one SM pull 128-byte-swizzled 64x128 fp16 boxes through TMA?  A persistent
// CTA per SM, one issuing thread, an S-stage ring whose stages are released as soon as they land (no
// consumer).  Used to tell a TMA / L2 delivery limit from contention inside the edge-MLP kernels
// (profiles/r01_notes.md).  Not part of the product path.
#include "common.cuh"
#include "sm100.cuh"

namespace dpvo {
namespace {
using namespace sm100;

__global__ void __launch_bounds__(64, 1)
tma_probe_kernel(const __grid_constant__ CUtensorMap tm, int64_t rows, int kblocks, int stages, int iters,
                 long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full[12];
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int64_t tiles = rows / 128;
    const long long t0 = clock64();
    uint32_t it = 0;
    for (int rep = 0; rep < iters; ++rep)
      for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x)
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % stages;
          if (it >= (uint32_t)stages) mbar_wait(&full[s], ((it / stages) - 1) & 1);   // previous use of the stage landed
          mbar_arrive_expect_tx(&full[s], 16384);
          tma_load_2d(smem + s * 16384, &tm, &full[s], kb * 64, static_cast<int32_t>(tile * 128));
        }
    for (int s = 0; s < stages && s < (int)it; ++s) {
      const uint32_t uses = (it - s + stages - 1) / stages;
      mbar_wait(&full[s], (uses - 1) & 1);
    }
    cycles[blockIdx.x] = clock64() - t0;
  }
}

}  // namespace
}  // namespace dpvo

using namespace dpvo;

typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

/* Diagnostics only: A fp16 [rows, K] (rows % 128 == 0, K % 64 == 0, contiguous).  Every CTA walks its
 * row tiles `iters` times loading all K chunks; cycles[grid] receives each CTA's elapsed clocks. */
extern "C" int dpvo_debug_tma_probe(const void* A, int64_t rows, int32_t K, int32_t stages, int32_t iters,
                                    int32_t grid, long long* cycles, void* stream) {
  DPVO_REQUIRE(A && cycles && rows > 0 && rows % 128 == 0 && K > 0 && K % 64 == 0 && stages >= 1 && stages <= 12 &&
                   grid >= 1 && iters >= 1, "tma_probe: bad arguments");
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  DPVO_REQUIRE(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
                   q == cudaDriverEntryPointSuccess, "tma_probe: no cuTensorMapEncodeTiled");
  CUtensorMap tm;
  const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)K * 2};
  const cuuint32_t box[2] = {64, 128};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = reinterpret_cast<EncodeTiledFn2>(sym)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(A),
                                                          gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DPVO_REQUIRE(r == CUDA_SUCCESS, "tma_probe: encode failed (%d)", (int)r);
  const int smem = stages * 16384 + 1024;
  cudaError_t e = cudaFuncSetAttribute(tma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  DPVO_REQUIRE(e == cudaSuccess, "tma_probe: smem: %s", cudaGetErrorString(e));
  tma_probe_kernel<<<grid, 64, smem, reinterpret_cast<cudaStream_t>(stream)>>>(tm, rows, K / 64, stages, iters, cycles);
  return check_launch("tma_probe");
}
