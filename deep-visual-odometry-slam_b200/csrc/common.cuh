// Shared helpers for libdpvo_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/dpvo_b200.h"

#ifndef __CUDA_ARCH_LIST__
#endif

namespace dpvo {

constexpr int kNumSMs = 148;  // B200

void set_error(const char* fmt, ...);
void note_launches(int n);  // kernel launches issued by this library (dpvo_launch_count)

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: CUDA launch failed: %s", what, cudaGetErrorString(e));
    return 1;
  }
  return 0;
}

// Development switches (DPVO_LINEAR_*, DPVO_CORR_NARROW, DPVO_ENC_MT, DPVO_DEBUG_SYNC) exist only in a library
// built with -DDPVO_DEBUG_HOOKS (DPVO_DEBUG_HOOKS=1 python build.py); the shipped library never reads the environment.
#ifdef DPVO_DEBUG_HOOKS
#include <stdlib.h>
inline const char* debug_env(const char* name) { return getenv(name); }
#else
inline const char* debug_env(const char*) { return nullptr; }
#endif

#define DPVO_REQUIRE(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      ::dpvo::set_error(__VA_ARGS__);      \
      return 1;                            \
    }                                      \
  } while (0)

__host__ __device__ inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

// floor(v) -> int with the reference's float->int conversion (cvt.rzi: NaN -> 0,
// +-inf saturate), then clamped so later +-radius arithmetic cannot wrap.  Any
// |value| beyond 2^24 is out of every feature map either way
// (correlation_kernel.cu:119-123 treats it as out of bounds -> 0).
__device__ __forceinline__ int floor_to_int(float v) {
  int f = static_cast<int>(floorf(v));
  return max(-(1 << 24), min(1 << 24, f));
}

// 16-byte read-only global load.
__device__ __forceinline__ uint4 ldg128(const void* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}

}  // namespace dpvo
