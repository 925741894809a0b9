// Shared device/host helpers for libm3r_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#include "../../include/m3r_abi.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libm3r_b200 is written for sm_100a (B200) only"
#endif

namespace m3r {

void set_last_error(const char* what, cudaError_t err);
void set_last_error_text(const char* what);

#define M3R_CUDA_CHECK(expr)                         \
  do {                                               \
    cudaError_t _e = (expr);                         \
    if (_e != cudaSuccess) {                         \
      ::m3r::set_last_error(#expr, _e);              \
      return M3R_ERR_CUDA;                           \
    }                                                \
  } while (0)

#define M3R_LAUNCH_CHECK()                           \
  do {                                               \
    cudaError_t _e = cudaPeekAtLastError();          \
    if (_e != cudaSuccess) {                         \
      ::m3r::set_last_error("kernel launch", _e);    \
      (void)cudaGetLastError();                      \
      return M3R_ERR_CUDA;                           \
    }                                                \
  } while (0)

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// The opt-in to more than 48 KB of dynamic shared memory is an attribute of (kernel, device): a process that drives
// several GPUs has to set it on each of them.  One of these per kernel (static storage); ensure() is cheap after the
// first call on a device.
struct DynSmemOptIn {
  int bytes_set[64] = {};
  template <typename Kernel>
  cudaError_t ensure(Kernel kernel, int bytes) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64 && bytes_set[dev] >= bytes) return cudaSuccess;  // (benign race: setting twice is harmless)
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess && dev >= 0 && dev < 64) bytes_set[dev] = bytes;
    return e;
  }
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace m3r
