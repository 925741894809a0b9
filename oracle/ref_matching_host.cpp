// TEST INFRASTRUCTURE — not product code.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load the library this file builds.
//
// Host build of the REFERENCE's own matching kernel bodies.  The kernel source is NOT in this
// repo: oracle/build_ref.sh extracts matching_kernels.cu:17-81 (inside_image, clamp,
// refine_matches_kernel<T>) and :119-275 (iter_proj_kernel) from /root/reference into a scratch
// file and passes its path as M3R_REF_KERNELS_INC.  This shim supplies just enough of the CUDA
// execution model (blockIdx/blockDim/threadIdx, __global__/__device__) for those bodies to
// compile as plain C++ and runs one "CUDA thread" per (b, n) in an OpenMP loop.  Neither kernel
// uses shared memory or barriers, so a sequential per-thread run is exact.
//
// This is BASELINE.json's "straightforward host C++ build of the reference kernel source"; it is
// what `bench.py --impl reference` times for the matching ops and what pins oracle/cpu_oracle.cpp.
#include <torch/torch.h>
#include <cuda/std/limits>
#include <cmath>
#include <cstdint>

#ifndef __global__
#define __global__
#endif
#ifndef __device__
#define __device__
#endif
#ifndef __forceinline__
#define __forceinline__ inline
#endif

// torch only defines RestrictPtrTraits under __CUDACC__ (torch/headeronly/core/TensorAccessor.h:22);
// the host build needs the same trait name, with the same meaning.
namespace torch {
template <typename T>
struct RestrictPtrTraits {
  typedef T* __restrict__ PtrType;
};
}  // namespace torch

namespace {
struct Idx3 { unsigned x, y, z; };
thread_local Idx3 blockIdx{0, 0, 0};
thread_local Idx3 threadIdx{0, 0, 0};
thread_local Idx3 blockDim{1, 1, 1};

#include M3R_REF_KERNELS_INC

template <typename T>
torch::Tensor wrap(const void* p, std::vector<int64_t> sizes, torch::ScalarType st) {
  return torch::from_blob(const_cast<void*>(p), sizes, torch::TensorOptions().dtype(st));
}
}  // namespace

extern "C" int ref_host_iter_proj(const float* rays, const float* pts, const float* p_init,
                                   float* p_new, uint8_t* converged, int b, int h, int w, int n,
                                   int max_iter, float lambda_init, float cost_thresh) {
  auto t_rays = wrap<float>(rays, {b, h, w, 9}, torch::kFloat32);
  auto t_pts = wrap<float>(pts, {b, n, 3}, torch::kFloat32);
  auto t_p0 = wrap<float>(p_init, {b, n, 2}, torch::kFloat32);
  auto t_p1 = wrap<float>(p_new, {b, n, 2}, torch::kFloat32);
  auto t_cv = wrap<bool>(converged, {b, n}, torch::kBool);
  auto a_rays = t_rays.packed_accessor32<float, 4, torch::RestrictPtrTraits>();
  auto a_pts = t_pts.packed_accessor32<float, 3, torch::RestrictPtrTraits>();
  auto a_p0 = t_p0.packed_accessor32<float, 3, torch::RestrictPtrTraits>();
  auto a_p1 = t_p1.packed_accessor32<float, 3, torch::RestrictPtrTraits>();
  auto a_cv = t_cv.packed_accessor32<bool, 2, torch::RestrictPtrTraits>();
  for (int bb = 0; bb < b; bb++) {
#pragma omp parallel for schedule(static)
    for (int k = 0; k < n; k++) {
      blockDim = {1, 1, 1};
      threadIdx = {0, 0, 0};
      blockIdx = {(unsigned)k, (unsigned)bb, 0};
      iter_proj_kernel(a_rays, a_pts, a_p0, a_p1, a_cv, max_iter, lambda_init, cost_thresh);
    }
  }
  return 0;
}

template <typename T>
static int run_refine(const void* D11, const void* D21, const int64_t* p1, int64_t* p1_new, int b,
                      int h, int w, int n, int fdim, int radius, int dilation_max,
                      torch::ScalarType st) {
  auto t_d11 = wrap<T>(D11, {b, h, w, fdim}, st);
  auto t_d21 = wrap<T>(D21, {b, n, fdim}, st);
  auto t_p1 = wrap<long>(p1, {b, n, 2}, torch::kInt64);
  auto t_p1n = wrap<long>(p1_new, {b, n, 2}, torch::kInt64);
  auto a_d11 = t_d11.template packed_accessor32<T, 4, torch::RestrictPtrTraits>();
  auto a_d21 = t_d21.template packed_accessor32<T, 3, torch::RestrictPtrTraits>();
  auto a_p1 = t_p1.packed_accessor32<long, 3, torch::RestrictPtrTraits>();
  auto a_p1n = t_p1n.packed_accessor32<long, 3, torch::RestrictPtrTraits>();
  for (int bb = 0; bb < b; bb++) {
#pragma omp parallel for schedule(static)
    for (int k = 0; k < n; k++) {
      blockDim = {1, 1, 1};
      threadIdx = {0, 0, 0};
      blockIdx = {(unsigned)k, (unsigned)bb, 0};
      refine_matches_kernel<T>(a_d11, a_d21, a_p1, a_p1n, radius, dilation_max);
    }
  }
  return 0;
}

extern "C" int ref_host_refine_matches_f16(const void* D11, const void* D21, const int64_t* p1,
                                            int64_t* p1_new, int b, int h, int w, int n, int fdim,
                                            int radius, int dilation_max) {
  return run_refine<c10::Half>(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max,
                               torch::kFloat16);
}

extern "C" int ref_host_refine_matches_f32(const void* D11, const void* D21, const int64_t* p1,
                                            int64_t* p1_new, int b, int h, int w, int n, int fdim,
                                            int radius, int dilation_max) {
  return run_refine<float>(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max,
                           torch::kFloat32);
}
