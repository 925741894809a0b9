// refine_matches — dilated-window descriptor argmax (sm_100a).
//
// Replaces mast3r_slam/backend/src/matching_kernels.cu:25-116.  Spec: SURVEY.md Appendix B.
// Parity rules that decide ties (all kept):
//   * scan order: dilation d = D..1; u-offset outer, v-offset inner (:54-57)
//   * strict `>` against a best score that is carried across levels (:47,:65), initial best =
//     cuda::std::numeric_limits<T>::min(): 0 for c10::Half, FLT_MIN / DBL_MIN for float / double
//   * fp16: score = fold_k( round16( score + round16(q_k * c_k) ) ), k ascending.  c10::Half does
//     this through float; IEEE half mul/add with a single RN rounding give the same bits (the
//     float product of two halves is exact; a float sum of two halves rounded again to half is
//     innocuous double rounding because 24 >= 2*11+2).  HMUL2/HADD2 `_rn` intrinsics are never
//     contracted to HFMA2.
//
// fp16 / fdim = 24 fast path (the shape every caller uses): one thread per query keeps the query
// descriptor in registers, duplicated per half2 lane, and scores TWO window candidates at a time,
// one per half2 lane (candidate words are lane-transposed with PRMT), so each pair of MACs for two
// candidates costs 2 PRMT + 2 HMUL2 + 2 HADD2.  A candidate descriptor is 48 contiguous bytes =
// three 16-byte read-only loads.  Neighbouring queries have neighbouring windows, so a warp's
// loads for one candidate offset are (nearly) contiguous and L1/L2-resident.
#include <cuda_fp16.h>

#include <cfloat>

#include "common.cuh"

namespace m3r {
namespace {

constexpr int kRefineThreads = 128;

__device__ __forceinline__ bool inside(int u, int v, int w, int h) {
  return v >= 0 && v < h && u >= 0 && u < w;
}

union Word4 {
  uint4 v;
  __half2 h2[4];
};

__global__ void __launch_bounds__(kRefineThreads)
refine_f16_fd24_kernel(const __half* __restrict__ D11, const __half* __restrict__ D21,
                       const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new, int h, int w,
                       int n, int radius, int dilation_max) {
  constexpr int FD = 24;
  const int k = blockIdx.x * kRefineThreads + threadIdx.x;
  const int b = blockIdx.y;
  if (k >= n) return;
  const size_t qi = static_cast<size_t>(b) * n + k;
  const uint4* img = reinterpret_cast<const uint4*>(D11 + static_cast<size_t>(b) * h * w * FD);

  // query descriptor, each element duplicated into both half2 lanes
  __half2 qd[FD];
  {
    const uint4* q4 = reinterpret_cast<const uint4*>(D21 + qi * FD);
#pragma unroll
    for (int t = 0; t < 3; t++) {
      Word4 wq;
      wq.v = __ldg(q4 + t);
#pragma unroll
      for (int c = 0; c < 4; c++) {
        qd[t * 8 + c * 2 + 0] = __low2half2(wq.h2[c]);
        qd[t * 8 + c * 2 + 1] = __high2half2(wq.h2[c]);
      }
    }
  }

  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + qi);
  int u0 = static_cast<int>(p.x), v0 = static_cast<int>(p.y);
  int ub = u0, vb = v0;
  float best = 0.0f;  // numeric_limits<c10::Half>::min() == Half() == 0
  const int cnt = 2 * radius + 1;
  const int total = cnt * cnt;

  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    int ia = 0, ja = 0;  // (u-offset index, v-offset index) of candidate A
    for (int c = 0; c < total; c += 2) {
      int ib = ia, jb = ja + 1;
      if (jb == cnt) {
        jb = 0;
        ib = ia + 1;
      }
      const int uA = u0 - rd + ia * d, vA = v0 - rd + ja * d;
      const int uB = u0 - rd + ib * d, vB = v0 - rd + jb * d;
      const bool inA = inside(uA, vA, w, h);
      const bool inB = (c + 1 < total) && inside(uB, vB, w, h);
      if (inA || inB) {
        const uint4* pa = img + (inA ? (static_cast<size_t>(vA) * w + uA) * 3 : 0);
        const uint4* pb = img + (inB ? (static_cast<size_t>(vB) * w + uB) * 3 : 0);
        Word4 a[3], bw[3];
#pragma unroll
        for (int t = 0; t < 3; t++) {
          a[t].v = __ldg(pa + t);
          bw[t].v = __ldg(pb + t);
        }
        __half2 s = __float2half2_rn(0.0f);
#pragma unroll
        for (int t = 0; t < 3; t++) {
#pragma unroll
          for (int cc = 0; cc < 4; cc++) {
            const __half2 lo = __lows2half2(a[t].h2[cc], bw[t].h2[cc]);
            const __half2 hi = __highs2half2(a[t].h2[cc], bw[t].h2[cc]);
            s = __hadd2_rn(s, __hmul2_rn(qd[t * 8 + cc * 2 + 0], lo));
            s = __hadd2_rn(s, __hmul2_rn(qd[t * 8 + cc * 2 + 1], hi));
          }
        }
        const float sA = __low2float(s), sB = __high2float(s);
        if (inA && sA > best) {
          best = sA;
          ub = uA;
          vb = vA;
        }
        if (inB && sB > best) {
          best = sB;
          ub = uB;
          vb = vB;
        }
      }
      // advance A by two candidates
      ja += 2;
      if (ja >= cnt) {
        ja -= cnt;
        ia += 1;
      }
    }
    u0 = ub;
    v0 = vb;
  }
  reinterpret_cast<longlong2*>(p1_new)[qi] = make_longlong2(ub, vb);
}

// ---- generic path: any fdim, fp16 / fp32 / fp64, one candidate at a time ----------------------
template <typename T>
struct Acc;
template <>
struct Acc<__half> {
  using S = __half;
  static __device__ __forceinline__ S zero() { return __float2half_rn(0.0f); }
  static __device__ __forceinline__ S mac(S s, __half a, __half b) { return __hadd_rn(s, __hmul_rn(a, b)); }
  static __device__ __forceinline__ float cmp(S s) { return __half2float(s); }
  static __device__ __forceinline__ float init() { return 0.0f; }
};
template <>
struct Acc<float> {
  using S = float;
  static __device__ __forceinline__ S zero() { return 0.0f; }
  static __device__ __forceinline__ S mac(S s, float a, float b) { return s + a * b; }  // FFMA like the ref
  static __device__ __forceinline__ float cmp(S s) { return s; }
  static __device__ __forceinline__ float init() { return FLT_MIN; }
};
template <>
struct Acc<double> {
  using S = double;
  static __device__ __forceinline__ S zero() { return 0.0; }
  static __device__ __forceinline__ S mac(S s, double a, double b) { return s + a * b; }
  static __device__ __forceinline__ double cmp(S s) { return s; }
  static __device__ __forceinline__ double init() { return DBL_MIN; }
};

template <typename T>
__global__ void __launch_bounds__(kRefineThreads)
refine_generic_kernel(const T* __restrict__ D11, const T* __restrict__ D21,
                      const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new, int h, int w,
                      int n, int fdim, int radius, int dilation_max) {
  const int k = blockIdx.x * kRefineThreads + threadIdx.x;
  const int b = blockIdx.y;
  if (k >= n) return;
  const size_t qi = static_cast<size_t>(b) * n + k;
  const T* img = D11 + static_cast<size_t>(b) * h * w * fdim;
  const T* q = D21 + qi * fdim;
  int u0 = static_cast<int>(p1[qi * 2 + 0]), v0 = static_cast<int>(p1[qi * 2 + 1]);
  int ub = u0, vb = v0;
  auto best = Acc<T>::init();
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      for (int j = 0; j < diam; j += d) {
        const int u = u0 - rd + i, v = v0 - rd + j;
        if (inside(u, v, w, h)) {
          const T* c = img + (static_cast<size_t>(v) * w + u) * fdim;
          typename Acc<T>::S s = Acc<T>::zero();
          for (int t = 0; t < fdim; t++) s = Acc<T>::mac(s, q[t], c[t]);
          if (Acc<T>::cmp(s) > best) {
            best = Acc<T>::cmp(s);
            ub = u;
            vb = v;
          }
        }
      }
    }
    u0 = ub;
    v0 = vb;
  }
  p1_new[qi * 2 + 0] = ub;
  p1_new[qi * 2 + 1] = vb;
}

template <typename T>
int launch_generic(const T* D11, const T* D21, const int64_t* p1, int64_t* p1_new, int b, int h, int w,
                   int n, int fdim, int radius, int dilation_max, void* stream) {
  dim3 grid((n + kRefineThreads - 1) / kRefineThreads, b);
  refine_generic_kernel<T><<<grid, kRefineThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      D11, D21, p1, p1_new, h, w, n, fdim, radius, dilation_max);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int check_args(const void* D11, const void* D21, const void* p1, const void* p1_new, int b, int h, int w,
               int n, int fdim, int radius, int dilation_max) {
  if (!D11 || !D21 || !p1 || !p1_new) return M3R_ERR_BAD_ARG;
  if (b < 0 || n < 0 || h < 1 || w < 1 || fdim < 1 || radius < 0 || dilation_max < 0) return M3R_ERR_BAD_ARG;
  if (b > 65535) return M3R_ERR_BAD_ARG;
  return M3R_OK;
}

}  // namespace
}  // namespace m3r

extern "C" int m3r_refine_matches_f16(const void* D11, const void* D21, const int64_t* p1,
                                      int64_t* p1_new, int b, int h, int w, int n, int fdim,
                                      int radius, int dilation_max, void* stream) {
  int rc = m3r::check_args(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max);
  if (rc != M3R_OK) return rc;
  if (b == 0 || n == 0) return M3R_OK;
  const bool aligned = ((reinterpret_cast<uintptr_t>(D11) | reinterpret_cast<uintptr_t>(D21) |
                         reinterpret_cast<uintptr_t>(p1) | reinterpret_cast<uintptr_t>(p1_new)) & 15) == 0;
  if (fdim == 24 && aligned) {
    dim3 grid((n + m3r::kRefineThreads - 1) / m3r::kRefineThreads, b);
    m3r::refine_f16_fd24_kernel<<<grid, m3r::kRefineThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __half*>(D11), static_cast<const __half*>(D21), p1, p1_new, h, w, n, radius,
        dilation_max);
    M3R_LAUNCH_CHECK();
    return M3R_OK;
  }
  return m3r::launch_generic<__half>(static_cast<const __half*>(D11), static_cast<const __half*>(D21), p1,
                                     p1_new, b, h, w, n, fdim, radius, dilation_max, stream);
}

extern "C" int m3r_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1,
                                      int64_t* p1_new, int b, int h, int w, int n, int fdim,
                                      int radius, int dilation_max, void* stream) {
  int rc = m3r::check_args(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max);
  if (rc != M3R_OK) return rc;
  if (b == 0 || n == 0) return M3R_OK;
  return m3r::launch_generic<float>(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max, stream);
}

extern "C" int m3r_refine_matches_f64(const double* D11, const double* D21, const int64_t* p1,
                                      int64_t* p1_new, int b, int h, int w, int n, int fdim,
                                      int radius, int dilation_max, void* stream) {
  int rc = m3r::check_args(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max);
  if (rc != M3R_OK) return rc;
  if (b == 0 || n == 0) return M3R_OK;
  return m3r::launch_generic<double>(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max, stream);
}
