// iter_proj — per-pixel Levenberg-Marquardt ray projection (sm_100a).
//
// Replaces mast3r_slam/backend/src/matching_kernels.cu:119-316 (iter_proj_kernel + launcher).
// Spec: SURVEY.md Appendix B.  Results must be identical to the reference's, so every float
// expression below keeps the reference's operand order (nvcc contracts them to the same FMAs) and
// its double promotions where they can change a bit (1.0/x, lambda*0.1, lambda*10.0); the
// (1.0-du) products are exact in float and stay float.
//
// What is different is the data movement.  The reference re-gathers all 9 channels x 4 taps at
// the current pixel and 3 x 4 at the candidate in every one of the max_iter iterations (48 scalar
// gathers).  Here the state at the current pixel (normalised-ray error, cost, undamped JtJ, Jtr)
// lives in registers:
//   * rejected step  -> pixel unchanged -> nothing at the current pixel is re-read or recomputed,
//                       only lambda changes;
//   * accepted step  -> the candidate's ray sample / error / cost (just computed, bit-identical to
//                       what the next iteration would recompute) become the current state and
//                       only the 6 gradient channels are gathered at the new pixel.
// After convergence almost every step is a reject, so an iteration costs 12 gathers instead of 48.
// One thread per target direction, 128-thread CTAs (4 warps), grid sized over all of b*n.
#include "common.cuh"

namespace m3r {
namespace {

constexpr int kIterProjThreads = 128;
constexpr int kCh = 9;

struct Taps {
  const float* __restrict__ p22;  // (v11, u11)       top-left texel
  const float* __restrict__ p21;  // (v11, u11+1)
  const float* __restrict__ p12;  // (v11+1, u11)
  const float* __restrict__ p11;  // (v11+1, u11+1)
  float w11, w12, w21, w22;
};

__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

__device__ __forceinline__ Taps make_taps(const float* __restrict__ img, int w, float u, float v) {
  Taps t;
  const int u11 = static_cast<int>(floorf(u));
  const int v11 = static_cast<int>(floorf(v));
  const float du = u - static_cast<float>(u11);
  const float dv = v - static_cast<float>(v11);
  t.w11 = du * dv;
  t.w12 = (1.0f - du) * dv;
  t.w21 = du * (1.0f - dv);
  t.w22 = (1.0f - du) * (1.0f - dv);
  t.p22 = img + (static_cast<size_t>(v11) * w + u11) * kCh;
  t.p21 = t.p22 + kCh;
  t.p12 = t.p22 + static_cast<size_t>(w) * kCh;
  t.p11 = t.p12 + kCh;
  return t;
}

__device__ __forceinline__ float tap(const Taps& t, int j) {
  return t.w11 * __ldg(t.p11 + j) + t.w12 * __ldg(t.p12 + j) + t.w21 * __ldg(t.p21 + j) +
         t.w22 * __ldg(t.p22 + j);
}

// normalised interpolated ray minus target -> err, returns squared error
__device__ __forceinline__ float ray_error(const Taps& t, const float tgt[3], float err[3]) {
  float r[3];
#pragma unroll
  for (int j = 0; j < 3; j++) r[j] = tap(t, j);
  const float r_norm = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  const float r_norm_inv = 1.0 / r_norm;  // double divide, rounded to float (reference :187)
#pragma unroll
  for (int j = 0; j < 3; j++) r[j] *= r_norm_inv;
#pragma unroll
  for (int j = 0; j < 3; j++) err[j] = r[j] - tgt[j];
  return err[0] * err[0] + err[1] * err[1] + err[2] * err[2];
}

struct Normal {
  float A00, A01, A11, b0, b1;
};

__device__ __forceinline__ Normal normal_eq(const Taps& t, const float err[3]) {
  float gx[3], gy[3];
#pragma unroll
  for (int j = 3; j < 6; j++) gx[j - 3] = tap(t, j);
#pragma unroll
  for (int j = 6; j < 9; j++) gy[j - 6] = tap(t, j);
  Normal s;
  s.A00 = gx[0] * gx[0] + gx[1] * gx[1] + gx[2] * gx[2];
  s.A01 = gx[0] * gy[0] + gx[1] * gy[1] + gx[2] * gy[2];
  s.A11 = gy[0] * gy[0] + gy[1] * gy[1] + gy[2] * gy[2];
  s.b0 = -(err[0] * gx[0] + err[1] * gx[1] + err[2] * gx[2]);
  s.b1 = -(err[0] * gy[0] + err[1] * gy[1] + err[2] * gy[2]);
  return s;
}

__global__ void __launch_bounds__(kIterProjThreads)
iter_proj_kernel(const float* __restrict__ rays_img, const float* __restrict__ pts3d_norm,
                 const float* __restrict__ p_init, float* __restrict__ p_new,
                 uint8_t* __restrict__ converged, int h, int w, int n, int max_iter,
                 float lambda_init, float cost_thresh) {
  const int k = blockIdx.x * kIterProjThreads + threadIdx.x;
  const int b = blockIdx.y;
  if (k >= n) return;
  const size_t pix = static_cast<size_t>(b) * n + k;
  const float* img = rays_img + static_cast<size_t>(b) * h * w * kCh;

  const float2 p0 = __ldg(reinterpret_cast<const float2*>(p_init) + pix);
  float u = clampf(p0.x, 1.0f, static_cast<float>(w - 2));
  float v = clampf(p0.y, 1.0f, static_cast<float>(h - 2));
  float tgt[3];
#pragma unroll
  for (int j = 0; j < 3; j++) tgt[j] = __ldg(pts3d_norm + pix * 3 + j);

  float lambda = lambda_init;
  bool conv = false;
  if (max_iter > 0) {
    float err[3];
    Taps t = make_taps(img, w, u, v);
    float cost = ray_error(t, tgt, err);
    Normal s = normal_eq(t, err);
    for (int it = 0; it < max_iter; it++) {
      const float A00 = s.A00 + lambda;
      const float A11 = s.A11 + lambda;
      const float det_inv = 1.0 / (A00 * A11 - s.A01 * s.A01);  // double divide (reference :213)
      const float delta_u = det_inv * (A11 * s.b0 - s.A01 * s.b1);
      const float delta_v = det_inv * (-s.A01 * s.b0 + A00 * s.b1);
      const float u_new = clampf(u + delta_u, 1.0f, static_cast<float>(w - 2));
      const float v_new = clampf(v + delta_v, 1.0f, static_cast<float>(h - 2));
      float err_new[3];
      const Taps tn = make_taps(img, w, u_new, v_new);
      const float new_cost = ray_error(tn, tgt, err_new);
      if (new_cost < cost) {
        u = u_new;
        v = v_new;
        lambda *= 0.1;  // double multiply (reference :262)
        conv = new_cost < cost_thresh;
        cost = new_cost;
#pragma unroll
        for (int j = 0; j < 3; j++) err[j] = err_new[j];
        if (it + 1 < max_iter) s = normal_eq(tn, err);
      } else {
        lambda *= 10.0;  // (reference :266)
        conv = cost < cost_thresh;
      }
    }
  }
  reinterpret_cast<float2*>(p_new)[pix] = make_float2(u, v);
  converged[pix] = conv ? 1 : 0;
}

}  // namespace
}  // namespace m3r

extern "C" int m3r_iter_proj(const float* rays_img, const float* pts3d_norm, const float* p_init,
                             float* p_new, uint8_t* converged, int b, int h, int w, int n,
                             int max_iter, float lambda_init, float cost_thresh, void* stream) {
  if (!rays_img || !pts3d_norm || !p_init || !p_new || !converged) return M3R_ERR_BAD_ARG;
  if (b < 0 || n < 0 || h < 3 || w < 3 || max_iter < 0) return M3R_ERR_BAD_ARG;
  if (b == 0 || n == 0) return M3R_OK;
  if (b > 65535) return M3R_ERR_BAD_ARG;
  dim3 grid((n + m3r::kIterProjThreads - 1) / m3r::kIterProjThreads, b);
  m3r::iter_proj_kernel<<<grid, m3r::kIterProjThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      rays_img, pts3d_norm, p_init, p_new, converged, h, w, n, max_iter, lambda_init, cost_thresh);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}
