// iter_proj — per-pixel Levenberg-Marquardt ray projection (sm_100a).
//
// Replaces mast3r_slam/backend/src/matching_kernels.cu:119-316 (iter_proj_kernel + launcher).
// Spec: SURVEY.md Appendix B.  Results must be identical to the reference's, so every float
// expression below keeps the reference's operand order (nvcc contracts them to the same FMAs) and
// its double promotions where they can change a bit (lambda*0.1, lambda*10.0); the (1.0-du) products
// are exact in float and `1.0 / x` is the IEEE single division (see recip_as_double).
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
#include <cuda_fp16.h>

#include <mutex>

#include "common.cuh"

namespace m3r {
namespace {

// CTA size / minimum CTAs per SM of the LM kernel are build parameters (tools/ab_iterproj.py, same box, 512x384:
// 128 threads at 59 registers = 32 resident warps per SM, 1.3 waves -> 27.6 us with the re-pack; forcing one wave
// with 48 / 40 registers (20-60 bytes of spills) -> 29.7 / 29.7-31.7 us; 64-thread CTAs -> 27.6 us).
#ifndef M3R_IP_THREADS
#define M3R_IP_THREADS 128
#endif
#ifndef M3R_IP_MINCTAS
#define M3R_IP_MINCTAS 1
#endif
constexpr int kIterProjThreads = M3R_IP_THREADS;
constexpr int kCh = 9;

struct Taps {
  const float* __restrict__ p22;  // (v11, u11)       top-left texel
  const float* __restrict__ p21;  // (v11, u11+1)
  const float* __restrict__ p12;  // (v11+1, u11)
  const float* __restrict__ p11;  // (v11+1, u11+1)
  float w11, w12, w21, w22;
};

// The reference writes `float y = 1.0 / x` with a float x: a double-precision divide rounded to float.
// Rounding the correctly rounded 53-bit quotient again to 24 bits equals rounding the exact quotient once
// (double rounding is innocuous for / when the wide format has >= 2*24 + 2 digits), so the IEEE single
// division gives the same bits for every input, NaN / inf / denormals included, at a third of the
// instructions and without touching the fp64 pipe.
__device__ __forceinline__ float recip_as_double(float x) { return __fdiv_rn(1.0f, x); }

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
  const float r_norm_inv = recip_as_double(r_norm);  // (float)(1.0 / (double)r_norm) (reference :187)
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
      const float det_inv = recip_as_double(A00 * A11 - s.A01 * s.A01);  // (reference :213)
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


// ---- padded-texel path ------------------------------------------------------------------------------
// The 36-byte texel of the caller's [b,h,w,9] image is only 4-byte aligned, so every tap costs nine
// scalar LDG and the kernel above is bound by the load/store queue (ncu: lg_throttle is its top stall).
// With a workspace the image is first re-packed into three PLANES over all b*h*w texels
//     A float4 {ray.x, ray.y, ray.z, gu.x}   B float4 {gu.y, gu.z, gv.x, gv.y}   C float {gv.z}
// (one coalesced pass, 7 MB in / 7 MB out at 512x384, L2-resident afterwards) and a tap is then one
// LDG.128 for the error sample and LDG.128 + LDG.32 for the gradients: 4 / 12 load instructions per
// iteration instead of 12 / 36, same values, same arithmetic, same results.  Planes rather than one
// 48-byte texel: neighbouring lanes sample neighbouring pixels, so a warp's 32 x 16 bytes of one tap are
// (nearly) contiguous -- 4-5 128-byte lines and 16-17 sectors per request instead of 12 lines / ~22 sectors
// with the interleaved texel; the L1 spends a wavefront per line touched.
constexpr int kPadF4 = 3;  // float4 of workspace per texel (A, B, and C's float in the third)

struct TexelPlanes {
  float4* A;
  float4* B;
  float* C;
};
__host__ __device__ __forceinline__ TexelPlanes texel_planes(float4* ws, size_t npix) {
  return TexelPlanes{ws, ws + npix, reinterpret_cast<float*>(ws + 2 * npix)};
}

// One warp re-packs 32 consecutive texels: 1152 contiguous bytes in (72 coalesced float4 loads), a
// transpose through shared memory (stride 9 floats: conflict-free), then 512 + 512 + 128 contiguous bytes
// out, one per plane.  `src` must be 16-byte aligned; the last partial group of < 32 texels goes scalar.
constexpr int kRepackWarps = 8;

__global__ void __launch_bounds__(kRepackWarps * 32)
iter_proj_repack_kernel(const float* __restrict__ src, float4* __restrict__ dst, size_t npix) {
  __shared__ __align__(16) float s_in[kRepackWarps][32 * kCh];
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  const size_t g0 = (static_cast<size_t>(blockIdx.x) * kRepackWarps + wp) * 32;  // first texel of this warp
  // the LM kernel is launched with programmatic stream serialisation: its CTAs may be scheduled (and read their
  // targets / initial pixels) as soon as every CTA here is running; they wait for this grid before sampling
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (g0 >= npix) return;
  const TexelPlanes P = texel_planes(dst, npix);
  float v[kCh];
  if (g0 + 32 <= npix) {
    const float4* in4 = reinterpret_cast<const float4*>(src + g0 * kCh);
    float4* sin4 = reinterpret_cast<float4*>(s_in[wp]);
    sin4[lane] = __ldg(in4 + lane);
    sin4[lane + 32] = __ldg(in4 + lane + 32);
    if (lane < 8) sin4[lane + 64] = __ldg(in4 + lane + 64);
    __syncwarp();
#pragma unroll
    for (int j = 0; j < kCh; j++) v[j] = s_in[wp][lane * kCh + j];
  } else if (g0 + lane < npix) {
    const float* sp = src + (g0 + lane) * kCh;
#pragma unroll
    for (int j = 0; j < kCh; j++) v[j] = __ldg(sp + j);
  } else {
    return;
  }
  const size_t i = g0 + lane;
  P.A[i] = make_float4(v[0], v[1], v[2], v[3]);
  P.B[i] = make_float4(v[4], v[5], v[6], v[7]);
  P.C[i] = v[8];
}

struct TapsP {
  const float4* __restrict__ a22;  // plane A at the top-left texel (v11, u11); (v11, u11+1) = +1, next row = +w
  const float* __restrict__ c22;   // plane C at the same texel
  size_t plane;                    // plane B = plane A + plane
  int row;                         // w
  float w11, w12, w21, w22;
};

__device__ __forceinline__ TapsP make_taps_p(const float4* __restrict__ A, const float* __restrict__ C, size_t plane,
                                             int w, float u, float v) {
  TapsP t;
  const int u11 = static_cast<int>(floorf(u));
  const int v11 = static_cast<int>(floorf(v));
  const float du = u - static_cast<float>(u11);
  const float dv = v - static_cast<float>(v11);
  t.w11 = du * dv;
  t.w12 = (1.0f - du) * dv;
  t.w21 = du * (1.0f - dv);
  t.w22 = (1.0f - du) * (1.0f - dv);
  t.row = w;
  t.plane = plane;
  const size_t i = static_cast<size_t>(v11) * w + u11;
  t.a22 = A + i;
  t.c22 = C + i;
  return t;
}

#define M3R_TAP(t, a11, a12, a21, a22) ((t).w11 * (a11) + (t).w12 * (a12) + (t).w21 * (a21) + (t).w22 * (a22))

// normalised interpolated ray minus target -> err, squared error returned; gu.x (4th lane of the same
// 16 bytes) comes along for free and is kept for the normal equations of an accepted step
__device__ __forceinline__ float ray_error_p(const TapsP& t, const float tgt[3], float err[3], float& gux) {
  const float4 a22 = __ldg(t.a22), a21 = __ldg(t.a22 + 1);
  const float4 a12 = __ldg(t.a22 + t.row), a11 = __ldg(t.a22 + t.row + 1);
  float r[3];
  r[0] = M3R_TAP(t, a11.x, a12.x, a21.x, a22.x);
  r[1] = M3R_TAP(t, a11.y, a12.y, a21.y, a22.y);
  r[2] = M3R_TAP(t, a11.z, a12.z, a21.z, a22.z);
  gux = M3R_TAP(t, a11.w, a12.w, a21.w, a22.w);
  const float r_norm = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  const float r_norm_inv = recip_as_double(r_norm);  // (float)(1.0 / (double)r_norm) (reference :187)
#pragma unroll
  for (int j = 0; j < 3; j++) r[j] *= r_norm_inv;
#pragma unroll
  for (int j = 0; j < 3; j++) err[j] = r[j] - tgt[j];
  return err[0] * err[0] + err[1] * err[1] + err[2] * err[2];
}

__device__ __forceinline__ Normal normal_eq_p(const TapsP& t, const float err[3], float gux) {
  const float4* pb = t.a22 + t.plane;
  const float4 b22 = __ldg(pb), b21 = __ldg(pb + 1);
  const float4 b12 = __ldg(pb + t.row), b11 = __ldg(pb + t.row + 1);
  const float c22 = __ldg(t.c22), c21 = __ldg(t.c22 + 1);
  const float c12 = __ldg(t.c22 + t.row), c11 = __ldg(t.c22 + t.row + 1);
  float gx[3], gy[3];
  gx[0] = gux;
  gx[1] = M3R_TAP(t, b11.x, b12.x, b21.x, b22.x);
  gx[2] = M3R_TAP(t, b11.y, b12.y, b21.y, b22.y);
  gy[0] = M3R_TAP(t, b11.z, b12.z, b21.z, b22.z);
  gy[1] = M3R_TAP(t, b11.w, b12.w, b21.w, b22.w);
  gy[2] = M3R_TAP(t, c11, c12, c21, c22);
  Normal s;
  s.A00 = gx[0] * gx[0] + gx[1] * gx[1] + gx[2] * gx[2];
  s.A01 = gx[0] * gy[0] + gx[1] * gy[1] + gx[2] * gy[2];
  s.A11 = gy[0] * gy[0] + gy[1] * gy[1] + gy[2] * gy[2];
  s.b0 = -(err[0] * gx[0] + err[1] * gx[1] + err[2] * gx[2]);
  s.b1 = -(err[0] * gy[0] + err[1] * gy[1] + err[2] * gy[2]);
  return s;
}

// POST = true (fused matching glue, matching.py:68-76): instead of the float pixel the kernel writes the
// truncated int64 pixel p1 = (long)(u, v) and valid = converged && |X11[p1] - X21[k]| < dist_thresh.
struct IterProjPost {
  const float* X11;   // [b,h,w,3]
  const float* X21;   // [b,n,3]
  int64_t* p1;        // [b,n,2]
  float dist_thresh;
};

template <bool POST>
__global__ void __launch_bounds__(kIterProjThreads, M3R_IP_MINCTAS)
iter_proj_padded_kernel(float4* __restrict__ rays_pad, size_t npix_all, const float* __restrict__ pts3d_norm,
                        const float* __restrict__ p_init, float* __restrict__ p_new,
                        uint8_t* __restrict__ converged, int h, int w, int n, int max_iter,
                        float lambda_init, float cost_thresh, IterProjPost post) {
  const int k = blockIdx.x * kIterProjThreads + threadIdx.x;
  const int b = blockIdx.y;
  if (k >= n) return;
  const size_t pix = static_cast<size_t>(b) * n + k;
  const TexelPlanes P = texel_planes(rays_pad, npix_all);
  const float4* imgA = P.A + static_cast<size_t>(b) * h * w;
  const float* imgC = P.C + static_cast<size_t>(b) * h * w;

  const float2 p0 = __ldg(reinterpret_cast<const float2*>(p_init) + pix);
  float u = clampf(p0.x, 1.0f, static_cast<float>(w - 2));
  float v = clampf(p0.y, 1.0f, static_cast<float>(h - 2));
  float tgt[3];
#pragma unroll
  for (int j = 0; j < 3; j++) tgt[j] = __ldg(pts3d_norm + pix * 3 + j);
  asm volatile("griddepcontrol.wait;" ::: "memory");  // planes written by the previous kernel (no-op without PDL)

  float lambda = lambda_init;
  bool conv = false;
  if (max_iter > 0) {
    float err[3], gux;
    TapsP t = make_taps_p(imgA, imgC, npix_all, w, u, v);
    float cost = ray_error_p(t, tgt, err, gux);
    Normal s = normal_eq_p(t, err, gux);
    for (int it = 0; it < max_iter; it++) {
      const float A00 = s.A00 + lambda;
      const float A11 = s.A11 + lambda;
      const float det_inv = recip_as_double(A00 * A11 - s.A01 * s.A01);  // (reference :213)
      const float delta_u = det_inv * (A11 * s.b0 - s.A01 * s.b1);
      const float delta_v = det_inv * (-s.A01 * s.b0 + A00 * s.b1);
      const float u_new = clampf(u + delta_u, 1.0f, static_cast<float>(w - 2));
      const float v_new = clampf(v + delta_v, 1.0f, static_cast<float>(h - 2));
      float err_new[3], gux_new;
      const TapsP tn = make_taps_p(imgA, imgC, npix_all, w, u_new, v_new);
      const float new_cost = ray_error_p(tn, tgt, err_new, gux_new);
      if (new_cost < cost) {
        u = u_new;
        v = v_new;
        lambda *= 0.1;  // double multiply (reference :262)
        conv = new_cost < cost_thresh;
        cost = new_cost;
#pragma unroll
        for (int j = 0; j < 3; j++) err[j] = err_new[j];
        if (it + 1 < max_iter) s = normal_eq_p(tn, err, gux_new);
      } else {
        lambda *= 10.0;  // (reference :266)
        conv = cost < cost_thresh;
      }
    }
  }
  if (POST) {
    const int pu = static_cast<int>(u), pv = static_cast<int>(v);  // .long(): truncation (u, v >= 1)
    const float* a = post.X11 + (static_cast<size_t>(b) * h * w + static_cast<size_t>(pv) * w + pu) * 3;
    const float* q = post.X21 + pix * 3;
    const float d0 = __ldg(a + 0) - __ldg(q + 0), d1 = __ldg(a + 1) - __ldg(q + 1), d2 = __ldg(a + 2) - __ldg(q + 2);
    const float dist = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);
    reinterpret_cast<longlong2*>(post.p1)[pix] = make_longlong2(pu, pv);
    converged[pix] = (conv && dist < post.dist_thresh) ? 1 : 0;
  } else {
    reinterpret_cast<float2*>(p_new)[pix] = make_float2(u, v);
    converged[pix] = conv ? 1 : 0;
  }
}

// ---- fused matching glue: prep (matching.py:25-49 + image.py:5-38) -------------------------------------
// One kernel replaces F.normalize, permute, reflect pad, two grouped 3x3 convolutions, cat, permute +
// contiguous, the second F.normalize and the arange / lin_to_pixel / float chain (~12 torch kernels that
// move 4-5x the bytes of iter_proj itself).  A CTA normalises a (32+2) x (8+2) patch of X11 into shared
// memory (reflect padding at the image border), applies the Scharr-like gradients
// 1/32 [[-3,0,3],[-10,0,10],[-3,0,3]] and its transpose, and writes the three texel planes the LM kernel reads;
// the same thread normalises X21[k] and emits p_init[k] for its own pixel index k.
constexpr int kPrepTW = 32, kPrepTH = 8;

__device__ __forceinline__ int reflect(int i, int n) { return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i); }

__global__ void __launch_bounds__(kPrepTW * kPrepTH)
match_prep_kernel(const float* __restrict__ X11, const float* __restrict__ X21, const int64_t* __restrict__ idx_init,
                  float4* __restrict__ rays_pad, size_t npix_all, float* __restrict__ pts3d_norm,
                  float* __restrict__ p_init, int h, int w) {
  __shared__ float s_r[3][kPrepTH + 2][kPrepTW + 2 + 1];
  const int b = blockIdx.z;
  const int x0 = blockIdx.x * kPrepTW, y0 = blockIdx.y * kPrepTH;
  const int tid = threadIdx.y * kPrepTW + threadIdx.x;
  const float* img = X11 + static_cast<size_t>(b) * h * w * 3;
  for (int t = tid; t < (kPrepTH + 2) * (kPrepTW + 2); t += kPrepTW * kPrepTH) {
    const int ly = t / (kPrepTW + 2), lx = t % (kPrepTW + 2);
    const int gy = reflect(min(y0 + ly - 1, h), h), gx = reflect(min(x0 + lx - 1, w), w);  // clamp keeps reflect in range
    const float* p = img + (static_cast<size_t>(gy) * w + gx) * 3;
    const float a0 = __ldg(p), a1 = __ldg(p + 1), a2 = __ldg(p + 2);
    const float nrm = fmaxf(sqrtf(a0 * a0 + a1 * a1 + a2 * a2), 1e-12f);  // F.normalize: x / max(|x|, eps)
    s_r[0][ly][lx] = a0 / nrm;
    s_r[1][ly][lx] = a1 / nrm;
    s_r[2][ly][lx] = a2 / nrm;
  }
  __syncthreads();
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  if (x >= w || y >= h) return;
  const int lx = threadIdx.x + 1, ly = threadIdx.y + 1;
  float r[3], gx[3], gy[3];
  const float k3 = 3.0f / 32.0f, k10 = 10.0f / 32.0f;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float (*s)[kPrepTW + 2 + 1] = s_r[c];
    r[c] = s[ly][lx];
    gx[c] = k3 * (s[ly - 1][lx + 1] - s[ly - 1][lx - 1]) + k10 * (s[ly][lx + 1] - s[ly][lx - 1]) +
            k3 * (s[ly + 1][lx + 1] - s[ly + 1][lx - 1]);
    gy[c] = k3 * (s[ly + 1][lx - 1] - s[ly - 1][lx - 1]) + k10 * (s[ly + 1][lx] - s[ly - 1][lx]) +
            k3 * (s[ly + 1][lx + 1] - s[ly - 1][lx + 1]);
  }
  const size_t k = static_cast<size_t>(y) * w + x;
  const size_t pix = static_cast<size_t>(b) * h * w + k;
  const TexelPlanes P = texel_planes(rays_pad, npix_all);
  P.A[pix] = make_float4(r[0], r[1], r[2], gx[0]);
  P.B[pix] = make_float4(gx[1], gx[2], gy[0], gy[1]);
  P.C[pix] = gy[2];
  {
    const float* q = X21 + pix * 3;
    const float a0 = __ldg(q), a1 = __ldg(q + 1), a2 = __ldg(q + 2);
    const float nrm = fmaxf(sqrtf(a0 * a0 + a1 * a1 + a2 * a2), 1e-12f);
    pts3d_norm[pix * 3 + 0] = a0 / nrm;
    pts3d_norm[pix * 3 + 1] = a1 / nrm;
    pts3d_norm[pix * 3 + 2] = a2 / nrm;
  }
  {
    const int64_t lin = idx_init ? idx_init[pix] : static_cast<int64_t>(k);
    reinterpret_cast<float2*>(p_init)[pix] = make_float2(static_cast<float>(lin % w), static_cast<float>(lin / w));
  }
}

// p1 [b,n,2] int64 (u, v) -> u + w * v  (matching.py:13-15), elementwise
__global__ void pixel_to_lin_kernel(const int64_t* __restrict__ p1, int64_t* __restrict__ lin, size_t count, int w) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + i);
  lin[i] = p.x + static_cast<int64_t>(w) * p.y;
}

// descriptors to fp16 (matching.py:80-81 `.half()`), both tensors in one launch; counts are multiples of 8
__global__ void f32_to_f16_pair_kernel(const float4* __restrict__ a, uint4* __restrict__ ah, size_t na8,
                                       const float4* __restrict__ c, uint4* __restrict__ ch, size_t nc8) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const float4* src;
  uint4* dst;
  size_t j;
  if (i < na8) {
    src = a; dst = ah; j = i;
  } else if (i < na8 + nc8) {
    src = c; dst = ch; j = i - na8;
  } else {
    return;
  }
  const float4 lo = __ldg(src + 2 * j), hi = __ldg(src + 2 * j + 1);
  const __half2 h0 = __floats2half2_rn(lo.x, lo.y), h1 = __floats2half2_rn(lo.z, lo.w);
  const __half2 h2 = __floats2half2_rn(hi.x, hi.y), h3 = __floats2half2_rn(hi.z, hi.w);
  uint4 o;
  o.x = *reinterpret_cast<const uint32_t*>(&h0);
  o.y = *reinterpret_cast<const uint32_t*>(&h1);
  o.z = *reinterpret_cast<const uint32_t*>(&h2);
  o.w = *reinterpret_cast<const uint32_t*>(&h3);
  dst[j] = o;
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

extern "C" size_t m3r_iter_proj_workspace_bytes(int b, int h, int w) {
  if (b < 1 || h < 1 || w < 1) return 0;
  return static_cast<size_t>(b) * h * w * m3r::kPadF4 * sizeof(float4);
}

extern "C" int m3r_iter_proj_ws(const float* rays_img, const float* pts3d_norm, const float* p_init,
                                float* p_new, uint8_t* converged, int b, int h, int w, int n,
                                int max_iter, float lambda_init, float cost_thresh, void* workspace,
                                size_t workspace_bytes, void* stream) {
  if (!rays_img || !pts3d_norm || !p_init || !p_new || !converged) return M3R_ERR_BAD_ARG;
  if (b < 0 || n < 0 || h < 3 || w < 3 || max_iter < 0) return M3R_ERR_BAD_ARG;
  if (b == 0 || n == 0) return M3R_OK;
  if (b > 65535) return M3R_ERR_BAD_ARG;
  const bool aligned = ((reinterpret_cast<uintptr_t>(workspace) | reinterpret_cast<uintptr_t>(rays_img)) & 15) == 0 &&
                       ((reinterpret_cast<uintptr_t>(p_init) | reinterpret_cast<uintptr_t>(p_new)) & 7) == 0;
  if (!workspace || !aligned || max_iter == 0)  // nothing to sample: the plain kernel is the cheaper one
    return m3r_iter_proj(rays_img, pts3d_norm, p_init, p_new, converged, b, h, w, n, max_iter, lambda_init,
                         cost_thresh, stream);
  if (workspace_bytes < m3r_iter_proj_workspace_bytes(b, h, w)) return M3R_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t npix = static_cast<size_t>(b) * h * w;
  float4* pad = static_cast<float4*>(workspace);
  const size_t per_cta = static_cast<size_t>(m3r::kRepackWarps) * 32;
  m3r::iter_proj_repack_kernel<<<static_cast<unsigned>((npix + per_cta - 1) / per_cta), m3r::kRepackWarps * 32, 0, st>>>(
      rays_img, pad, npix);
  M3R_LAUNCH_CHECK();
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((n + m3r::kIterProjThreads - 1) / m3r::kIterProjThreads, b);
  cfg.blockDim = dim3(m3r::kIterProjThreads);
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  M3R_CUDA_CHECK(cudaLaunchKernelEx(&cfg, m3r::iter_proj_padded_kernel<false>, pad, npix, pts3d_norm, p_init, p_new,
                                    converged, h, w, n, max_iter, lambda_init, cost_thresh, m3r::IterProjPost{}));
  return M3R_OK;
}

// ---- iter_proj with library-owned static scratch (no allocation anywhere on the call path) ---------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops).  The re-packed ray image lives in a file-scope device array
// of M3R_ITER_PROJ_STATIC_MAX_TEXELS texels x 48 B = 9 437 184 B, held by every GPU that loads the library until exit.
// Every call shares the array: calls that overlap on the GPU (two streams, two host threads) RACE on it.  Nothing in
// it has to start at zero -- the re-pack kernel rewrites texels [0, b*h*w) before the projection kernel reads exactly
// that range -- so the last call's contents are simply overwritten.
namespace m3r {
namespace {
static __device__ __align__(256) float4 g_iter_proj_scratch[(size_t)M3R_ITER_PROJ_STATIC_MAX_TEXELS * kPadF4];
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
float4* g_scratch_ptr[kMaxDevices] = {};  // device address of the array, per device (the host-side symbol is NOT one)

// set-up, once per device, with that device current
int iter_proj_static_scratch(float4** out) {
  int dev = -1;
  M3R_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_last_error_text("m3r_iter_proj_static: device ordinal out of range");
    return M3R_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (!g_scratch_ptr[dev]) {
    float4* p = nullptr;
    M3R_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_iter_proj_scratch));
    if (!p) {
      set_last_error_text("m3r_iter_proj_static: cudaGetSymbolAddress returned a null address");
      return M3R_ERR_CUDA;
    }
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return M3R_OK;
}
}  // namespace
}  // namespace m3r

extern "C" size_t m3r_iter_proj_static_max_texels(void) { return M3R_ITER_PROJ_STATIC_MAX_TEXELS; }

extern "C" int m3r_iter_proj_static(const float* rays_img, const float* pts3d_norm, const float* p_init,
                                    float* p_new, uint8_t* converged, int b, int h, int w, int n, int max_iter,
                                    float lambda_init, float cost_thresh, void* stream) {
  if (!rays_img || !pts3d_norm || !p_init || !p_new || !converged) return M3R_ERR_BAD_ARG;
  if (b < 0 || n < 0 || h < 3 || w < 3 || max_iter < 0) return M3R_ERR_BAD_ARG;
  if (b == 0 || n == 0) return M3R_OK;
  if (b > 65535) return M3R_ERR_BAD_ARG;
  const size_t npix = static_cast<size_t>(b) * h * w;
  if (npix > static_cast<size_t>(M3R_ITER_PROJ_STATIC_MAX_TEXELS)) {
    char msg[160];
    snprintf(msg, sizeof(msg), "m3r_iter_proj_static: b*h*w = %zu texels exceeds the static scratch of %d texels (use m3r_iter_proj_ws)",
             npix, M3R_ITER_PROJ_STATIC_MAX_TEXELS);
    m3r::set_last_error_text(msg);
    return M3R_ERR_WORKSPACE;
  }
  if ((reinterpret_cast<uintptr_t>(rays_img) & 15) != 0 ||
      ((reinterpret_cast<uintptr_t>(p_init) | reinterpret_cast<uintptr_t>(p_new)) & 7) != 0) {
    m3r::set_last_error_text("m3r_iter_proj_static: rays_img must be 16-byte aligned, p_init / p_new 8-byte aligned");
    return M3R_ERR_BAD_ARG;
  }
  if (max_iter == 0)  // nothing to sample: the plain kernel (no scratch at all) is the cheaper one, as in m3r_iter_proj_ws
    return m3r_iter_proj(rays_img, pts3d_norm, p_init, p_new, converged, b, h, w, n, max_iter, lambda_init, cost_thresh, stream);
  float4* pad = nullptr;
  if (int rc = m3r::iter_proj_static_scratch(&pad)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t per_cta = static_cast<size_t>(m3r::kRepackWarps) * 32;
  m3r::iter_proj_repack_kernel<<<static_cast<unsigned>((npix + per_cta - 1) / per_cta), m3r::kRepackWarps * 32, 0, st>>>(
      rays_img, pad, npix);
  M3R_LAUNCH_CHECK();
  // plain stream order between the two launches (no programmatic overlap: nothing else may touch the shared array)
  m3r::iter_proj_padded_kernel<false><<<dim3((n + m3r::kIterProjThreads - 1) / m3r::kIterProjThreads, b),
                                        m3r::kIterProjThreads, 0, st>>>(pad, npix, pts3d_norm, p_init, p_new, converged, h, w, n,
                                                                        max_iter, lambda_init, cost_thresh, m3r::IterProjPost{});
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

// ---- fused matching: matching.match() in five launches ------------------------------------------------
namespace {
struct MatchWs {
  size_t off_rays, off_pts, off_p0, off_p1, off_p1r, off_d11h, off_d21h, total;
};
MatchWs match_ws_layout(int b, int h, int w, int fdim, int desc_is_f32) {
  MatchWs L{};
  const size_t n = static_cast<size_t>(b) * h * w;
  size_t o = 0;
  auto take = [&](size_t bytes) {
    const size_t r = o;
    o = m3r::align_up(o + bytes, 256);
    return r;
  };
  L.off_rays = take(n * m3r::kPadF4 * sizeof(float4));
  L.off_pts = take(n * 3 * sizeof(float));
  L.off_p0 = take(n * 2 * sizeof(float));
  L.off_p1 = take(n * 2 * sizeof(int64_t));
  L.off_p1r = take(n * 2 * sizeof(int64_t));
  L.off_d11h = take(desc_is_f32 ? n * fdim * 2 : 0);
  L.off_d21h = take(desc_is_f32 ? n * fdim * 2 : 0);
  L.total = o;
  return L;
}
}  // namespace

extern "C" size_t m3r_match_workspace_bytes(int b, int h, int w, int fdim, int desc_is_f32) {
  if (b < 1 || h < 1 || w < 1 || fdim < 1) return 0;
  return match_ws_layout(b, h, w, fdim, desc_is_f32).total;
}

extern "C" int m3r_match_fused(const float* X11, const float* X21, const void* D11, const void* D21, int fdim,
                               int desc_is_f32, const int64_t* idx_init, int b, int h, int w, int max_iter,
                               float lambda_init, float cost_thresh, float dist_thresh, int radius,
                               int dilation_max, int64_t* idx_out, uint8_t* valid_out, void* workspace,
                               size_t workspace_bytes, void* stream) {
  if (!X11 || !X21 || !idx_out || !valid_out || !workspace) return M3R_ERR_BAD_ARG;
  if (radius > 0 && (!D11 || !D21)) return M3R_ERR_BAD_ARG;
  if (b < 0 || h < 3 || w < 3 || fdim < 1 || max_iter < 0 || radius < 0 || dilation_max < 0) return M3R_ERR_BAD_ARG;
  if (b == 0) return M3R_OK;
  if (b > 65535) return M3R_ERR_BAD_ARG;
  const size_t n_all = static_cast<size_t>(b) * h * w;
  if (static_cast<size_t>(h) * w > 0x7fffffffull) return M3R_ERR_UNSUPPORTED;
  if (radius > 0 && desc_is_f32 && (n_all * fdim) % 8 != 0) return M3R_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(workspace) & 255) != 0) return M3R_ERR_WORKSPACE;
  const MatchWs L = match_ws_layout(b, h, w, fdim, desc_is_f32);
  if (workspace_bytes < L.total) return M3R_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  char* ws = static_cast<char*>(workspace);
  float4* rays = reinterpret_cast<float4*>(ws + L.off_rays);
  float* pts = reinterpret_cast<float*>(ws + L.off_pts);
  float* p0 = reinterpret_cast<float*>(ws + L.off_p0);
  int64_t* p1 = reinterpret_cast<int64_t*>(ws + L.off_p1);
  int64_t* p1r = reinterpret_cast<int64_t*>(ws + L.off_p1r);
  const int n = h * w;
  {
    dim3 grid((w + m3r::kPrepTW - 1) / m3r::kPrepTW, (h + m3r::kPrepTH - 1) / m3r::kPrepTH, b);
    m3r::match_prep_kernel<<<grid, dim3(m3r::kPrepTW, m3r::kPrepTH), 0, st>>>(X11, X21, idx_init, rays, n_all, pts, p0, h, w);
    M3R_LAUNCH_CHECK();
  }
  {
    m3r::IterProjPost post{X11, X21, p1, dist_thresh};
    dim3 grid((n + m3r::kIterProjThreads - 1) / m3r::kIterProjThreads, b);
    m3r::iter_proj_padded_kernel<true><<<grid, m3r::kIterProjThreads, 0, st>>>(
        rays, n_all, pts, p0, nullptr, valid_out, h, w, n, max_iter, lambda_init, cost_thresh, post);
    M3R_LAUNCH_CHECK();
  }
  const int64_t* p_final = p1;
  if (radius > 0) {
    const void* d11h = D11;
    const void* d21h = D21;
    if (desc_is_f32) {
      const size_t n8 = n_all * fdim / 8;
      uint4* a = reinterpret_cast<uint4*>(ws + L.off_d11h);
      uint4* c = reinterpret_cast<uint4*>(ws + L.off_d21h);
      m3r::f32_to_f16_pair_kernel<<<static_cast<unsigned>((2 * n8 + 255) / 256), 256, 0, st>>>(
          static_cast<const float4*>(D11), a, n8, static_cast<const float4*>(D21), c, n8);
      M3R_LAUNCH_CHECK();
      d11h = a;
      d21h = c;
    }
    const int rc = m3r_refine_matches_f16(d11h, d21h, p1, p1r, b, h, w, n, fdim, radius, dilation_max, stream);
    if (rc != M3R_OK) return rc;
    p_final = p1r;
  }
  m3r::pixel_to_lin_kernel<<<static_cast<unsigned>((n_all + 255) / 256), 256, 0, st>>>(p_final, idx_out, n_all, w);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}
