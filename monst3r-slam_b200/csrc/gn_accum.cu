// Gauss-Newton accumulate: per-edge Sim(3) residual / Jacobian / Hessian reduction (sm_100a).
//
// Replaces the three accumulate kernels of mast3r_slam/backend/src/gn_kernels.cu
// (point_align_kernel :455-723, ray_align_kernel :813-1138, calib_proj_kernel :1231-1543).
// Math spec: SURVEY.md Appendix A.  This is a re-derivation, not a translation:
//
//  * J_i = -J_j and J_j = J_loc * Ad(T_i^-1) with Ad constant per edge, so per point only the
//    LOCAL normal equations are accumulated: the 28 upper-triangle terms of sum w J_loc^T J_loc and
//    the 7 terms of sum w r J_loc (35 fp32 registers per thread, with the structural zeros of every
//    Jacobian row removed at compile time).  The reference accumulates 105 + 14 terms per point in
//    local memory and applies the adjoint per point (~1.5 kflop/pt); here it is ~0.2 kflop/pt and
//    the adjoint is applied once per edge in fp64: A_e = G^T M G, b_e = G^T v.
//  * the grid is (point-chunks x edges), not one CTA per edge, so a 2-edge tracking problem still
//    fills the 148 SMs; each thread streams 4 consecutive points per step with 16-byte loads
//    (idx as 2x longlong2, Xs[j] as 3x float4, Q / Cs[j] as float4, valid as uchar4).
//  * reduction: warp shuffles -> shared memory -> one fp64 partial per CTA, then the LAST CTA of
//    an edge (ticket counter) sums the partials in a fixed order, applies the adjoint and writes
//    the edge block EB[e] = {A_e (7x7), b_e (7)} in fp64.  Per-pixel Jacobians never touch HBM and
//    the result is bit-reproducible from run to run (no floating-point atomics).
//
// H_ii = H_jj = A_e, H_ij = H_ji = -A_e, g_i = -b_e, g_j = +b_e (gn_kernels.cu:1096-1137).
#include "gn_common.cuh"

namespace m3r {
namespace {

__host__ __device__ constexpr int tri(int a, int b) { return a * 7 - a * (a - 1) / 2 + (b - a); }  // a <= b

// acc[0..27] += w * J J^T (upper triangle), acc[28..34] += w * r * J ; MASK = non-zero columns of J
template <unsigned MASK>
__device__ __forceinline__ void rank1(float (&acc)[kAcc], const float (&J)[7], float w, float r) {
  float wJ[7];
#pragma unroll
  for (int a = 0; a < 7; a++)
    if ((MASK >> a) & 1u) wJ[a] = w * J[a];
#pragma unroll
  for (int a = 0; a < 7; a++) {
    if (!((MASK >> a) & 1u)) continue;
#pragma unroll
    for (int b = a; b < 7; b++) {
      if (!((MASK >> b) & 1u)) continue;
      acc[tri(a, b)] = fmaf(wJ[a], J[b], acc[tri(a, b)]);
    }
  }
  const float wr = w * r;
#pragma unroll
  for (int a = 0; a < 7; a++)
    if ((MASK >> a) & 1u) acc[28 + a] = fmaf(wr, J[a], acc[28 + a]);
}

__device__ __forceinline__ float huber(float x) {
  const float a = fabsf(x);
  return a < 1.345f ? 1.0f : __fdividef(1.345f, a);
}

struct EdgeXform {  // P = sR * X + t   (T_ij = T_i^-1 T_j), fp32 copy of an fp64 computation
  float m[9];
  float t[3];
};

struct AccumParams {
  float inv_sigma_a, inv_sigma_b, C_thresh, Q_thresh;
  float fx, fy, cx, cy;
  float u_lo, u_hi, v_lo, v_hi, z_eps;
  int width;
};

template <int KIND>
__device__ __forceinline__ void accum_point(float (&acc)[kAcc], const EdgeXform& X, const AccumParams& prm,
                                            float xjx, float xjy, float xjz, float cj, float q, bool vm,
                                            long long m, const float* __restrict__ Xi_base,
                                            const float* __restrict__ Ci_base) {
  // gather the matched point of frame i (index 0 for an invalid match, like the reference :914)
  const long long mi = vm ? m : 0;
  const float xix = __ldg(Xi_base + mi * 3 + 0);
  const float xiy = __ldg(Xi_base + mi * 3 + 1);
  const float xiz = __ldg(Xi_base + mi * 3 + 2);
  const float ci = __ldg(Ci_base + mi);
  bool ok = vm & (q > prm.Q_thresh) & (ci > prm.C_thresh) & (cj > prm.C_thresh);

  const float px = fmaf(X.m[0], xjx, fmaf(X.m[1], xjy, fmaf(X.m[2], xjz, X.t[0])));
  const float py = fmaf(X.m[3], xjx, fmaf(X.m[4], xjy, fmaf(X.m[5], xjz, X.t[1])));
  const float pz = fmaf(X.m[6], xjx, fmaf(X.m[7], xjy, fmaf(X.m[8], xjz, X.t[2])));
  const float sq = sqrtf(q);

  if (KIND == M3R_GN_POINTS) {
    if (!ok) return;
    const float sw = prm.inv_sigma_a * sq;
    const float wc = sw * sw;
    const float e0 = px - xix, e1 = py - xiy, e2 = pz - xiz;
    {
      const float J[7] = {1.f, 0.f, 0.f, 0.f, pz, -py, px};
      rank1<0x71u>(acc, J, huber(sw * e0) * wc, e0);
    }
    {
      const float J[7] = {0.f, 1.f, 0.f, -pz, 0.f, px, py};
      rank1<0x6Au>(acc, J, huber(sw * e1) * wc, e1);
    }
    {
      const float J[7] = {0.f, 0.f, 1.f, py, -px, 0.f, pz};
      rank1<0x5Cu>(acc, J, huber(sw * e2) * wc, e2);
    }
  } else if (KIND == M3R_GN_RAYS) {
    if (!ok) return;
    const float n2j = fmaf(px, px, fmaf(py, py, pz * pz));
    const float inv_j = rsqrtf(n2j);
    const float rho = n2j * inv_j;
    const float rx = px * inv_j, ry = py * inv_j, rz = pz * inv_j;
    const float n2i = fmaf(xix, xix, fmaf(xiy, xiy, xiz * xiz));
    const float inv_i = rsqrtf(n2i);
    const float e0 = rx - xix * inv_i, e1 = ry - xiy * inv_i, e2 = rz - xiz * inv_i;
    const float e3 = rho - n2i * inv_i;
    const float swr = prm.inv_sigma_a * sq, swd = prm.inv_sigma_b * sq;
    const float wr = swr * swr, wd = swd * swd;
    // d(P/|P|)/dP = (I - r r^T)/|P|
    const float xx = fmaf(-rx * rx, inv_j, inv_j), yy = fmaf(-ry * ry, inv_j, inv_j),
                zz = fmaf(-rz * rz, inv_j, inv_j);
    const float xy = -rx * ry * inv_j, xz = -rx * rz * inv_j, yz = -ry * rz * inv_j;
    {
      const float J[7] = {xx, xy, xz, 0.f, rz, -ry, 0.f};
      rank1<0x37u>(acc, J, huber(swr * e0) * wr, e0);
    }
    {
      const float J[7] = {xy, yy, yz, -rz, 0.f, rx, 0.f};
      rank1<0x2Fu>(acc, J, huber(swr * e1) * wr, e1);
    }
    {
      const float J[7] = {xz, yz, zz, ry, -rx, 0.f, 0.f};
      rank1<0x1Fu>(acc, J, huber(swr * e2) * wr, e2);
    }
    {
      const float J[7] = {rx, ry, rz, 0.f, 0.f, 0.f, rho};
      rank1<0x47u>(acc, J, huber(swd * e3) * wd, e3);
    }
  } else {  // calib
    const bool vz = (pz > prm.z_eps) && (xiz > prm.z_eps);
    const float zinv = vz ? __frcp_rn(pz) : 0.f;
    const float xz_ = px * zinv, yz_ = py * zinv;
    const float u = fmaf(prm.fx, xz_, prm.cx), v = fmaf(prm.fy, yz_, prm.cy);
    ok = ok & vz & (u > prm.u_lo) & (u < prm.u_hi) & (v > prm.v_lo) & (v < prm.v_hi);
    if (!ok) return;
    const int mi32 = static_cast<int>(mi);
    const int ut = mi32 % prm.width, vt = mi32 / prm.width;
    const float e0 = u - static_cast<float>(ut), e1 = v - static_cast<float>(vt);
    const float e2 = logf(pz) - logf(xiz);
    const float swp = prm.inv_sigma_a * sq, swd = prm.inv_sigma_b * sq;
    const float wp = swp * swp, wd = swd * swd;
    {
      const float J[7] = {prm.fx * zinv, 0.f, -prm.fx * xz_ * zinv, -prm.fx * xz_ * yz_,
                          prm.fx * fmaf(xz_, xz_, 1.f), -prm.fx * yz_, 0.f};
      rank1<0x3Du>(acc, J, huber(swp * e0) * wp, e0);
    }
    {
      const float J[7] = {0.f, prm.fy * zinv, -prm.fy * yz_ * zinv, -prm.fy * fmaf(yz_, yz_, 1.f),
                          prm.fy * xz_ * yz_, prm.fy * xz_, 0.f};
      rank1<0x3Eu>(acc, J, huber(swp * e1) * wp, e1);
    }
    {
      const float J[7] = {0.f, 0.f, zinv, yz_, -xz_, 0.f, 1.f};
      rank1<0x5Cu>(acc, J, huber(swd * e2) * wd, e2);
    }
  }
}

struct AccumArgs {
  const float* Twc;
  const float* Xs;
  const float* Cs;
  const float* K;
  const int64_t* idx;   // local slabs [E_local, n]
  const uint8_t* valid;
  const float* Q;
  const int* ci;
  const int* cj;
  const int* ctrl;
  unsigned* counters;
  double* partials;
  double* eb;
  int n, e_lo, chunks, groups_per_chunk;  // group = VEC points
  AccumParams prm;
};

// relative pose T_ij = T_i^-1 T_j in fp64 -> fp32 3x4
__device__ __forceinline__ void edge_xform(const float* __restrict__ Ti, const float* __restrict__ Tj, EdgeXform& X) {
  const double tix = Ti[0], tiy = Ti[1], tiz = Ti[2];
  const double qix = -(double)Ti[3], qiy = -(double)Ti[4], qiz = -(double)Ti[5], qiw = Ti[6];  // q_i^-1
  const double si_inv = 1.0 / (double)Ti[7];
  const double qjx = Tj[3], qjy = Tj[4], qjz = Tj[5], qjw = Tj[6];
  // q = q_i^-1 * q_j
  const double x = qiw * qjx + qix * qjw + qiy * qjz - qiz * qjy;
  const double y = qiw * qjy - qix * qjz + qiy * qjw + qiz * qjx;
  const double z = qiw * qjz + qix * qjy - qiy * qjx + qiz * qjw;
  const double w = qiw * qjw - qix * qjx - qiy * qjy - qiz * qjz;
  const double s = si_inv * (double)Tj[7];
  // R(q) as the linear map of actSO3 (no normalisation, like the reference): I + 2w[q]x + 2[q]x^2
  double R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - z * w),     2 * (x * z + y * w),
                 2 * (x * y + z * w),     1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                 2 * (x * z - y * w),     2 * (y * z + x * w),     1 - 2 * (x * x + y * y)};
#pragma unroll
  for (int k = 0; k < 9; k++) X.m[k] = (float)(s * R[k]);
  // t = R(q_i^-1) (t_j - t_i) / s_i
  const double dx = (double)Tj[0] - tix, dy = (double)Tj[1] - tiy, dz = (double)Tj[2] - tiz;
  const double Ri[9] = {1 - 2 * (qiy * qiy + qiz * qiz), 2 * (qix * qiy - qiz * qiw),     2 * (qix * qiz + qiy * qiw),
                        2 * (qix * qiy + qiz * qiw),     1 - 2 * (qix * qix + qiz * qiz), 2 * (qiy * qiz - qix * qiw),
                        2 * (qix * qiz - qiy * qiw),     2 * (qiy * qiz + qix * qiw),     1 - 2 * (qix * qix + qiy * qiy)};
  X.t[0] = (float)((Ri[0] * dx + Ri[1] * dy + Ri[2] * dz) * si_inv);
  X.t[1] = (float)((Ri[3] * dx + Ri[4] * dy + Ri[5] * dz) * si_inv);
  X.t[2] = (float)((Ri[6] * dx + Ri[7] * dy + Ri[8] * dz) * si_inv);
}

template <int KIND, int VEC>
__global__ void __launch_bounds__(kAccumThreads)
gn_accum_kernel(const AccumArgs a) {
  if (a.ctrl[CTRL_DONE]) return;
  const int el = blockIdx.y;          // local edge
  const int e = a.e_lo + el;          // global edge
  const int chunk = blockIdx.x;
  const int i = a.ci[e], j = a.cj[e];
  const int n = a.n;

  EdgeXform X;
  edge_xform(a.Twc + (size_t)i * 8, a.Twc + (size_t)j * 8, X);
  AccumParams prm = a.prm;
  if (KIND == M3R_GN_CALIB) {  // K[0][0], K[1][1], K[0][2], K[1][2] (gn_kernels.cu:1274-1279)
    prm.fx = __ldg(a.K + 0);
    prm.fy = __ldg(a.K + 4);
    prm.cx = __ldg(a.K + 2);
    prm.cy = __ldg(a.K + 5);
  }

  const float* __restrict__ Xi = a.Xs + (size_t)i * n * 3;
  const float* __restrict__ Ci = a.Cs + (size_t)i * n;
  const float* __restrict__ Xj = a.Xs + (size_t)j * n * 3;
  const float* __restrict__ Cj = a.Cs + (size_t)j * n;
  const int64_t* __restrict__ idx = a.idx + (size_t)el * n;
  const uint8_t* __restrict__ valid = a.valid + (size_t)el * n;
  const float* __restrict__ Q = a.Q + (size_t)el * n;

  float acc[kAcc];
#pragma unroll
  for (int t = 0; t < kAcc; t++) acc[t] = 0.f;

  const int total_groups = (n + VEC - 1) / VEC;
  const int g_lo = chunk * a.groups_per_chunk;
  int g_hi = g_lo + a.groups_per_chunk;
  if (g_hi > total_groups) g_hi = total_groups;

  if (i != j) {
    for (int g = g_lo + threadIdx.x; g < g_hi; g += kAccumThreads) {
      if (VEC == 4) {
        const size_t k0 = (size_t)g * 4;
        const uchar4 vm = __ldg(reinterpret_cast<const uchar4*>(valid + k0));
        const longlong2 m01 = __ldg(reinterpret_cast<const longlong2*>(idx + k0));
        const longlong2 m23 = __ldg(reinterpret_cast<const longlong2*>(idx + k0) + 1);
        const float4 q4 = __ldg(reinterpret_cast<const float4*>(Q + k0));
        const float4 c4 = __ldg(reinterpret_cast<const float4*>(Cj + k0));
        const float4 xa = __ldg(reinterpret_cast<const float4*>(Xj + k0 * 3));
        const float4 xb = __ldg(reinterpret_cast<const float4*>(Xj + k0 * 3) + 1);
        const float4 xc = __ldg(reinterpret_cast<const float4*>(Xj + k0 * 3) + 2);
        accum_point<KIND>(acc, X, prm, xa.x, xa.y, xa.z, c4.x, q4.x, vm.x != 0, m01.x, Xi, Ci);
        accum_point<KIND>(acc, X, prm, xa.w, xb.x, xb.y, c4.y, q4.y, vm.y != 0, m01.y, Xi, Ci);
        accum_point<KIND>(acc, X, prm, xb.z, xb.w, xc.x, c4.z, q4.z, vm.z != 0, m23.x, Xi, Ci);
        accum_point<KIND>(acc, X, prm, xc.y, xc.z, xc.w, c4.w, q4.w, vm.w != 0, m23.y, Xi, Ci);
      } else {
        const size_t k0 = (size_t)g;
        accum_point<KIND>(acc, X, prm, __ldg(Xj + k0 * 3), __ldg(Xj + k0 * 3 + 1), __ldg(Xj + k0 * 3 + 2),
                          __ldg(Cj + k0), __ldg(Q + k0), valid[k0] != 0, idx[k0], Xi, Ci);
      }
    }
  }

  // ---- CTA reduction: shuffles -> smem -> fp64 partial ----
  __shared__ float s_warp[kAccumThreads / 32][kAcc];
  __shared__ double s_mv[kAcc];
  __shared__ double s_G[49];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int t = 0; t < kAcc; t++) {
    float v = acc[t];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0) s_warp[warp][t] = v;
  }
  __syncthreads();
  double* part = a.partials + ((size_t)el * a.chunks + chunk) * kAcc;
  if (threadIdx.x < kAcc) {
    double s = 0.0;
#pragma unroll
    for (int wv = 0; wv < kAccumThreads / 32; wv++) s += (double)s_warp[wv][threadIdx.x];
    part[threadIdx.x] = s;
  }
  if (a.chunks > 1) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned ticket = atomicAdd(a.counters + e, 1u);
      s_last = (ticket == (unsigned)a.chunks - 1u);
      if (s_last) a.counters[e] = 0;  // re-arm for the next iteration
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
  }

  // ---- last CTA of the edge: fixed-order fp64 sum, adjoint, edge block ----
  if (threadIdx.x < kAcc) {
    const double* p0 = a.partials + (size_t)el * a.chunks * kAcc + threadIdx.x;
    double s = 0.0;
    for (int c = 0; c < a.chunks; c++) s += __ldcg(p0 + (size_t)c * kAcc);
    s_mv[threadIdx.x] = s;
  }
  if (threadIdx.x >= 64 && threadIdx.x < 64 + 7) {
    // G: Y = X * G is the row-vector adjoint of gn_kernels.cu:277-297 for pose i; row m = image of e_m
    const int m = threadIdx.x - 64;
    const float* Ti = a.Twc + (size_t)i * 8;
    const double tx = Ti[0], ty = Ti[1], tz = Ti[2], x = Ti[3], y = Ti[4], z = Ti[5], w = Ti[6];
    const double s_inv = 1.0 / (double)Ti[7];
    const double R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - z * w),     2 * (x * z + y * w),
                         2 * (x * y + z * w),     1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                         2 * (x * z - y * w),     2 * (y * z + x * w),     1 - 2 * (x * x + y * y)};
    double row[7] = {0, 0, 0, 0, 0, 0, 0};
    if (m < 3) {
      const double c0 = R[0 + m], c1 = R[3 + m], c2 = R[6 + m];  // R e_m
      row[0] = s_inv * c0; row[1] = s_inv * c1; row[2] = s_inv * c2;
      row[3] = s_inv * (ty * c2 - tz * c1);
      row[4] = s_inv * (tz * c0 - tx * c2);
      row[5] = s_inv * (tx * c1 - ty * c0);
      row[6] = s_inv * (tx * c0 + ty * c1 + tz * c2);
    } else if (m < 6) {
      row[3] = R[0 + (m - 3)]; row[4] = R[3 + (m - 3)]; row[5] = R[6 + (m - 3)];
    } else {
      row[6] = 1.0;
    }
#pragma unroll
    for (int c = 0; c < 7; c++) s_G[m * 7 + c] = row[c];
  }
  __syncthreads();
  double* out = a.eb + (size_t)e * kEdgeBlock;
  if (threadIdx.x < 49) {
    const int r = threadIdx.x / 7, c = threadIdx.x % 7;
    double s = 0.0;
    for (int p = 0; p < 7; p++) {
      double t = 0.0;
      for (int q = 0; q < 7; q++) {
        const double mpq = (p <= q) ? s_mv[tri(p, q)] : s_mv[tri(q, p)];
        t += mpq * s_G[q * 7 + c];
      }
      s += s_G[p * 7 + r] * t;
    }
    out[threadIdx.x] = s;
  } else if (threadIdx.x < 56) {
    const int r = threadIdx.x - 49;
    double s = 0.0;
    for (int p = 0; p < 7; p++) s += s_G[p * 7 + r] * s_mv[28 + p];
    out[threadIdx.x] = s;
  }
}

template <int KIND>
int launch_kind(const AccumArgs& a, int e_local, bool vec4, cudaStream_t s) {
  dim3 grid(a.chunks, e_local);
  if (vec4)
    gn_accum_kernel<KIND, 4><<<grid, kAccumThreads, 0, s>>>(a);
  else
    gn_accum_kernel<KIND, 1><<<grid, kAccumThreads, 0, s>>>(a);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

// Hs [4,E,7,7] f32 / gs [2,E,7] f32 in the reference layout, from the fp64 edge blocks
__global__ void export_hs_gs_kernel(const double* __restrict__ eb, int E, float* __restrict__ Hs,
                                    float* __restrict__ gs) {
  const int e = blockIdx.x;
  const int t = threadIdx.x;
  const double* b = eb + (size_t)e * kEdgeBlock;
  if (t < 49) {
    const float v = (float)b[t];
    Hs[((size_t)0 * E + e) * 49 + t] = v;
    Hs[((size_t)1 * E + e) * 49 + t] = -v;
    Hs[((size_t)2 * E + e) * 49 + t] = -v;
    Hs[((size_t)3 * E + e) * 49 + t] = v;
  } else if (t < 56) {
    const float v = (float)b[t];
    gs[((size_t)0 * E + e) * 7 + (t - 49)] = -v;
    gs[((size_t)1 * E + e) * 7 + (t - 49)] = v;
  }
}

}  // namespace

int launch_gn_accumulate(const float* Twc, const float* Xs, const float* Cs, const float* K,
                         const int64_t* idx, const uint8_t* valid, const float* Q, const GnLayout& L,
                         const GnPtrs& P, const m3r_gn_params& prm, int e_lo, int e_hi, cudaStream_t s) {
  const int e_local = e_hi - e_lo;
  if (e_local <= 0) return M3R_OK;
  if (e_local > 65535) return M3R_ERR_UNSUPPORTED;
  AccumArgs a{};
  a.Twc = Twc; a.Xs = Xs; a.Cs = Cs; a.K = K; a.idx = idx; a.valid = valid; a.Q = Q;
  a.ci = P.ci; a.cj = P.cj; a.ctrl = P.ctrl; a.counters = P.counters; a.partials = P.partials; a.eb = P.eb;
  a.n = L.num_pts;
  a.e_lo = e_lo;
  a.chunks = gn_chunks(e_local, L.num_pts);
  const bool vec4 = (L.num_pts % 4 == 0) &&
                    ((reinterpret_cast<uintptr_t>(Xs) | reinterpret_cast<uintptr_t>(Cs) |
                      reinterpret_cast<uintptr_t>(idx) | reinterpret_cast<uintptr_t>(Q) |
                      reinterpret_cast<uintptr_t>(valid)) % 16 == 0);
  const int vec = vec4 ? 4 : 1;
  const int total_groups = (L.num_pts + vec - 1) / vec;
  a.groups_per_chunk = (total_groups + a.chunks - 1) / a.chunks;
  a.prm.inv_sigma_a = (float)(1.0 / (double)prm.sigma_a);
  a.prm.inv_sigma_b = prm.kind == M3R_GN_POINTS ? 0.f : (float)(1.0 / (double)prm.sigma_b);
  a.prm.C_thresh = prm.C_thresh;
  a.prm.Q_thresh = prm.Q_thresh;
  a.prm.z_eps = prm.z_eps;
  a.prm.width = prm.width > 0 ? prm.width : 1;
  a.prm.u_lo = (float)prm.pixel_border;
  a.prm.u_hi = (float)(prm.width - 1 - prm.pixel_border);
  a.prm.v_lo = (float)prm.pixel_border;
  a.prm.v_hi = (float)(prm.height - 1 - prm.pixel_border);
  if (prm.kind == M3R_GN_CALIB) {
    if (!K) return M3R_ERR_BAD_ARG;  // read on the device by the kernel: the call stays asynchronous
  }
  switch (prm.kind) {
    case M3R_GN_POINTS: return launch_kind<M3R_GN_POINTS>(a, e_local, vec4, s);
    case M3R_GN_RAYS: return launch_kind<M3R_GN_RAYS>(a, e_local, vec4, s);
    case M3R_GN_CALIB: return launch_kind<M3R_GN_CALIB>(a, e_local, vec4, s);
    default: return M3R_ERR_BAD_ARG;
  }
}

int launch_export_hs_gs(const GnLayout& L, const GnPtrs& P, float* Hs, float* gs, cudaStream_t s) {
  if (L.num_edges <= 0) return M3R_OK;
  export_hs_gs_kernel<<<L.num_edges, 64, 0, s>>>(P.eb, L.num_edges, Hs, gs);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r
