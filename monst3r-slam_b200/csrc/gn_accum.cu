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
//  * everything that does not depend on the poses is folded ONCE per solve by gn_pack_kernel into an
//    8-byte record per (edge, point): {int32 m, float sqrt(Q)} with m = -1 when
//    !(valid_match && Q > Q_thresh && C_i[m] > C_thresh && C_j[k] > C_thresh).  The max_iter
//    iterations then stream 8 B/pt/edge instead of 13 B + two confidence reads, and never touch Cs.
//  * the grid is (point-chunks x edges).  A CTA stages its chunk of the two contiguous streams
//    (packed records, Xs[j]) through shared memory with TMA bulk copies (cp.async.bulk + mbarrier),
//    all tiles issued up-front, so the only long-latency accesses left in the loop are the Xs[i][m]
//    gathers, issued for all of a thread's points of the tile before any arithmetic.  Consecutive
//    CTAs work on the same edge and the two directions of a keyframe pair are adjacent in the edge
//    list, so Xs[i] / Xs[j] are served by the L2 after their first touch.
//  * reduction: warp shuffles -> shared memory -> one fp64 partial per CTA, then the LAST CTA of
//    an edge (ticket counter) sums the partials in a fixed order, applies the adjoint and writes
//    the edge block EB[e] = {A_e (7x7), b_e (7)} in fp64.  Per-pixel Jacobians never touch HBM and
//    the result is bit-reproducible from run to run (no floating-point atomics).
//  * small systems (tracking): the last CTA of the last edge also assembles, factors and solves the
//    system and retracts the poses (gn_device.cuh: small_solve), i.e. ONE launch per GN iteration.
//
// H_ii = H_jj = A_e, H_ij = H_ji = -A_e, g_i = -b_e, g_j = +b_e (gn_kernels.cu:1096-1137).
#include <cooperative_groups.h>

#include <algorithm>
#include <atomic>

#include "gn_device.cuh"
#include "tma.cuh"

namespace m3r {
namespace {

__host__ __device__ constexpr int tri(int a, int b) { return a * 7 - a * (a - 1) / 2 + (b - a); }  // a <= b

// acc[0..27] += w * J J^T (upper triangle), acc[28..34] += w * r * J ; MASK = non-zero columns of J
template <unsigned MASK>
__device__ __forceinline__ void rank1(float (&acc)[kAcc], const float (&J)[7], float w, float r,
                                      float* cost = nullptr) {
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
  if (cost) *cost = fmaf(wr, r, *cost);  // sum of (robust sqrt-info * r)^2: the frame tracker's cost (x2)
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
                                            float xjx, float xjy, float xjz, float sq, int mi,
                                            float xix, float xiy, float xiz, float* cost = nullptr) {
  bool ok = true;
  const float px = fmaf(X.m[0], xjx, fmaf(X.m[1], xjy, fmaf(X.m[2], xjz, X.t[0])));
  const float py = fmaf(X.m[3], xjx, fmaf(X.m[4], xjy, fmaf(X.m[5], xjz, X.t[1])));
  const float pz = fmaf(X.m[6], xjx, fmaf(X.m[7], xjy, fmaf(X.m[8], xjz, X.t[2])));

  if (KIND == M3R_GN_POINTS) {
    const float sw = prm.inv_sigma_a * sq;
    const float wc = sw * sw;
    const float e0 = px - xix, e1 = py - xiy, e2 = pz - xiz;
    {
      const float J[7] = {1.f, 0.f, 0.f, 0.f, pz, -py, px};
      rank1<0x71u>(acc, J, huber(sw * e0) * wc, e0, cost);
    }
    {
      const float J[7] = {0.f, 1.f, 0.f, -pz, 0.f, px, py};
      rank1<0x6Au>(acc, J, huber(sw * e1) * wc, e1, cost);
    }
    {
      const float J[7] = {0.f, 0.f, 1.f, py, -px, 0.f, pz};
      rank1<0x5Cu>(acc, J, huber(sw * e2) * wc, e2, cost);
    }
  } else if (KIND == M3R_GN_RAYS) {
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
      rank1<0x37u>(acc, J, huber(swr * e0) * wr, e0, cost);
    }
    {
      const float J[7] = {xy, yy, yz, -rz, 0.f, rx, 0.f};
      rank1<0x2Fu>(acc, J, huber(swr * e1) * wr, e1, cost);
    }
    {
      const float J[7] = {xz, yz, zz, ry, -rx, 0.f, 0.f};
      rank1<0x1Fu>(acc, J, huber(swr * e2) * wr, e2, cost);
    }
    {
      const float J[7] = {rx, ry, rz, 0.f, 0.f, 0.f, rho};
      rank1<0x47u>(acc, J, huber(swd * e3) * wd, e3, cost);
    }
  } else {  // calib
    const bool vz = (pz > prm.z_eps) && (xiz > prm.z_eps);
    const float zinv = vz ? __frcp_rn(pz) : 0.f;
    const float xz_ = px * zinv, yz_ = py * zinv;
    const float u = fmaf(prm.fx, xz_, prm.cx), v = fmaf(prm.fy, yz_, prm.cy);
    ok = ok & vz & (u > prm.u_lo) & (u < prm.u_hi) & (v > prm.v_lo) & (v < prm.v_hi);
    if (!ok) return;
    const int ut = mi % prm.width, vt = mi / prm.width;
    const float e0 = u - static_cast<float>(ut), e1 = v - static_cast<float>(vt);
    const float e2 = logf(pz) - logf(xiz);
    const float swp = prm.inv_sigma_a * sq, swd = prm.inv_sigma_b * sq;
    const float wp = swp * swp, wd = swd * swd;
    {
      const float J[7] = {prm.fx * zinv, 0.f, -prm.fx * xz_ * zinv, -prm.fx * xz_ * yz_,
                          prm.fx * fmaf(xz_, xz_, 1.f), -prm.fx * yz_, 0.f};
      rank1<0x3Du>(acc, J, huber(swp * e0) * wp, e0, cost);
    }
    {
      const float J[7] = {0.f, prm.fy * zinv, -prm.fy * yz_ * zinv, -prm.fy * fmaf(yz_, yz_, 1.f),
                          prm.fy * xz_ * yz_, prm.fy * xz_, 0.f};
      rank1<0x3Eu>(acc, J, huber(swp * e1) * wp, e1, cost);
    }
    {
      const float J[7] = {0.f, 0.f, zinv, yz_, -xz_, 0.f, 1.f};
      rank1<0x5Cu>(acc, J, huber(swd * e2) * wd, e2, cost);
    }
  }
}

struct AccumArgs {
  const float* Twc;
  const float* Xs;
  const float* K;
  const uint2* packed;  // local slabs [E_local, n]: per chunk, `counts` compacted {int32 m, float sqrt(Q)}
  const uint16_t* kloc; // chunk-local point index of each record
  const int* counts;    // [E_local, chunks]
  const int* ci;
  const int* cj;
  const int* ctrl;
  const int* order;     // launch slot -> local edge (NULL: identity); see gn_prepare_kernel
  unsigned* counters;   // [E] per-edge tickets, [E] (last word) = edges done in this launch
  double* partials;
  double* eb;
  int n, e_lo, e_local, chunks, chunk_pts;
  AccumParams prm;
  int ss_counter;       // index in counters of the "edges done" ticket
  int fuse;             // 1: the last CTA of the launch solves + retracts (small systems)
  SmallSolveArgs ss;
};

// relative pose T_ij = T_i^-1 T_j in fp64 -> fp32 3x4
// element k of the edge transform: k < 9 -> m[k] = s R(q_i^-1 q_j)[k], k >= 9 -> t[k-9]; fp64 arithmetic.
// One element per call so that the persistent kernel can spread the 12 elements over 12 warps (the serial
// version cost ~1 us per iteration on its critical path); with a compile-time k the common terms fold.
__device__ __forceinline__ float edge_xform_elem(const float* __restrict__ Ti, const float* __restrict__ Tj, int k) {
  const double qix = -(double)Ti[3], qiy = -(double)Ti[4], qiz = -(double)Ti[5], qiw = Ti[6];  // q_i^-1
  const double si_inv = 1.0 / (double)Ti[7];
  if (k < 9) {
    const double qjx = Tj[3], qjy = Tj[4], qjz = Tj[5], qjw = Tj[6];
    // q = q_i^-1 * q_j
    const double x = qiw * qjx + qix * qjw + qiy * qjz - qiz * qjy;
    const double y = qiw * qjy - qix * qjz + qiy * qjw + qiz * qjx;
    const double z = qiw * qjz + qix * qjy - qiy * qjx + qiz * qjw;
    const double w = qiw * qjw - qix * qjx - qiy * qjy - qiz * qjz;
    const double s = si_inv * (double)Tj[7];
    // R(q) as the linear map of actSO3 (no normalisation, like the reference): I + 2w[q]x + 2[q]x^2
    double r;
    switch (k) {
      case 0: r = 1 - 2 * (y * y + z * z); break;
      case 1: r = 2 * (x * y - z * w); break;
      case 2: r = 2 * (x * z + y * w); break;
      case 3: r = 2 * (x * y + z * w); break;
      case 4: r = 1 - 2 * (x * x + z * z); break;
      case 5: r = 2 * (y * z - x * w); break;
      case 6: r = 2 * (x * z - y * w); break;
      case 7: r = 2 * (y * z + x * w); break;
      default: r = 1 - 2 * (x * x + y * y); break;
    }
    return (float)(s * r);
  }
  // t = R(q_i^-1) (t_j - t_i) / s_i
  const double dx = (double)Tj[0] - (double)Ti[0], dy = (double)Tj[1] - (double)Ti[1], dz = (double)Tj[2] - (double)Ti[2];
  double a0, a1, a2;
  switch (k) {
    case 9: a0 = 1 - 2 * (qiy * qiy + qiz * qiz); a1 = 2 * (qix * qiy - qiz * qiw); a2 = 2 * (qix * qiz + qiy * qiw); break;
    case 10: a0 = 2 * (qix * qiy + qiz * qiw); a1 = 1 - 2 * (qix * qix + qiz * qiz); a2 = 2 * (qiy * qiz - qix * qiw); break;
    default: a0 = 2 * (qix * qiz - qiy * qiw); a1 = 2 * (qiy * qiz + qix * qiw); a2 = 1 - 2 * (qix * qix + qiy * qiy); break;
  }
  return (float)((a0 * dx + a1 * dy + a2 * dz) * si_inv);
}

__device__ __forceinline__ void edge_xform(const float* __restrict__ Ti, const float* __restrict__ Tj, EdgeXform& X) {
#pragma unroll
  for (int k = 0; k < 9; k++) X.m[k] = edge_xform_elem(Ti, Tj, k);
#pragma unroll
  for (int k = 0; k < 3; k++) X.t[k] = edge_xform_elem(Ti, Tj, 9 + k);
}

// ---- pose-independent fold + compaction, once per solve -----------------------------------------
// Per (edge, chunk): the points that pass  valid_match && Q > Q_thresh && C_i[m] > C_thresh &&
// C_j[k] > C_thresh  (gn_kernels.cu:953-957; none of it depends on the poses) are compacted, in
// order, into {int32 m, float sqrt(Q)} records + their chunk-local index k (u16) + a count.  The
// iterations then stream ~10 B per SURVIVING point instead of 13 B + two confidence gathers per point
// and spend no instruction on rejected points.
struct PackArgs {
  const float* Cs;
  const int64_t* idx;   // local slabs [E_local, n]
  const uint8_t* valid;
  const float* Q;
  const int* ci;
  const int* cj;
  uint2* packed;
  uint16_t* kloc;
  int* counts;
  int n, e_lo, chunk_pts, chunks;
  float C_thresh, Q_thresh;
};

__device__ __forceinline__ bool pack_one(long long m, bool vm, float q, float cj, const float* __restrict__ Ci,
                                         float C_thresh, float Q_thresh, uint2& rec) {
  // invalid matches read C_i[0] in the reference (gn_kernels.cu:913-914); the result is discarded
  // (Ci == nullptr: frame-tracker mode, the caller's valid mask already holds every confidence test)
  const float ci = Ci ? __ldg(Ci + (vm ? m : 0)) : INFINITY;
  rec = make_uint2((unsigned)(int)m, __float_as_uint(sqrtf(q)));
  return vm & (q > Q_thresh) & (ci > C_thresh) & (cj > C_thresh);
}

template <bool VEC4>
__global__ void __launch_bounds__(256)
gn_pack_kernel(const PackArgs a) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int el = blockIdx.y, chunk = blockIdx.x;
  const int e = a.e_lo + el;
  const int n = a.n;
  const int ic = a.ci[e], jc = a.cj[e];
  const bool track = a.Cs == nullptr;  // frame-tracker mode: identity index, no confidence maps
  const float* __restrict__ Ci = track ? nullptr : a.Cs + (size_t)ic * n;
  const float* __restrict__ Cj = track ? nullptr : a.Cs + (size_t)jc * n;
  const int64_t* __restrict__ idx = track ? nullptr : a.idx + (size_t)el * n;
  const uint8_t* __restrict__ valid = a.valid + (size_t)el * n;
  const float* __restrict__ Q = a.Q + (size_t)el * n;
  const int p_lo = chunk * a.chunk_pts;
  const int p_hi = min(n, p_lo + a.chunk_pts);
  uint2* __restrict__ out_rec = a.packed + (size_t)el * n + p_lo;
  uint16_t* __restrict__ out_k = a.kloc + (size_t)el * n + p_lo;
  __shared__ int s_wsum[8];
  int base = 0;
  if (ic != jc) {  // a self-edge contributes nothing: count 0
    for (int pass_lo = p_lo; pass_lo < p_hi; pass_lo += 1024) {
      const int k0 = pass_lo + tid * 4;
      uint2 rec[4];
      bool ok[4] = {false, false, false, false};
      if (VEC4 && k0 + 3 < p_hi) {
        const uchar4 vm = __ldcs(reinterpret_cast<const uchar4*>(valid + k0));
        longlong2 m01 = make_longlong2(k0, k0 + 1), m23 = make_longlong2(k0 + 2, k0 + 3);
        float4 c4 = make_float4(INFINITY, INFINITY, INFINITY, INFINITY);
        if (!track) {
          m01 = __ldcs(reinterpret_cast<const longlong2*>(idx + k0));
          m23 = __ldcs(reinterpret_cast<const longlong2*>(idx + k0) + 1);
          c4 = __ldg(reinterpret_cast<const float4*>(Cj + k0));
        }
        const float4 q4 = __ldcs(reinterpret_cast<const float4*>(Q + k0));
        ok[0] = pack_one(m01.x, vm.x != 0, q4.x, c4.x, Ci, a.C_thresh, a.Q_thresh, rec[0]);
        ok[1] = pack_one(m01.y, vm.y != 0, q4.y, c4.y, Ci, a.C_thresh, a.Q_thresh, rec[1]);
        ok[2] = pack_one(m23.x, vm.z != 0, q4.z, c4.z, Ci, a.C_thresh, a.Q_thresh, rec[2]);
        ok[3] = pack_one(m23.y, vm.w != 0, q4.w, c4.w, Ci, a.C_thresh, a.Q_thresh, rec[3]);
      } else {
#pragma unroll
        for (int r = 0; r < 4; r++) {
          const int k = k0 + r;
          if (k < p_hi)
            ok[r] = pack_one(track ? (long long)k : idx[k], valid[k] != 0, Q[k], track ? INFINITY : __ldg(Cj + k), Ci,
                             a.C_thresh, a.Q_thresh, rec[r]);
        }
      }
      const int cnt = (int)ok[0] + (int)ok[1] + (int)ok[2] + (int)ok[3];
      int incl = cnt;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += v;
      }
      if (lane == 31) s_wsum[warp] = incl;
      __syncthreads();
      int woff = 0, total = 0;
#pragma unroll
      for (int wv = 0; wv < 8; wv++) {
        const int t = s_wsum[wv];
        if (wv < warp) woff += t;
        total += t;
      }
      int off = base + woff + incl - cnt;
#pragma unroll
      for (int r = 0; r < 4; r++) {
        if (ok[r]) {
          out_rec[off] = rec[r];
          out_k[off] = (uint16_t)(k0 + r - p_lo);
          off++;
        }
      }
      base += total;
      __syncthreads();
    }
  }
  if (tid == 0) a.counts[(size_t)el * a.chunks + chunk] = base;
}

template <int KIND, bool STAGED>
__global__ void __launch_bounds__(kAccumThreads, kAccumCtasPerSM)
gn_accum_kernel(const AccumArgs a) {
  // Programmatic dependent launch: iteration k+1 of the fused GN loop is allowed on the SMs while
  // iteration k is still in its serial tail.  Everything up to griddepcontrol.wait touches only data
  // that is constant during a solve (launch order, compact indices, packed records, Xs): the TMA
  // staging of this CTA's chunk overlaps the previous iteration's reduction / solve / retraction.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ EdgeXform s_X;
  const int tid = threadIdx.x;
  const int el = a.order ? a.order[blockIdx.y] : blockIdx.y;  // local edge
  const int e = a.e_lo + el;                                   // global edge
  const int chunk = blockIdx.x;
  const int i = a.ci[e], j = a.cj[e];
  const int n = a.n;
  const int p_lo = chunk * a.chunk_pts;
  const int npts = min(n, p_lo + a.chunk_pts) - p_lo;
  const int cnt = a.counts[(size_t)el * a.chunks + chunk];  // surviving points of this chunk

  const float* __restrict__ Xi = a.Xs + (size_t)i * n * 3;
  const float* __restrict__ Xj = a.Xs + ((size_t)j * n + p_lo) * 3;
  const uint2* __restrict__ grec = a.packed + (size_t)el * n + p_lo;
  const uint16_t* __restrict__ gk = a.kloc + (size_t)el * n + p_lo;
  // smem: Xs[j] chunk (12 B/pt) | records (8 B) | local indices (2 B); chunk_pts % 8 == 0 keeps all 16-B aligned
  const uint32_t off_rec = (uint32_t)a.chunk_pts * 12u, off_k = (uint32_t)a.chunk_pts * 20u;

  if (STAGED) {
    if (tid == 0) {
      mbar_init(smem_u32(&s_bar), 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      if (cnt > 0) {
        const uint64_t pol_stream = l2_policy_evict_first();  // records are read once per iteration
        const uint64_t pol_keep = l2_policy_evict_last();     // Xs[j] is re-read by the other edges of j
        const uint32_t cnt8 = ((uint32_t)cnt + 7u) & ~7u;     // <= chunk_pts; the pad is never read back
        const uint32_t bar = smem_u32(&s_bar);
        const uint32_t dst = smem_u32(dsm);
        mbar_expect_tx(bar, (uint32_t)npts * 12u + cnt8 * 10u);
        bulk_g2s(dst, Xj, (uint32_t)npts * 12u, bar, pol_keep);
        bulk_g2s(dst + off_rec, grec, cnt8 * 8u, bar, pol_stream);
        bulk_g2s(dst + off_k, gk, cnt8 * 2u, bar, pol_stream);
      }
    }
  }
  // from here on: poses and the stop flag written by the previous iteration
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int done = a.ctrl[CTRL_DONE];
  if (tid == 0 && !done) edge_xform(a.Twc + (size_t)i * 8, a.Twc + (size_t)j * 8, s_X);
  __syncthreads();  // s_X, and the mbarrier init
  if (done) {
    if (STAGED && cnt > 0) mbar_wait(smem_u32(&s_bar), 0);  // never exit with bulk copies still landing here
    return;
  }
  EdgeXform X;
#pragma unroll
  for (int t = 0; t < 9; t++) X.m[t] = s_X.m[t];
#pragma unroll
  for (int t = 0; t < 3; t++) X.t[t] = s_X.t[t];
  AccumParams prm = a.prm;
  if (KIND == M3R_GN_CALIB) {  // K[0][0], K[1][1], K[0][2], K[1][2] (gn_kernels.cu:1274-1279)
    prm.fx = __ldg(a.K + 0);
    prm.fy = __ldg(a.K + 4);
    prm.cx = __ldg(a.K + 2);
    prm.cy = __ldg(a.K + 5);
  }

  float acc[kAcc];
#pragma unroll
  for (int t = 0; t < kAcc; t++) acc[t] = 0.f;

  if (cnt > 0) {
    const float* sx;
    const uint2* srec;
    const uint16_t* sk;
    if (STAGED) {
      mbar_wait(smem_u32(&s_bar), 0);
      sx = reinterpret_cast<const float*>(dsm);
      srec = reinterpret_cast<const uint2*>(dsm + off_rec);
      sk = reinterpret_cast<const uint16_t*>(dsm + off_k);
    } else {
      sx = Xj;
      srec = grec;
      sk = gk;
    }
    for (int k0 = tid; k0 < cnt; k0 += 4 * kAccumThreads) {
      int m[4], kl[4];
      float sq[4], xi[4][3];
#pragma unroll
      for (int r = 0; r < 4; r++) {
        const int k = k0 + r * kAccumThreads;
        m[r] = -1;
        if (k < cnt) {
          const uint2 rec = srec[k];
          m[r] = (int)rec.x;
          sq[r] = __uint_as_float(rec.y);
          kl[r] = sk[k];
        }
      }
      // all gathers of this pass are in flight before any arithmetic
#pragma unroll
      for (int r = 0; r < 4; r++) {
        if (m[r] >= 0) {
          const float* g = Xi + (size_t)m[r] * 3;
          xi[r][0] = __ldg(g + 0);
          xi[r][1] = __ldg(g + 1);
          xi[r][2] = __ldg(g + 2);
        }
      }
#pragma unroll
      for (int r = 0; r < 4; r++) {
        if (m[r] >= 0) {
          const float* xj = sx + kl[r] * 3;
          accum_point<KIND>(acc, X, prm, xj[0], xj[1], xj[2], sq[r], m[r], xi[r][0], xi[r][1], xi[r][2]);
        }
      }
    }
  }

  // ---- CTA reduction: shuffles -> smem -> fp64 partial ----
  __shared__ float s_warp[kAccumThreads / 32][kAcc];
  __shared__ double s_part[7][kAcc];
  __shared__ double s_mv[kAcc];
  __shared__ double s_G[49];
  __shared__ double s_T[56];
  __shared__ bool s_last;
  const int lane = tid & 31, warp = tid >> 5;
  // terms 0..31: transposing butterfly (31 shuffles instead of 160): after the step with offset
  // `off`, slot i of a lane holds the sum over its xor-group of term i | (lane & off); lane l ends
  // up with the warp total of term l.  Terms 32..34: plain butterflies.
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int t = 0; t < off; t++) {
      const float send = up ? acc[t] : acc[t + off];
      const float keep = up ? acc[t + off] : acc[t];
      acc[t] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  s_warp[warp][lane] = acc[0];
#pragma unroll
  for (int t = 32; t < kAcc; t++) {
    float v = acc[t];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0) s_warp[warp][t] = v;
  }
  __syncthreads();
  if (tid < kAcc) {
    double s = 0.0;
#pragma unroll
    for (int wv = 0; wv < kAccumThreads / 32; wv++) s += (double)s_warp[wv][tid];
    if (a.chunks > 1)
      a.partials[((size_t)el * a.chunks + chunk) * kAcc + tid] = s;
    else
      s_mv[tid] = s;
  }
  if (a.chunks > 1) {
    __threadfence();
    __syncthreads();
    if (tid == 0) {
      const unsigned ticket = atomicAdd(a.counters + e, 1u);
      s_last = (ticket == (unsigned)a.chunks - 1u);
      if (s_last) a.counters[e] = 0;  // re-arm for the next iteration
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // last CTA of the edge: fixed-order fp64 sum of the partials (7 interleaved slices, then 7 -> 1)
    if (tid < 7 * kAcc) {
      const int t = tid % kAcc, sl = tid / kAcc;
      const double* p0 = a.partials + (size_t)el * a.chunks * kAcc + t;
      double s = 0.0;
      int c = sl;
      for (; c + 7 * 7 < a.chunks; c += 7 * 8) {  // 8 loads in flight, summed in a fixed order
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) v[u] = __ldcg(p0 + (size_t)(c + 7 * u) * kAcc);
#pragma unroll
        for (int u = 0; u < 8; u++) s += v[u];
      }
      for (; c < a.chunks; c += 7) s += __ldcg(p0 + (size_t)c * kAcc);
      s_part[sl][t] = s;
    }
    __syncthreads();
    if (tid < kAcc) {
      double s = s_part[0][tid];
#pragma unroll
      for (int sl = 1; sl < 7; sl++) s += s_part[sl][tid];
      s_mv[tid] = s;
    }
  }

  // ---- adjoint: A_e = G^T M G, b_e = G^T v in fp64 ----
  if (tid >= 64 && tid < 64 + 7) {
    // G: Y = X * G is the row-vector adjoint of gn_kernels.cu:277-297 for pose i; row m = image of e_m
    const int m = tid - 64;
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
  if (tid < 49) {  // T = M G
    const int p = tid / 7, c = tid % 7;
    double t = 0.0;
#pragma unroll
    for (int q = 0; q < 7; q++) t += ((p <= q) ? s_mv[tri(p, q)] : s_mv[tri(q, p)]) * s_G[q * 7 + c];
    s_T[tid] = t;
  }
  __syncthreads();
  double* out = a.eb + (size_t)e * kEdgeBlock;
  if (tid < 49) {
    const int r = tid / 7, c = tid % 7;
    double s = 0.0;
#pragma unroll
    for (int p = 0; p < 7; p++) s += s_G[p * 7 + r] * s_T[p * 7 + c];
    out[tid] = s;
  } else if (tid < 56) {
    const int r = tid - 49;
    double s = 0.0;
#pragma unroll
    for (int p = 0; p < 7; p++) s += s_G[p * 7 + r] * s_mv[28 + p];
    out[tid] = s;
  }
  if (!a.fuse) return;

  // ---- fused small-system solve: the CTA that completes the LAST edge of the launch ----
  if (a.e_local > 1) {
    __threadfence();
    __syncthreads();
    if (tid == 0) {
      unsigned* done = a.counters + a.ss_counter;
      const unsigned ticket = atomicAdd(done, 1u);
      s_last = (ticket == (unsigned)a.e_local - 1u);
      if (s_last) *done = 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
  } else {
    __threadfence();
    __syncthreads();
  }
  small_solve(a.ss, reinterpret_cast<double*>(dsm));
}


// ---- persistent tracking kernel ------------------------------------------------------------------
// Tracking-sized problems (<= kTrackMaxEdges directed edges, all of a CTA's points fit in shared
// memory): the whole Gauss-Newton loop is ONE cooperative launch with one CTA per SM.  A CTA stages
// its share of one edge ONCE (TMA bulk copies of the packed records and Xs[j], then a gather of the
// pose-independent Xs[i][m] into shared memory), so the iterations never touch HBM or L2 for point
// data.  Per iteration: accumulate from shared memory -> one fp64 partial per CTA -> grid barrier ->
// EVERY CTA sums the partials in the same fixed order, applies the adjoints, solves the small system
// and retracts its own shared-memory copy of the poses (bit-identical in all CTAs, so there is no
// second barrier and no broadcast); CTA 0 writes dx / Twc / info.  Replaces max_iter launches of
// gn_accum_kernel (each ~20 us of launch + staging + last-CTA serial tail at 2 x 196 608 points).
#ifdef M3R_TRACK_PROF
__device__ unsigned long long g_dbg[1024];
// phase stamps of iteration 5 in the solver CTA (globaltimer, ns): cheap enough not to disturb the timeline
#define TRK(slot)                                                     \
  do {                                                                \
    if (blockIdx.x == 0 && tid == 0 && it == 5) {                     \
      unsigned long long gt_;                                         \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_));         \
      g_dbg[700 + (slot)] = gt_;                                      \
    }                                                                 \
    (void)trk_t;                                                      \
  } while (0)
#else
#define TRK(slot) do { } while (0)
#endif
constexpr int kTrackThreads = 512;
// (kTrackMaxEdges = 8: gn_common.cuh)
constexpr int kTrackPtBytes = 34;  // Xs[j] 12 | record 8 | local index 2 | gathered Xs[i][m] 12
constexpr int kTrackMaxPts = 4800; // x 34 B = 163 KB staged + 33 KB solve + 33 KB static <= 227 KB per CTA

struct TrackArgs {
  AccumArgs a;
  int cpe, upc;       // CTAs per edge, point-chunks ("units") per CTA
  int max_iter;
  int stage_pts;      // upc * chunk_pts
  unsigned epoch;     // per-launch nonce (20 bits) of the partial all-gather flags
  // frame-tracker mode (COST): one directed edge whose two point sets are separate arrays, and the stop test of
  // mast3r_slam/nonlinear_optimizer.py:5-26 on the relative decrease of the robust cost
  const float* Xi_base;  // target points (keyframe), or nullptr: a.Xs + i*n*3
  const float* Xj_base;  // moving points (frame, already gathered), or nullptr: a.Xs + j*n*3
  float rel_error;
};

template <int KIND, bool COST>
__global__ void __launch_bounds__(kTrackThreads, 1)
gn_track_kernel(const TrackArgs ta) {
  const AccumArgs& a = ta.a;
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ EdgeXform s_X;
  __shared__ float s_pose[(kSmallDim / 7 + 1) * 8];
  __shared__ float s_warp[kTrackThreads / 32][kAcc];
  __shared__ double s_part[8][kTrackMaxEdges * kAcc];
  __shared__ double s_mvE[kTrackMaxEdges * kAcc];
  __shared__ double s_G[kTrackThreads / 64][49];
  __shared__ double s_TT[kTrackThreads / 64][49];
  __shared__ double s_eb[kTrackMaxEdges * kEdgeBlock];
  __shared__ int s_cnt[64];
  __shared__ int s_pref[65];  // exclusive prefix of s_cnt: the accumulate loop runs over the CTA's points as one range
  __shared__ TrackGraph s_g;
  __shared__ float s_dx[64];
  __shared__ int s_flags[2];
  __shared__ float s_costw[kTrackThreads / 32];
  __shared__ double s_cost[2];  // [0] this iteration's sum, [1] previous cost

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int E = a.e_local, n = a.n, N = a.ss.N;
  const int e = blockIdx.x / ta.cpe, sub = blockIdx.x % ta.cpe;
  const int i = a.ci[e], j = a.cj[e];
  const int u_lo = min(sub * ta.upc, a.chunks), u_hi = min(a.chunks, u_lo + ta.upc);
  const int nu = u_hi - u_lo;
  if (a.ctrl[CTRL_DONE]) return;  // uniform over the grid: written before this launch
#ifdef M3R_TRACK_PROF
  long long trk_t = clock64();
  if (blockIdx.x == 0 && tid == 0) {
    unsigned long long gt_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_));
    g_dbg[720] = gt_;  // kernel start
  }
#endif

  // shared-memory layout: per unit u a slab of chunk_pts points: Xj | rec | kloc | Xi
  const uint32_t cp = (uint32_t)a.chunk_pts;
  const uint32_t slab = cp * kTrackPtBytes;
  const uint32_t off_rec = cp * 12u, off_k = cp * 20u, off_xi = cp * 22u;
  double* solve_sm = reinterpret_cast<double*>(dsm + (size_t)ta.stage_pts * kTrackPtBytes);

  if (tid < nu) s_cnt[tid] = a.counts[(size_t)e * a.chunks + u_lo + tid];
  if (tid == 0) {
    mbar_init(smem_u32(&s_bar), 1);
    mbar_fence_init();
  }
  for (int t = tid; t < N * 8; t += kTrackThreads) s_pose[t] = a.Twc[t];
  if (tid < E) {
    s_g.ci[tid] = a.ci[tid];
    s_g.cj[tid] = a.cj[tid];
  }
  if (tid >= 32 && tid < 32 + N + 1) s_g.inc_off[tid - 32] = a.ss.inc_off[tid - 32];
  if (tid >= 64 && tid < 64 + 2 * E) s_g.inc_list[tid - 64] = a.ss.inc_list[tid - 64];
  const int bad_ids = a.ctrl[CTRL_BAD];
  __syncthreads();
  if (tid == 0) {
    const uint64_t pol = l2_policy_evict_first();
    const uint32_t bar = smem_u32(&s_bar);
    uint32_t total = 0;
    for (int u = 0; u < nu; u++) {
      const int cnt = s_cnt[u];
      if (cnt <= 0) continue;
      const int p_lo = (u_lo + u) * a.chunk_pts;
      const int npts = min(n, p_lo + a.chunk_pts) - p_lo;
      const uint32_t cnt8 = ((uint32_t)cnt + 7u) & ~7u;
      total += (uint32_t)npts * 12u + cnt8 * 10u;
    }
    mbar_expect_tx(bar, total);
    for (int u = 0; u < nu; u++) {
      const int cnt = s_cnt[u];
      if (cnt <= 0) continue;
      const int p_lo = (u_lo + u) * a.chunk_pts;
      const int npts = min(n, p_lo + a.chunk_pts) - p_lo;
      const uint32_t cnt8 = ((uint32_t)cnt + 7u) & ~7u;
      const uint32_t dst = smem_u32(dsm) + (uint32_t)u * slab;
      const float* Xj = ta.Xj_base ? ta.Xj_base : a.Xs + (size_t)j * n * 3;
      bulk_g2s(dst, Xj + (size_t)p_lo * 3, (uint32_t)npts * 12u, bar, pol);
      bulk_g2s(dst + off_rec, a.packed + (size_t)e * n + p_lo, cnt8 * 8u, bar, pol);
      bulk_g2s(dst + off_k, a.kloc + (size_t)e * n + p_lo, cnt8 * 2u, bar, pol);
    }
  }
  if (tid == 0) {
    int run = 0;
    for (int u = 0; u < nu; u++) {
      s_pref[u] = run;
      run += max(s_cnt[u], 0);
    }
    s_pref[nu] = run;
  }
  mbar_wait(smem_u32(&s_bar), 0);
  // gather the matched points of frame i once (they do not depend on the poses)
  {
    const float* __restrict__ Xi = ta.Xi_base ? ta.Xi_base : a.Xs + (size_t)i * n * 3;
    for (int u = 0; u < nu; u++) {
      const int cnt = s_cnt[u];
      const uint2* srec = reinterpret_cast<const uint2*>(dsm + (size_t)u * slab + off_rec);
      float* sxi = reinterpret_cast<float*>(dsm + (size_t)u * slab + off_xi);
      for (int k = tid; k < cnt; k += kTrackThreads) {
        const float* g = Xi + (size_t)srec[k].x * 3;
        sxi[k * 3 + 0] = __ldg(g + 0);
        sxi[k * 3 + 1] = __ldg(g + 1);
        sxi[k * 3 + 2] = __ldg(g + 2);
      }
    }
  }
#ifdef M3R_TRACK_PROF
  if (blockIdx.x == 0 && tid == 0) {
    unsigned long long gt_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_));
    g_dbg[721] = gt_;  // staging + gather done
  }
#endif
  AccumParams prm = a.prm;
  if (KIND == M3R_GN_CALIB) {
    prm.fx = __ldg(a.K + 0);
    prm.fy = __ldg(a.K + 4);
    prm.cx = __ldg(a.K + 2);
    prm.cy = __ldg(a.K + 5);
  }
  const bool writer = blockIdx.x == 0;
  const int pairs = E * kAcc;
  int nslice = kTrackThreads / pairs;
  nslice = nslice > 8 ? 8 : nslice;
  int nfail = 0;

  for (int it = 0; it < ta.max_iter; it++) {
    if (lane == 0 && warp < 12) {  // one element per warp: no divergence, ~0.1 us instead of ~1 us
      const float v = edge_xform_elem(s_pose + (size_t)i * 8, s_pose + (size_t)j * 8, warp);
      if (warp < 9) s_X.m[warp] = v;
      else s_X.t[warp - 9] = v;
    }
    __syncthreads();  // s_X (and, first iteration, the gathered Xi)
    TRK(1);  // edge xform
    EdgeXform X;
#pragma unroll
    for (int t = 0; t < 9; t++) X.m[t] = s_X.m[t];
#pragma unroll
    for (int t = 0; t < 3; t++) X.t[t] = s_X.t[t];
    float acc[kAcc];
#pragma unroll
    for (int t = 0; t < kAcc; t++) acc[t] = 0.f;
    float cost = 0.f;
    // one flat range over the surviving points of all units: ceil(total / 512) rounds instead of one partly filled
    // last round per unit
    {
      const int total = s_pref[nu];
      int u = 0;
      for (int f = tid; f < total; f += kTrackThreads) {
        while (f >= s_pref[u + 1]) u++;
        const int k = f - s_pref[u];
        const unsigned char* base = dsm + (size_t)u * slab;
        const float* sx = reinterpret_cast<const float*>(base);
        const uint2 rec = reinterpret_cast<const uint2*>(base + off_rec)[k];
        const float* xj = sx + (int)reinterpret_cast<const uint16_t*>(base + off_k)[k] * 3;
        const float* xi = reinterpret_cast<const float*>(base + off_xi) + k * 3;
        accum_point<KIND>(acc, X, prm, xj[0], xj[1], xj[2], __uint_as_float(rec.y), (int)rec.x, xi[0], xi[1], xi[2],
                          COST ? &cost : nullptr);
      }
    }
    TRK(2);  // accumulate
    if (COST) {
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) cost += __shfl_xor_sync(0xffffffffu, cost, off);
      if (lane == 0) s_costw[warp] = cost;
    }
    // CTA reduction (same transposing butterfly as gn_accum_kernel) -> fp64 partial of this CTA
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const bool up = (lane & off) != 0;
#pragma unroll
      for (int t = 0; t < off; t++) {
        const float send = up ? acc[t] : acc[t + off];
        const float keep = up ? acc[t + off] : acc[t];
        acc[t] = keep + __shfl_xor_sync(0xffffffffu, send, off);
      }
    }
    s_warp[warp][lane] = acc[0];
#pragma unroll
    for (int t = 32; t < kAcc; t++) {
      float v = acc[t];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
      if (lane == 0) s_warp[warp][t] = v;
    }
    __syncthreads();
    // All-gather of the per-CTA partials without a barrier (the NCCL "LL" idea): every fp64 partial is
    // published as two 8-byte words {flag : 32 | half of the value : 32}; 8-byte stores are single-copy
    // atomic, so a reader that sees this iteration's flag in a word also sees its payload and simply
    // polls the words it needs.  One store + one L2 round trip replace [fence, atomic arrive, spin,
    // load].  Two buffers alternate: a CTA can only be two iterations ahead of the slowest reader
    // after that reader has consumed the older buffer.  The launch is cooperative (co-resident CTAs).
    ulonglong2* ll = reinterpret_cast<ulonglong2*>(a.partials) + (size_t)(it & 1) * gridDim.x * kAcc;
    const unsigned long long flag = ((unsigned long long)(ta.epoch << 12 | (unsigned)(it + 1))) << 32;
    // (COST) the CTA's cost partial travels the same way, in its own words behind the pose words
    ulonglong2* cost_ll = reinterpret_cast<ulonglong2*>(reinterpret_cast<unsigned long long*>(
                              reinterpret_cast<ulonglong2*>(a.partials) + (size_t)2 * gridDim.x * kAcc) + 192) +
                          (size_t)(it & 1) * gridDim.x;
    if (COST && tid == kAcc) {
      double s = 0.0;
#pragma unroll
      for (int wv = 0; wv < kTrackThreads / 32; wv++) s += (double)s_costw[wv];
      const unsigned long long bits = (unsigned long long)__double_as_longlong(s);
      asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(cost_ll + blockIdx.x),
                   "l"(flag | (bits & 0xffffffffull)), "l"(flag | (bits >> 32))
                   : "memory");
    }
    if (tid < kAcc) {
      double s = 0.0;
#pragma unroll
      for (int wv = 0; wv < kTrackThreads / 32; wv++) s += (double)s_warp[wv][tid];
      const unsigned long long bits = (unsigned long long)__double_as_longlong(s);
      ulonglong2 w;
      w.x = flag | (bits & 0xffffffffull);
      w.y = flag | (bits >> 32);
      asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(ll + (size_t)blockIdx.x * kAcc + tid), "l"(w.x),
                   "l"(w.y)
                   : "memory");
    }
    TRK(3);  // CTA reduction + partial store
#ifdef M3R_TRACK_PROF
    if (it == 5 && tid == 0) {
      unsigned long long gt;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
      g_dbg[blockIdx.x] = gt;
    }
#endif
    // the poses of this iteration travel back the same way: N*8 floats + the stop flag, one word each
    unsigned long long* pose_ll = reinterpret_cast<unsigned long long*>(
                                      reinterpret_cast<ulonglong2*>(a.partials) + (size_t)2 * gridDim.x * kAcc) +
                                  (size_t)(it & 1) * 96;
    if (!writer) {
      // every other CTA just waits for the new poses (81 words at most, polled with a short back-off)
      if (tid <= N * 8) {
        unsigned long long wv;
        while (true) {
          asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(wv) : "l"(pose_ll + tid) : "memory");
          if ((wv >> 32) == (flag >> 32)) break;
          __nanosleep(64);
        }
        if (tid < N * 8) s_pose[tid] = __uint_as_float((unsigned)wv);
        else s_flags[0] = (int)(unsigned)wv;
      }
#ifdef M3R_TRACK_PROF
      if (it == 4 && tid == 0) {
        unsigned long long gt;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
        g_dbg[300 + blockIdx.x] = gt;
      }
#endif
      __syncthreads();
      if (s_flags[0]) break;
      __syncthreads();  // s_flags is rewritten by the next iteration
      continue;
    }
    // CTA 0 (the solver).  The adjoint rows G_e depend on the poses only: the threads that do not poll compute
    // them while the partials are still arriving.
    {
      const int used = pairs * nslice;
      const int nfree = kTrackThreads - used - (COST ? 32 : 0);
      if (tid >= used && tid < used + nfree) {
        for (int row = tid - used; row < 7 * E; row += nfree) {
          const int ee = row / 7, m = row % 7;
          const float* Ti = s_pose + (size_t)s_g.ci[ee] * 8;
          const double tx = Ti[0], ty = Ti[1], tz = Ti[2], x = Ti[3], y = Ti[4], z = Ti[5], w = Ti[6];
          const double s_inv = 1.0 / (double)Ti[7];
          const int mm = m < 3 ? m : m - 3;  // column of R used by this row (selects, no indexed local array)
          const double c0 = mm == 0 ? 1 - 2 * (y * y + z * z) : mm == 1 ? 2 * (x * y - z * w) : 2 * (x * z + y * w);
          const double c1 = mm == 0 ? 2 * (x * y + z * w) : mm == 1 ? 1 - 2 * (x * x + z * z) : 2 * (y * z - x * w);
          const double c2 = mm == 0 ? 2 * (x * z - y * w) : mm == 1 ? 2 * (y * z + x * w) : 1 - 2 * (x * x + y * y);
          double rowv[7] = {0, 0, 0, 0, 0, 0, 0};
          if (m < 3) {
            rowv[0] = s_inv * c0; rowv[1] = s_inv * c1; rowv[2] = s_inv * c2;
            rowv[3] = s_inv * (ty * c2 - tz * c1);
            rowv[4] = s_inv * (tz * c0 - tx * c2);
            rowv[5] = s_inv * (tx * c1 - ty * c0);
            rowv[6] = s_inv * (tx * c0 + ty * c1 + tz * c2);
          } else if (m < 6) {
            rowv[3] = c0; rowv[4] = c1; rowv[5] = c2;
          } else {
            rowv[6] = 1.0;
          }
#pragma unroll
          for (int c = 0; c < 7; c++) s_G[ee][m * 7 + c] = rowv[c];
        }
      }
    }
    // fixed-order sum of the cpe partials of each edge
    if (tid < pairs * nslice) {
      const int pr = tid % pairs, sl = tid / pairs;
      const int ee = pr / kAcc, t = pr % kAcc;
      const ulonglong2* p0 = ll + (size_t)ee * ta.cpe * kAcc + t;
      double s = 0.0;
      for (int c = sl; c < ta.cpe; c += 12 * nslice) {  // 12 polls in flight, summed in a fixed order
        ulonglong2 w[12];
        unsigned pending = 0;
#pragma unroll
        for (int q = 0; q < 12; q++)
          if (c + q * nslice < ta.cpe) pending |= 1u << q;
        while (pending) {
#pragma unroll
          for (int q = 0; q < 12; q++) {
            if (pending & (1u << q))
              asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];"
                           : "=l"(w[q].x), "=l"(w[q].y)
                           : "l"(p0 + (size_t)(c + q * nslice) * kAcc)
                           : "memory");
          }
#pragma unroll
          for (int q = 0; q < 12; q++) {
            if ((pending & (1u << q)) && (w[q].x >> 32) == (flag >> 32) && (w[q].y >> 32) == (flag >> 32))
              pending &= ~(1u << q);
          }
        }
#pragma unroll
        for (int q = 0; q < 12; q++) {
          if (c + q * nslice < ta.cpe)
            s += __longlong_as_double((long long)((w[q].x & 0xffffffffull) | (w[q].y << 32)));
        }
      }
      s_part[sl][pr] = s;
    }
    if (COST && warp == kTrackThreads / 32 - 1) {  // (E == 1: the last warp is free) cost partials, fixed order
      double s = 0.0;
      for (int c = lane; c < (int)gridDim.x; c += 32) {
        ulonglong2 w;
        do {
          asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(w.x), "=l"(w.y) : "l"(cost_ll + c) : "memory");
        } while ((w.x >> 32) != (flag >> 32) || (w.y >> 32) != (flag >> 32));
        s += __longlong_as_double((long long)((w.x & 0xffffffffull) | (w.y << 32)));
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
      if (lane == 0) s_cost[0] = s;
    }
    TRK(4);  // all-gather
#ifdef M3R_TRACK_PROF
    if (it == 5 && tid == 0) {
      unsigned long long gt;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
      g_dbg[600] = gt;
    }
#endif
    __syncthreads();
    if (tid < pairs) {
      double s = s_part[0][tid];
      for (int sl = 1; sl < nslice; sl++) s += s_part[sl][tid];
      s_mvE[tid] = s;
    }
    __syncthreads();
    TRK(5);  // partial sums
    // adjoints: A_e = G^T M G, b_e = G^T v (groups of 64 threads, one edge each; G_e = s_G[e] is ready)
    for (int e0 = 0; e0 < E; e0 += kTrackThreads / 64) {
      const int grp = tid >> 6, lt = tid & 63, ee = e0 + grp;
      const bool act = ee < E;
      const double* mv = s_mvE + (act ? ee : 0) * kAcc;
      if (act && lt < 49) {
        const int p = lt / 7, c = lt % 7;
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < 7; q++) t += ((p <= q) ? mv[tri(p, q)] : mv[tri(q, p)]) * s_G[ee][q * 7 + c];
        s_TT[grp][lt] = t;
      }
      __syncthreads();
      if (act && lt < 49) {
        const int r = lt / 7, c = lt % 7;
        double s = 0.0;
#pragma unroll
        for (int p = 0; p < 7; p++) s += s_G[ee][p * 7 + r] * s_TT[grp][p * 7 + c];
        s_eb[ee * kEdgeBlock + lt] = s;
      } else if (act && lt < 56) {
        const int r = lt - 49;
        double s = 0.0;
#pragma unroll
        for (int p = 0; p < 7; p++) s += s_G[ee][p * 7 + r] * mv[28 + p];
        s_eb[ee * kEdgeBlock + lt] = s;
      }
      __syncthreads();
    }
    TRK(6);  // adjoints
    track_solve(s_g, s_eb, s_pose, N, a.ss.dim, a.ss.delta_thresh, solve_sm, s_dx, s_flags);
    TRK(7);  // solve + retraction
    if (COST) {
      // nonlinear_optimizer.py:14-18 — after the update: |old - new| / old < rel_error (old = +inf at step 0: no stop)
      if (tid == 0) {
        const double new_cost = 0.5 * s_cost[0];
        if (it > 0) {
          const double rel_dec = fabs((s_cost[1] - new_cost) / s_cost[1]);
          if (rel_dec < (double)ta.rel_error) s_flags[0] = 1;
        }
        s_cost[1] = new_cost;
      }
      __syncthreads();
    }
    nfail += s_flags[1];
    const int stop = s_flags[0];
    {  // publish first (the other CTAs are waiting), then the plain stores of the outputs
      if (tid <= N * 8) {
        const unsigned payload = tid < N * 8 ? __float_as_uint(s_pose[tid]) : (unsigned)stop;
        asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(pose_ll + tid), "l"(flag | (unsigned long long)payload)
                     : "memory");
      }
#ifdef M3R_TRACK_PROF
      if (tid == 0) {
        unsigned long long gt;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
        if (it == 4) g_dbg[601] = gt;
        g_dbg[730 + (it & 63)] = gt;  // pose of iteration `it` published
      }
#endif
      for (int t = tid; t < N * 8; t += kTrackThreads) a.ss.Twc[t] = s_pose[t];
      for (int t = tid; t < a.ss.dim; t += kTrackThreads) a.ss.dx_out[t] = s_dx[t];
      if (a.eb)
        for (int t = tid; t < E * kEdgeBlock; t += kTrackThreads) a.eb[t] = s_eb[t];
      if (tid == 0) {
        a.ss.ctrl[CTRL_ITERS] = it + 1;
        a.ss.ctrl[CTRL_FAILS] = nfail;
        if (stop) {
          a.ss.ctrl[CTRL_DONE] = 1;
          a.ss.ctrl[CTRL_EARLY] = 1;
        }
        if (a.ss.info_out) {
          a.ss.info_out[0] = it + 1;
          a.ss.info_out[1] = nfail;
          a.ss.info_out[2] = stop;
          a.ss.info_out[3] = bad_ids;
        }
      }
    }
    if (stop) break;  // the other CTAs read the same flag from the published words
    __syncthreads();   // s_flags / s_dx are rewritten by the next iteration
  }
}

template <int KIND, bool COST = false>
int launch_track(const TrackArgs& ta, size_t dyn, cudaStream_t s) {
  static DynSmemOptIn opt_in;  // per kernel instantiation
  M3R_CUDA_CHECK(opt_in.ensure(gn_track_kernel<KIND, COST>, kTrackMaxPts * kTrackPtBytes + (int)small_solve_smem_bytes(kSmallDim)));
  void* params[] = {const_cast<TrackArgs*>(&ta)};
  M3R_CUDA_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(gn_track_kernel<KIND, COST>),
                                             dim3(ta.a.e_local * ta.cpe), dim3(kTrackThreads), params, dyn, s));
  return M3R_OK;
}

template <int KIND, bool STAGED>
int launch_one(const AccumArgs& a, size_t dyn, bool pdl, cudaStream_t s) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(a.chunks, a.e_local);
  cfg.blockDim = dim3(kAccumThreads);
  cfg.dynamicSmemBytes = dyn;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  M3R_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gn_accum_kernel<KIND, STAGED>, a));
  return M3R_OK;
}

template <int KIND>
int launch_kind(const AccumArgs& a, bool staged, size_t dyn, bool pdl, cudaStream_t s) {
  if (staged) {
    static DynSmemOptIn opt_in;  // per kernel instantiation
    M3R_CUDA_CHECK(opt_in.ensure(gn_accum_kernel<KIND, true>, kMaxChunkPts * 22));
    return launch_one<KIND, true>(a, dyn, pdl, s);
  }
  return launch_one<KIND, false>(a, dyn, pdl, s);
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

int launch_gn_pack(const float* Cs, const int64_t* idx, const uint8_t* valid, const float* Q, const GnLayout& L,
                   const GnPtrs& P, const m3r_gn_params& prm, int e_lo, int e_hi, cudaStream_t s) {
  const int e_local = e_hi - e_lo;
  if (e_local <= 0) return M3R_OK;
  if (e_local > 65535 || e_local > L.edges_local) return M3R_ERR_UNSUPPORTED;
  PackArgs a{};
  a.Cs = Cs; a.idx = idx; a.valid = valid; a.Q = Q; a.ci = P.ci; a.cj = P.cj; a.packed = P.packed;
  a.n = L.num_pts; a.e_lo = e_lo; a.C_thresh = prm.C_thresh; a.Q_thresh = prm.Q_thresh;
  const bool vec4 = (L.num_pts % 4 == 0) &&
                    ((reinterpret_cast<uintptr_t>(Cs) | reinterpret_cast<uintptr_t>(idx) |
                      reinterpret_cast<uintptr_t>(Q) | reinterpret_cast<uintptr_t>(valid)) % 16 == 0);
  const GnChunking ck = gn_chunking(L.num_edges, L.num_pts);
  a.kloc = P.kloc; a.counts = P.counts; a.chunk_pts = ck.chunk_pts; a.chunks = ck.chunks;
  dim3 grid(ck.chunks, e_local);
  if (vec4)
    gn_pack_kernel<true><<<grid, 256, 0, s>>>(a);
  else
    gn_pack_kernel<false><<<grid, 256, 0, s>>>(a);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

// AccumArgs shared by the per-iteration launch and the persistent tracking launch
static int fill_accum_args(AccumArgs& a, bool& staged, size_t& dyn, const float* Twc, const float* Xs, const float* K,
                           const GnLayout& L, const GnPtrs& P, const m3r_gn_params& prm, int e_lo, int e_hi,
                           const GnFuse& fuse) {
  const int e_local = e_hi - e_lo;
  a = AccumArgs{};
  a.Twc = Twc; a.Xs = Xs; a.K = K; a.packed = P.packed;
  a.ci = P.ci; a.cj = P.cj; a.ctrl = P.ctrl; a.counters = P.counters; a.partials = P.partials; a.eb = P.eb;
  a.n = L.num_pts;
  a.e_lo = e_lo;
  a.e_local = e_local;
  // chunking depends on the FULL edge count, so a shard's blocks are bit-identical to the unsharded run's
  const GnChunking ck = gn_chunking(L.num_edges, L.num_pts);
  a.chunks = ck.chunks; a.chunk_pts = ck.chunk_pts; a.kloc = P.kloc; a.counts = P.counts;
  a.ss_counter = L.num_edges;
  a.order = (e_lo == 0 && e_hi == L.num_edges) ? P.order : nullptr;  // shards are pre-ordered by the host
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
  if (prm.kind == M3R_GN_CALIB && !K) return M3R_ERR_BAD_ARG;  // read on the device: the call stays asynchronous
  // TMA bulk copies need 16-byte aligned sources and sizes: n % 8 == 0 makes every chunk of all three streams so
  staged = (L.num_pts % 8 == 0) && (reinterpret_cast<uintptr_t>(Xs) % 16 == 0);
  dyn = staged ? (size_t)ck.chunk_pts * 22 : 0;
  a.fuse = 0;
  if (fuse.enabled && L.dim <= kSmallDim && e_lo == 0 && e_hi == L.num_edges) {
    a.fuse = 1;
    a.ss.Twc = fuse.Twc; a.ss.eb = P.eb; a.ss.ci = P.ci; a.ss.cj = P.cj; a.ss.inc_off = P.inc_off;
    a.ss.inc_list = P.inc_list; a.ss.ctrl = P.ctrl; a.ss.dx_out = fuse.dx_out; a.ss.info_out = fuse.info_out;
    a.ss.N = L.num_kf; a.ss.dim = L.dim; a.ss.delta_thresh = fuse.delta_thresh;
    const size_t need = small_solve_smem_bytes(L.dim);
    if (need > dyn) dyn = need;
  }
  return M3R_OK;
}

int launch_gn_accumulate(const float* Twc, const float* Xs, const float* K, const GnLayout& L, const GnPtrs& P,
                         const m3r_gn_params& prm, int e_lo, int e_hi, const GnFuse& fuse, cudaStream_t s) {
  const int e_local = e_hi - e_lo;
  if (e_local <= 0) return M3R_OK;
  if (e_local > 65535 || e_local > L.edges_local) return M3R_ERR_UNSUPPORTED;
  AccumArgs a;
  bool staged;
  size_t dyn;
  int rc = fill_accum_args(a, staged, dyn, Twc, Xs, K, L, P, prm, e_lo, e_hi, fuse);
  if (rc != M3R_OK) return rc;
  switch (prm.kind) {
    case M3R_GN_POINTS: return launch_kind<M3R_GN_POINTS>(a, staged, dyn, fuse.pdl != 0, s);
    case M3R_GN_RAYS: return launch_kind<M3R_GN_RAYS>(a, staged, dyn, fuse.pdl != 0, s);
    case M3R_GN_CALIB: return launch_kind<M3R_GN_CALIB>(a, staged, dyn, fuse.pdl != 0, s);
    default: return M3R_ERR_BAD_ARG;
  }
}

// Whole GN loop in one cooperative launch (gn_track_kernel).  *launched = 0 when the problem is not
// tracking-sized (the caller then runs the per-iteration path).
int launch_gn_track(const float* Twc, const float* Xs, const float* K, const GnLayout& L, const GnPtrs& P,
                    const m3r_gn_params& prm, const GnFuse& fuse, int* launched, cudaStream_t s, const GnFrameTrack* ft,
                    unsigned ll_epoch) {
  *launched = 0;
  if (ft && L.num_edges != 1) return M3R_ERR_BAD_ARG;
  const int E = L.num_edges;
  if (!fuse.enabled || E < 1 || E > kTrackMaxEdges || L.dim > kSmallDim || prm.max_iter < 1) return M3R_OK;
  int num_sms = 0, coop = 0;  // (per call: two attribute reads are cheaper than a cache that goes stale across devices)
  {
    int dev = 0;
    M3R_CUDA_CHECK(cudaGetDevice(&dev));
    M3R_CUDA_CHECK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    M3R_CUDA_CHECK(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
  if (!coop || num_sms < E) return M3R_OK;
  AccumArgs a;
  bool staged;
  size_t dyn;
  int rc = fill_accum_args(a, staged, dyn, Twc, ft ? ft->Xj : Xs, K, L, P, prm, 0, E, fuse);
  if (rc != M3R_OK) return rc;
  if (ft && (reinterpret_cast<uintptr_t>(ft->Xi) % 4 != 0)) return M3R_ERR_BAD_ARG;
  if (!staged || !a.fuse) return M3R_OK;
  TrackArgs ta{};
  if (ft) {
    ta.Xi_base = ft->Xi;
    ta.Xj_base = ft->Xj;
    ta.rel_error = ft->rel_error;
  }
  ta.a = a;
  ta.cpe = std::min(a.chunks, num_sms / E);
  ta.upc = (a.chunks + ta.cpe - 1) / ta.cpe;
  ta.stage_pts = ta.upc * a.chunk_pts;
  ta.max_iter = prm.max_iter;
  if (ta.upc > 64 || ta.stage_pts > kTrackMaxPts || prm.max_iter > 4095) return M3R_OK;
  // two gather buffers of grid x kAcc x 16 B + two pose buffers of 96 words live in P.partials
  // (gn_layout reserves >= (4 * 148 + 8) x kAcc doubles)
  // (+ two buffers of grid x 16 B of cost partials in frame-tracker mode)
  if ((size_t)4 * E * ta.cpe + 32 > std::max((size_t)E * a.chunks, (size_t)4 * kNumSMs + 32)) return M3R_OK;
  // per-launch nonce in the flag tag, drawn by the caller before gn_prepare (which clears this workspace's flag words
  // on first use and whenever the 20-bit nonce has wrapped since: gn_common.cuh)
  if (ll_epoch == 0) return M3R_ERR_BAD_ARG;
  ta.epoch = ll_epoch & 0xfffffu;
  const size_t dyn_track = (size_t)ta.stage_pts * kTrackPtBytes + small_solve_smem_bytes(L.dim);
  if (ft) {
    switch (prm.kind) {
      case M3R_GN_RAYS: rc = launch_track<M3R_GN_RAYS, true>(ta, dyn_track, s); break;
      case M3R_GN_CALIB: rc = launch_track<M3R_GN_CALIB, true>(ta, dyn_track, s); break;
      default: return M3R_ERR_BAD_ARG;
    }
  } else {
    switch (prm.kind) {
      case M3R_GN_POINTS: rc = launch_track<M3R_GN_POINTS>(ta, dyn_track, s); break;
      case M3R_GN_RAYS: rc = launch_track<M3R_GN_RAYS>(ta, dyn_track, s); break;
      case M3R_GN_CALIB: rc = launch_track<M3R_GN_CALIB>(ta, dyn_track, s); break;
      default: return M3R_ERR_BAD_ARG;
    }
  }
  if (rc == M3R_OK) *launched = 1;
  return rc;
}

int launch_export_hs_gs(const GnLayout& L, const GnPtrs& P, float* Hs, float* gs, cudaStream_t s) {
  if (L.num_edges <= 0) return M3R_OK;
  export_hs_gs_kernel<<<L.num_edges, 64, 0, s>>>(P.eb, L.num_edges, Hs, gs);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r

#ifdef M3R_TRACK_PROF
extern "C" int m3r_debug_read(void* out, size_t bytes) {
  return (int)cudaMemcpyFromSymbol(out, m3r::g_dbg, bytes);
}
#endif
