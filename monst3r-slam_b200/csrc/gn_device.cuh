// Device functions shared by the accumulate and solve kernels of the Gauss-Newton path (sm_100a):
// the Sim(3) retraction (gn_kernels.cu:299-413) and the in-CTA solve of small systems.
#pragma once
#include "gn_common.cuh"

namespace m3r {

__device__ __forceinline__ void quat_mul(const float* a, const float* b, float* o) {
  o[0] = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
  o[1] = a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0];
  o[2] = a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3];
  o[3] = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
}

__device__ __forceinline__ void rot_act(const float* q, const float* X, float* Y) {
  const float u0 = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  const float u1 = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  const float u2 = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  const float y0 = X[0] + q[3] * u0 + (q[1] * u2 - q[2] * u1);
  const float y1 = X[1] + q[3] * u1 + (q[2] * u0 - q[0] * u2);
  const float y2 = X[2] + q[3] * u2 + (q[0] * u1 - q[1] * u0);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}

__device__ __forceinline__ void cross_into(const float* a, float* b) {
  const float x0 = a[1] * b[2] - a[2] * b[1];
  const float x1 = a[2] * b[0] - a[0] * b[2];
  const float x2 = a[0] * b[1] - a[1] * b[0];
  b[0] = x0; b[1] = x1; b[2] = x2;
}

// T <- Exp(xi) * T, fp32, same closed form and the same 1e-6 branches as gn_kernels.cu:299-413
__device__ inline void sim3_retract(const float* xi, float* T) {
  float tau[3] = {xi[0], xi[1], xi[2]};
  const float phi[3] = {xi[3], xi[4], xi[5]};
  const float sigma = xi[6];
  const float scale = expf(sigma);
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float dq[4];
  {
    float imag, real;
    if (theta_sq < 1e-6f) {
      const float p4 = theta_sq * theta_sq;
      imag = 0.5f - (1.0f / 48.0f) * theta_sq + (1.0f / 3840.0f) * p4;
      real = 1.0f - (1.0f / 8.0f) * theta_sq + (1.0f / 384.0f) * p4;
    } else {
      const float th = sqrtf(theta_sq);
      imag = sinf(0.5f * th) / th;
      real = cosf(0.5f * th);
    }
    dq[0] = imag * phi[0]; dq[1] = imag * phi[1]; dq[2] = imag * phi[2]; dq[3] = real;
  }
  const float theta = sqrtf(theta_sq);
  float A, B, C;
  if (fabsf(sigma) < 1e-6f) {
    C = 1.0f;
    if (fabsf(theta) < 1e-6f) {
      A = 0.5f;
      B = 1.0f / 6.0f;
    } else {
      A = (1.0f - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - 1.0f) / sigma;
    if (fabsf(theta) < 1e-6f) {
      const float s2 = sigma * sigma;
      A = ((sigma - 1.0f) * scale + 1.0f) / s2;
      B = (scale * 0.5f * s2 + scale - 1.0f - sigma * scale) / (s2 * sigma);
    } else {
      const float a = scale * sinf(theta);
      const float b = scale * cosf(theta);
      const float c = theta_sq + sigma * sigma;
      A = (a * sigma + (1.0f - b) * theta) / (theta * c);
      B = (C - ((b - 1.0f) * sigma + a * theta) / c) / theta_sq;
    }
  }
  float dt[3] = {C * tau[0], C * tau[1], C * tau[2]};
  cross_into(phi, tau);
  dt[0] += A * tau[0]; dt[1] += A * tau[1]; dt[2] += A * tau[2];
  cross_into(phi, tau);
  dt[0] += B * tau[0]; dt[1] += B * tau[1]; dt[2] += B * tau[2];
  float q1[4], t1[3];
  quat_mul(dq, T + 3, q1);
  rot_act(dq, T, t1);
  T[0] = t1[0] * scale + dt[0];
  T[1] = t1[1] * scale + dt[1];
  T[2] = t1[2] * scale + dt[2];
  T[3] = q1[0]; T[4] = q1[1]; T[5] = q1[2]; T[6] = q1[3];
  T[7] = scale * T[7];
}

// ---------------------------------------------------------------------------------------------
// In-CTA solve for small systems (dim = 7 (N-1) <= kSmallDim): assemble from the fp64 edge blocks,
// dense Cholesky in shared memory, two triangular solves, dx = -x, retraction, stop test.
// Replaces SparseBlock::update_lhs / update_rhs / solve (gn_kernels.cu:57-159), pose_retr_kernel
// (:415-453) and the host-side norm test (:1219-1222) for tracking-sized problems.  Called by ALL
// threads of one CTA (blockDim.x >= 64).  `sm` needs small_solve_smem_bytes(dim) bytes.
struct SmallSolveArgs {
  float* Twc;
  const double* eb;
  const int* ci;
  const int* cj;
  const int* inc_off;
  const int* inc_list;
  int* ctrl;
  float* dx_out;
  int32_t* info_out;
  int N, dim;
  float delta_thresh;
};

__host__ __device__ inline size_t small_solve_smem_bytes(int dim) {
  return sizeof(double) * ((size_t)dim * (dim + 1) + 2 * (size_t)dim + 2);
}

// LOCAL = true (persistent tracking kernel): every CTA of the grid runs the same solve on its own
// shared-memory copies (a.eb and a.Twc point to shared memory); only the `writer` CTA touches the
// global outputs (dx_out, ctrl, info_out), and the stop decision is returned through *stop_out (smem).
template <bool LOCAL = false>
__device__ inline void small_solve(const SmallSolveArgs& a, double* sm, bool writer = true, int* stop_out = nullptr) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const int dim = a.dim, ld = dim + 1, nfree = a.N - 1;
  double* H = sm;                    // [dim][ld], lower triangle; H[j][j] keeps the raw pivot
  double* x = sm + (size_t)dim * ld; // rhs / solution
  double* dg = x + dim;              // sqrt of the pivots
  int* s_fail = reinterpret_cast<int*>(dg + dim);

  for (int t = tid; t < dim * ld; t += nt) H[t] = 0.0;
  if (tid == 0) *s_fail = 0;
  __syncthreads();
  // every (pose p, r, c) entry of block row p is owned by one thread: deterministic edge order
  for (int t = tid; t < nfree * 49; t += nt) {
    const int p = t / 49, rc = t % 49, r = rc / 7, c = rc % 7;
    double diag = 0.0;
    for (int q = a.inc_off[p + 1]; q < a.inc_off[p + 2]; q++) {
      const int ent = a.inc_list[q];
      const int e = ent >> 1;
      const int other = ((ent & 1) ? a.ci[e] : a.cj[e]) - 1;
      const double v = LOCAL ? a.eb[(size_t)e * kEdgeBlock + rc] : __ldcg(a.eb + (size_t)e * kEdgeBlock + rc);
      diag += v;
      if (other >= 0 && other < p) H[(7 * p + r) * ld + 7 * other + c] -= v;
    }
    H[(7 * p + r) * ld + 7 * p + c] = diag;
  }
  for (int t = tid; t < dim; t += nt) {
    const int p = t / 7, k = t % 7;
    double g = 0.0;
    for (int q = a.inc_off[p + 1]; q < a.inc_off[p + 2]; q++) {
      const int ent = a.inc_list[q];
      const double bv = LOCAL ? a.eb[(size_t)(ent >> 1) * kEdgeBlock + 49 + k]
                              : __ldcg(a.eb + (size_t)(ent >> 1) * kEdgeBlock + 49 + k);
      g += (ent & 1) ? bv : -bv;  // g_i = -b_e, g_j = +b_e
    }
    x[t] = g;
  }
  __syncthreads();
  // left-looking Cholesky, thread i owns row i
  for (int j = 0; j < dim; j++) {
    const int i = tid;
    double s = 0.0;
    if (i >= j && i < dim) {
      s = H[i * ld + j];
      for (int k = 0; k < j; k++) s -= H[i * ld + k] * H[j * ld + k];
      if (i == j) H[j * ld + j] = s;
    }
    __syncthreads();
    if (i >= j && i < dim) {
      double d = H[j * ld + j];
      if (!(d > 0.0)) {
        d = 1.0;
        if (i == j) *s_fail = 1;
      }
      const double rt = sqrt(d);
      if (i == j) dg[j] = rt; else H[i * ld + j] = s / rt;
    }
    __syncthreads();
  }
  // L y = g, then L^T x = y (warp 0; lanes own rows lane and lane + 32)
  if (tid < 32) {
    const int lane = tid;
    for (int j = 0; j < dim; j++) {
      const double yj = x[j] / dg[j];
      __syncwarp();
      if (lane == 0) x[j] = yj;
      for (int i = lane; i < dim; i += 32)
        if (i > j) x[i] -= H[i * ld + j] * yj;
      __syncwarp();
    }
    for (int j = dim - 1; j >= 0; j--) {
      const double xj = x[j] / dg[j];
      __syncwarp();
      if (lane == 0) x[j] = xj;
      for (int i = lane; i < j; i += 32) x[i] -= H[j * ld + i] * xj;
      __syncwarp();
    }
  }
  __syncthreads();
  const bool failed = *s_fail != 0;  // non-SPD => dx = 0 (gn_kernels.cu:147-150)
  if (writer)
    for (int t = tid; t < dim; t += nt) a.dx_out[t] = failed ? 0.0f : (float)(-x[t]);
  for (int k = 1 + tid; k < a.N; k += nt) {
    float T[8], xi[7];
#pragma unroll
    for (int c = 0; c < 8; c++) T[c] = a.Twc[(size_t)k * 8 + c];
#pragma unroll
    for (int c = 0; c < 7; c++) xi[c] = failed ? 0.0f : (float)(-x[(k - 1) * 7 + c]);
    sim3_retract(xi, T);
#pragma unroll
    for (int c = 0; c < 8; c++) a.Twc[(size_t)k * 8 + c] = T[c];
  }
  if (tid == 0) {
    double tot = 0.0;
    for (int t = 0; t < dim; t++) {
      const float d = failed ? 0.0f : (float)(-x[t]);
      tot += (double)d * (double)d;
    }
    const float nrm = (float)sqrt(tot);
    if (LOCAL) *stop_out = (nrm < a.delta_thresh) ? 1 : 0;
    if (!writer) return;
    a.ctrl[CTRL_ITERS] += 1;
    if (failed) a.ctrl[CTRL_FAILS] += 1;
    if (nrm < a.delta_thresh) {
      a.ctrl[CTRL_DONE] = 1;
      a.ctrl[CTRL_EARLY] = 1;
    }
    if (a.info_out) {
      a.info_out[0] = a.ctrl[CTRL_ITERS];
      a.info_out[1] = a.ctrl[CTRL_FAILS];
      a.info_out[2] = a.ctrl[CTRL_EARLY];
      a.info_out[3] = a.ctrl[CTRL_BAD];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Solve of the persistent tracking kernel: same mathematics as small_solve, laid out for latency.
// Everything it reads lives in shared memory (edge blocks, poses, the cached graph structure), the
// Cholesky / substitutions / retraction run in warp 0 with __syncwarp only (lanes own rows lane and
// lane + 32; reciprocals of the pivots' roots are kept so the substitutions multiply), and nothing is
// read from global memory.  Called by ALL threads of the CTA; ends with a __syncthreads.
struct TrackGraph {      // shared-memory copy of what gn_prepare produced (E <= 8, N <= 10)
  int ci[8], cj[8];
  int inc_off[12];
  int inc_list[16];
};

// Two keyframes (dim == 7, the frame-to-keyframe case): warp 0 alone assembles the 7x7 system, factors it, solves
// and retracts with shuffles — lane r owns row r of H / L, pivots travel by shuffle, the only shared-memory traffic is
// one transpose of L for the back-substitution.  Same operations in the same order as the general path below (except the
// back-substitution, which subtracts in descending k), but ~1/2 the latency of a single-thread version
// (~1400 dependent instructions) that used to sit on every iteration's critical path.  Call with threadIdx.x < 32.
__device__ __forceinline__ void track_solve7_warp(const TrackGraph& g, const double* eb, float* pose, float delta_thresh,
                                                  double* sm, float* dx_sm, int* flags) {
  const int lane = threadIdx.x & 31;
  const int r = lane < 7 ? lane : 0;
  double a[7] = {0, 0, 0, 0, 0, 0, 0};
  double y = 0.0;
  for (int q = g.inc_off[1]; q < g.inc_off[2]; q++) {  // every edge of a two-keyframe graph touches the free pose
    const int ent = g.inc_list[q];
    const double* b = eb + (ent >> 1) * kEdgeBlock;
#pragma unroll
    for (int c = 0; c < 7; c++) a[c] += b[r * 7 + c];
    const double bv = b[49 + r];
    y += (ent & 1) ? bv : -bv;  // g_i = -b_e, g_j = +b_e
  }
  bool failed = false;
  double my_rinv = 1.0;
#pragma unroll
  for (int j = 0; j < 7; j++) {
    double d = __shfl_sync(0xffffffffu, a[j], j);
    if (!(d > 0.0)) {  // non-SPD => dx = 0 (gn_kernels.cu:147-150)
      d = 1.0;
      failed = true;
    }
    const double rinv = rsqrt(d);
    const double l = a[j] * rinv;  // lane j: sqrt(d); lanes > j: L[lane][j]
    a[j] = l;
    if (lane == j) my_rinv = rinv;
#pragma unroll
    for (int c = j + 1; c < 7; c++) a[c] -= l * __shfl_sync(0xffffffffu, l, c);
  }
  // L y = g
#pragma unroll
  for (int j = 0; j < 7; j++) {
    const double yj = __shfl_sync(0xffffffffu, y, j) * __shfl_sync(0xffffffffu, my_rinv, j);
    if (lane == j) y = yj;
    if (lane > j) y -= a[j] * yj;
  }
  // L^T x = y: lane r needs column r of L, i.e. a[r] of the lanes below it -> one transpose through shared memory
  if (lane < 7) {
#pragma unroll
    for (int c = 0; c < 7; c++) sm[lane * 8 + c] = a[c];
  }
  __syncwarp();
  double col[7];
#pragma unroll
  for (int j = 0; j < 7; j++) col[j] = sm[j * 8 + r];  // L[j][r]
#pragma unroll
  for (int j = 6; j >= 0; j--) {
    const double xj = __shfl_sync(0xffffffffu, y, j) * __shfl_sync(0xffffffffu, my_rinv, j);
    if (lane == j) y = xj;
    if (lane < j) y -= col[j] * xj;
  }
  const float mine = failed ? 0.0f : (float)(-y);
  float xi[7];
  double tot = 0.0;
#pragma unroll
  for (int c = 0; c < 7; c++) {
    xi[c] = __shfl_sync(0xffffffffu, mine, c);
    tot += (double)xi[c] * (double)xi[c];
  }
  if (lane == 0) {
    float T[8];
#pragma unroll
    for (int c = 0; c < 7; c++) dx_sm[c] = xi[c];
#pragma unroll
    for (int c = 0; c < 8; c++) T[c] = pose[8 + c];
    sim3_retract(xi, T);
#pragma unroll
    for (int c = 0; c < 8; c++) pose[8 + c] = T[c];
    flags[0] = ((float)sqrt(tot) < delta_thresh) ? 1 : 0;
    flags[1] = failed ? 1 : 0;
  }
}

// returns (in *flags, shared): flags[0] = stop (|dx| < delta_thresh), flags[1] = factorisation failed
__device__ inline void track_solve(const TrackGraph& g, const double* eb, float* pose, int N, int dim,
                                   float delta_thresh, double* sm, float* dx_sm, int* flags) {
  const int tid = threadIdx.x, nt = blockDim.x;
  if (dim == 7) {
    if (tid < 32) track_solve7_warp(g, eb, pose, delta_thresh, sm, dx_sm, flags);
    __syncthreads();
    return;
  }
  const int ld = dim + 1, nfree = N - 1;
  double* H = sm;                     // [dim][ld] lower triangle
  double* x = sm + (size_t)dim * ld;  // rhs / solution
  double* dinv = x + dim;             // 1 / sqrt(pivot)
  // assemble: every (pose p, r, c) entry of the block lower triangle is owned by one thread
  for (int t = tid; t < nfree * nfree * 49; t += nt) {
    const int blk = t / 49, rc = t % 49, r = rc / 7, c = rc % 7;
    const int p = blk / nfree, o = blk % nfree;
    if (o > p) continue;
    double v = 0.0;
    for (int q = g.inc_off[p + 1]; q < g.inc_off[p + 2]; q++) {
      const int ent = g.inc_list[q];
      const int e = ent >> 1;
      const int other = ((ent & 1) ? g.ci[e] : g.cj[e]) - 1;
      const double b = eb[e * kEdgeBlock + rc];
      if (o == p) v += b;
      else if (other == o) v -= b;
    }
    H[(7 * p + r) * ld + 7 * o + c] = v;
  }
  for (int t = tid; t < dim; t += nt) {
    const int p = t / 7, k = t % 7;
    double gsum = 0.0;
    for (int q = g.inc_off[p + 1]; q < g.inc_off[p + 2]; q++) {
      const int ent = g.inc_list[q];
      const double bv = eb[(ent >> 1) * kEdgeBlock + 49 + k];
      gsum += (ent & 1) ? bv : -bv;  // g_i = -b_e, g_j = +b_e
    }
    x[t] = gsum;
  }
  __syncthreads();
  if (tid < 32) {
    const int lane = tid;
    const int r0 = lane, r1 = lane + 32;
    bool failed = false;
    for (int j = 0; j < dim; j++) {
      double s0 = 0.0, s1 = 0.0;
      if (r0 >= j && r0 < dim) {
        s0 = H[r0 * ld + j];
        for (int k = 0; k < j; k++) s0 -= H[r0 * ld + k] * H[j * ld + k];
      }
      if (r1 >= j && r1 < dim) {
        s1 = H[r1 * ld + j];
        for (int k = 0; k < j; k++) s1 -= H[r1 * ld + k] * H[j * ld + k];
      }
      double d = __shfl_sync(0xffffffffu, j < 32 ? s0 : s1, j & 31);
      if (!(d > 0.0)) {  // non-SPD => dx = 0 (gn_kernels.cu:147-150)
        d = 1.0;
        failed = true;
      }
      const double inv = 1.0 / sqrt(d);
      if (r0 > j && r0 < dim) H[r0 * ld + j] = s0 * inv;
      if (r1 > j && r1 < dim) H[r1 * ld + j] = s1 * inv;
      if (lane == 0) dinv[j] = inv;
      __syncwarp();
    }
    // L y = g, then L^T x = y
    for (int j = 0; j < dim; j++) {
      const double yj = x[j] * dinv[j];
      __syncwarp();
      if (lane == 0) x[j] = yj;
      if (r0 > j && r0 < dim) x[r0] -= H[r0 * ld + j] * yj;
      if (r1 > j && r1 < dim) x[r1] -= H[r1 * ld + j] * yj;
      __syncwarp();
    }
    for (int j = dim - 1; j >= 0; j--) {
      const double xj = x[j] * dinv[j];
      __syncwarp();
      if (lane == 0) x[j] = xj;
      if (r0 < j) x[r0] -= H[j * ld + r0] * xj;
      if (r1 < j) x[r1] -= H[j * ld + r1] * xj;
      __syncwarp();
    }
    float d0 = 0.f, d1 = 0.f;
    if (r0 < dim) d0 = failed ? 0.0f : (float)(-x[r0]);
    if (r1 < dim) d1 = failed ? 0.0f : (float)(-x[r1]);
    if (r0 < dim) dx_sm[r0] = d0;
    if (r1 < dim) dx_sm[r1] = d1;
    double tot = (double)d0 * (double)d0 + (double)d1 * (double)d1;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, off);
    __syncwarp();
    if (lane >= 1 && lane < N) {
      float T[8], xi[7];
#pragma unroll
      for (int c = 0; c < 8; c++) T[c] = pose[lane * 8 + c];
#pragma unroll
      for (int c = 0; c < 7; c++) xi[c] = dx_sm[(lane - 1) * 7 + c];
      sim3_retract(xi, T);
#pragma unroll
      for (int c = 0; c < 8; c++) pose[lane * 8 + c] = T[c];
    }
    if (lane == 0) {
      flags[0] = ((float)sqrt(tot) < delta_thresh) ? 1 : 0;
      flags[1] = failed ? 1 : 0;
    }
  }
  __syncthreads();
}

}  // namespace m3r
