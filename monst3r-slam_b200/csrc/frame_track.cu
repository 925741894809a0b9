// Frame tracker pose optimisation (sm_100a): SURVEY.md §8(f) rank 3.
//
// Replaces FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 + solve
// (mast3r_slam/tracker.py:141-154 solve, :156-214 rays, :216-270 calib) — a torch loop of up to 50 iterations
// that materialises J [n,4,7] (22 MB), builds A^T A with a GEMM, factorises on the device and synchronises with
// `.item()` every iteration — by ONE cooperative launch of the persistent Gauss-Newton kernel in its frame-tracker
// mode (gn_accum.cu: gn_track_kernel<KIND, COST=true>): points staged once in shared memory, per iteration a
// fused residual / Huber / normal-equation pass, an LL all-gather of 36 fp64 partials, a register 7x7 Cholesky
// (fp64) and the Sim(3) retraction, with both stop tests of nonlinear_optimizer.py:5-26 evaluated on the device.
//
// Why the GN kernel fits: the tracker minimises over the relative pose T_CkCf with the keyframe fixed, i.e. a graph
// of two keyframes {k: identity (fixed), f: T_CkCf} and ONE directed edge whose moving points are the frame points
// already gathered by idx_f2k (tracker.py:139 `Xf[idx_f2k]`) and whose targets are the keyframe points at the same
// index.  Residuals, sqrt-information (valid * sqrt(Q) / sigma), Huber weight (k = 1.345) and the left-perturbation
// Jacobians are the ones of gauss_newton_rays / gauss_newton_calib with T_i = identity.
#include <cfloat>
#include <cmath>

#include "gn_common.cuh"

namespace m3r {
namespace {

struct Sim3d {
  double t[3], q[4], s;  // q = (x, y, z, w), lietorch layout [t, q, s]
};

__device__ Sim3d load_sim3(const float* p) {
  Sim3d T;
  for (int k = 0; k < 3; k++) T.t[k] = p[k];
  for (int k = 0; k < 4; k++) T.q[k] = p[3 + k];
  // Unit quaternion restored on entry: inv() below is the conjugate, so a norm 1 + e in the keyframe pose would come
  // back as (1 + e)^2 in every tracked pose and compound frame after frame (seen as an exponential drift of |q| over a
  // 60-frame sequence with 8 keyframes, bench.py --sequence).  A no-op to within one fp32 ulp for unit input.
  const double nq = 1.0 / sqrt(T.q[0] * T.q[0] + T.q[1] * T.q[1] + T.q[2] * T.q[2] + T.q[3] * T.q[3]);
  for (int k = 0; k < 4; k++) T.q[k] *= nq;
  T.s = p[7];
  return T;
}
__device__ void store_sim3(float* p, const Sim3d& T) {
  for (int k = 0; k < 3; k++) p[k] = (float)T.t[k];
  for (int k = 0; k < 4; k++) p[3 + k] = (float)T.q[k];
  p[7] = (float)T.s;
}
__device__ void rotate(const double* q, const double* v, double* out) {
  const double ux = q[0], uy = q[1], uz = q[2], w = q[3];
  const double cx = uy * v[2] - uz * v[1], cy = uz * v[0] - ux * v[2], cz = ux * v[1] - uy * v[0];  // u x v
  out[0] = v[0] + 2.0 * (w * cx + (uy * cz - uz * cy));
  out[1] = v[1] + 2.0 * (w * cy + (uz * cx - ux * cz));
  out[2] = v[2] + 2.0 * (w * cz + (ux * cy - uy * cx));
}
__device__ Sim3d sim3_inverse(const Sim3d& T) {
  Sim3d I;
  I.q[0] = -T.q[0]; I.q[1] = -T.q[1]; I.q[2] = -T.q[2]; I.q[3] = T.q[3];
  I.s = 1.0 / T.s;
  double r[3];
  rotate(I.q, T.t, r);
  for (int k = 0; k < 3; k++) I.t[k] = -I.s * r[k];
  return I;
}
__device__ Sim3d sim3_compose(const Sim3d& A, const Sim3d& B) {  // A * B
  Sim3d C;
  double r[3];
  rotate(A.q, B.t, r);
  for (int k = 0; k < 3; k++) C.t[k] = A.s * r[k] + A.t[k];
  const double x1 = A.q[0], y1 = A.q[1], z1 = A.q[2], w1 = A.q[3];
  const double x2 = B.q[0], y2 = B.q[1], z2 = B.q[2], w2 = B.q[3];
  C.q[0] = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2;
  C.q[1] = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2;
  C.q[2] = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2;
  C.q[3] = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2;
  C.s = A.s * B.s;
  return C;
}

// tracker.py:163 / :229: T_CkCf = T_WCk.inv() * T_WCf; poses of the 2-keyframe graph = [identity, T_CkCf]
__global__ void track_init_kernel(const float* __restrict__ T_WCk, const float* __restrict__ T_WCf,
                                  float* __restrict__ Twc2, int64_t* __restrict__ ii, int64_t* __restrict__ jj) {
  if (threadIdx.x != 0) return;
  const Sim3d rel = sim3_compose(sim3_inverse(load_sim3(T_WCk)), load_sim3(T_WCf));
  const float ident[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 1.f, 1.f};
  for (int k = 0; k < 8; k++) Twc2[k] = ident[k];
  store_sim3(Twc2 + 8, rel);
  ii[0] = 0;
  jj[0] = 1;
}

// tracker.py:192 / :267: T_WCf = T_WCk * T_CkCf
__global__ void track_finish_kernel(const float* __restrict__ T_WCk, const float* __restrict__ Twc2,
                                    float* __restrict__ T_WCf, float* __restrict__ T_CkCf) {
  if (threadIdx.x != 0) return;
  const Sim3d rel = load_sim3(Twc2 + 8);
  if (T_CkCf) store_sim3(T_CkCf, rel);
  store_sim3(T_WCf, sim3_compose(load_sim3(T_WCk), rel));
}

constexpr size_t kTrackExtraBytes = 256;  // Twc2 [2,8] f32 | ii | jj

}  // namespace
}  // namespace m3r

using namespace m3r;

extern "C" size_t m3r_track_workspace_bytes(int num_pts) {
  if (num_pts < 1) return 0;
  return gn_layout(2, 1, num_pts).off_total + kTrackExtraBytes;
}

extern "C" int m3r_track_frame(float* T_WCf, float* T_CkCf_out, const float* T_WCk, const float* Xf, const float* Xk,
                               const uint8_t* valid, const float* Q, const float* K, int num_pts,
                               const m3r_track_params* p, float* dx_out, int32_t* info_out, void* workspace,
                               size_t workspace_bytes, void* stream) {
  if (!T_WCf || !T_WCk || !Xf || !Xk || !valid || !Q || !p || !dx_out || !info_out) return M3R_ERR_BAD_ARG;
  if (num_pts < 1 || p->max_iter < 1) return M3R_ERR_BAD_ARG;
  if (p->kind != M3R_GN_RAYS && p->kind != M3R_GN_CALIB) return M3R_ERR_BAD_ARG;
  if (p->kind == M3R_GN_CALIB && (!K || p->height * p->width != num_pts)) return M3R_ERR_BAD_ARG;
  if (std::fabs(p->huber - 1.345f) > 1e-6f) {
    set_last_error_text("m3r_track_frame: only the Huber threshold 1.345 (config/base.yaml tracking.huber) is built");
    return M3R_ERR_UNSUPPORTED;
  }
  if (!workspace || (reinterpret_cast<uintptr_t>(workspace) & 255)) return M3R_ERR_WORKSPACE;
  const GnLayout L = gn_layout(2, 1, num_pts);
  if (workspace_bytes < L.off_total + kTrackExtraBytes) return M3R_ERR_WORKSPACE;
  const GnPtrs P = gn_ptrs(L, workspace);
  char* extra = static_cast<char*>(workspace) + L.off_total;
  float* Twc2 = reinterpret_cast<float*>(extra);
  int64_t* ii = reinterpret_cast<int64_t*>(extra + 64);
  int64_t* jj = reinterpret_cast<int64_t*>(extra + 72);
  cudaStream_t s = static_cast<cudaStream_t>(stream);

  m3r_gn_params prm{};
  prm.kind = p->kind;
  prm.sigma_a = p->sigma_a;
  prm.sigma_b = p->sigma_b;
  prm.C_thresh = -FLT_MAX;  // every confidence test is already in `valid` (tracker.py:62-67)
  prm.Q_thresh = -FLT_MAX;
  prm.max_iter = p->max_iter;
  prm.delta_thresh = p->delta_norm;
  prm.height = p->height;
  prm.width = p->width;
  prm.pixel_border = p->pixel_border;
  prm.z_eps = p->z_eps;

  track_init_kernel<<<1, 32, 0, s>>>(T_WCk, T_WCf, Twc2, ii, jj);
  M3R_LAUNCH_CHECK();
  const unsigned ll_epoch = gn_next_epoch();
  int rc = launch_gn_prepare(ii, jj, L, P, s, dx_out, kPose, info_out, ll_epoch);  // also zeroes dx_out / info_out
  if (rc != M3R_OK) return rc;
  rc = launch_gn_pack(nullptr, nullptr, valid, Q, L, P, prm, 0, 1, s);
  if (rc != M3R_OK) return rc;
  GnFuse fuse{};
  fuse.Twc = Twc2;
  fuse.dx_out = dx_out;
  fuse.info_out = info_out;
  fuse.delta_thresh = p->delta_norm;
  fuse.enabled = 1;
  GnFrameTrack ft{Xk, Xf, p->rel_error};
  int launched = 0;
  rc = launch_gn_track(Twc2, nullptr, K, L, P, prm, fuse, &launched, s, &ft, ll_epoch);
  if (rc != M3R_OK) return rc;
  if (!launched) {
    set_last_error_text("m3r_track_frame: the persistent kernel cannot hold this frame (num_pts % 8 != 0, Xf not "
                        "16-byte aligned, more than 148 x 4800 points, or no cooperative launch)");
    return M3R_ERR_UNSUPPORTED;
  }
  track_finish_kernel<<<1, 32, 0, s>>>(T_WCk, Twc2, T_WCf, T_CkCf_out);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}
