// Python module `mast3r_slam_backends` — drop-in for the reference's extension of the same name.
//
// Mirrors mast3r_slam/backend/src/gn.cpp:3-123 (the five pybind functions, their positional
// signatures, the "must be contiguous" RuntimeErrors, in-place Twc update, list return values) and
// forwards to the C ABI of libm3r_b200.so (include/m3r_abi.h).  torch is used only for what the
// reference uses it for at this boundary: argument tensors, output allocation, current stream.
// Supersets of the reference (SURVEY §8g "S"): device guard, current (not legacy) stream, dtype /
// device / shape validation.  There is no CPU path: CPU tensors raise.
//
// Additive API (multi-GPU): nccl_unique_id(), init_distributed(rank, world, id),
// set_edge_shard(lo, hi), clear_edge_shard(), set_sync_checks(bool), last_info().
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include <string>
#include <vector>

#include "../../include/m3r_abi.h"

#define CHECK_CONTIGUOUS(x) TORCH_CHECK(x.is_contiguous(), #x " must be contiguous")
#define CHECK_CUDA(x) TORCH_CHECK(x.is_cuda(), #x " must be a CUDA tensor (this backend has no CPU path)")
#define CHECK_DTYPE(x, dt) TORCH_CHECK(x.scalar_type() == dt, #x " has the wrong dtype")

namespace {

struct DistState {
  void* comm = nullptr;
  bool sharded = false;
  int lo = 0, hi = 0;
  int exchange = M3R_EXCHANGE_ALLREDUCE;  // same value on every rank (m3r_abi.h: m3r_gn_shard)
  bool sync_checks = true;
  std::vector<int64_t> last_info{0, 0, 0, 0};
} g_dist;

void check_rc(int rc, const char* what) {
  if (rc == M3R_OK) return;
  TORCH_CHECK(false, what, " failed: code ", rc, " (", m3r_last_error(), ")");
}

void* cur_stream() { return (void*)at::cuda::getCurrentCUDAStream().stream(); }

std::vector<torch::Tensor> gn_common(int kind, torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs,
                                     const torch::Tensor* K, torch::Tensor ii, torch::Tensor jj,
                                     torch::Tensor idx_ii2jj, torch::Tensor valid_match, torch::Tensor Q,
                                     m3r_gn_params prm) {
  CHECK_CONTIGUOUS(Twc);
  CHECK_CONTIGUOUS(Xs);
  CHECK_CONTIGUOUS(Cs);
  CHECK_CONTIGUOUS(ii);
  CHECK_CONTIGUOUS(jj);
  CHECK_CONTIGUOUS(idx_ii2jj);
  CHECK_CONTIGUOUS(valid_match);
  CHECK_CONTIGUOUS(Q);
  CHECK_CUDA(Twc);
  CHECK_CUDA(Xs);
  CHECK_CUDA(Cs);
  CHECK_CUDA(ii);
  CHECK_CUDA(jj);
  CHECK_CUDA(idx_ii2jj);
  CHECK_CUDA(valid_match);
  CHECK_CUDA(Q);
  CHECK_DTYPE(Twc, torch::kFloat32);
  CHECK_DTYPE(Xs, torch::kFloat32);
  CHECK_DTYPE(Cs, torch::kFloat32);
  CHECK_DTYPE(ii, torch::kInt64);
  CHECK_DTYPE(jj, torch::kInt64);
  CHECK_DTYPE(idx_ii2jj, torch::kInt64);
  CHECK_DTYPE(valid_match, torch::kBool);
  CHECK_DTYPE(Q, torch::kFloat32);
  const float* Kp = nullptr;
  if (K) {
    CHECK_CONTIGUOUS((*K));
    CHECK_CUDA((*K));
    CHECK_DTYPE((*K), torch::kFloat32);
    TORCH_CHECK(K->numel() == 9, "K must be 3x3");
    Kp = K->data_ptr<float>();
  }
  TORCH_CHECK(Twc.dim() == 2 && Twc.size(1) == 8, "Twc must be [N,8]");
  TORCH_CHECK(Xs.dim() == 3 && Xs.size(2) == 3, "Xs must be [N,n,3]");
  const int num_kf = (int)Xs.size(0);
  const int n = (int)Xs.size(1);
  const int E = (int)ii.size(0);
  TORCH_CHECK(Twc.size(0) == num_kf, "Twc and Xs disagree on the number of keyframes");
  TORCH_CHECK(Cs.numel() == (int64_t)num_kf * n, "Cs must be [N,n,1]");
  TORCH_CHECK(jj.size(0) == E, "ii and jj must have the same length");
  int lo = 0, hi = E;
  if (g_dist.sharded) {
    lo = g_dist.lo;
    hi = g_dist.hi;
    TORCH_CHECK(lo >= 0 && hi <= E && lo <= hi, "edge shard out of range");
  }
  const int64_t e_local = hi - lo;
  TORCH_CHECK(idx_ii2jj.numel() == e_local * n, "idx_ii2jj must be [E,n] (local edges when sharded)");
  TORCH_CHECK(valid_match.numel() == e_local * n, "valid_match must be [E,n,1]");
  TORCH_CHECK(Q.numel() == e_local * n, "Q must be [E,n,1]");

  const c10::cuda::CUDAGuard guard(Twc.device());
  auto opts = Twc.options();
  torch::Tensor dx = torch::zeros({std::max(num_kf - 1, 0), 7}, opts);
  if (num_kf < 2 || E < 1 || n < 1) return {dx};
  // a rank keeps packed records for its own edges only (8 + 2 B per point per edge: GBs at 2000 edges)
  const size_t ws_bytes = m3r_gn_workspace_bytes_sharded(num_kf, E, n, (int)e_local);
  torch::Tensor ws = torch::empty({(int64_t)ws_bytes}, opts.dtype(torch::kUInt8));
  torch::Tensor info = torch::zeros({4}, opts.dtype(torch::kInt32));
  m3r_gn_shard shard{lo, hi, g_dist.comm, g_dist.exchange};
  // no shard set -> plain single-GPU call even if a communicator exists (every rank would hold ALL edges: an
  // exchange would count each block `world` times)
  const bool use_shard = g_dist.sharded;
  int rc = m3r_gauss_newton(Twc.data_ptr<float>(), Xs.data_ptr<float>(), Cs.data_ptr<float>(), Kp,
                            ii.data_ptr<int64_t>(), jj.data_ptr<int64_t>(), idx_ii2jj.data_ptr<int64_t>(),
                            reinterpret_cast<const uint8_t*>(valid_match.data_ptr<bool>()),
                            Q.data_ptr<float>(), num_kf, n, E, &prm, dx.data_ptr<float>(),
                            info.data_ptr<int32_t>(), ws.data_ptr(), ws_bytes, use_shard ? &shard : nullptr,
                            cur_stream());
  check_rc(rc, "m3r_gauss_newton");
  if (g_dist.sync_checks) {
    // the reference blocks the host twice per iteration; one read-back per call keeps bad keyframe
    // ids from passing silently (set_sync_checks(False) makes the op fully asynchronous)
    auto h = info.cpu();
    auto a = h.accessor<int32_t, 1>();
    for (int k = 0; k < 4; k++) g_dist.last_info[k] = a[k];
    TORCH_CHECK(a[3] == 0, "gauss_newton: keyframe ids in ii/jj must lie in [0, ", (int)M3R_MAX_KF_ID,
                ") and have exactly Xs.size(0) distinct values (flag ", a[3], ")");
  }
  return {dx};
}

}  // namespace

std::vector<torch::Tensor> gauss_newton_points(torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs,
                                               torch::Tensor ii, torch::Tensor jj, torch::Tensor idx_ii2jj,
                                               torch::Tensor valid_match, torch::Tensor Q,
                                               const float sigma_point, const float C_thresh,
                                               const float Q_thresh, const int max_iter,
                                               const float delta_thresh) {
  m3r_gn_params p{};
  p.kind = M3R_GN_POINTS;
  p.sigma_a = sigma_point;
  p.sigma_b = 1.f;
  p.C_thresh = C_thresh;
  p.Q_thresh = Q_thresh;
  p.max_iter = max_iter;
  p.delta_thresh = delta_thresh;
  return gn_common(M3R_GN_POINTS, Twc, Xs, Cs, nullptr, ii, jj, idx_ii2jj, valid_match, Q, p);
}

std::vector<torch::Tensor> gauss_newton_rays(torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs,
                                             torch::Tensor ii, torch::Tensor jj, torch::Tensor idx_ii2jj,
                                             torch::Tensor valid_match, torch::Tensor Q, const float sigma_ray,
                                             const float sigma_dist, const float C_thresh,
                                             const float Q_thresh, const int max_iter,
                                             const float delta_thresh) {
  m3r_gn_params p{};
  p.kind = M3R_GN_RAYS;
  p.sigma_a = sigma_ray;
  p.sigma_b = sigma_dist;
  p.C_thresh = C_thresh;
  p.Q_thresh = Q_thresh;
  p.max_iter = max_iter;
  p.delta_thresh = delta_thresh;
  return gn_common(M3R_GN_RAYS, Twc, Xs, Cs, nullptr, ii, jj, idx_ii2jj, valid_match, Q, p);
}

std::vector<torch::Tensor> gauss_newton_calib(torch::Tensor Twc, torch::Tensor Xs, torch::Tensor Cs,
                                              torch::Tensor K, torch::Tensor ii, torch::Tensor jj,
                                              torch::Tensor idx_ii2jj, torch::Tensor valid_match,
                                              torch::Tensor Q, const int height, const int width,
                                              const int pixel_border, const float z_eps,
                                              const float sigma_pixel, const float sigma_depth,
                                              const float C_thresh, const float Q_thresh, const int max_iter,
                                              const float delta_thresh) {
  m3r_gn_params p{};
  p.kind = M3R_GN_CALIB;
  p.sigma_a = sigma_pixel;
  p.sigma_b = sigma_depth;
  p.C_thresh = C_thresh;
  p.Q_thresh = Q_thresh;
  p.max_iter = max_iter;
  p.delta_thresh = delta_thresh;
  p.height = height;
  p.width = width;
  p.pixel_border = pixel_border;
  p.z_eps = z_eps;
  return gn_common(M3R_GN_CALIB, Twc, Xs, Cs, &K, ii, jj, idx_ii2jj, valid_match, Q, p);
}

std::vector<torch::Tensor> iter_proj(torch::Tensor rays_img_with_grad, torch::Tensor pts_3d_norm,
                                     torch::Tensor p_init, const int max_iter, const float lambda_init,
                                     const float cost_thresh) {
  CHECK_CONTIGUOUS(rays_img_with_grad);
  CHECK_CONTIGUOUS(pts_3d_norm);
  CHECK_CONTIGUOUS(p_init);
  CHECK_CUDA(rays_img_with_grad);
  CHECK_CUDA(pts_3d_norm);
  CHECK_CUDA(p_init);
  CHECK_DTYPE(rays_img_with_grad, torch::kFloat32);
  CHECK_DTYPE(pts_3d_norm, torch::kFloat32);
  CHECK_DTYPE(p_init, torch::kFloat32);
  TORCH_CHECK(rays_img_with_grad.dim() == 4 && rays_img_with_grad.size(3) == 9,
              "rays_img_with_grad must be [b,h,w,9]");
  TORCH_CHECK(p_init.dim() == 3 && p_init.size(2) == 2, "p_init must be [b,n,2]");
  const int b = (int)p_init.size(0), n = (int)p_init.size(1);
  const int h = (int)rays_img_with_grad.size(1), w = (int)rays_img_with_grad.size(2);
  TORCH_CHECK(rays_img_with_grad.size(0) == b, "batch mismatch");
  TORCH_CHECK(pts_3d_norm.numel() == (int64_t)b * n * 3, "pts_3d_norm must be [b,n,3]");
  const c10::cuda::CUDAGuard guard(p_init.device());
  auto opts = p_init.options();
  torch::Tensor p_new = torch::empty({b, n, 2}, opts);
  torch::Tensor converged = torch::empty({b, n}, opts.dtype(torch::kBool));
  // scratch for the texel-plane re-pack (torch's caching allocator: no cudaMalloc in steady state)
  const size_t ws_bytes = m3r_iter_proj_workspace_bytes(b, h, w);
  torch::Tensor ws = torch::empty({(int64_t)std::max<size_t>(ws_bytes, 16)}, opts.dtype(torch::kUInt8));
  int rc = m3r_iter_proj_ws(rays_img_with_grad.data_ptr<float>(), pts_3d_norm.data_ptr<float>(),
                            p_init.data_ptr<float>(), p_new.data_ptr<float>(),
                            reinterpret_cast<uint8_t*>(converged.data_ptr<bool>()), b, h, w, n, max_iter,
                            lambda_init, cost_thresh, ws.data_ptr(), ws_bytes, cur_stream());
  check_rc(rc, "m3r_iter_proj_ws");
  return {p_new, converged};
}

std::vector<torch::Tensor> refine_matches(torch::Tensor D11, torch::Tensor D21, torch::Tensor p1,
                                          const int radius, const int dilation_max) {
  CHECK_CONTIGUOUS(D11);
  CHECK_CONTIGUOUS(D21);
  CHECK_CONTIGUOUS(p1);
  CHECK_CUDA(D11);
  CHECK_CUDA(D21);
  CHECK_CUDA(p1);
  CHECK_DTYPE(p1, torch::kInt64);
  TORCH_CHECK(D11.dim() == 4, "D11 must be [b,h,w,f]");
  TORCH_CHECK(p1.dim() == 3 && p1.size(2) == 2, "p1 must be [b,n,2]");
  TORCH_CHECK(D21.scalar_type() == D11.scalar_type(), "D11 and D21 must share a dtype");
  const int b = (int)p1.size(0), n = (int)p1.size(1);
  const int h = (int)D11.size(1), w = (int)D11.size(2), fdim = (int)D11.size(3);
  TORCH_CHECK(D11.size(0) == b, "batch mismatch");
  TORCH_CHECK(D21.numel() == (int64_t)b * n * fdim, "D21 must be [b,n,f]");
  const c10::cuda::CUDAGuard guard(p1.device());
  torch::Tensor p1_new = torch::empty({b, n, 2}, p1.options());
  int rc;
  switch (D11.scalar_type()) {
    case torch::kFloat16:
      rc = m3r_refine_matches_f16(D11.data_ptr(), D21.data_ptr(), p1.data_ptr<int64_t>(),
                                  p1_new.data_ptr<int64_t>(), b, h, w, n, fdim, radius, dilation_max,
                                  cur_stream());
      break;
    case torch::kFloat32:
      rc = m3r_refine_matches_f32(D11.data_ptr<float>(), D21.data_ptr<float>(), p1.data_ptr<int64_t>(),
                                  p1_new.data_ptr<int64_t>(), b, h, w, n, fdim, radius, dilation_max,
                                  cur_stream());
      break;
    case torch::kFloat64:
      rc = m3r_refine_matches_f64(D11.data_ptr<double>(), D21.data_ptr<double>(), p1.data_ptr<int64_t>(),
                                  p1_new.data_ptr<int64_t>(), b, h, w, n, fdim, radius, dilation_max,
                                  cur_stream());
      break;
    default:
      TORCH_CHECK(false, "refine_matches: D11 must be float16, float32 or float64");
  }
  check_rc(rc, "m3r_refine_matches");
  return {p1_new};
}

// ------------------------------------------------------------------------------- additive API
py::bytes nccl_unique_id() {
  char id[M3R_NCCL_UNIQUE_ID_BYTES];
  check_rc(m3r_comm_unique_id(id), "m3r_comm_unique_id");
  return py::bytes(id, M3R_NCCL_UNIQUE_ID_BYTES);
}

void init_distributed(int rank, int world, py::bytes id) {
  std::string s = id;
  TORCH_CHECK((int)s.size() == M3R_NCCL_UNIQUE_ID_BYTES, "unique id must be 128 bytes");
  if (g_dist.comm) {
    m3r_comm_destroy(g_dist.comm);
    g_dist.comm = nullptr;
  }
  check_rc(m3r_comm_init(rank, world, s.data(), &g_dist.comm), "m3r_comm_init");
}

void shutdown_distributed() {
  if (g_dist.comm) m3r_comm_destroy(g_dist.comm);
  g_dist.comm = nullptr;
  g_dist.sharded = false;
}

// exchange: "allreduce" (any contiguous partition) or "allgather" (equal partition only) — the SAME on every rank;
// derive it from the global partition (dist.exchange_for_partition), never from the rank's own range.
void set_edge_shard(int lo, int hi, const std::string& exchange) {
  TORCH_CHECK(exchange == "allreduce" || exchange == "allgather", "exchange must be 'allreduce' or 'allgather'");
  g_dist.sharded = true;
  g_dist.lo = lo;
  g_dist.hi = hi;
  g_dist.exchange = exchange == "allgather" ? M3R_EXCHANGE_ALLGATHER : M3R_EXCHANGE_ALLREDUCE;
}
void clear_edge_shard() { g_dist.sharded = false; }
void set_sync_checks(bool on) { g_dist.sync_checks = on; }
std::vector<int64_t> last_info() { return g_dist.last_info; }


// Additive op (not in the reference module): matching.match() fused into five launches (m3r_match_fused).
// Returns [idx_1_to_2 i64 [b,n], valid_match2 bool [b,n,1]] like mast3r_slam/matching.py:8-10.
std::vector<torch::Tensor> match_fused(torch::Tensor X11, torch::Tensor X21, torch::Tensor D11, torch::Tensor D21,
                                       c10::optional<torch::Tensor> idx_1_to_2_init, const int max_iter,
                                       const float lambda_init, const float convergence_thresh,
                                       const float dist_thresh, const int radius, const int dilation_max) {
  CHECK_CONTIGUOUS(X11);
  CHECK_CONTIGUOUS(X21);
  CHECK_CONTIGUOUS(D11);
  CHECK_CONTIGUOUS(D21);
  CHECK_CUDA(X11);
  CHECK_CUDA(X21);
  CHECK_CUDA(D11);
  CHECK_CUDA(D21);
  CHECK_DTYPE(X11, torch::kFloat32);
  CHECK_DTYPE(X21, torch::kFloat32);
  TORCH_CHECK(X11.dim() == 4 && X11.size(3) == 3 && X21.sizes() == X11.sizes(), "X11 / X21 must be [b,h,w,3]");
  TORCH_CHECK(D11.scalar_type() == D21.scalar_type() &&
                  (D11.scalar_type() == torch::kFloat32 || D11.scalar_type() == torch::kFloat16),
              "D11 / D21 must both be float32 or float16");
  const int b = (int)X11.size(0), h = (int)X11.size(1), w = (int)X11.size(2);
  const int fdim = (int)D11.size(-1);
  TORCH_CHECK(D11.numel() == (int64_t)b * h * w * fdim && D21.numel() == D11.numel(), "descriptor shape mismatch");
  const int is_f32 = D11.scalar_type() == torch::kFloat32 ? 1 : 0;
  const int64_t* init = nullptr;
  if (idx_1_to_2_init.has_value()) {
    CHECK_CONTIGUOUS((*idx_1_to_2_init));
    CHECK_CUDA((*idx_1_to_2_init));
    CHECK_DTYPE((*idx_1_to_2_init), torch::kInt64);
    TORCH_CHECK(idx_1_to_2_init->numel() == (int64_t)b * h * w, "idx_1_to_2_init must be [b,h*w]");
    init = idx_1_to_2_init->data_ptr<int64_t>();
  }
  const c10::cuda::CUDAGuard guard(X11.device());
  auto opts = X11.options();
  torch::Tensor idx = torch::empty({b, (int64_t)h * w}, opts.dtype(torch::kInt64));
  torch::Tensor valid = torch::empty({b, (int64_t)h * w, 1}, opts.dtype(torch::kBool));
  const size_t ws_bytes = m3r_match_workspace_bytes(b, h, w, fdim, is_f32);
  torch::Tensor ws = torch::empty({(int64_t)std::max<size_t>(ws_bytes, 256)}, opts.dtype(torch::kUInt8));
  int rc = m3r_match_fused(X11.data_ptr<float>(), X21.data_ptr<float>(), D11.data_ptr(), D21.data_ptr(), fdim, is_f32,
                           init, b, h, w, max_iter, lambda_init, convergence_thresh, dist_thresh, radius, dilation_max,
                           idx.data_ptr<int64_t>(), reinterpret_cast<uint8_t*>(valid.data_ptr<bool>()), ws.data_ptr(),
                           ws_bytes, cur_stream());
  check_rc(rc, "m3r_match_fused");
  return {idx, valid};
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("gauss_newton_points", &gauss_newton_points, "gauss newton points");
  m.def("gauss_newton_rays", &gauss_newton_rays, "gauss newton rays");
  m.def("gauss_newton_calib", &gauss_newton_calib, "gauss newton calib");
  m.def("iter_proj", &iter_proj, "iterative projection matching");
  m.def("refine_matches", &refine_matches, "refine matches");
  m.def("match_fused", &match_fused, "fused matching.match (additive)", py::arg("X11"), py::arg("X21"), py::arg("D11"),
        py::arg("D21"), py::arg("idx_1_to_2_init") = py::none(), py::arg("max_iter") = 10,
        py::arg("lambda_init") = 1e-8f, py::arg("convergence_thresh") = 1e-6f, py::arg("dist_thresh") = 1e-1f,
        py::arg("radius") = 3, py::arg("dilation_max") = 5);
  m.def("nccl_unique_id", &nccl_unique_id);
  m.def("init_distributed", &init_distributed);
  m.def("shutdown_distributed", &shutdown_distributed);
  m.def("set_edge_shard", &set_edge_shard, py::arg("lo"), py::arg("hi"), py::arg("exchange") = "allreduce");
  m.def("clear_edge_shard", &clear_edge_shard);
  m.def("set_sync_checks", &set_sync_checks);
  m.def("last_info", &last_info);
  m.attr("__m3r_b200__") = true;
  m.attr("abi_version") = m3r_abi_version();
}
