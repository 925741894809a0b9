// Workspace layout and shared declarations of the Gauss-Newton path (sm_100a).
#pragma once
#include "common.cuh"

namespace m3r {

constexpr int kPose = 7;        // Sim(3) tangent dimension
constexpr int kAcc = 35;        // 28 upper-triangle terms of the local 7x7 + 7 gradient terms
constexpr int kEdgeBlock = 56;  // per directed edge: A_e (49, row-major) + b_e (7), fp64
constexpr int kTile = 32;       // Cholesky tile
constexpr int kAccumThreads = 256;
constexpr int kSolveThreads = 256;
constexpr int kMaxChunks = 256;

// ctrl words (device int32)
enum { CTRL_DONE = 0, CTRL_ITERS = 1, CTRL_FAILS = 2, CTRL_EARLY = 3, CTRL_BAD = 4, CTRL_FAIL_NOW = 5, CTRL_WORDS = 16 };

// How many point-chunks (CTAs) each edge is split into.  Pure function of (E, n) so that the
// workspace size and every launch agree.  Target: >= 8 CTAs per SM in flight overall.
inline int gn_chunks(int num_edges_local, int n) {
  const int want = (kNumSMs * 8 + num_edges_local - 1) / (num_edges_local > 0 ? num_edges_local : 1);
  int max_chunks = (n + 1023) / 1024;
  if (max_chunks < 1) max_chunks = 1;
  int c = want < 1 ? 1 : want;
  if (c > max_chunks) c = max_chunks;
  if (c > kMaxChunks) c = kMaxChunks;
  return c;
}

struct GnLayout {
  int num_kf, num_edges, num_pts;
  int dim, nblk, ld, rows;  // system dim 7(N-1); tiles; leading dimension; rows incl. rhs tile row
  size_t off_ctrl, off_counters, off_ci, off_cj, off_bitmap, off_prefix, off_inc_off, off_inc_list,
      off_partials, off_eb, off_hd, off_linv, off_total;
};

inline GnLayout gn_layout(int num_kf, int num_edges, int num_pts) {
  GnLayout L{};
  L.num_kf = num_kf;
  L.num_edges = num_edges;
  L.num_pts = num_pts;
  L.dim = kPose * (num_kf - 1);
  L.nblk = (L.dim + kTile - 1) / kTile;
  if (L.nblk < 1) L.nblk = 1;
  L.ld = L.nblk * kTile;
  L.rows = (L.nblk + 1) * kTile;
  size_t o = 0;
  auto take = [&](size_t bytes) {
    size_t r = o;
    o = align_up(o + bytes, 256);
    return r;
  };
  L.off_ctrl = take(sizeof(int) * CTRL_WORDS);
  L.off_counters = take(sizeof(unsigned) * (size_t)num_edges);
  L.off_ci = take(sizeof(int) * (size_t)num_edges);
  L.off_cj = take(sizeof(int) * (size_t)num_edges);
  L.off_bitmap = take(sizeof(unsigned) * (M3R_MAX_KF_ID / 32));
  L.off_prefix = take(sizeof(unsigned) * (M3R_MAX_KF_ID / 32));
  L.off_inc_off = take(sizeof(int) * ((size_t)num_kf + 1));
  L.off_inc_list = take(sizeof(int) * 2 * (size_t)num_edges);
  // chunks(el) * el <= kNumSMs*8 + el for el <= kNumSMs*8, and == el above it (chunks == 1)
  size_t worst = (size_t)2 * kNumSMs * 8 + 256;
  if ((size_t)num_edges > worst) worst = (size_t)num_edges;
  L.off_partials = take(sizeof(double) * kAcc * worst);
  L.off_eb = take(sizeof(double) * kEdgeBlock * (size_t)num_edges);
  L.off_hd = take(sizeof(double) * (size_t)L.rows * L.ld);
  L.off_linv = take(sizeof(double) * (size_t)L.nblk * kTile * kTile);
  L.off_total = o;
  return L;
}

struct GnPtrs {
  int* ctrl;
  unsigned* counters;
  int* ci;
  int* cj;
  unsigned* bitmap;
  unsigned* prefix;
  int* inc_off;
  int* inc_list;
  double* partials;
  double* eb;
  double* hd;
  double* linv;
};

inline GnPtrs gn_ptrs(const GnLayout& L, void* ws) {
  char* b = static_cast<char*>(ws);
  GnPtrs p;
  p.ctrl = reinterpret_cast<int*>(b + L.off_ctrl);
  p.counters = reinterpret_cast<unsigned*>(b + L.off_counters);
  p.ci = reinterpret_cast<int*>(b + L.off_ci);
  p.cj = reinterpret_cast<int*>(b + L.off_cj);
  p.bitmap = reinterpret_cast<unsigned*>(b + L.off_bitmap);
  p.prefix = reinterpret_cast<unsigned*>(b + L.off_prefix);
  p.inc_off = reinterpret_cast<int*>(b + L.off_inc_off);
  p.inc_list = reinterpret_cast<int*>(b + L.off_inc_list);
  p.partials = reinterpret_cast<double*>(b + L.off_partials);
  p.eb = reinterpret_cast<double*>(b + L.off_eb);
  p.hd = reinterpret_cast<double*>(b + L.off_hd);
  p.linv = reinterpret_cast<double*>(b + L.off_linv);
  return p;
}

// entry points implemented in gn_accum.cu / gn_solve.cu (host launchers)
int launch_gn_prepare(const int64_t* ii, const int64_t* jj, const GnLayout& L, const GnPtrs& P, cudaStream_t s);
int launch_gn_accumulate(const float* Twc, const float* Xs, const float* Cs, const float* K,
                         const int64_t* idx, const uint8_t* valid, const float* Q, const GnLayout& L,
                         const GnPtrs& P, const m3r_gn_params& prm, int e_lo, int e_hi, cudaStream_t s);
int launch_gn_solve(float* Twc, const GnLayout& L, const GnPtrs& P, float delta_thresh, float* dx_out,
                    int32_t* info_out, cudaStream_t s);
int launch_export_hs_gs(const GnLayout& L, const GnPtrs& P, float* Hs, float* gs, cudaStream_t s);
int launch_pose_retr(float* Twc, const float* dx, int num_kf, int num_fix, cudaStream_t s);

}  // namespace m3r
