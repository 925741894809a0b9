// Workspace layout and shared declarations of the Gauss-Newton path (sm_100a).
#pragma once
#include <algorithm>

#include "common.cuh"

namespace m3r {

constexpr int kPose = 7;        // Sim(3) tangent dimension
constexpr int kAcc = 35;        // 28 upper-triangle terms of the local 7x7 + 7 gradient terms
constexpr int kEdgeBlock = 56;  // per directed edge: A_e (49, row-major) + b_e (7), fp64
constexpr int kTile = 32;       // Cholesky tile
constexpr int kLlKinds = 5;     // L_kk, L_(k+1,k), tile (k+2,k+1), tile (k+2,k+2), L_(k+2,k)
constexpr int kLlRing = 4;      // steps a slot lives before it is reused
constexpr int kAccumThreads = 256;
constexpr int kSolveThreads = 256;

// ctrl words (device int32)
enum { CTRL_DONE = 0, CTRL_ITERS = 1, CTRL_FAILS = 2, CTRL_EARLY = 3, CTRL_BAD = 4, CTRL_FAIL_NOW = 5, CTRL_LL_GEN = 15, CTRL_WORDS = 16 };
// (CTRL_LL_GEN survives gn_prepare: generation under which the tracking kernel's flag words were last cleared)

// Point-chunking of one edge: a pure function of (E, n) with E the FULL edge count, so that the
// workspace size, every launch and every shard of a multi-GPU run agree (a shard's edge blocks are
// then bit-identical to the unsharded run's).  A CTA owns `chunk_pts` consecutive points of one edge.
//  * big graphs: 3072-point chunks -> few edges in flight at a time, so both directions of a
//    keyframe pair find Xs[i] / Xs[j] in the 126 MB L2;
//  * small graphs (tracking: 2 edges): the chunk shrinks until the grid is one full wave
//    (kAccumCtasPerSM CTAs on each of the 148 SMs).
constexpr int kAccumCtasPerSM = 3;
constexpr int kMaxChunkPts = 3072;  // x 22 B (Xs[j] 12 + record 8 + local index 2) = 66 KB of smem per CTA
constexpr int kSmallDim = 63;       // systems up to 9 free poses are solved inside the accumulate kernel
constexpr int kTrackMaxEdges = 8;   // directed edges the persistent tracking kernel takes
struct GnChunking {
  int chunk_pts, chunks;
};
inline GnChunking gn_chunking(int num_edges, int n) {
  GnChunking c;
  const long long e = num_edges > 0 ? num_edges : 1;
  const long long slots = (long long)kNumSMs * kAccumCtasPerSM;
  c.chunk_pts = kMaxChunkPts;
  c.chunks = (n + c.chunk_pts - 1) / c.chunk_pts;
  if (e * c.chunks < slots) {
    const long long per_edge = slots / e > 0 ? slots / e : 1;
    long long chunk = (n + per_edge - 1) / per_edge;
    chunk = (chunk + 7) / 8 * 8;  // 16-byte granularity for every staged stream
    if (chunk < 256) chunk = 256;
    if (chunk > kMaxChunkPts) chunk = kMaxChunkPts;
    c.chunk_pts = (int)chunk;
    c.chunks = (n + c.chunk_pts - 1) / c.chunk_pts;
  }
  if (c.chunks < 1) c.chunks = 1;
  return c;
}

struct GnLayout {
  int num_kf, num_edges, num_pts;
  int dim, nblk, ld, rows;  // system dim 7(N-1); tiles; leading dimension; rows incl. rhs tile row
  size_t off_ctrl, off_counters, off_ci, off_cj, off_bitmap, off_prefix, off_inc_off, off_inc_list,
      off_order, off_partials, off_eb, off_hd, off_linv, off_flags, off_ll, off_counts, off_kloc, off_packed, off_total;
  int edges_local;  // packed slabs are sized for this many edges (== num_edges unless sharded)
};

inline GnLayout gn_layout(int num_kf, int num_edges, int num_pts, int edges_local = -1) {
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
  L.off_counters = take(sizeof(unsigned) * ((size_t)num_edges + 1));  // per-edge tickets + [E] = edges done
  L.off_ci = take(sizeof(int) * (size_t)num_edges);
  L.off_cj = take(sizeof(int) * (size_t)num_edges);
  L.off_bitmap = take(sizeof(unsigned) * (M3R_MAX_KF_ID / 32));
  L.off_prefix = take(sizeof(unsigned) * (M3R_MAX_KF_ID / 32));
  L.off_inc_off = take(sizeof(int) * ((size_t)num_kf + 1));
  L.off_inc_list = take(sizeof(int) * 2 * (size_t)num_edges);
  L.off_order = take(sizeof(int) * (size_t)num_edges);
  // (the persistent tracking kernel keeps two all-gather buffers of 16 B x kAcc per CTA of a <= 148-CTA grid here)
  const size_t worst = std::max((size_t)num_edges * (size_t)gn_chunking(num_edges, num_pts).chunks, (size_t)4 * kNumSMs + 32);
  L.off_partials = take(sizeof(double) * kAcc * worst);
  L.off_eb = take(sizeof(double) * kEdgeBlock * (size_t)num_edges);
  L.off_hd = take(sizeof(double) * (size_t)L.rows * L.ld);
  L.off_linv = take(sizeof(double) * (size_t)L.nblk * kTile * kTile);
  // tile flags of the dataflow factorisation, three arrays of (nblk+1) x nblk: done (tile holds its final L values),
  // rdy (helper -> chain hand-off of a diagonal / sub-diagonal tile), rdyh (cyclic owner -> helper hand-off of a band tile)
  L.off_flags = take(sizeof(int) * 3 * (size_t)(L.nblk + 1) * L.nblk);
  // LL transfer buffers of the chain <-> helper loop: kLlKinds tiles x ring of kLlRing steps x 2048 tagged 8-byte words
  L.off_ll = take(sizeof(unsigned long long) * kLlKinds * kLlRing * 2 * kTile * kTile);
  // compacted per-point records of the locally owned edges (last regions): per (edge, chunk) a count,
  // then for each surviving point its chunk-local index (u16) and {int32 m, float sqrt(Q)}
  L.edges_local = edges_local < 0 ? num_edges : edges_local;
  L.off_counts = take(sizeof(int) * (size_t)L.edges_local * (size_t)gn_chunking(num_edges, num_pts).chunks);
  L.off_kloc = take(sizeof(uint16_t) * ((size_t)L.edges_local * (size_t)num_pts + 8));
  L.off_packed = take(sizeof(uint2) * ((size_t)L.edges_local * (size_t)num_pts + 8));
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
  int* order;  // launch order of the edges: both directions of a keyframe pair adjacent, pairs sorted
  double* partials;
  double* eb;
  double* hd;
  double* linv;
  int* flags;
  unsigned long long* ll;
  uint2* packed;
  uint16_t* kloc;
  int* counts;
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
  p.order = reinterpret_cast<int*>(b + L.off_order);
  p.partials = reinterpret_cast<double*>(b + L.off_partials);
  p.eb = reinterpret_cast<double*>(b + L.off_eb);
  p.hd = reinterpret_cast<double*>(b + L.off_hd);
  p.linv = reinterpret_cast<double*>(b + L.off_linv);
  p.flags = reinterpret_cast<int*>(b + L.off_flags);
  p.ll = reinterpret_cast<unsigned long long*>(b + L.off_ll);
  p.packed = reinterpret_cast<uint2*>(b + L.off_packed);
  p.kloc = reinterpret_cast<uint16_t*>(b + L.off_kloc);
  p.counts = reinterpret_cast<int*>(b + L.off_counts);
  return p;
}

// fused tail of the accumulate kernel (dim <= kSmallDim, no collective): solve + retract in the last CTA
struct GnFuse {
  float* Twc;        // written by the fused solve
  float* dx_out;
  int32_t* info_out;
  float delta_thresh;
  int enabled;
  int pdl;  // launch with programmatic stream serialization (the previous op on the stream is the same kernel)
};

// entry points implemented in gn_accum.cu / gn_solve.cu (host launchers)
// ll_epoch: nonce of the persistent tracking kernel's flag words for the solve that follows (gn_next_epoch(); 0 = no such
// kernel will run on this workspace).  The prepare kernel clears the flag words whenever the workspace was last
// cleared under another generation (epoch >> 20) of the 20-bit nonce, i.e. on first use and once per wrap.
unsigned gn_next_epoch();
int launch_gn_prepare(const int64_t* ii, const int64_t* jj, const GnLayout& L, const GnPtrs& P, cudaStream_t s,
                      float* zero_dx = nullptr, int n_dx = 0, int32_t* zero_info = nullptr, unsigned ll_epoch = 0);
int launch_gn_pack(const float* Cs, const int64_t* idx, const uint8_t* valid, const float* Q, const GnLayout& L,
                   const GnPtrs& P, const m3r_gn_params& prm, int e_lo, int e_hi, cudaStream_t s);
int launch_gn_accumulate(const float* Twc, const float* Xs, const float* K, const GnLayout& L, const GnPtrs& P,
                         const m3r_gn_params& prm, int e_lo, int e_hi, const GnFuse& fuse, cudaStream_t s);
// frame-tracker mode of the persistent kernel: one directed edge, target / moving points in separate arrays,
// extra stop test on the relative decrease of the robust cost (tracker.py:156-214)
struct GnFrameTrack {
  const float* Xi;  // [n,3] keyframe points (targets)
  const float* Xj;  // [n,3] frame points, already gathered by idx_f2k (moved by T_CkCf)
  float rel_error;
};
int launch_gn_track(const float* Twc, const float* Xs, const float* K, const GnLayout& L, const GnPtrs& P,
                    const m3r_gn_params& prm, const GnFuse& fuse, int* launched, cudaStream_t s,
                    const GnFrameTrack* ft = nullptr, unsigned ll_epoch = 0);
int launch_gn_solve(float* Twc, const GnLayout& L, const GnPtrs& P, float delta_thresh, float* dx_out,
                    int32_t* info_out, cudaStream_t s);
// shared-memory-resident single-CTA solve (gn_solve_resident.cu): systems of up to kResidentMaxTiles tiles
constexpr int kResidentMaxTiles = 7;
struct ResidentSolveArgs {
  float* Twc;
  const double* eb;
  const int* ci;
  const int* cj;
  const int* inc_off;
  const int* inc_list;
  int* ctrl;
  float* dx_out;
  int32_t* info_out;
  int N, dim, nblk;
  float delta_thresh;
};
int launch_gn_solve_resident(const ResidentSolveArgs& a, cudaStream_t s);
// peer-memory all-gather of the edge blocks (peer_exchange.cu)
int peer_xchg_check(const void* xchg, int num_edges, int lo, int hi);
unsigned peer_xchg_next_seq(void* xchg);
double* peer_xchg_buffer(void* xchg, int parity);
int launch_peer_allgather(void* xchg, int lo, int hi, int it, unsigned seq, int* ctrl, cudaStream_t s);
int launch_export_hs_gs(const GnLayout& L, const GnPtrs& P, float* Hs, float* gs, cudaStream_t s);
int launch_pose_retr(float* Twc, const float* dx, int num_kf, int num_fix, cudaStream_t s);

}  // namespace m3r
