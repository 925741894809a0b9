// C-ABI entry points of the Gauss-Newton path + NCCL plumbing (sm_100a build, host code).
//
// Replaces the three host loops gauss_newton_{points,rays,calib}_cuda of
// mast3r_slam/backend/src/gn_kernels.cu (:725-811, :1140-1228, :1546-1638).  One GN iteration is
//     accumulate -> [exchange of the edge blocks: in-place ncclAllGather, or memset + ncclAllReduce] -> solve
// all enqueued on the caller's stream; the stop test lives on the device (ctrl[CTRL_DONE]) and later
// iterations turn into no-ops, so there is no host synchronisation anywhere in the loop.
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <mutex>
#include <string>

#include "gn_common.cuh"

namespace m3r {

static thread_local std::string g_last_error;

void set_last_error(const char* what, cudaError_t err) {
  g_last_error = std::string(what) + ": " + cudaGetErrorString(err);
}
void set_last_error_text(const char* what) { g_last_error = what; }

namespace {

// ---- NCCL through dlopen: torch already maps libnccl.so.2 into the process; no link dependency
typedef struct { char internal[M3R_NCCL_UNIQUE_ID_BYTES]; } NcclUniqueId;
typedef int (*fn_GetUniqueId)(NcclUniqueId*);
typedef int (*fn_CommInitRank)(void**, int, NcclUniqueId, int);
typedef int (*fn_CommDestroy)(void*);
typedef int (*fn_AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_AllGather)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*fn_CommQuery)(const void*, int*);
typedef const char* (*fn_GetErrorString)(int);

struct NcclApi {
  void* handle = nullptr;
  fn_GetUniqueId GetUniqueId = nullptr;
  fn_CommInitRank CommInitRank = nullptr;
  fn_CommDestroy CommDestroy = nullptr;
  fn_AllReduce AllReduce = nullptr;
  fn_AllGather AllGather = nullptr;   // optional (exchange falls back to the all-reduce without them)
  fn_CommQuery CommCount = nullptr;
  fn_CommQuery CommUserRank = nullptr;
  fn_GetErrorString GetErrorString = nullptr;
  bool ok = false;
};

NcclApi& nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
    }
    if (!api.handle) return;
    api.GetUniqueId = (fn_GetUniqueId)dlsym(api.handle, "ncclGetUniqueId");
    api.CommInitRank = (fn_CommInitRank)dlsym(api.handle, "ncclCommInitRank");
    api.CommDestroy = (fn_CommDestroy)dlsym(api.handle, "ncclCommDestroy");
    api.AllReduce = (fn_AllReduce)dlsym(api.handle, "ncclAllReduce");
    api.AllGather = (fn_AllGather)dlsym(api.handle, "ncclAllGather");
    api.CommCount = (fn_CommQuery)dlsym(api.handle, "ncclCommCount");
    api.CommUserRank = (fn_CommQuery)dlsym(api.handle, "ncclCommUserRank");
    api.GetErrorString = (fn_GetErrorString)dlsym(api.handle, "ncclGetErrorString");
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce;
  });
  return api;
}

constexpr int kNcclFloat64 = 8;  // ncclDouble
constexpr int kNcclSum = 0;      // ncclSum

int nccl_fail(const char* what, int rc) {
  NcclApi& api = nccl();
  std::string msg = std::string(what) + ": ";
  msg += (api.GetErrorString ? api.GetErrorString(rc) : "nccl error");
  set_last_error_text(msg.c_str());
  return M3R_ERR_NCCL;
}

// the packed slabs (last workspace region) are sized for the edges this rank owns
int check_common(int num_kf, int num_pts, int num_edges, const void* ws, size_t ws_bytes, GnLayout* L,
                 const m3r_gn_shard* shard = nullptr, bool needs_packed = false) {
  if (num_kf < 2 || num_pts < 1 || num_edges < 1) return M3R_ERR_BAD_ARG;
  if (!ws || (reinterpret_cast<uintptr_t>(ws) & 255)) return M3R_ERR_WORKSPACE;
  int e_local = num_edges;
  if (shard) {
    if (shard->edge_lo < 0 || shard->edge_hi > num_edges || shard->edge_lo > shard->edge_hi) return M3R_ERR_BAD_ARG;
    e_local = shard->edge_hi - shard->edge_lo;
  }
  *L = gn_layout(num_kf, num_edges, num_pts, e_local);
  // calls without shard information (prepare / solve / export) never touch the packed slabs
  if (ws_bytes < (needs_packed ? L->off_total : L->off_counts)) return M3R_ERR_WORKSPACE;
  return M3R_OK;
}

int shard_range(const m3r_gn_shard* shard, int num_edges, int* lo, int* hi) {
  *lo = 0;
  *hi = num_edges;
  if (shard) {
    *lo = shard->edge_lo;
    *hi = shard->edge_hi;
    if (*lo < 0 || *hi > num_edges || *lo > *hi) return M3R_ERR_BAD_ARG;
  }
  return M3R_OK;
}

// The exchange of the edge blocks is an all-gather over disjoint slots.  Which collective carries it is part of the
// shard description (m3r_gn_shard::exchange) and therefore the same on every rank by contract: it is derived from
// the GLOBAL partition (dist.exchange_for_partition), never from a rank's own range or environment, so ranks cannot
// end up in different collectives.  ALLGATHER (equal contiguous ranges in rank order): one in-place ncclAllGather,
// each rank sends only its own blocks and the two memsets of the foreign slots disappear.  ALLREDUCE (any
// partition): zero the foreign slots and sum-all-reduce (exact, a sum with zeros).
// -> 1 all-gather, 0 all-reduce, < 0 error (an ALLGATHER request this rank's range does not satisfy)
int exchange_mode(const m3r_gn_shard* shard, int num_edges, int lo, int hi) {
  if (shard && shard->exchange == M3R_EXCHANGE_PEER) {  // -> 2: the library's own all-gather over peer memory
    const int rc = peer_xchg_check(shard->xchg, num_edges, lo, hi);
    return rc == M3R_OK ? 2 : rc;
  }
  if (!shard || !shard->comm) return 0;
  if (shard->exchange == M3R_EXCHANGE_ALLREDUCE) return 0;
  if (shard->exchange != M3R_EXCHANGE_ALLGATHER) return M3R_ERR_BAD_ARG;
  NcclApi& api = nccl();
  if (!api.AllGather || !api.CommCount || !api.CommUserRank) return M3R_ERR_NCCL;
  int world = 0, rank = -1;
  if (api.CommCount(shard->comm, &world) != 0 || api.CommUserRank(shard->comm, &rank) != 0) return M3R_ERR_NCCL;
  if (world < 1 || num_edges % world != 0) return M3R_ERR_BAD_ARG;
  const int cnt = num_edges / world;
  if (lo != rank * cnt || hi != lo + cnt) return M3R_ERR_BAD_ARG;
  return 1;
}

int accumulate_step(const float* Twc, const float* Xs, const float* K, const GnLayout& L, const GnPtrs& P,
                    const m3r_gn_params& prm, const m3r_gn_shard* shard, int lo, int hi, const GnFuse& fuse,
                    cudaStream_t s, int it = 0, unsigned xchg_seq = 0) {
  const int mode = exchange_mode(shard, L.num_edges, lo, hi);
  if (mode < 0) {
    set_last_error_text("m3r_gn_shard: exchange = ALLGATHER / PEER needs the equal partition (world | num_edges, rank-ordered ranges) and, for PEER, a connected exchange of sufficient capacity");
    return mode;
  }
  const bool gather = mode == 1;
  if (!gather && shard && (lo > 0 || hi < L.num_edges)) {
    // non-owned slots must be exact zeros so that the sum all-reduce acts as an all-gather
    if (lo > 0) M3R_CUDA_CHECK(cudaMemsetAsync(P.eb, 0, sizeof(double) * kEdgeBlock * (size_t)lo, s));
    if (hi < L.num_edges)
      M3R_CUDA_CHECK(cudaMemsetAsync(P.eb + (size_t)hi * kEdgeBlock, 0,
                                     sizeof(double) * kEdgeBlock * (size_t)(L.num_edges - hi), s));
  }
  int rc = launch_gn_accumulate(Twc, Xs, K, L, P, prm, lo, hi, fuse, s);
  if (rc != M3R_OK) return rc;
  if (mode == 2) return launch_peer_allgather(shard->xchg, lo, hi, it, xchg_seq, P.ctrl, s);  // (P.eb is this rank's exchange buffer)
  if (gather) {
    const size_t count = (size_t)(hi - lo) * kEdgeBlock;
    const int nrc = nccl().AllGather(P.eb + (size_t)lo * kEdgeBlock, P.eb, count, kNcclFloat64, shard->comm, s);
    if (nrc != 0) return nccl_fail("ncclAllGather", nrc);
  } else if (shard && shard->comm) {
    rc = m3r_comm_allreduce_f64(shard->comm, P.eb, (size_t)L.num_edges * kEdgeBlock, s);
  }
  return rc;
}

}  // namespace
}  // namespace m3r

using namespace m3r;

extern "C" int m3r_abi_version(void) { return M3R_ABI_VERSION; }
extern "C" const char* m3r_last_error(void) { return g_last_error.c_str(); }

extern "C" size_t m3r_gn_workspace_bytes(int num_kf, int num_edges, int num_pts) {
  if (num_kf < 2 || num_edges < 1 || num_pts < 1) return 0;
  return gn_layout(num_kf, num_edges, num_pts).off_total;
}

extern "C" size_t m3r_gn_workspace_bytes_sharded(int num_kf, int num_edges, int num_pts, int edges_local) {
  if (num_kf < 2 || num_edges < 1 || num_pts < 1 || edges_local < 0 || edges_local > num_edges) return 0;
  return gn_layout(num_kf, num_edges, num_pts, edges_local).off_total;
}

extern "C" int m3r_gn_prepare(const int64_t* ii, const int64_t* jj, int num_kf, int num_pts, int num_edges,
                              void* workspace, size_t workspace_bytes, void* stream) {
  if (!ii || !jj) return M3R_ERR_BAD_ARG;
  GnLayout L;
  int rc = check_common(num_kf, num_pts, num_edges, workspace, workspace_bytes, &L);
  if (rc != M3R_OK) return rc;
  return launch_gn_prepare(ii, jj, L, gn_ptrs(L, workspace), static_cast<cudaStream_t>(stream));
}

extern "C" int m3r_gn_pack(const float* Cs, const int64_t* idx, const uint8_t* valid, const float* Q, int num_kf,
                          int num_pts, int num_edges, const m3r_gn_params* params, void* workspace,
                          size_t workspace_bytes, const m3r_gn_shard* shard, void* stream) {
  if (!Cs || !idx || !valid || !Q || !params) return M3R_ERR_BAD_ARG;
  GnLayout L;
  int rc = check_common(num_kf, num_pts, num_edges, workspace, workspace_bytes, &L, shard, true);
  if (rc != M3R_OK) return rc;
  int lo, hi;
  rc = shard_range(shard, num_edges, &lo, &hi);
  if (rc != M3R_OK) return rc;
  return launch_gn_pack(Cs, idx, valid, Q, L, gn_ptrs(L, workspace), *params, lo, hi,
                        static_cast<cudaStream_t>(stream));
}

extern "C" int m3r_gn_accumulate(const float* Twc, const float* Xs, const float* K, int num_kf, int num_pts,
                                 int num_edges, const m3r_gn_params* params, void* workspace,
                                 size_t workspace_bytes, const m3r_gn_shard* shard, void* stream) {
  if (!Twc || !Xs || !params) return M3R_ERR_BAD_ARG;
  GnLayout L;
  int rc = check_common(num_kf, num_pts, num_edges, workspace, workspace_bytes, &L, shard, true);
  if (rc != M3R_OK) return rc;
  int lo, hi;
  rc = shard_range(shard, num_edges, &lo, &hi);
  if (rc != M3R_OK) return rc;
  if (shard && shard->exchange == M3R_EXCHANGE_PEER) {
    set_last_error_text("m3r_gn_accumulate: the peer exchange keeps its edge blocks in the exchange buffer; use m3r_gauss_newton");
    return M3R_ERR_UNSUPPORTED;
  }
  GnFuse nofuse{};
  return accumulate_step(Twc, Xs, K, L, gn_ptrs(L, workspace), *params, shard, lo, hi, nofuse,
                         static_cast<cudaStream_t>(stream));
}

extern "C" int m3r_gn_edge_blocks(const void* workspace, int num_kf, int num_edges, int num_pts,
                                  double** blocks_out) {
  if (!workspace || !blocks_out || num_kf < 2 || num_edges < 1 || num_pts < 1) return M3R_ERR_BAD_ARG;
  GnLayout L = gn_layout(num_kf, num_edges, num_pts);
  *blocks_out = gn_ptrs(L, const_cast<void*>(workspace)).eb;
  return M3R_OK;
}

extern "C" int m3r_gn_solve_retract(float* Twc, int num_kf, int num_pts, int num_edges, float delta_thresh,
                                    float* dx_out, int32_t* info_out, void* workspace, size_t workspace_bytes,
                                    void* stream) {
  if (!Twc || !dx_out) return M3R_ERR_BAD_ARG;
  GnLayout L;
  int rc = check_common(num_kf, num_pts, num_edges, workspace, workspace_bytes, &L);
  if (rc != M3R_OK) return rc;
  return launch_gn_solve(Twc, L, gn_ptrs(L, workspace), delta_thresh, dx_out, info_out,
                         static_cast<cudaStream_t>(stream));
}

extern "C" int m3r_gn_export_hs_gs(const void* workspace, int num_kf, int num_edges, int num_pts, float* Hs,
                                   float* gs, void* stream) {
  if (!workspace || !Hs || !gs || num_kf < 2 || num_edges < 1 || num_pts < 1) return M3R_ERR_BAD_ARG;
  GnLayout L = gn_layout(num_kf, num_edges, num_pts);
  return launch_export_hs_gs(L, gn_ptrs(L, const_cast<void*>(workspace)), Hs, gs, static_cast<cudaStream_t>(stream));
}

extern "C" int m3r_pose_retr(float* Twc, const float* dx, int num_kf, int num_fix, void* stream) {
  if (!Twc || !dx || num_kf < 0 || num_fix < 0) return M3R_ERR_BAD_ARG;
  return launch_pose_retr(Twc, dx, num_kf, num_fix, static_cast<cudaStream_t>(stream));
}

extern "C" int m3r_gauss_newton(float* Twc, const float* Xs, const float* Cs, const float* K,
                                const int64_t* ii, const int64_t* jj, const int64_t* idx,
                                const uint8_t* valid, const float* Q, int num_kf, int num_pts, int num_edges,
                                const m3r_gn_params* params, float* dx_out, int32_t* info_out, void* workspace,
                                size_t workspace_bytes, const m3r_gn_shard* shard, void* stream) {
  if (!Twc || !Xs || !Cs || !ii || !jj || !idx || !valid || !Q || !params || !dx_out) return M3R_ERR_BAD_ARG;
  if (params->kind == M3R_GN_CALIB && !K) return M3R_ERR_BAD_ARG;
  if (params->max_iter < 0) return M3R_ERR_BAD_ARG;
  GnLayout L;
  int rc = check_common(num_kf, num_pts, num_edges, workspace, workspace_bytes, &L, shard, true);
  if (rc != M3R_OK) return rc;
  int lo, hi;
  rc = shard_range(shard, num_edges, &lo, &hi);
  if (rc != M3R_OK) return rc;
  const int xmode = exchange_mode(shard, num_edges, lo, hi);
  if (xmode < 0) {  // before anything is enqueued
    set_last_error_text("m3r_gn_shard: exchange = ALLGATHER / PEER needs the equal partition (world | num_edges, rank-ordered ranges) and, for PEER, a connected exchange of sufficient capacity");
    return M3R_ERR_BAD_ARG;
  }
  const unsigned xchg_seq = xmode == 2 ? peer_xchg_next_seq(shard->xchg) : 0u;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  GnPtrs P = gn_ptrs(L, workspace);
  // small systems without a collective: solve + retract ride in the accumulate kernel (1 launch / iteration)
  GnFuse fuse{};
  fuse.Twc = Twc; fuse.dx_out = dx_out; fuse.info_out = info_out; fuse.delta_thresh = params->delta_thresh;
  fuse.enabled = (L.dim <= kSmallDim) && !(shard && (shard->comm || xmode == 2)) && lo == 0 && hi == num_edges;
  const unsigned ll_epoch = (params->max_iter > 0 && fuse.enabled) ? gn_next_epoch() : 0u;
  rc = launch_gn_prepare(ii, jj, L, P, s, dx_out, kPose * (num_kf - 1), info_out, ll_epoch);  // also zeroes dx_out / info_out
  if (rc != M3R_OK) return rc;
  if (params->max_iter > 0) {
    rc = launch_gn_pack(Cs, idx, valid, Q, L, P, *params, lo, hi, s);
    if (rc != M3R_OK) return rc;
  }
  if (params->max_iter > 0 && fuse.enabled) {
    // tracking-sized: the whole loop is one cooperative launch (gn_track_kernel)
    int launched = 0;
    rc = launch_gn_track(Twc, Xs, K, L, P, *params, fuse, &launched, s, nullptr, ll_epoch);
    if (rc != M3R_OK) return rc;
    if (launched) return M3R_OK;
  }
  for (int it = 0; it < params->max_iter; it++) {
    fuse.pdl = fuse.enabled && it > 0;  // iteration 0 follows pack: full serialisation
    GnPtrs Pit = P;
    if (xmode == 2) Pit.eb = peer_xchg_buffer(shard->xchg, it);  // the edge blocks of this iteration live in the exchange buffer
    rc = accumulate_step(Twc, Xs, K, L, Pit, *params, shard, lo, hi, fuse, s, it, xchg_seq);
    if (rc != M3R_OK) return rc;
    if (!fuse.enabled) {
      rc = launch_gn_solve(Twc, L, Pit, params->delta_thresh, dx_out, info_out, s);
      if (rc != M3R_OK) return rc;
    }
  }
  return M3R_OK;
}

// ---------------------------------------------------------------------------------- NCCL plumbing
extern "C" int m3r_comm_unique_id(void* id_out) {
  NcclApi& api = nccl();
  if (!api.ok) {
    set_last_error_text("libnccl.so.2 not loadable");
    return M3R_ERR_NCCL;
  }
  if (!id_out) return M3R_ERR_BAD_ARG;
  NcclUniqueId id;
  int rc = api.GetUniqueId(&id);
  if (rc != 0) return nccl_fail("ncclGetUniqueId", rc);
  std::memcpy(id_out, &id, sizeof(id));
  return M3R_OK;
}

extern "C" int m3r_comm_init(int rank, int world, const void* id, void** comm_out) {
  NcclApi& api = nccl();
  if (!api.ok) {
    set_last_error_text("libnccl.so.2 not loadable");
    return M3R_ERR_NCCL;
  }
  if (!id || !comm_out || world < 1 || rank < 0 || rank >= world) return M3R_ERR_BAD_ARG;
  NcclUniqueId uid;
  std::memcpy(&uid, id, sizeof(uid));
  void* comm = nullptr;
  int rc = api.CommInitRank(&comm, world, uid, rank);
  if (rc != 0) return nccl_fail("ncclCommInitRank", rc);
  *comm_out = comm;
  return M3R_OK;
}

extern "C" int m3r_comm_destroy(void* comm) {
  NcclApi& api = nccl();
  if (!api.ok || !comm) return M3R_ERR_BAD_ARG;
  int rc = api.CommDestroy(comm);
  if (rc != 0) return nccl_fail("ncclCommDestroy", rc);
  return M3R_OK;
}

extern "C" int m3r_comm_allreduce_f64(void* comm, double* buf, size_t count, void* stream) {
  NcclApi& api = nccl();
  if (!api.ok) {
    set_last_error_text("libnccl.so.2 not loadable");
    return M3R_ERR_NCCL;
  }
  if (!comm || !buf) return M3R_ERR_BAD_ARG;
  int rc = api.AllReduce(buf, buf, count, kNcclFloat64, kNcclSum, comm, static_cast<cudaStream_t>(stream));
  if (rc != 0) return nccl_fail("ncclAllReduce", rc);
  return M3R_OK;
}
