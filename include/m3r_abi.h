/* m3r_abi.h — C ABI of libm3r_b200.so, the B200 (sm_100a) geometric backend that replaces the
 * reference's `mast3r_slam_backends` CUDA extension.
 *
 * Every entry point takes plain DEVICE pointers + sizes + a cudaStream_t (passed as void*).
 * Nothing here allocates, synchronises the host or throws: all work is enqueued on `stream`
 * (graph-capturable) and an int status is returned (0 = ok, <0 = M3R_ERR_*).
 * The torch-facing shim (csrc/bindings.cpp → python module `mast3r_slam_backends`) and the
 * ctypes wrapper (monst3r-slam_b200/abi.py) are the only callers.
 *
 * Reference interface each entry point replaces (paths relative to
 * /root/reference/MASt3R-SLAM/mast3r_slam/backend/):
 *   m3r_iter_proj            include/gn.h:89-95   src/gn.cpp:84-99    src/matching_kernels.cu:119-316
 *   m3r_refine_matches_*     include/gn.h:112-117 src/gn.cpp:101-114  src/matching_kernels.cu:25-116
 *   m3r_gauss_newton_rays    include/gn.h:45-55   src/gn.cpp:28-52    src/gn_kernels.cu:813-1228
 *   m3r_gauss_newton_calib   include/gn.h:74-87   src/gn.cpp:54-82    src/gn_kernels.cu:1231-1638
 *   m3r_gauss_newton_points  include/gn.h:22-31   src/gn.cpp:3-26     src/gn_kernels.cu:455-811
 *   (index prep gn_kernels.cu:161-170, SparseBlock assemble+solve :57-159 and pose_retr_kernel
 *    :415-453 run on the device inside the m3r_gauss_newton_* calls.)
 */
#ifndef M3R_ABI_H
#define M3R_ABI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define M3R_ABI_VERSION 4

/* status codes */
#define M3R_OK 0
#define M3R_ERR_BAD_ARG (-1)      /* null pointer / non-positive size / unsupported parameter   */
#define M3R_ERR_WORKSPACE (-2)    /* workspace too small or misaligned                          */
#define M3R_ERR_CUDA (-3)         /* a CUDA runtime call failed (see m3r_last_error)            */
#define M3R_ERR_NCCL (-4)         /* NCCL unavailable or a collective failed                    */
#define M3R_ERR_UNSUPPORTED (-5)

/* residual kinds of the Gauss-Newton ops */
#define M3R_GN_POINTS 0
#define M3R_GN_RAYS 1
#define M3R_GN_CALIB 2

/* largest keyframe id accepted in ii/jj (device-side compaction uses a presence bitmap) */
#define M3R_MAX_KF_ID (1 << 20)

int m3r_abi_version(void);
/* last CUDA / NCCL error text seen by this thread ("" if none) */
const char* m3r_last_error(void);

/* ---- matching -------------------------------------------------------------------------- */

/* iter_proj: rays_img [b,h,w,9] f32, pts3d_norm [b,n,3] f32, p_init [b,n,2] f32 (u,v)
 *   -> p_new [b,n,2] f32, converged [b,n] u8 (0/1; torch.bool layout).  Every output element is
 *   written (n need not be a multiple of the block size; the reference has no guard). */
int m3r_iter_proj(const float* rays_img, const float* pts3d_norm, const float* p_init,
                  float* p_new, uint8_t* converged, int b, int h, int w, int n, int max_iter,
                  float lambda_init, float cost_thresh, void* stream);

/* Same op with a caller-provided scratch buffer of m3r_iter_proj_workspace_bytes(b,h,w) bytes
 * (16-byte aligned): the ray image is first re-packed into float4/float planes (48 bytes of scratch per texel) so that every bilinear tap is
 * a 128-bit load.  Results are identical to m3r_iter_proj; workspace == NULL falls back to it. */
size_t m3r_iter_proj_workspace_bytes(int b, int h, int w);
int m3r_iter_proj_ws(const float* rays_img, const float* pts3d_norm, const float* p_init, float* p_new,
                     uint8_t* converged, int b, int h, int w, int n, int max_iter, float lambda_init,
                     float cost_thresh, void* workspace, size_t workspace_bytes, void* stream);

/* Same op again, for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  The 48-byte-per-texel
 * re-pack scratch of m3r_iter_proj_ws lives in a file-scope `static __device__` array inside the library, sized for
 * M3R_ITER_PROJ_STATIC_MAX_TEXELS texels (b*h*w <= 512*384 = one frame at the reference's working resolution;
 * 9 437 184 bytes that every GPU loading the library holds until exit).  The call allocates nothing (no cudaMalloc /
 * cudaFree, no library or torch allocation) and takes no workspace argument; results are identical to m3r_iter_proj.
 *   - b*h*w beyond the maximum -> M3R_ERR_WORKSPACE, with the sizes in m3r_last_error(); nothing is enqueued.
 *   - rays_img must be 16-byte aligned and p_init / p_new 8-byte aligned (M3R_ERR_BAD_ARG otherwise).
 *   - ALL calls on a device share the one array: calls that overlap on the GPU (different streams or host threads)
 *     race on it; serialise them on one stream.  Nothing in it needs clearing between calls (texels [0, b*h*w) are
 *     rewritten by the call's first kernel before its second kernel reads exactly that range). */
#define M3R_ITER_PROJ_STATIC_MAX_TEXELS (512 * 384)
size_t m3r_iter_proj_static_max_texels(void);
int m3r_iter_proj_static(const float* rays_img, const float* pts3d_norm, const float* p_init, float* p_new,
                         uint8_t* converged, int b, int h, int w, int n, int max_iter, float lambda_init,
                         float cost_thresh, void* stream);

/* refine_matches: D11 [b,h,w,fdim], D21 [b,n,fdim], p1 [b,n,2] i64 (u,v) -> p1_new [b,n,2] i64.
 * The f16 variant reproduces c10::Half arithmetic (every * and + rounded to fp16, k ascending,
 * best score starts at 0); f32/f64 start at FLT_MIN/DBL_MIN like the reference template. */
int m3r_refine_matches_f16(const void* D11, const void* D21, const int64_t* p1, int64_t* p1_new,
                           int b, int h, int w, int n, int fdim, int radius, int dilation_max,
                           void* stream);
int m3r_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1, int64_t* p1_new,
                           int b, int h, int w, int n, int fdim, int radius, int dilation_max,
                           void* stream);
int m3r_refine_matches_f64(const double* D11, const double* D21, const int64_t* p1,
                           int64_t* p1_new, int b, int h, int w, int n, int fdim, int radius,
                           int dilation_max, void* stream);

/* Fused matching.match() (additive; the five reference ops above are unchanged): matching.py:25-90 in
 * five launches -- ray normalisation + Scharr gradients + p_init (one kernel), LM projection with the
 * occlusion test |X11[p] - X21| < dist_thresh fused into its epilogue, fp32->fp16 of both descriptor
 * maps (skipped when they are fp16 already), refine_matches, pixel_to_lin.
 *   X11, X21 [b,h,w,3] f32; D11, D21 [b,h,w,fdim] f32 (desc_is_f32 = 1) or f16; idx_init [b,h*w] i64 or NULL
 *   -> idx_out [b,h*w] i64 (u + w v), valid_out [b,h*w] u8.  workspace: m3r_match_workspace_bytes(), 256-B aligned. */
size_t m3r_match_workspace_bytes(int b, int h, int w, int fdim, int desc_is_f32);
int m3r_match_fused(const float* X11, const float* X21, const void* D11, const void* D21, int fdim,
                    int desc_is_f32, const int64_t* idx_init, int b, int h, int w, int max_iter,
                    float lambda_init, float cost_thresh, float dist_thresh, int radius, int dilation_max,
                    int64_t* idx_out, uint8_t* valid_out, void* workspace, size_t workspace_bytes,
                    void* stream);

/* ---- Gauss-Newton ---------------------------------------------------------------------- */

/* Edge sharding for multi-GPU global optimisation (additive; NULL = single GPU, all edges).
 * A rank owns directed edges [edge_lo, edge_hi) of the FULL edge list.  idx/valid/Q passed to
 * the op then hold ONLY those edges' slabs ([edge_hi-edge_lo, n]); ii/jj stay full-length so
 * every rank derives the same compact indices.  `comm` is an ncclComm_t (from m3r_comm_init);
 * once per iteration the per-edge 7x7 blocks + 7-vectors (fp64) are exchanged.  HOW is part of the
 * shard description and must be THE SAME VALUE ON EVERY RANK (it selects the NCCL collective; ranks
 * that disagree would hang):
 *   M3R_EXCHANGE_ALLREDUCE  any contiguous partition: foreign slots zeroed, sum all-reduce (a sum with
 *                           zeros is exact).  The safe default.
 *   M3R_EXCHANGE_ALLGATHER  in-place ncclAllGather of the own slots (half the bytes, no memsets).  Legal
 *                           only for the equal partition: world | num_edges and rank r owns
 *                           [r * num_edges / world, (r + 1) * num_edges / world).  A rank whose range is not
 *                           that returns M3R_ERR_BAD_ARG before anything is enqueued.
 *   M3R_EXCHANGE_PEER       no NCCL: the library's own all-gather kernel over NVLink peer memory (m3r_xchg_*
 *                           below; `xchg` must be a connected exchange, `comm` is ignored).  Same equal-partition
 *                           rule as ALLGATHER.  The edge blocks of iteration it live in parity it & 1 of the
 *                           exchange buffer instead of the workspace (m3r_gn_edge_blocks / export do not see them).
 * dist.exchange_for_partition() derives the value from the global partition, which every rank knows. */
#define M3R_EXCHANGE_ALLREDUCE 0
#define M3R_EXCHANGE_ALLGATHER 1
#define M3R_EXCHANGE_PEER 2
typedef struct m3r_gn_shard {
  int edge_lo;
  int edge_hi;
  void* comm;
  int exchange; /* M3R_EXCHANGE_*; ignored when comm == NULL and xchg == NULL */
  void* xchg;   /* M3R_EXCHANGE_PEER: from m3r_xchg_create + m3r_xchg_connect; NULL otherwise (ABI v4) */
} m3r_gn_shard;

typedef struct m3r_gn_params {
  int kind;            /* M3R_GN_POINTS / RAYS / CALIB                                        */
  float sigma_a;       /* points: sigma_point | rays: sigma_ray   | calib: sigma_pixel        */
  float sigma_b;       /* points: unused      | rays: sigma_dist  | calib: sigma_depth        */
  float C_thresh;
  float Q_thresh;
  int max_iter;
  float delta_thresh;
  /* calib only */
  int height, width, pixel_border;
  float z_eps;
} m3r_gn_params;

/* bytes of device workspace needed for a problem with num_kf keyframes and num_edges directed
 * edges (full list) of num_pts points. 256-byte aligned pointer required. */
size_t m3r_gn_workspace_bytes(int num_kf, int num_edges, int num_pts);
/* same for a rank that owns only `edges_local` of the edges (the 8 B/pt packed slabs, the bulk of
 * the workspace, are kept for owned edges only) */
size_t m3r_gn_workspace_bytes_sharded(int num_kf, int num_edges, int num_pts, int edges_local);

/* Full GN loop, asynchronous on `stream`, zero host round-trips.
 *   Twc  [num_kf,8] f32  IN/OUT (t, q=xyzw, s), pose 0 fixed (num_fix = 1, gn_kernels.cu:1157)
 *   Xs   [num_kf,n,3] f32, Cs [num_kf,n] f32, K [3,3] f32 row-major (calib only, else NULL)
 *   ii,jj [num_edges] i64 keyframe ids (any ids < M3R_MAX_KF_ID; rows of Twc/Xs/Cs are in
 *         sorted-unique id order exactly as the reference expects)
 *   idx  [E_local,n] i64, valid [E_local,n] u8, Q [E_local,n] f32
 *   dx_out [num_kf-1,7] f32 : dx of the last executed iteration (zeros if max_iter == 0)
 *   info_out (nullable) device int32[4]: {iterations executed, cholesky failures, stopped early,
 *         bad-index flag}
 * Non-SPD system => dx = 0 for that iteration, poses untouched (gn_kernels.cu:147-150). */
int m3r_gauss_newton(float* Twc, const float* Xs, const float* Cs, const float* K,
                     const int64_t* ii, const int64_t* jj, const int64_t* idx,
                     const uint8_t* valid, const float* Q, int num_kf, int num_pts,
                     int num_edges, const m3r_gn_params* params, float* dx_out, int32_t* info_out,
                     void* workspace, size_t workspace_bytes, const m3r_gn_shard* shard,
                     void* stream);

/* Building blocks of one GN iteration, exposed for tests / profiling (same stream semantics).
 * m3r_gn_prepare must be called once per problem before the others (index compaction + the
 * per-pose incidence lists); it also resets the iteration/stop state. */
int m3r_gn_prepare(const int64_t* ii, const int64_t* jj, int num_kf, int num_pts, int num_edges,
                   void* workspace, size_t workspace_bytes, void* stream);
/* pack: folds everything that does not depend on the poses (valid_match, Q > Q_thresh, C_i[m] and
 * C_j[k] > C_thresh, sqrt(Q)) into one 8-byte record per (owned edge, point) in the workspace.
 * Once per problem, after m3r_gn_prepare; the iterations then stream 8 B/pt/edge instead of 13 B
 * plus two confidence reads. */
int m3r_gn_pack(const float* Cs, const int64_t* idx, const uint8_t* valid, const float* Q,
                int num_kf, int num_pts, int num_edges, const m3r_gn_params* params,
                void* workspace, size_t workspace_bytes, const m3r_gn_shard* shard, void* stream);
/* accumulate (after m3r_gn_pack): writes the per-edge fp64 blocks EB[num_edges][56] =
 * {A_e (7x7 row-major), b_e (7)} for edges [edge_lo,edge_hi) and zeros for the others
 * (shard == NULL: all edges); exchanges them (see m3r_gn_shard) when shard->comm is set. */
int m3r_gn_accumulate(const float* Twc, const float* Xs, const float* K, int num_kf, int num_pts,
                      int num_edges, const m3r_gn_params* params, void* workspace,
                      size_t workspace_bytes, const m3r_gn_shard* shard, void* stream);
/* copy the per-edge blocks out (device double[num_edges*56]) */
int m3r_gn_edge_blocks(const void* workspace, int num_kf, int num_edges, int num_pts,
                       double** blocks_out);
/* assemble + Cholesky + solve + retract + stop test, from the blocks in the workspace */
int m3r_gn_solve_retract(float* Twc, int num_kf, int num_pts, int num_edges, float delta_thresh,
                         float* dx_out, int32_t* info_out, void* workspace,
                         size_t workspace_bytes, void* stream);
/* reference-layout export of the blocks: Hs [4,E,7,7] f32, gs [2,E,7] f32 (gn_kernels.cu:1096-1137) */
int m3r_gn_export_hs_gs(const void* workspace, int num_kf, int num_edges, int num_pts, float* Hs,
                        float* gs, void* stream);

/* pose retraction alone: T_k <- exp(dx[k-num_fix]) * T_k for k >= num_fix (gn_kernels.cu:415-453) */
int m3r_pose_retr(float* Twc, const float* dx, int num_kf, int num_fix, void* stream);

/* ---- factor-graph staging (additive) -------------------------
 * m3r_edge_admit replaces the tensor arithmetic of FactorGraph.add_factors that follows the symmetric
 * matcher (mast3r_slam/global_opt.py:54-67) and the slab growth by torch.cat (:89-96, :104-110).
 * For each of `batch` new keyframe pairs b and both directions it computes
 *     Q[k] = sqrt(Q_src[b, idx[b,k]] * Q_dst[b,k])      dir 0: idx_i2j, Qii, Qji   dir 1: idx_j2i, Qjj, Qij
 * writes {idx, valid, Q} into the caller's edge-store slabs (row-major [2*capacity, num_pts]) at
 * directed slot 2*(slot0+b)+dir, and counts_out[2*b+dir] = #{k: valid && Q > Q_conf} (the numerator
 * of match_frac; the accept/reject decision of :69-76 stays with the host).  The slabs are laid out
 * exactly as m3r_gauss_newton's idx / valid / Q arguments, so a solve reads them in place.
 * All inputs [batch, num_pts] contiguous; int32 counts_out[2*batch] is zeroed by the call. */
int m3r_edge_admit(const int64_t* idx_i2j, const int64_t* idx_j2i, const uint8_t* valid_j, const uint8_t* valid_i,
                   const float* Qii, const float* Qjj, const float* Qji, const float* Qij, int batch, int num_pts,
                   float Q_conf, int64_t* store_idx, uint8_t* store_valid, float* store_Q, int64_t slot0,
                   int32_t* counts_out, void* stream);

/* constrain_points_to_ray (mast3r_slam/geometry.py:37-42, :107-114): out[kf,k] = z * ((u-cx)/fx, (v-cy)/fy, 1)
 * with z = Xs[kf,k,2], u = k mod width, v = k div width, K row-major 3x3 on the device.  out may alias Xs. */
int m3r_constrain_points_to_ray(const float* Xs, float* out, const float* K, int num_kf, int height, int width,
                                void* stream);

/* ---- frame tracker pose optimisation (additive) -------------------------
 * Replaces FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 and FrameTracker.solve
 * (mast3r_slam/tracker.py:141-154, :156-214, :216-270) with one persistent cooperative launch.
 *   T_WCf  [8] in/out  frame pose (t, q=xyzw, s); on return T_WCk * T_CkCf        (:192, :267)
 *   T_CkCf_out [8]     optimised relative pose (may be NULL)
 *   Xf [n,3]           frame points ALREADY gathered by idx_f2k (what get_points_poses returns, :139)
 *   Xk [n,3]           keyframe points; valid [n] = valid_opt (:67); Q [n] = Qk (:41)
 *   kind               M3R_GN_RAYS (sigma_a = sigma_ray, sigma_b = sigma_dist) or
 *                      M3R_GN_CALIB (sigma_a = sigma_pixel, sigma_b = sigma_depth; K, height, width,
 *                      pixel_border, z_eps = depth_eps; Xf / Xk constrained to their rays by the caller)
 *   stop tests         after each update: |tau| < delta_norm  or  |old_cost - new_cost| / old_cost < rel_error
 *                      (nonlinear_optimizer.py:5-26), at most max_iter iterations
 *   dx_out [7]         last update; info_out int32[4] = {iterations, cholesky failures, stopped early, 0}
 * A Cholesky failure leaves the pose of that iteration unchanged and is counted in info_out[1] (the reference
 * raises and skips the frame, tracker.py:92-94: the host mirror does the same from info_out). */
typedef struct m3r_track_params {
  int kind;
  float sigma_a, sigma_b;
  int max_iter;
  float rel_error, delta_norm, huber; /* huber must be 1.345 */
  int height, width, pixel_border;
  float z_eps;
} m3r_track_params;
size_t m3r_track_workspace_bytes(int num_pts);
int m3r_track_frame(float* T_WCf, float* T_CkCf_out, const float* T_WCk, const float* Xf, const float* Xk,
                    const uint8_t* valid, const float* Q, const float* K, int num_pts, const m3r_track_params* params,
                    float* dx_out, int32_t* info_out, void* workspace, size_t workspace_bytes, void* stream);

/* ---- NCCL plumbing (resolved with dlopen; no link-time dependency) ------------------------ */
#define M3R_NCCL_UNIQUE_ID_BYTES 128
int m3r_comm_unique_id(void* id_out /* 128 bytes, host */);
int m3r_comm_init(int rank, int world, const void* id /* 128 bytes, host */, void** comm_out);
int m3r_comm_destroy(void* comm);
/* in-place sum all-reduce of `count` doubles (used by the GN loop; exposed for tests) */
int m3r_comm_allreduce_f64(void* comm, double* buf, size_t count, void* stream);

/* Peer-memory exchange (no counterpart in the reference, which is single-GPU; replaces the NCCL call of the sharded
 * Gauss-Newton loop by ONE kernel of this library: 16-byte stores into every peer's buffer over NVLink, a
 * release/acquire flag per peer, csrc/peer_exchange.cu).  One process per GPU on one node:
 *   m3r_xchg_create   allocates this rank's exchange buffer (2 x capacity_bytes + flags, cudaMalloc -- set-up time,
 *                     not the hot path) on the current device and returns its 64-byte CUDA IPC handle (host memory);
 *                     capacity_bytes >= num_edges * 56 * 8 of the largest graph to be solved.
 *   m3r_xchg_connect  takes the `world` handles in rank order (exchanged by the caller, e.g. torch.distributed
 *                     all_gather) and maps the peers' buffers (needs P2P access between the GPUs).
 *   m3r_xchg_destroy  after a barrier across the ranks (a peer may still be writing otherwise). */
#define M3R_IPC_HANDLE_BYTES 64
int m3r_xchg_create(size_t capacity_bytes, void** xchg_out, void* ipc_handle_out /* 64 bytes, host */);
int m3r_xchg_connect(void* xchg, int rank, int world, const void* all_handles /* world x 64 bytes, host */);
int m3r_xchg_destroy(void* xchg);

#ifdef __cplusplus
}
#endif
#endif /* M3R_ABI_H */
