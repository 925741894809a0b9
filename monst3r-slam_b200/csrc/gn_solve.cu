// On-device linear algebra of the Gauss-Newton loop (sm_100a): index compaction, block-sparse
// assemble, dense fp64 Cholesky, triangular solves, Sim(3) retraction and the stop test.
//
// Replaces HOST code of the reference: get_unique_kf_idx / create_inds (gn_kernels.cu:161-170),
// SparseBlock::update_lhs / update_rhs / solve with Eigen::SimplicialLLT<double> (:57-159), plus
// pose_retr_kernel (:415-453) and the linalg_norm().item() stop test (:1219-1222).  The reference
// does two host round-trips per iteration; here one iteration is  accumulate -> (all-reduce) ->
// gn_solve_kernel  with no host involvement, so the whole loop can sit in a CUDA graph.
//
// gn_solve_kernel (cooperative when the system has more than a few tiles, one CTA otherwise):
//   A. assemble   each CTA owns pose rows; it zero-fills its 7-row strip of the dense lower
//                 triangle and adds the incident edge blocks in edge order (deterministic, no atomics)
//                 H_pp += A_e, H_pq -= A_e (q < p), g_p -+= b_e.  g is stored as one extra row
//                 below H, so the factorisation also produces y = L^-1 g (bordered-matrix trick).
//   B. Cholesky   right-looking, 32x32 tiles, fp64.  From 8 tiles on (round 2, "v4"): ONE grid barrier per tile
//                 step.  A dedicated chain CTA factors tile (k+1,k+1) (one-warp LDL-ordered potrf), inverts it by
//                 16x16 blocks and publishes L^-1; every other CTA owns a contiguous chunk of the trailing tiles
//                 and forms the panel tiles it needs itself as DMMA products with L_kk^-1 (operands streamed
//                 through a cp.async ring), writing the final L_Ik to the mirror slot in the upper triangle.
//                 "v2" (M3R_SOLVE_NOV4=1) keeps a separate panel phase with forward substitution and two barriers
//                 per step; below 8 tiles (and on the round-1 fallback path) every CTA factors + inverts the
//                 diagonal tile redundantly.  Systems of up to 7 tiles never get here: gn_solve_resident.cu.
//   C. back-substitution  L^T x = y as a barrier-free column-block dataflow over the CTAs with the L_kk^-1 tiles
//                 (v2 forms them after the factorisation, all tiles in parallel).
//   D. dx = -x (0 if a pivot was <= 0: gn_kernels.cu:142-150), retraction T <- Exp(dx) T in fp32
//      following :299-413, ||dx|| stop test, next-iteration bookkeeping.
#include <cooperative_groups.h>

#include <cstdlib>

#include <atomic>

#include "gn_tile.cuh"

namespace cg = cooperative_groups;

namespace m3r {
namespace {

// =============================================================================== prepare kernel
// compact index of a keyframe id = number of distinct ids below it (== searchsorted(unique(ids)))
__global__ void __launch_bounds__(1024)
gn_prepare_kernel(const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int E, int N,
                  int* __restrict__ ci, int* __restrict__ cj, unsigned* __restrict__ bitmap,
                  unsigned* __restrict__ prefix, int* __restrict__ inc_off, int* __restrict__ inc_list,
                  int* __restrict__ order, unsigned* __restrict__ counters, int* __restrict__ ctrl,
                  float* __restrict__ zero_dx, int n_dx, int32_t* __restrict__ zero_info,
                  uint4* __restrict__ zero_ll, int n_ll16, int ll_gen) {
  // presence bitmap over [id_min, id_max] only: keyframe ids are a few hundred apart at most, so the
  // scan touches a handful of words instead of the whole M3R_MAX_KF_ID range
  const int tid = threadIdx.x, nt = blockDim.x;
  __shared__ unsigned s_scan[1024];
  __shared__ int s_bad;
  __shared__ long long s_min, s_max;
  if (tid == 0) {
    s_bad = 0;
    s_min = M3R_MAX_KF_ID;
    s_max = -1;
  }
  for (int k = tid; k <= E; k += nt) counters[k] = 0u;
  const bool clear_ll = n_ll16 > 0 && ctrl[CTRL_LL_GEN] != ll_gen;  // (every thread reads before anyone writes: barrier below)
  __syncthreads();
  if (tid < CTRL_WORDS && tid != CTRL_LL_GEN) ctrl[tid] = 0;
  if (clear_ll && tid == 0) ctrl[CTRL_LL_GEN] = ll_gen;
  // outputs of the solve that follows start from zero (max_iter == 0, or a solve that never runs): folded in
  // here instead of two cudaMemsetAsync launches in front of every Gauss-Newton call
  if (zero_dx)
    for (int k = tid; k < n_dx; k += nt) zero_dx[k] = 0.f;
  if (zero_info && tid < 4) zero_info[tid] = 0;
  // The persistent tracking kernel's flag words ({tag | payload}, gn_accum.cu) live in caller memory that is never
  // initialised and outlives the launch; the tag carries a 20-bit launch nonce.  They are cleared (tag 0 is never
  // waited for) when this workspace is first used and whenever the nonce has wrapped since the last clear, so a stale
  // word can never carry the tag a later launch waits for.
  if (clear_ll)
    for (int k = tid; k < n_ll16; k += nt) zero_ll[k] = make_uint4(0u, 0u, 0u, 0u);
  if (2 * E <= 32 && N <= 32) {
    // Small graphs (tracking: 2 edges): everything in warp 0 with shuffles, no block barrier, no pass through global
    // memory — same outputs as the general path below (which costs ~10 us of barriers for a two-edge graph).
    if (tid >= 32) return;
    __syncwarp();  // the zeroing stores above (lanes < 16 own ctrl) before lane 0 writes ctrl[CTRL_BAD]
    const unsigned full = 0xffffffffu;
    const int lane = tid;
    const bool has = lane < 2 * E;
    const long long id = has ? (lane < E ? ii[lane] : jj[lane - E]) : -1;
    const bool ok = has && id >= 0 && id < M3R_MAX_KF_ID;
    int bad = __any_sync(full, has && !ok) ? 1 : 0;
    bool first = ok;
    for (int m = 0; m < 2 * E; m++) {
      const long long om = __shfl_sync(full, id, m);
      const bool okm = __shfl_sync(full, (int)ok, m) != 0;
      if (okm && m < lane && om == id) first = false;
    }
    const unsigned firstmask = __ballot_sync(full, first);
    int c = 0;
    for (int m = 0; m < 2 * E; m++) {
      const long long om = __shfl_sync(full, id, m);
      if (((firstmask >> m) & 1u) && om < id) c++;
    }
    if (!ok) c = 0;
    const bool over = ok && c >= N;
    if (over) c = 0;
    bad |= __any_sync(full, over) ? 1 : 0;
    if (has) (lane < E ? ci : cj)[lane < E ? lane : lane - E] = c;
    if (lane == 0) ctrl[CTRL_BAD] = bad | ((__popc(firstmask) != N) ? 2 : 0);
    // per edge (lanes e < E): compact endpoints, launch order by (min, max, e)
    const int ea = c, eb2 = __shfl_sync(full, c, (lane + E) & 31);
    const long long key = (long long)min(ea, eb2) * N + max(ea, eb2);
    int rank = 0;
    for (int f = 0; f < E; f++) {
      const long long kf = __shfl_sync(full, key, f);
      rank += (kf < key) || (kf == key && f < lane);
    }
    if (lane < E) order[rank] = lane;
    // incidence lists per compact pose (lanes c < N)
    int cnt = 0;
    for (int e = 0; e < E; e++) {
      const int a = __shfl_sync(full, ea, e), b = __shfl_sync(full, eb2, e);
      if (a != b) cnt += (a == lane) + (b == lane);
    }
    if (lane >= N) cnt = 0;
    int incl = cnt;
    for (int off = 1; off < 32; off <<= 1) {
      const int v = __shfl_up_sync(full, incl, off);
      if (lane >= off) incl += v;
    }
    if (lane == 0) inc_off[0] = 0;
    if (lane < N) inc_off[lane + 1] = incl;
    int o = incl - cnt;
    for (int e = 0; e < E; e++) {
      const int a = __shfl_sync(full, ea, e), b = __shfl_sync(full, eb2, e);
      if (a == b || lane >= N) continue;
      if (a == lane) inc_list[o++] = (e << 1) | 0;
      if (b == lane) inc_list[o++] = (e << 1) | 1;
    }
    return;
  }
  __syncthreads();
  {
    long long lo = M3R_MAX_KF_ID, hi = -1;
    for (int k = tid; k < 2 * E; k += nt) {
      const int64_t id = k < E ? ii[k] : jj[k - E];
      if (id < 0 || id >= M3R_MAX_KF_ID) {
        s_bad = 1;
      } else {
        lo = id < lo ? id : lo;
        hi = id > hi ? id : hi;
      }
    }
    for (int off = 16; off > 0; off >>= 1) {
      const long long l2 = __shfl_xor_sync(0xffffffffu, lo, off), h2 = __shfl_xor_sync(0xffffffffu, hi, off);
      lo = l2 < lo ? l2 : lo;
      hi = h2 > hi ? h2 : hi;
    }
    if ((tid & 31) == 0) {
      atomicMin(&s_min, lo);
      atomicMax(&s_max, hi);
    }
  }
  __syncthreads();
  const long long base = s_max >= 0 ? (s_min & ~31LL) : 0;  // first id of bitmap word 0
  const int W = s_max >= 0 ? (int)((s_max - base) / 32 + 1) : 1;
  for (int k = tid; k < W; k += nt) bitmap[k] = 0u;
  __syncthreads();
  for (int k = tid; k < 2 * E; k += nt) {
    const int64_t id = k < E ? ii[k] : jj[k - E];
    if (id >= 0 && id < M3R_MAX_KF_ID) atomicOr(bitmap + ((id - base) >> 5), 1u << ((id - base) & 31));
  }
  __syncthreads();
  // exclusive prefix of popcounts over W words: each thread owns `per` consecutive words
  const int per = (W + nt - 1) / nt;
  unsigned local = 0;
  for (int k = 0; k < per; k++) {
    const int wd = tid * per + k;
    if (wd < W) local += __popc(bitmap[wd]);
  }
  s_scan[tid] = local;
  __syncthreads();
  for (int off = 1; off < nt; off <<= 1) {
    unsigned v = tid >= off ? s_scan[tid - off] : 0u;
    __syncthreads();
    s_scan[tid] += v;
    __syncthreads();
  }
  unsigned run = s_scan[tid] - local;
  for (int k = 0; k < per; k++) {
    const int wd = tid * per + k;
    if (wd < W) {
      prefix[wd] = run;
      run += __popc(bitmap[wd]);
    }
  }
  __syncthreads();
  const unsigned total = s_scan[nt - 1];
  for (int k = tid; k < E; k += nt) {
    const int64_t a = ii[k], b = jj[k];
    int ca = 0, cb = 0;
    if (a >= 0 && a < M3R_MAX_KF_ID) {
      const long long r = a - base;
      ca = prefix[r >> 5] + __popc(bitmap[r >> 5] & ((1u << (r & 31)) - 1u));
    }
    if (b >= 0 && b < M3R_MAX_KF_ID) {
      const long long r = b - base;
      cb = prefix[r >> 5] + __popc(bitmap[r >> 5] & ((1u << (r & 31)) - 1u));
    }
    if (ca >= N) { ca = 0; s_bad = 1; }
    if (cb >= N) { cb = 0; s_bad = 1; }
    ci[k] = ca;
    cj[k] = cb;
  }
  __syncthreads();
  if (tid == 0) ctrl[CTRL_BAD] = s_bad | ((total != (unsigned)N) ? 2 : 0);
  // launch order for the accumulate kernel: edges sorted by (min, max) keyframe, ties by edge id, so
  // that (i,j) and (j,i) are adjacent and share Xs[i] / Xs[j] in L2 (rank sort, once per solve)
  for (int e = tid; e < E; e += nt) {
    const long long key = (long long)min(ci[e], cj[e]) * N + max(ci[e], cj[e]);
    int rank = 0;
    for (int f = 0; f < E; f++) {
      const long long kf = (long long)min(ci[f], cj[f]) * N + max(ci[f], cj[f]);
      rank += (kf < key) || (kf == key && f < e);
    }
    order[rank] = e;
  }
  // incidence lists per compact pose c: entries (e << 1 | role) in ascending e; role 0: c == ci[e]
  // two passes (count, fill); self-edges contribute nothing (H_ii + H_jj - H_ij - H_ji = 0).
  for (int c = tid; c < N; c += nt) {
    int cnt = 0;
    for (int e = 0; e < E; e++) {
      const int a = ci[e], b = cj[e];
      if (a != b) cnt += (a == c) + (b == c);
    }
    inc_off[c + 1] = cnt;
  }
  __syncthreads();
  if (tid == 0) {
    inc_off[0] = 0;
    for (int c = 0; c < N; c++) inc_off[c + 1] += inc_off[c];
  }
  __syncthreads();
  for (int c = tid; c < N; c += nt) {
    int o = inc_off[c];
    for (int e = 0; e < E; e++) {
      const int a = ci[e], b = cj[e];
      if (a == b) continue;
      if (a == c) inc_list[o++] = (e << 1) | 0;
      if (b == c) inc_list[o++] = (e << 1) | 1;
    }
  }
}

// =============================================================================== Sim(3) retraction
// (sim3_retract lives in gn_device.cuh: it is shared with the fused tail of the accumulate kernel)
__global__ void pose_retr_kernel(float* __restrict__ Twc, const float* __restrict__ dx, int num_kf, int num_fix) {
  for (int k = num_fix + blockIdx.x * blockDim.x + threadIdx.x; k < num_kf; k += gridDim.x * blockDim.x) {
    float T[8], xi[7];
#pragma unroll
    for (int c = 0; c < 8; c++) T[c] = Twc[(size_t)k * 8 + c];
#pragma unroll
    for (int c = 0; c < 7; c++) xi[c] = dx[(size_t)(k - num_fix) * 7 + c];
    sim3_retract(xi, T);
#pragma unroll
    for (int c = 0; c < 8; c++) Twc[(size_t)k * 8 + c] = T[c];
  }
}

// =============================================================================== solve kernel
struct SolveArgs {
  float* Twc;
  const double* eb;
  const int* ci;
  const int* cj;
  const int* inc_off;
  const int* inc_list;
  double* hd;
  double* linv;
  int* ctrl;
  float* dx_out;
  int32_t* info_out;
  int N, dim, nblk, ld;
  float delta_thresh;
  // dataflow factorisation (gn_solve_df_kernel): tile flags and the P x Q grid of worker CTAs
  int* done;
  int* rdy;
  unsigned long long* ll;
  int P, Q;
  int v4;        // gn_solve_kernel: one barrier per tile step (panel tiles formed inside the update with L_kk^-1)
  double* ybuf;  // v4: y = L^-1 g, one block per step
  int pipe_off;  // gn_solve_kernel: offset (doubles) of the trailing update's cp.async tile pipeline in dynamic smem; 0 = none
};

// ---- pieces of the pipelined trailing update (gn_solve_kernel, B2) ------------------------------------------------
#ifndef M3R_UPD_STAGES
#define M3R_UPD_STAGES 3
#endif
constexpr int kUpdStages = M3R_UPD_STAGES;          // tiles in flight per CTA (24 KB each)
constexpr int kUpdStageDoubles = 3 * kTile * kTile; // L_Ik, L_Jk and the destination tile A_IJ
__device__ __forceinline__ void cp_async16(double* smem_dst, const double* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <bool COOP>
__device__ __forceinline__ void grid_barrier() {
  if (COOP) {
    cg::this_grid().sync();
  } else {
    __syncthreads();
  }
}

// Factor the 32x32 tile held in Ls (lower, stride 33) and invert the factor, all by WARP 0 with the tile in registers
// (lane r owns row r): no block barriers inside.  Ls <- L (upper part zeroed), Li <- L^-1.
//  * factorisation: right-looking in LDL^T order; per pivot the unscaled column goes through a double-buffered 32-double
//    shared array (1 STS + 16 broadcast LDS.128) and the chain carries one reciprocal; the square roots are applied to
//    whole columns at the end.  Measured in isolation (tools/ubench_solve.cu, profiles/r02a_ubench_solve.log): 7.4k
//    cycles against 12.5k for the round-1 version that broadcast the scaled column with 62 shuffles per pivot.
//  * inverse: row r of L^-T by forward substitution from the unit vector e_r (lane = row, broadcast reads of L).
// Returns (to every thread, through s_flag) false on a pivot <= 0.
__device__ __forceinline__ bool tile_potrf_invert(double (*Ls)[kTile + 1], double (*Li)[kTile + 1], double* s_dinv, int* s_flag,
                                                  double* s_colbuf /* 2 x 32 doubles, 16-byte aligned */) {
  const int lane = threadIdx.x;
  if (threadIdx.x < 32) {
    double a[kTile];
#pragma unroll
    for (int c = 0; c < kTile; c++) a[c] = Ls[lane][c];
    bool bad = false;
    double my_d = 1.0;
#pragma unroll
    for (int j = 0; j < kTile; j++) {
      double* cb = s_colbuf + (j & 1) * kTile;
      cb[lane] = a[j];
      __syncwarp();
      double d = cb[j];
      if (!(d > 0.0)) {
        bad = true;
        d = 1.0;
      }
      if (lane == j) my_d = d;
      const double tj = a[j] * (1.0 / d);
#pragma unroll
      for (int c = (j + 1) & ~1; c < kTile; c += 2) {
        const double2 v = *reinterpret_cast<const double2*>(cb + c);
        if (c > j) a[c] -= tj * v.x;
        a[c + 1] -= tj * v.y;
      }
    }
    const double rs = rsqrt(my_d);  // 1 / L[lane][lane]
    s_colbuf[lane] = rs;            // (lower half: last used by pivot 30, every lane is past pivot 31's barrier)
    s_dinv[lane] = rs;
    if (lane == 0) *s_flag = bad ? 1 : 0;
    __syncwarp();
#pragma unroll
    for (int c = 0; c < kTile; c++) Ls[lane][c] = (c <= lane) ? a[c] * s_colbuf[c] : 0.0;
    __syncwarp();
    // x = row `lane` of L^-T (= column `lane` of L^-1): x L^T = e_lane
    double x[kTile];
#pragma unroll
    for (int c = 0; c < kTile; c++) x[c] = (c == lane) ? 1.0 : 0.0;
#pragma unroll
    for (int c = 0; c < kTile; c++) {
      const double xc = x[c] * s_dinv[c];
      x[c] = xc;
#pragma unroll
      for (int c2 = c + 1; c2 < kTile; c2++) x[c2] -= xc * Ls[c2][c];
    }
#pragma unroll
    for (int c = 0; c < kTile; c++) Li[c][lane] = x[c];
  }
  __syncthreads();
  return *s_flag == 0;
}

// measured (tools/solve_prof.py): dim 1785 (56 tiles) 1.90 -> 1.66 ms with the look-ahead, dim 217 (7 tiles) 139 -> 160 us
constexpr int kLookaheadMinTiles = 8;

// One CTA per SM (round 2): the one-warp substitution / factorisation of the v2 chain keep a tile row in registers
// (64 + operands); at the 128 registers of two CTAs per SM they spilled (508 B) and the solve ran at 990 us, at 255
// registers it runs at 796 us (profiles/r02s).  The trailing update hides its L2 latency with cp.async instead of a
// second resident CTA.
#ifndef M3R_SOLVE_MINCTAS
#define M3R_SOLVE_MINCTAS 1
#endif
template <bool COOP>
__global__ void __launch_bounds__(kSolveThreads, M3R_SOLVE_MINCTAS)
gn_solve_kernel(const SolveArgs a) {
  if (a.ctrl[CTRL_DONE]) return;  // uniform over the grid: set only by a previous launch
  __shared__ double sA[kTile][kTile + 1];
  __shared__ double sB[kTile][kTile + 1];
  __shared__ int s_flag;
  __shared__ double s_dinv[kTile];
  __shared__ __align__(16) double s_colbuf[2 * kTile];
  __shared__ double s_red[kSolveThreads / 32];
  extern __shared__ double s_x[];  // dimP doubles (CTA 0, phase C)

  const int tid = threadIdx.x;
  const int nblk = a.nblk, ld = a.ld, dim = a.dim, dimP = a.nblk * kTile;
  double* H = a.hd;
  const int nfree = a.N - 1;

  if (blockIdx.x == 0 && tid == 0) a.ctrl[CTRL_FAIL_NOW] = 0;  // set by the look-ahead factorisation (phase B)
  // ---------------------------------------------------------------- A. assemble
  // strips: 0..nfree-1 pose strips (7 rows), then padding rows [dim, dimP), then the rhs tile row
  const int nstrips = nfree + 2;
  for (int sidx = blockIdx.x; sidx < nstrips; sidx += gridDim.x) {
    if (sidx < nfree) {
      const int p = sidx;
      const int r0 = 7 * p, ncol = 7 * p + 7;
      for (int t = tid; t < 7 * ncol; t += kSolveThreads) H[(size_t)(r0 + t / ncol) * ld + (t % ncol)] = 0.0;
      __syncthreads();
      const int c = p + 1;  // compact pose index
      const int lo = a.inc_off[c], hi = a.inc_off[c + 1];
      for (int q = lo; q < hi; q++) {
        const int ent = a.inc_list[q];
        const int e = ent >> 1, role = ent & 1;
        const int other = (role == 0 ? a.cj[e] : a.ci[e]) - 1;
        const double* blk = a.eb + (size_t)e * kEdgeBlock;
        if (tid < 49) {
          const int r = tid / 7, cc = tid % 7;
          const double v = blk[tid];
          H[(size_t)(r0 + r) * ld + (r0 + cc)] += v;
          if (other >= 0 && other < p) H[(size_t)(r0 + r) * ld + (7 * other + cc)] -= v;
        }
      }
      __syncthreads();
    } else if (sidx == nfree) {
      for (int t = tid; t < (dimP - dim) * ld; t += kSolveThreads) {
        const int r = dim + t / ld, cc = t % ld;
        H[(size_t)r * ld + cc] = (r == cc) ? 1.0 : 0.0;
      }
    } else {
      // rhs tile row: row dimP holds g, rows dimP+1.. are zero
      for (int t = tid; t < kTile * ld; t += kSolveThreads) H[(size_t)dimP * ld + t] = 0.0;
      __syncthreads();
      for (int p = tid; p < nfree; p += kSolveThreads) {
        const int c = p + 1;
        double g[7] = {0, 0, 0, 0, 0, 0, 0};
        for (int q = a.inc_off[c]; q < a.inc_off[c + 1]; q++) {
          const int ent = a.inc_list[q];
          const double* bv = a.eb + (size_t)(ent >> 1) * kEdgeBlock + 49;
          const double sgn = (ent & 1) ? 1.0 : -1.0;  // g_i = -b_e, g_j = +b_e
#pragma unroll
          for (int k = 0; k < 7; k++) g[k] += sgn * bv[k];
        }
#pragma unroll
        for (int k = 0; k < 7; k++) H[(size_t)dimP * ld + 7 * p + k] = g[k];
      }
    }
  }
  grid_barrier<COOP>();

#ifdef M3R_TRACK_PROF
  long long trk_t = clock64();
#define SLV(slot) do { if (blockIdx.x == 0 && tid == 0) { const long long n_ = clock64(); a.ctrl[6 + (slot)] += (int)((n_ - trk_t) >> 4); trk_t = n_; } } while (0)
#else
#define SLV(slot) do { } while (0)
#endif
  // ---------------------------------------------------------------- B. Cholesky
  bool failed = false;
  // Look-ahead: from step 1 on the diagonal tile is factored and inverted DURING the previous step's trailing
  // update by one dedicated CTA (the last one), which updates tile (kb+1, kb+1) first and takes no other tile,
  // so the ~8 us serial potrf + inverse leaves the critical path whenever the update has more than a few tiles
  // per CTA.  The other CTAs then only read L_kk^-1 (8 KB) at the start of the step.
  const bool lookahead = COOP && gridDim.x > 1 && nblk >= kLookaheadMinTiles;
  const int diag_cta = (int)gridDim.x - 1;
  const bool is_diag = lookahead && (int)blockIdx.x == diag_cta;
  const int n_work = lookahead ? (int)gridDim.x - 1 : (int)gridDim.x;  // CTAs that share the bulk work
  // v2 (round 2) of the look-ahead path: no tile inverse on the chain.  The diagonal CTA factors tile (kb+1, kb+1) with
  // the LDL-ordered one-warp potrf of gn_tile.cuh (3.9 us instead of 7 for potrf + inverse), its symmetric update is 16
  // DMMA per warp on swizzled shared-memory tiles with the destination tile prefetched before the barrier, and every
  // panel tile is solved by forward substitution against L_kk (one warp, the tile's rows in registers) instead of a
  // product with L_kk^-1.  The inverses the back-substitution wants are formed afterwards, all tiles in parallel.
  const bool v2 = lookahead && a.pipe_off > 0;
  const bool v4 = v2 && a.v4 != 0 && nblk <= (int)gridDim.x;  // (the back-substitution of v4 wants one CTA per tile column)
  const int lane = tid & 31, wrp = tid >> 5;
  double* const pipe = s_x + a.pipe_off;
  double* Lcur = pipe;                      // L_kk, swizzled (diagonal CTA: alternates with Dnext)
  double* Dnext = pipe + 2 * kTileEl;       // diagonal CTA: tile (kb+1, kb+1) being updated / factored
  double* const Xs = pipe + kTileEl;        // diagonal CTA: L_(kb+1, kb)
  double* const colb = pipe + 3 * kTileEl;  // potrf column buffer (2 x 32), then 1 / diag (32)
  double* const rinv = colb + 2 * kTile;
  // one warp: rows of panel tile (I, kb) <- rows * L_kk^-T, lane = row; optionally also into a swizzled smem tile
  auto load_rows = [&](double (&x)[kTile], int I, int kb_) {
    const double* row = &H[(size_t)(I * kTile + lane) * ld + kb_ * kTile];
#pragma unroll
    for (int c = 0; c < kTile; c += 2) {
      const double2 v = __ldcg(reinterpret_cast<const double2*>(row + c));
      x[c] = v.x;
      x[c + 1] = v.y;
    }
  };
  auto trsm_loaded = [&](double (&x)[kTile], int I, int kb_, const double* Ltile, double* smem_out) {
    double* row = &H[(size_t)(I * kTile + lane) * ld + kb_ * kTile];
    trsm_rows_sw(x, Ltile, rinv);
#pragma unroll
    for (int c = 0; c < kTile; c += 2) {
      *reinterpret_cast<double2*>(row + c) = make_double2(x[c], x[c + 1]);
      if (smem_out) *reinterpret_cast<double2*>(smem_out + sw(lane, c)) = make_double2(x[c], x[c + 1]);
    }
  };
  auto trsm_tile = [&](int I, int kb_, const double* Ltile, double* smem_out) {
    double x[kTile];
    load_rows(x, I, kb_);
    trsm_rows_sw(x, Ltile, rinv);
    double* row = &H[(size_t)(I * kTile + lane) * ld + kb_ * kTile];
#pragma unroll
    for (int c = 0; c < kTile; c += 2) {
      *reinterpret_cast<double2*>(row + c) = make_double2(x[c], x[c + 1]);
      if (smem_out) *reinterpret_cast<double2*>(smem_out + sw(lane, c)) = make_double2(x[c], x[c + 1]);
    }
  };
  if (v4) {
    // ---- v4: ONE grid barrier per tile step -----------------------------------------------------------------------
    // The panel phase is gone: a CTA that updates tile (I,J) with column kb forms the two panel tiles it needs itself,
    // L_Ik = A_Ik L_kk^-T and L_Jk = A_Jk L_kk^-T, as DMMA products with the inverse of the diagonal tile (L_Ik is
    // reused along a row of its contiguous chunk of tiles).  The raw panel tiles are never overwritten -- concurrent
    // CTAs read them -- so the final L_Ik goes to the mirror slot (kb, I) in the unused upper triangle of H (written
    // by the CTA that owns the first tile of row I) and y_kb to ybuf.  The chain CTA meanwhile forms L_(kb+1,kb),
    // updates + factors + inverts tile (kb+1, kb+1) and publishes L^-1 for the next step.
#ifdef M3R_TRACK_PROF
    long long cpt[6] = {0, 0, 0, 0, 0, 0}, cp0 = 0;
#define CPS() do { cp0 = clock64(); } while (0)
#define CPT(i) do { const long long n_ = clock64(); cpt[i] += n_ - cp0; cp0 = n_; } while (0)
#else
#define CPS() do { } while (0)
#define CPT(i) do { } while (0)
#endif
    double* const LinvS = pipe + kUpdStages * kUpdStageDoubles;   // L_kk^-1, swizzled
    double* const XI = LinvS + kTileEl;                           // L_Ik of the current row
    double* const XJ = XI + kTileEl;                              // L_Jk
    double* const colb4 = XJ + kTileEl;                           // chain: potrf column buffer (64) + 1/diag (32)
    double* const rinv4 = colb4 + 2 * kTile;
    // chain CTA (it does not use the ring): its tiles of this step, Ak = (kb+1,kb) and Dk = (kb+1,kb+1), are in shared
    // memory already; R = raw (kb+2,kb), T = (kb+2,kb+1) and D2 = (kb+2,kb+2) are updated with column kb by warps 4-7
    // WHILE warp 0 factors Dk, and become the next step's Ak / Dk -- no load on the chain's critical path
    double* Ak = pipe;
    double* Dk = pipe + 2 * kTileEl;
    double* const Rr = pipe + 3 * kTileEl;
    double* Tn = pipe + 4 * kTileEl;
    double* D2 = pipe + 5 * kTileEl;
    double* const X2 = pipe + 6 * kTileEl;
    const int g = lane >> 2, t4 = lane & 3, mi = wrp >> 1, ni0 = (wrp & 1) * 2;
    auto g2s = [&](double* dst, const double* src, int ldsrc) {  // 32 x 32 tile -> swizzled smem, all threads
      for (int e = tid; e < kTileEl / 2; e += kSolveThreads) {
        const int r = e >> 4, c2 = (e & 15) * 2;
        *reinterpret_cast<double2*>(dst + sw(r, c2)) = __ldcg(reinterpret_cast<const double2*>(src + (size_t)r * ldsrc + c2));
      }
    };
    auto s2g = [&](double* dst, int lddst, const double* src) {  // swizzled smem tile -> global, all threads
      for (int e = tid; e < kTileEl / 2; e += kSolveThreads) {
        const int r = e >> 4, c2 = (e & 15) * 2;
        *reinterpret_cast<double2*>(dst + (size_t)r * lddst + c2) = *reinterpret_cast<const double2*>(src + sw(r, c2));
      }
    };
    // OUT (=|-=) A B^T on swizzled smem tiles, 8 warps x two 8x8 blocks
    auto cta_gemm = [&](double* OUT, const double* A, const double* B, bool subtract) {
      double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
      for (int kk = 0; kk < 8; kk++) {
        const int pc = (kk * 4 + t4) ^ (g << 2);
        const double av = A[(mi * 8 + g) * kTile + pc];
        dmma884(acc[0][0], acc[0][1], av, B[(ni0 * 8 + g) * kTile + pc]);
        dmma884(acc[1][0], acc[1][1], av, B[((ni0 + 1) * 8 + g) * kTile + pc]);
      }
#pragma unroll
      for (int n2 = 0; n2 < 2; n2++) {
        double2* d = reinterpret_cast<double2*>(OUT + (mi * 8 + g) * kTile + (((ni0 + n2) * 8 + 2 * t4) ^ (g << 2)));
        if (subtract) {
          double2 v = *d;
          v.x -= acc[n2][0];
          v.y -= acc[n2][1];
          *d = v;
        } else {
          *d = make_double2(acc[n2][0], acc[n2][1]);
        }
      }
    };
    // chain CTA: factor the tile in D (swizzled), write L to H(kk,kk), invert into LinvS, publish a.linv[kk]
    auto factor_invert_publish = [&](double* D, int kk, bool factor) {
      if (factor) {
        if (wrp == 0) {
          const bool bad = potrf_inplace(D, rinv4, colb4, lane);
          if (lane == 0) s_flag = bad ? 1 : 0;
        }
        __syncthreads();
      }
      CPT(2);
      for (int t = tid; t < kTileEl; t += kSolveThreads) {
        const int r = t >> 5, c = t & 31;
        if (c <= r) H[(size_t)(kk * kTile + r) * ld + kk * kTile + c] = D[sw(r, c)];
      }
      // L^-1 by 16 x 16 blocks, L = [A 0; C B]:  [A^-1 0; -B^-1 C A^-1  B^-1].  Warps 0 and 1 invert A and B side by
      // side (lane j < 16 = column j, forward substitution, four partial sums), then all threads form the lower-left
      // block with two 16^3 products.  (One warp doing the whole tile took 2.5 us of the chain's 8.7 us per step.)
      if (wrp < 2 && lane < 16) {
        const int o = wrp * 16;
        double x[16];
#pragma unroll
        for (int i = 0; i < 16; i++) {
          double s4[4] = {(i == lane) ? 1.0 : 0.0, 0.0, 0.0, 0.0};
#pragma unroll
          for (int m = 0; m < i; m++) s4[m & 3] -= D[sw(o + i, o + m)] * x[m];
          x[i] = ((s4[0] + s4[1]) + (s4[2] + s4[3])) * rinv4[o + i];
          LinvS[sw(o + i, o + lane)] = x[i];
        }
      }
      {  // upper-right block of the inverse is zero
        const int r = tid >> 4, c = tid & 15;
        LinvS[sw(r, 16 + c)] = 0.0;
      }
      __syncthreads();
      {
        const int r = tid >> 4, c = tid & 15;  // M = C A^-1 (into XJ, plain 16 x 16), then -B^-1 M
        double m = 0.0;
#pragma unroll
        for (int k = 0; k < 16; k++) m += D[sw(16 + r, k)] * LinvS[sw(k, c)];
        XJ[r * 16 + c] = m;
        __syncthreads();
        double v = 0.0;
#pragma unroll
        for (int k = 0; k < 16; k++) v += LinvS[sw(16 + r, 16 + k)] * XJ[k * 16 + c];
        LinvS[sw(16 + r, c)] = -v;
      }
      __syncthreads();
      CPT(3);
      for (int t = tid; t < kTileEl; t += kSolveThreads) a.linv[(size_t)kk * kTileEl + t] = LinvS[sw(t >> 5, t & 31)];
      if (s_flag) {
        failed = true;
        if (tid == 0) a.ctrl[CTRL_FAIL_NOW] = 1;
      }
    };
    // ---- prologue: the chain CTA factors and inverts tile (0,0)
    if (is_diag) {
      for (int t = tid; t < kTileEl; t += kSolveThreads) {
        const int r = t >> 5, c = t & 31;
        Dnext[sw(r, c)] = (c <= r) ? __ldcg(&H[(size_t)r * ld + c]) : 0.0;
      }
      __syncthreads();
      factor_invert_publish(Dnext, 0, true);
      if (nblk > 1) {  // the chain's tiles of step 0
        g2s(Ak, &H[(size_t)kTile * ld], ld);
        g2s(Dk, &H[(size_t)kTile * ld + kTile], ld);
      }
      __syncthreads();
    }
    SLV(0);
    grid_barrier<COOP>();
    SLV(2);
    for (int kb = 0; kb < nblk; kb++) {
      const int kn = kb + 1;
      if (is_diag) {
        // ---- chain (L_kk^-1 is still in LinvS from this CTA's own inversion)
        const int kq = kb + 2;  // kq == nblk: the rhs row
        const int w4 = wrp - 4;
        // OUT (=|-=) A B^T by warps 4-7 (8 rows x 32 columns each)
        auto gemm4 = [&](double* OUT, const double* A, const double* B, bool subtract) {
          double acc[4][2];
#pragma unroll
          for (int n2 = 0; n2 < 4; n2++) acc[n2][0] = acc[n2][1] = 0.0;
#pragma unroll
          for (int kk = 0; kk < 8; kk++) {
            const int pc = (kk * 4 + t4) ^ (g << 2);
            const double av = A[(w4 * 8 + g) * kTile + pc];
#pragma unroll
            for (int n2 = 0; n2 < 4; n2++) dmma884(acc[n2][0], acc[n2][1], av, B[(n2 * 8 + g) * kTile + pc]);
          }
#pragma unroll
          for (int n2 = 0; n2 < 4; n2++) {
            double2* d = reinterpret_cast<double2*>(OUT + (w4 * 8 + g) * kTile + ((n2 * 8 + 2 * t4) ^ (g << 2)));
            if (subtract) {
              double2 v = *d;
              v.x -= acc[n2][0];
              v.y -= acc[n2][1];
              *d = v;
            } else {
              *d = make_double2(acc[n2][0], acc[n2][1]);
            }
          }
        };
        auto bar4 = [&]() { asm volatile("bar.sync 1, 128;" ::: "memory"); };
        if (kn < nblk) {
          CPS();
          const bool has_q = kq <= nblk, has_d2 = kq < nblk;
          if (wrp >= 4 && has_q) {  // look-ahead tiles: in flight during the two products below
            for (int e = tid - 128; e < kTileEl / 2; e += 128) {
              const int r = e >> 4, c2 = (e & 15) * 2, so = sw(r, c2);
              cp_async16(Rr + so, &H[(size_t)(kq * kTile + r) * ld + kb * kTile + c2]);
              cp_async16(Tn + so, &H[(size_t)(kq * kTile + r) * ld + kn * kTile + c2]);
              if (has_d2) cp_async16(D2 + so, &H[(size_t)(kq * kTile + r) * ld + kq * kTile + c2]);
            }
            cp_async_commit();
          }
          CPT(0);
          cta_gemm(XI, Ak, LinvS, false);  // L_(kn,kb)
          __syncthreads();
          cta_gemm(Dk, XI, XI, true);
          __syncthreads();
          CPT(1);
          if (wrp == 0) {
            const bool bad = potrf_inplace(Dk, rinv4, colb4, lane);
            if (lane == 0) s_flag = bad ? 1 : 0;
          } else if (wrp < 4) {  // mirror slot (kb, kn) <- L_(kn,kb)
            for (int e = tid - 32; e < kTileEl / 2; e += 96) {
              const int r = e >> 4, c2 = (e & 15) * 2;
              *reinterpret_cast<double2*>(&H[(size_t)(kb * kTile + r) * ld + kn * kTile + c2]) = *reinterpret_cast<const double2*>(XI + sw(r, c2));
            }
          } else if (has_q) {
            cp_async_wait<0>();
            bar4();
            gemm4(X2, Rr, LinvS, false);  // L_(kq,kb)
            bar4();
            gemm4(Tn, X2, XI, true);      // (kq,kn) -= L_(kq,kb) L_(kn,kb)^T
            if (has_d2) gemm4(D2, X2, X2, true);
            bar4();
            for (int e = tid - 128; e < kTileEl / 2; e += 128) {
              const int r = e >> 4, c2 = (e & 15) * 2;
              // the workers of the next step read the raw tile (kq,kn) as a panel tile of column kn
              *reinterpret_cast<double2*>(&H[(size_t)(kq * kTile + r) * ld + kn * kTile + c2]) = *reinterpret_cast<const double2*>(Tn + sw(r, c2));
              if (has_d2)
                *reinterpret_cast<double2*>(&H[(size_t)(kb * kTile + r) * ld + kq * kTile + c2]) = *reinterpret_cast<const double2*>(X2 + sw(r, c2));
            }
            if (!has_d2 && tid - 128 < kTile) a.ybuf[kb * kTile + (tid - 128)] = X2[sw(0, tid - 128)];
          }
          __syncthreads();
          factor_invert_publish(Dk, kn, false);
          CPT(5);
          { double* tmp = Ak; Ak = Tn; Tn = tmp; }
          { double* tmp = Dk; Dk = D2; D2 = tmp; }
        } else {  // last step: y of the last block (row nblk has no tile left to hang it on)
          cta_gemm(XI, Ak, LinvS, false);
          __syncthreads();
          if (tid < kTile) a.ybuf[kb * kTile + tid] = XI[sw(0, tid)];
        }
        __syncthreads();
      } else {
        // ---- worker: a contiguous chunk of the tiles (I,J), kb < J <= I <= nblk, J < nblk, minus (kn,kn)
        const int rem = nblk - kb;
        const int ncols = rem - 1;
        const int ntri = ncols * (ncols + 1) / 2;
        const int ntiles = ntri + ncols;
        // tiles 0 = (kn,kn), 1 = (kb+2,kn) and (if it exists) 2 = (kb+2,kb+2) are the chain's
        const int nchain = 1 + (kb + 2 <= nblk ? 1 : 0) + (kb + 2 < nblk ? 1 : 0);
        const long long nw = ntiles > nchain ? ntiles - nchain : 0;
        const int t_lo = nchain + (int)(nw * blockIdx.x / n_work), t_hi = nchain + (int)(nw * (blockIdx.x + 1) / n_work);
        const int nmine = ncols > 0 ? t_hi - t_lo : 0;
        auto decode = [&](int t0, int& I, int& J) {
          int i, j;
          if (t0 < ntri) {
            i = (int)((sqrt(8.0 * (double)t0 + 1.0) - 1.0) * 0.5);
            while ((i + 1) * (i + 2) / 2 <= t0) i++;
            while (i * (i + 1) / 2 > t0) i--;
            j = t0 - i * (i + 1) / 2;
          } else {
            i = ncols;
            j = t0 - ntri;
          }
          I = kb + 1 + i;
          J = kb + 1 + j;
        };
        int I_iss = -1;  // row of the tile issued last
        auto issue = [&](int n) {
          if (n < nmine) {
            int I, J;
            decode(t_lo + n, I, J);
            double* st = pipe + (n % kUpdStages) * kUpdStageDoubles;
            const bool newrow = I != I_iss;
            I_iss = I;
#pragma unroll
            for (int q = 0; q < 2; q++) {
              const int e = tid + q * kSolveThreads, row = e >> 4, c2 = (e & 15) * 2;
              const int so = sw(row, c2);
              if (newrow) cp_async16(st + so, &H[(size_t)(I * kTile + row) * ld + kb * kTile + c2]);
              if (J != I) cp_async16(st + kTileEl + so, &H[(size_t)(J * kTile + row) * ld + kb * kTile + c2]);
              cp_async16(st + 2 * kTileEl + so, &H[(size_t)(I * kTile + row) * ld + J * kTile + c2]);
              if (n == 0) cp_async16(LinvS + so, &a.linv[(size_t)kb * kTileEl + row * kTile + c2]);
            }
          }
          cp_async_commit();
        };
        for (int n = 0; n < kUpdStages - 1; n++) issue(n);
        int I_cur = -1;
        for (int n = 0; n < nmine; n++) {
          CPS();
          issue(n + kUpdStages - 1);
          cp_async_wait<kUpdStages - 1>();
          __syncthreads();
          CPT(0);
          int I, J;
          decode(t_lo + n, I, J);
          double* st = pipe + (n % kUpdStages) * kUpdStageDoubles;
          if (I != I_cur) {
            I_cur = I;
            cta_gemm(XI, st, LinvS, false);  // L_Ik
            if (J == kn) {  // this CTA owns the first tile of the row: it publishes L_Ik
              __syncthreads();
              if (I < nblk) {
                s2g(&H[(size_t)(kb * kTile) * ld + I * kTile], ld, XI);
              } else if (tid < kTile) {
                a.ybuf[kb * kTile + tid] = XI[sw(0, tid)];
              }
            }
          }
          CPT(1);
          const double* xj = XI;
          if (J != I) {
            cta_gemm(XJ, st + kTileEl, LinvS, false);  // L_Jk
            xj = XJ;
          }
          __syncthreads();
          CPT(2);
          {
            const double* sd = st + 2 * kTileEl;
            double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
            for (int kk = 0; kk < 8; kk++) {
              const int pc = (kk * 4 + t4) ^ (g << 2);
              const double av = XI[(mi * 8 + g) * kTile + pc];
              dmma884(acc[0][0], acc[0][1], av, xj[(ni0 * 8 + g) * kTile + pc]);
              dmma884(acc[1][0], acc[1][1], av, xj[((ni0 + 1) * 8 + g) * kTile + pc]);
            }
            const int row = mi * 8 + g;
#pragma unroll
            for (int n2 = 0; n2 < 2; n2++) {
              const int col = (ni0 + n2) * 8 + 2 * t4;
              double2 d = *reinterpret_cast<const double2*>(sd + row * kTile + (col ^ (g << 2)));
              d.x -= acc[n2][0];
              d.y -= acc[n2][1];
              *reinterpret_cast<double2*>(&H[(size_t)(I * kTile + row) * ld + J * kTile + col]) = d;
            }
          }
          __syncthreads();
          CPT(3);
#ifdef M3R_TRACK_PROF
          cpt[4] += 1;
#endif
        }
        cp_async_wait<0>();
        if (__ldcg(&a.ctrl[CTRL_FAIL_NOW])) failed = true;  // (a step late is early enough: read again after the loop)
      }
      SLV(3);
      grid_barrier<COOP>();
      SLV(4);
    }
    if (__ldcg(&a.ctrl[CTRL_FAIL_NOW])) failed = true;
#ifdef M3R_TRACK_PROF
    if (blockIdx.x == 0 && tid == 0)
      printf("V4WORKER per tile (us) over %lld tiles: wait %.2f, L_Ik (row changes) %.2f, L_Jk %.2f, update + write %.2f\n", cpt[4], cpt[0] / 1965.0 / cpt[4],
             cpt[1] / 1965.0 / cpt[4], cpt[2] / 1965.0 / cpt[4], cpt[3] / 1965.0 / cpt[4]);
    if (is_diag && tid == 0)
      printf("V4CHAIN per step (us): load %.2f, X + SYRK %.2f, potrf %.2f, L write + invert %.2f, publish %.2f\n", cpt[0] / 1965.0 / nblk,
             cpt[1] / 1965.0 / nblk, cpt[2] / 1965.0 / nblk, cpt[3] / 1965.0 / nblk, cpt[5] / 1965.0 / nblk);
#endif
  } else
  for (int kb = 0; kb < nblk; kb++) {
    if (v2) {
      // B1 (v2): L_kk into shared memory, substitution on this CTA's panel tile
      if (kb == 0) {
        for (int t = tid; t < kTileEl; t += kSolveThreads) {
          const int r = t >> 5, c = t & 31;
          Lcur[sw(r, c)] = (c <= r) ? __ldcg(&H[(size_t)r * ld + c]) : 0.0;
        }
        __syncthreads();
        if (wrp == 0) {
          const bool bad = potrf_inplace(Lcur, rinv, colb, lane);  // redundant per CTA: no barrier before the first panel
          if (lane == 0) s_flag = bad ? 1 : 0;
        }
        __syncthreads();
        if (s_flag) failed = true;
        if (blockIdx.x == 0) {
          for (int t = tid; t < kTileEl; t += kSolveThreads) {
            const int r = t >> 5, c = t & 31;
            const double v = Lcur[sw(r, c)];
            a.linv[t] = v;
            if (c <= r) H[(size_t)r * ld + c] = v;
          }
        }
      } else if (!is_diag) {
        // this CTA's (first) panel tile travels together with L_kk: one exposed L2 round trip instead of two
        const int I0 = kb + 1 + ((kb + 1 < nblk) ? 1 : 0) + blockIdx.x;
        double x0[kTile];
        if (wrp == 0 && I0 <= nblk) load_rows(x0, I0, kb);
        __syncthreads();
        for (int t = tid; t < kTileEl; t += kSolveThreads) Lcur[sw(t >> 5, t & 31)] = __ldcg(&a.linv[(size_t)kb * kTileEl + t]);
        if (__ldcg(&a.ctrl[CTRL_FAIL_NOW])) failed = true;
        __syncthreads();
        if (tid < kTile) rinv[tid] = 1.0 / Lcur[sw(tid, tid)];
        __syncthreads();
        SLV(0);
        if (wrp == 0) {
          if (I0 <= nblk) trsm_loaded(x0, I0, kb, Lcur, nullptr);
          for (int I = I0 + n_work; I <= nblk; I += n_work) trsm_tile(I, kb, Lcur, nullptr);
        }
      }  // (the diagonal CTA holds L_kk and 1 / diag from its own factorisation)
      if (kb == 0 || is_diag) SLV(0);
      const int kn = kb + 1;
      if (is_diag) {
        if (kn < nblk) {
          if (wrp == 0) {
            trsm_tile(kn, kb, Lcur, Xs);
          } else {  // meanwhile: the tile the chain updates next (final since the previous step's update)
            for (int e = tid - 32; e < kTileEl / 2; e += kSolveThreads - 32) {
              const int r = e >> 4, c2 = (e & 15) * 2;
              *reinterpret_cast<double2*>(Dnext + sw(r, c2)) =
                  __ldcg(reinterpret_cast<const double2*>(&H[(size_t)(kn * kTile + r) * ld + kn * kTile + c2]));
            }
          }
        }
      } else if (wrp == 0 && kb == 0) {
        for (int I = kb + 1 + ((kn < nblk) ? 1 : 0) + blockIdx.x; I <= nblk; I += n_work) trsm_tile(I, kb, Lcur, nullptr);
      }
      SLV(1);
      grid_barrier<COOP>();
      SLV(2);
    } else {
    // B1: diagonal tile, inverse, TRSM on this CTA's panel tiles
    if (kb == 0 || !lookahead) {
      for (int t = tid; t < kTile * kTile; t += kSolveThreads) {
        const int r = t >> 5, c = t & 31;
        sA[r][c] = (c <= r) ? __ldcg(&H[(size_t)(kb * kTile + r) * ld + kb * kTile + c]) : 0.0;
      }
      __syncthreads();
      if (!tile_potrf_invert(sA, sB, s_dinv, &s_flag, s_colbuf)) failed = true;  // sA = L_kk, sB = L_kk^-1 (redundant per CTA)
      if (blockIdx.x == 0) {
        for (int t = tid; t < kTile * kTile; t += kSolveThreads) {
          const int r = t >> 5, c = t & 31;
          a.linv[(size_t)kb * kTile * kTile + t] = sB[r][c];
          if (c <= r) H[(size_t)(kb * kTile + r) * ld + kb * kTile + c] = sA[r][c];
        }
      }
    } else if (!is_diag) {
      __syncthreads();
      for (int t = tid; t < kTile * kTile; t += kSolveThreads)
        sB[t >> 5][t & 31] = __ldcg(&a.linv[(size_t)kb * kTile * kTile + t]);
      if (__ldcg(&a.ctrl[CTRL_FAIL_NOW])) failed = true;
      __syncthreads();
    }  // (the diagonal CTA still holds L_kk^-1 in sB from its own factorisation)
    SLV(0);
    if (is_diag) {
      // The serial chain of the factorisation lives in this CTA alone: panel tile (kb+1, kb) -> X = A L_kk^-T
      // here; after the barrier, while the others update the trailing matrix, diagonal tile (kb+1, kb+1) -= X X^T
      // is factored + inverted and L_(kb+1)^-1 published for the next step.
      const int kn = kb + 1;
      if (kn < nblk) {
        __syncthreads();
        for (int t = tid; t < kTile * kTile; t += kSolveThreads)
          sA[t >> 5][t & 31] = __ldcg(&H[(size_t)(kn * kTile + (t >> 5)) * ld + kb * kTile + (t & 31)]);
        __syncthreads();
        double x[kTile * kTile / kSolveThreads];
#pragma unroll
        for (int q = 0; q < kTile * kTile / kSolveThreads; q++) {
          const int t = tid + q * kSolveThreads, r = t >> 5, c = t & 31;
          double acc = 0.0;
          for (int m = 0; m <= c; m++) acc += sA[r][m] * sB[c][m];
          x[q] = acc;
          H[(size_t)(kn * kTile + r) * ld + kb * kTile + c] = acc;
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < kTile * kTile / kSolveThreads; q++) {
          const int t = tid + q * kSolveThreads;
          sA[t >> 5][t & 31] = x[q];
        }
        __syncthreads();
      }
    } else {
      // panel tiles of this CTA (with look-ahead, tile (kb+1, kb) belongs to the diagonal CTA unless it is the rhs row)
      for (int I = kb + 1 + ((lookahead && kb + 1 < nblk) ? 1 : 0) + blockIdx.x; I <= nblk; I += n_work) {
        // X = A_Ik * L_kk^-T :  X[r][c] = sum_{m <= c} A[r][m] * Linv[c][m]
        __syncthreads();
        for (int t = tid; t < kTile * kTile; t += kSolveThreads)
          sA[t >> 5][t & 31] = __ldcg(&H[(size_t)(I * kTile + (t >> 5)) * ld + kb * kTile + (t & 31)]);
        __syncthreads();
        {  // 4 outputs per thread, all 32 terms (L^-1 carries an explicit zero upper triangle): unrolled, 4-way ILP
          const int c = tid & 31, r0 = tid >> 5;
          double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll 8
          for (int m = 0; m < kTile; m++) {
            const double l = sB[c][m];
            s0 += sA[r0][m] * l; s1 += sA[r0 + 8][m] * l; s2 += sA[r0 + 16][m] * l; s3 += sA[r0 + 24][m] * l;
          }
          double* o = &H[(size_t)(I * kTile) * ld + kb * kTile + c];
          o[(size_t)r0 * ld] = s0; o[(size_t)(r0 + 8) * ld] = s1; o[(size_t)(r0 + 16) * ld] = s2; o[(size_t)(r0 + 24) * ld] = s3;
        }
      }
    }
    SLV(1);
    grid_barrier<COOP>();
    SLV(2);
    }  // (!v2)
    // B2: trailing update  A_IJ -= L_Ik * L_Jk^T  for kb < J <= I <= nblk (J < nblk)
    const int rem = nblk - kb;          // tile rows below kb, including the rhs row
    const int ncols = rem - 1;          // tile columns kb+1 .. nblk-1
    // tiles enumerated row by row: row i (1..rem) has min(i, ncols) tiles
    const int ntiles = ncols * (ncols + 1) / 2 + ncols;  // triangle + rhs row
    if (v2 && is_diag && kb + 1 < nblk) {
      const int kn = kb + 1;
      __syncthreads();  // Xs (warp 0) and Dnext (warps 1-7) are complete
      {
        const int g = lane >> 2, t4 = lane & 3, mi = wrp >> 1, ni0 = (wrp & 1) * 2;
        double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
        for (int kk = 0; kk < 8; kk++) {
          const int pc = (kk * 4 + t4) ^ (g << 2);
          const double av = Xs[(mi * 8 + g) * kTile + pc];
          dmma884(acc[0][0], acc[0][1], av, Xs[(ni0 * 8 + g) * kTile + pc]);
          dmma884(acc[1][0], acc[1][1], av, Xs[((ni0 + 1) * 8 + g) * kTile + pc]);
        }
#pragma unroll
        for (int n2 = 0; n2 < 2; n2++) {
          double2* d = reinterpret_cast<double2*>(Dnext + (mi * 8 + g) * kTile + (((ni0 + n2) * 8 + 2 * t4) ^ (g << 2)));
          double2 v = *d;
          v.x -= acc[n2][0];
          v.y -= acc[n2][1];
          *d = v;
        }
      }
      __syncthreads();
      if (wrp == 0) {
        const bool bad = potrf_inplace(Dnext, rinv, colb, lane);
        if (lane == 0) s_flag = bad ? 1 : 0;
      }
      __syncthreads();
      for (int t = tid; t < kTileEl; t += kSolveThreads) {
        const int r = t >> 5, c = t & 31;
        const double v = Dnext[sw(r, c)];
        a.linv[(size_t)kn * kTileEl + t] = v;  // L_(kn,kn) for the other CTAs' substitutions (inverted after the loop)
        if (c <= r) H[(size_t)(kn * kTile + r) * ld + kn * kTile + c] = v;
      }
      if (s_flag) {
        failed = true;
        if (tid == 0) a.ctrl[CTRL_FAIL_NOW] = 1;
      }
      double* tmp = Lcur;
      Lcur = Dnext;
      Dnext = tmp;
    } else
    if (is_diag && kb + 1 < nblk) {
      // while the other CTAs run the trailing update: diagonal tile (kb+1, kb+1) -= X X^T (X is still in sA),
      // factor + invert it and publish L^-1 for the next step
      const int kn = kb + 1;
        {  // D = A_(kn,kn) - X X^T into sB (L_kk^-1 is no longer needed here)
          const int r = tid >> 4, c = tid & 15;
          // tile (kn, kn) was last written in the previous step's update: its L2 round trip overlaps the products
          const double* o = &H[(size_t)(kn * kTile) * ld + kn * kTile];
          const double o00 = __ldcg(&o[(size_t)r * ld + c]), o01 = __ldcg(&o[(size_t)r * ld + c + 16]);
          const double o10 = __ldcg(&o[(size_t)(r + 16) * ld + c]), o11 = __ldcg(&o[(size_t)(r + 16) * ld + c + 16]);
          double acc00 = 0, acc01 = 0, acc10 = 0, acc11 = 0;
#pragma unroll 8
          for (int m = 0; m < kTile; m++) {
            const double a0 = sA[r][m], a1 = sA[r + 16][m], b0 = sA[c][m], b1 = sA[c + 16][m];
            acc00 += a0 * b0; acc01 += a0 * b1; acc10 += a1 * b0; acc11 += a1 * b1;
          }
          const double d00 = o00 - acc00;
          const double d01 = o01 - acc01;
          const double d10 = o10 - acc10;
          const double d11 = o11 - acc11;
          __syncthreads();
          sB[r][c] = (c <= r) ? d00 : 0.0;
          sB[r][c + 16] = (c + 16 <= r) ? d01 : 0.0;
          sB[r + 16][c] = d10;  // c <= 15 < r + 16
          sB[r + 16][c + 16] = (c <= r) ? d11 : 0.0;
        }
        __syncthreads();
        const bool ok = tile_potrf_invert(sB, sA, s_dinv, &s_flag, s_colbuf);  // sB = L, sA = L^-1
        for (int t = tid; t < kTile * kTile; t += kSolveThreads) {
          const int r = t >> 5, c = t & 31;
          a.linv[(size_t)kn * kTile * kTile + t] = sA[r][c];
          if (c <= r) H[(size_t)(kn * kTile + r) * ld + kn * kTile + c] = sB[r][c];
        }
        if (!ok) {
          failed = true;
          if (tid == 0) a.ctrl[CTRL_FAIL_NOW] = 1;
        }
        __syncthreads();
        for (int t = tid; t < kTile * kTile; t += kSolveThreads) sB[t >> 5][t & 31] = sA[t >> 5][t & 31];
        __syncthreads();
    }
    // with look-ahead tile 0 = (kb+1, kb+1) is updated (and factored) by the diagonal CTA
    const int t_first = lookahead ? (is_diag ? ntiles : 1 + (int)blockIdx.x) : (int)blockIdx.x;
    const int t_step = n_work;
    if (a.pipe_off > 0) {
      // Pipelined update (round 2): the three tiles of an update travel global -> shared memory by cp.async (L2 only,
      // like __ldcg) kUpdStages tiles ahead, so their L2 round trips overlap the products of the tiles before them
      // instead of being exposed once per tile; the product is 16 DMMA (mma.m8n8k4.f64) per warp on swizzled tiles
      // instead of 128 DFMA per thread fed by four shared-memory loads each.  Was 3.7 us per tile (profiles/r02s).
      double* pipe = s_x + a.pipe_off;
      const int nmine = t_first < ntiles ? (ntiles - t_first + t_step - 1) / t_step : 0;
      auto decode = [&](int t0, int& I, int& J) {
        int i, j;
        if (t0 < ncols * (ncols + 1) / 2) {
          i = (int)((sqrt(8.0 * (double)t0 + 1.0) - 1.0) * 0.5);
          while ((i + 1) * (i + 2) / 2 <= t0) i++;
          while (i * (i + 1) / 2 > t0) i--;
          j = t0 - i * (i + 1) / 2;
        } else {
          i = ncols;  // rhs row
          j = t0 - ncols * (ncols + 1) / 2;
        }
        I = kb + 1 + i;
        J = kb + 1 + j;
      };
      auto issue = [&](int n) {
        if (n < nmine) {
          int I, J;
          decode(t_first + n * t_step, I, J);
          double* st = pipe + (n % kUpdStages) * kUpdStageDoubles;
#pragma unroll
          for (int q = 0; q < 2; q++) {
            const int e = tid + q * kSolveThreads, row = e >> 4, c2 = (e & 15) * 2;
            const int so = sw(row, c2);
            cp_async16(st + so, &H[(size_t)(I * kTile + row) * ld + kb * kTile + c2]);
            cp_async16(st + kTile * kTile + so, &H[(size_t)(J * kTile + row) * ld + kb * kTile + c2]);
            cp_async16(st + 2 * kTile * kTile + so, &H[(size_t)(I * kTile + row) * ld + J * kTile + c2]);
          }
        }
        cp_async_commit();  // (an empty group keeps the group count uniform)
      };
      for (int n = 0; n < kUpdStages - 1; n++) issue(n);
      const int lane = tid & 31, wrp = tid >> 5, g = lane >> 2, t4 = lane & 3;
      const int mi = wrp >> 1, ni0 = (wrp & 1) * 2;
      for (int n = 0; n < nmine; n++) {
        issue(n + kUpdStages - 1);
        cp_async_wait<kUpdStages - 1>();
        __syncthreads();
        int I, J;
        decode(t_first + n * t_step, I, J);
        const double* sa = pipe + (n % kUpdStages) * kUpdStageDoubles;
        const double* sb = sa + kTile * kTile;
        const double* sd = sb + kTile * kTile;
        double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
        for (int kk = 0; kk < 8; kk++) {
          const int pc = (kk * 4 + t4) ^ (g << 2);  // rows 8*m + g: (row & 7) == g
          const double av = sa[(mi * 8 + g) * kTile + pc];
          const double b0 = sb[(ni0 * 8 + g) * kTile + pc], b1 = sb[((ni0 + 1) * 8 + g) * kTile + pc];
          dmma884(acc[0][0], acc[0][1], av, b0);
          dmma884(acc[1][0], acc[1][1], av, b1);
        }
        const int row = mi * 8 + g;
#pragma unroll
        for (int n2 = 0; n2 < 2; n2++) {
          const int col = (ni0 + n2) * 8 + 2 * t4;
          double2 d = *reinterpret_cast<const double2*>(sd + row * kTile + (col ^ (g << 2)));
          d.x -= acc[n2][0];
          d.y -= acc[n2][1];
          *reinterpret_cast<double2*>(&H[(size_t)(I * kTile + row) * ld + J * kTile + col]) = d;
        }
        __syncthreads();  // the stage is refilled by the next iteration's issue
      }
      cp_async_wait<0>();
    } else
    for (int t0 = t_first; t0 < ntiles; t0 += t_step) {
      int i, j;
      if (t0 < ncols * (ncols + 1) / 2) {
        // invert the triangular number: i = row (0-based within triangle), j <= i
        i = (int)((sqrt(8.0 * (double)t0 + 1.0) - 1.0) * 0.5);
        while ((i + 1) * (i + 2) / 2 <= t0) i++;
        while (i * (i + 1) / 2 > t0) i--;
        j = t0 - i * (i + 1) / 2;
      } else {
        i = ncols;  // rhs row
        j = t0 - ncols * (ncols + 1) / 2;
      }
      const int I = kb + 1 + i, J = kb + 1 + j;
      __syncthreads();
      // the destination tile travels with the operands: one exposed L2 round trip per tile instead of two
      const int r = tid >> 4, c = tid & 15;  // thread -> 2x2 outputs: rows r, r+16 ; cols c, c+16
      double* o = &H[(size_t)(I * kTile) * ld + J * kTile];
      const double o00 = __ldcg(&o[(size_t)r * ld + c]), o01 = __ldcg(&o[(size_t)r * ld + c + 16]);
      const double o10 = __ldcg(&o[(size_t)(r + 16) * ld + c]), o11 = __ldcg(&o[(size_t)(r + 16) * ld + c + 16]);
      for (int t = tid; t < kTile * kTile; t += kSolveThreads) {
        const int rr = t >> 5, cc = t & 31;
        sA[rr][cc] = __ldcg(&H[(size_t)(I * kTile + rr) * ld + kb * kTile + cc]);
        sB[rr][cc] = __ldcg(&H[(size_t)(J * kTile + rr) * ld + kb * kTile + cc]);
      }
      __syncthreads();
      {
        double acc00 = 0, acc01 = 0, acc10 = 0, acc11 = 0;
#pragma unroll 8
        for (int m = 0; m < kTile; m++) {
          const double a0 = sA[r][m], a1 = sA[r + 16][m], b0 = sB[c][m], b1 = sB[c + 16][m];
          acc00 += a0 * b0; acc01 += a0 * b1; acc10 += a1 * b0; acc11 += a1 * b1;
        }
        o[(size_t)r * ld + c] = o00 - acc00;
        o[(size_t)r * ld + c + 16] = o01 - acc01;
        o[(size_t)(r + 16) * ld + c] = o10 - acc10;
        o[(size_t)(r + 16) * ld + c + 16] = o11 - acc11;
      }
    }
    SLV(3);
    grid_barrier<COOP>();
    SLV(4);
  }
  if (v2 && !v4) {
    // the back-substitution multiplies by L_kk^-1: invert the diagonal tiles now, one per CTA, lane j = column j by
    // forward substitution (a.linv holds L_kk until here)
    for (int kb2 = blockIdx.x; kb2 < nblk; kb2 += gridDim.x) {
      __syncthreads();
      for (int t = tid; t < kTileEl; t += kSolveThreads) pipe[sw(t >> 5, t & 31)] = __ldcg(&a.linv[(size_t)kb2 * kTileEl + t]);
      __syncthreads();
      if (tid < kTile) rinv[tid] = 1.0 / pipe[sw(tid, tid)];
      __syncthreads();
      if (wrp == 0) {
        double x[kTile];
#pragma unroll
        for (int i = 0; i < kTile; i++) {
          double sacc = (i == lane) ? 1.0 : 0.0;
#pragma unroll
          for (int m = 0; m < i; m++) sacc -= pipe[sw(i, m)] * x[m];
          x[i] = sacc * rinv[i];
          a.linv[(size_t)kb2 * kTileEl + i * kTile + lane] = x[i];
        }
      }
    }
    grid_barrier<COOP>();
  }
  // ---------------------------------------------------------------- C. back-substitution  L^T x = y
  const bool par_back = COOP && nblk >= 2 && nblk <= (int)gridDim.x;
  if (par_back) {
    // Column-block dataflow, no grid barriers: CTA kb owns x_kb = L_kk^-T (y_kb - sum_{j>kb} L_jk^T x_j).  The x_j
    // travel through global memory as {flag:32 | half of the double:32} words (8-byte stores are single-copy
    // atomic, so a word with the flag also has its payload): a CTA polls x_j while the tile L_jk it needs next is
    // already in its registers.  Chain = one publish + one 32x32 GEMV per block instead of CTA 0 streaming the
    // whole factor through one SM.  The words live in rows dimP+2.. of the rhs tile row, which hold zeros after
    // the factorisation (zero rows stay zero under TRSM / updates).  CTA 0 ends up with all of x in s_x.
    const int kb = blockIdx.x;
    if (kb >= nblk) return;
    unsigned long long* xw = reinterpret_cast<unsigned long long*>(&H[(size_t)(dimP + 2) * ld]);
    const unsigned long long flag = 0x5A5A0001ull << 32;
    double* red = &sA[0][0];                 // [8][32] partial sums
    double* sLi = &sB[0][0];                 // L_kk^-1, row-major stride 33
    for (int t = tid; t < kTile * kTile; t += kSolveThreads)
      sB[t >> 5][t & 31] = __ldcg(&a.linv[(size_t)kb * kTile * kTile + t]);
    if (tid < kTile)  // y_kb (v4: from ybuf; the factor's sub-diagonal tiles from their mirror slots)
      s_x[kb * kTile + tid] = v4 ? __ldcg(&a.ybuf[kb * kTile + tid]) : __ldcg(&H[(size_t)dimP * ld + kb * kTile + tid]);
    auto Ltile = [&](int j, int r, int cc) -> const double* {
      return v4 ? &H[(size_t)(kb * kTile + r) * ld + j * kTile + cc] : &H[(size_t)(j * kTile + r) * ld + kb * kTile + cc];
    };
    const int c = tid & 31, rg = tid >> 5;   // column of the tile, group of 4 rows
    double tile[4];
    if (kb + 1 < nblk) {
#pragma unroll
      for (int q = 0; q < 4; q++) tile[q] = __ldcg(Ltile(nblk - 1, rg * 4 + q, c));
    }
    __syncthreads();
    for (int j = nblk - 1; j > kb; j--) {
      double cur[4];
#pragma unroll
      for (int q = 0; q < 4; q++) cur[q] = tile[q];
      if (j - 1 > kb) {  // next tile in flight while we wait for x_j
#pragma unroll
        for (int q = 0; q < 4; q++) tile[q] = __ldcg(Ltile(j - 1, rg * 4 + q, c));
      }
      if (tid < 2 * kTile) {
        unsigned long long wv;
        const unsigned long long* src = xw + (size_t)j * 2 * kTile + tid;
        while (true) {
          asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(wv) : "l"(src) : "memory");
          if ((wv >> 32) == (flag >> 32)) break;
        }
        reinterpret_cast<unsigned*>(&s_x[j * kTile])[tid] = (unsigned)wv;  // word 2r = low half, 2r+1 = high half
      }
      __syncthreads();
      double part = 0.0;
#pragma unroll
      for (int q = 0; q < 4; q++) part += cur[q] * s_x[j * kTile + rg * 4 + q];
      red[rg * kTile + c] = part;
      __syncthreads();
      if (tid < kTile) {
        double sum = 0.0;
#pragma unroll
        for (int g = 0; g < kSolveThreads / 32; g++) sum += red[g * kTile + tid];
        s_x[kb * kTile + tid] -= sum;
      }
      __syncthreads();
    }
    double xv = 0.0;
    if (tid < kTile)
      for (int m = tid; m < kTile; m++) xv += sLi[m * (kTile + 1) + tid] * s_x[kb * kTile + m];
    __syncthreads();
    if (tid < kTile) {
      s_x[kb * kTile + tid] = xv;
      const unsigned long long bits = (unsigned long long)__double_as_longlong(xv);
      unsigned long long* dst = xw + (size_t)kb * 2 * kTile + 2 * tid;
      asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(dst), "l"(flag | (bits & 0xffffffffull)) : "memory");
      asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(dst + 1), "l"(flag | (bits >> 32)) : "memory");
    }
    __syncthreads();
    if (blockIdx.x != 0) return;
  } else {
    if (blockIdx.x != 0) return;
    for (int t = tid; t < dimP; t += kSolveThreads) s_x[t] = __ldcg(&H[(size_t)dimP * ld + t]);  // y = L^-1 g
    __syncthreads();
    for (int kb = nblk - 1; kb >= 0; kb--) {
      // x_kb = L_kk^-T s_kb : x[c] = sum_{m >= c} Linv[m][c] * s[m]
      double xv = 0.0;
      if (tid < kTile) {
        const double* Li = a.linv + (size_t)kb * kTile * kTile;
        for (int m = tid; m < kTile; m++) xv += __ldcg(&Li[m * kTile + tid]) * s_x[kb * kTile + m];
      }
      __syncthreads();
      if (tid < kTile) s_x[kb * kTile + tid] = xv;
      __syncthreads();
      // s[col] -= sum_r L[kb*32 + r][col] * x_kb[r]   for col < kb*32
      for (int col = tid; col < kb * kTile; col += kSolveThreads) {
        double s = 0.0;
#pragma unroll 8
        for (int r = 0; r < kTile; r++) s += __ldcg(&H[(size_t)(kb * kTile + r) * ld + col]) * s_x[kb * kTile + r];
        s_x[col] -= s;
      }
      __syncthreads();
    }
  }

  SLV(5);
  // ---------------------------------------------------------------- D. dx, retraction, stop test
  double ss = 0.0;
  for (int t = tid; t < dim; t += kSolveThreads) {
    const float d = failed ? 0.0f : (float)(-s_x[t]);
    a.dx_out[t] = d;
    ss += (double)d * (double)d;
  }
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  if ((tid & 31) == 0) s_red[tid >> 5] = ss;
  __syncthreads();
  for (int k = 1 + tid; k < a.N; k += kSolveThreads) {
    float T[8], xi[7];
#pragma unroll
    for (int c = 0; c < 8; c++) T[c] = a.Twc[(size_t)k * 8 + c];
#pragma unroll
    for (int c = 0; c < 7; c++) xi[c] = failed ? 0.0f : (float)(-s_x[(k - 1) * 7 + c]);
    sim3_retract(xi, T);
#pragma unroll
    for (int c = 0; c < 8; c++) a.Twc[(size_t)k * 8 + c] = T[c];
  }
  if (tid == 0) {
    double tot = 0.0;
    for (int w = 0; w < kSolveThreads / 32; w++) tot += s_red[w];
    const float nrm = (float)sqrt(tot);
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


// =============================================================================== dataflow solver
// gn_solve_df_kernel: the same four phases, but the factorisation (B) is a flag-driven dataflow over 32x32 tiles
// instead of grid-wide steps.  Measured on B200 (profiles/r02a_ubench_solve.log): a grid barrier costs 1.24 us,
// a flag hand-off through L2 0.4 us one way, a 32x32 potrf in one warp >= 3.75 us, and mma.sync.m8n8k4.f64 (DMMA)
// runs at the full fp64 rate (36.9 TFLOP/s) with 4 warps per SM where DFMA code is bound by shared-memory operand
// loads.  So:
//   * CTA 0 is the CHAIN: it alone walks the critical path  potrf(k,k) -> TRSM (k+1,k) -> SYRK (k+1,k+1) -> potrf ...
//     with every operand in its shared memory; while warp 0 factors, warps 1-3 prefetch the two tiles of the next
//     step.  It never waits for a round trip: the only remote inputs are tiles whose owner finished them a step ago.
//   * the other CTAs are WORKERS on a P x Q block-cyclic grid; worker (i mod P, j mod Q) OWNS tile (i,j) and applies
//     every update to it, in column order m = 0,1,... (fixed order: bit-deterministic), then its TRSM against L_mm
//     (forward substitution, lane = row; no explicit inverse on the path).  The last update of a diagonal tile and
//     the TRSM of a sub-diagonal tile belong to the chain.
//   * tile products are DMMA (4 warps x 16x16 quadrants, operands in padded shared memory: stride 36 doubles makes
//     the 8x4 fragment loads conflict-free).
//   * done[i][j] = tile (i,j) holds final L values; rdy[i][j] = diagonal / sub-diagonal tile has received all the
//     updates its owner applies.  Writers: stores, __syncthreads / __syncwarp, fence, st.release.gpu; readers: one
//     thread spins with ld.acquire.gpu, barrier, data through __ldcg (L2).  All CTAs are co-resident (cooperative
//     launch) and every CTA consumes its tasks in (m, phase) order, so every wait points strictly backwards.
constexpr int kDfThreads = 128;
constexpr int kTS = 36;               // shared-memory tile row stride in doubles
constexpr int kTileSm = kTile * kTS;  // doubles per shared-memory tile
constexpr int kDfTiles = 7;           // tile buffers per CTA (dynamic shared memory)
constexpr int kHelpers = 8;           // row-helper CTAs (a helper holds one row for about three steps)
constexpr int kBand = 3;              // tiles (i,j) with i - j <= kBand are finished by helper(i) / the chain

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, const double a, const double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}
// whole CTA
__device__ __forceinline__ void cta_wait(const int* f) {
  if (threadIdx.x == 0)
    while (ld_acquire_gpu(f) == 0) {}
  __syncthreads();
}
__device__ __forceinline__ void cta_wait2(const int* f, const int* g) {
  if (threadIdx.x == 0) {
    while (ld_acquire_gpu(f) == 0) {}
    while (ld_acquire_gpu(g) == 0) {}
  }
  __syncthreads();
}
__device__ __forceinline__ void cta_signal(int* f) {
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) st_release_gpu(f, 1);
}

// 32x32 tile, global (row stride ld) <-> shared (row stride kTS); thread t of a group of NT
template <int NT>
__device__ __forceinline__ void tile_g2s(double* s, const double* g, int ld, int t) {
  for (int q = t; q < kTile * kTile / 2; q += NT) {
    const int r = q >> 4, c2 = (q & 15) * 2;
    const double2 v = __ldcg(reinterpret_cast<const double2*>(g + (size_t)r * ld + c2));
    *reinterpret_cast<double2*>(s + r * kTS + c2) = v;
  }
}
template <int NT>
__device__ __forceinline__ void tile_g2s_transposed(double* s, const double* g, int ld, int t) {
  for (int q = t; q < kTile * kTile / 2; q += NT) {
    const int r = q >> 4, c2 = (q & 15) * 2;
    const double2 v = __ldcg(reinterpret_cast<const double2*>(g + (size_t)r * ld + c2));
    s[c2 * kTS + r] = v.x;
    s[(c2 + 1) * kTS + r] = v.y;
  }
}
template <int NT>
__device__ __forceinline__ void tile_s2g(double* g, const double* s, int ld, int t) {
  for (int q = t; q < kTile * kTile / 2; q += NT) {
    const int r = q >> 4, c2 = (q & 15) * 2;
    __stcg(reinterpret_cast<double2*>(g + (size_t)r * ld + c2), *reinterpret_cast<const double2*>(s + r * kTS + c2));
  }
}

// LL transfer of a tile through L2 WITHOUT a release fence (a fence costs the writer 1-1.5 us, r02c): every 8-byte word
// carries a 32-bit tag next to 32 bits of payload; 8-byte accesses are single-copy atomic, so a word that shows the tag
// holds its payload.  The buffers are zeroed at the start of every solve and a slot is reused only kLlRing steps later
// with a different tag, so a stale word can never match.  dst/src: 2048 words; element e of the 32x32 tile = words 2e, 2e+1.
template <int NT>
__device__ __forceinline__ void ll_store_tile(unsigned long long* dst, const double* sT, unsigned tag, int t) {
  const unsigned long long hi = (unsigned long long)tag << 32;
  for (int q = t; q < kTile * kTile / 2; q += NT) {
    const int r = q >> 4, c2 = (q & 15) * 2;
    const double2 v = *reinterpret_cast<const double2*>(sT + r * kTS + c2);
    const unsigned long long bx = (unsigned long long)__double_as_longlong(v.x), by = (unsigned long long)__double_as_longlong(v.y);
    unsigned long long* w = dst + 2 * (r * kTile + c2);
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(w), "l"(hi | (bx & 0xffffffffull)), "l"(hi | (bx >> 32)) : "memory");
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(w + 2), "l"(hi | (by & 0xffffffffull)), "l"(hi | (by >> 32)) : "memory");
  }
}
// All chunks of a thread are requested before any tag is looked at (a check per chunk would serialise one L2 round
// trip per chunk: 16 per lane for a one-warp reader, measured 5-7 us per tile); only missing chunks are re-requested.
template <int NT, bool TRANSPOSED>
__device__ __forceinline__ void ll_load_tile(double* sT, const unsigned long long* src, unsigned tag, int t) {
  // Gate: ONE thread polls one word until the tag shows, then the group reads the tile (and re-reads stragglers).  With
  // every thread polling its own chunks the writer's stores queued behind ~16 KB of polling loads per round on the same
  // L2 lines (r02c: a 0.3 us store took 2 us).  Group barrier: id 0 = the whole CTA, NT == 96 = warps 1-3 (id 7).
  if (t == 0) {
    unsigned long long w;
    const unsigned long long* p = src + 2 * kTile * kTile - 1;
    do {
      asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    } while ((unsigned)(w >> 32) != tag);
  }
  if (NT == kDfThreads) __syncthreads();
  else if (NT == 96) asm volatile("bar.sync 7, 96;" ::: "memory");
  else __syncwarp();
  constexpr int kChunks = kTile * kTile / 2;                // 32-byte chunks (two doubles) per tile
  constexpr int kPer = (kChunks + NT - 1) / NT;              // chunks per thread
  constexpr int kBatch = kPer < 4 ? kPer : 4;                // in flight per thread (register budget)
  for (int b0 = 0; b0 < kPer; b0 += kBatch) {
    unsigned long long w[kBatch][4];
    unsigned pending = 0;
#pragma unroll
    for (int c = 0; c < kBatch; c++)
      if (b0 + c < kPer && t + (b0 + c) * NT < kChunks) pending |= 1u << c;
    while (pending) {
#pragma unroll
      for (int c = 0; c < kBatch; c++)
        if (pending & (1u << c)) {
          const int q = t + (b0 + c) * NT;
          const unsigned long long* p = src + 2 * ((q >> 4) * kTile + (q & 15) * 2);
          asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[c][0]), "=l"(w[c][1]) : "l"(p) : "memory");
          asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[c][2]), "=l"(w[c][3]) : "l"(p + 2) : "memory");
        }
#pragma unroll
      for (int c = 0; c < kBatch; c++)
        if ((pending & (1u << c)) && (unsigned)(w[c][0] >> 32) == tag && (unsigned)(w[c][1] >> 32) == tag &&
            (unsigned)(w[c][2] >> 32) == tag && (unsigned)(w[c][3] >> 32) == tag)
          pending &= ~(1u << c);
    }
#pragma unroll
    for (int c = 0; c < kBatch; c++) {
      const int q = t + (b0 + c) * NT;
      if (b0 + c < kPer && q < kChunks) {
        const int r = q >> 4, c2 = (q & 15) * 2;
        const double x = __longlong_as_double((long long)((w[c][1] << 32) | (w[c][0] & 0xffffffffull)));
        const double y = __longlong_as_double((long long)((w[c][3] << 32) | (w[c][2] & 0xffffffffull)));
        if (TRANSPOSED) {
          sT[c2 * kTS + r] = x;
          sT[(c2 + 1) * kTS + r] = y;
        } else {
          *reinterpret_cast<double2*>(sT + r * kTS + c2) = make_double2(x, y);
        }
      }
    }
  }
}

// acc (16x16 quadrant of this warp: rows r0.., cols c0..) += A * B^T, A and B 32x32 row-major in shared memory.
// Fragments of mma.m8n8k4.f64: A[m][k]: lane (g = lane/4, t = lane%4) holds row g, column t; B[k][n] = Bmat[n][k]: the
// same access pattern on Bmat; C: row g, columns 2t and 2t+1.
__device__ __forceinline__ void tile_mma(const double* sA, const double* sB, int r0, int c0, int g, int t,
                                         double (&acc)[2][2][2]) {
#pragma unroll
  for (int kk = 0; kk < 8; kk++) {
    const double a0 = sA[(r0 + g) * kTS + kk * 4 + t], a1 = sA[(r0 + 8 + g) * kTS + kk * 4 + t];
    const double b0 = sB[(c0 + g) * kTS + kk * 4 + t], b1 = sB[(c0 + 8 + g) * kTS + kk * 4 + t];
    dmma_m8n8k4(acc[0][0][0], acc[0][0][1], a0, b0);
    dmma_m8n8k4(acc[0][1][0], acc[0][1][1], a0, b1);
    dmma_m8n8k4(acc[1][0][0], acc[1][0][1], a1, b0);
    dmma_m8n8k4(acc[1][1][0], acc[1][1][1], a1, b1);
  }
}

// X L^T = A by forward substitution, one warp, lane = row: x holds row `lane` of A on entry and of X on exit.
// sLT[c][c2] = L[c2][c] (row c of sLT = column c of L, contiguous), rinv[c] = 1 / L[c][c].
__device__ __forceinline__ void trsm_rows(double (&x)[kTile], const double* sLT, const double* rinv) {
#pragma unroll
  for (int c = 0; c < kTile; c++) {
    const double xc = x[c] * rinv[c];
    x[c] = xc;
#pragma unroll
    for (int c2 = (c + 1) & ~1; c2 < kTile; c2 += 2) {
      const double2 v = *reinterpret_cast<const double2*>(sLT + c * kTS + c2);
      if (c2 > c) x[c2] -= xc * v.x;
      x[c2 + 1] -= xc * v.y;
    }
  }
}

// Cholesky of the 32x32 tile in sD (lower part valid) by ONE warp, lane = row, rows in registers.  Right-looking in
// LDL^T order: per pivot the unscaled column goes through a double-buffered 32-double shared array (1 STS + 16
// broadcast LDS.128 instead of 62 shuffles: 12.5k -> 7.4k cycles, r02a_ubench_solve.log) and the chain carries one
// reciprocal; the square roots are applied to whole columns at the end.  Writes sL (row-major, zero upper part),
// sLT (its transpose) and rinv[c] = 1 / L[c][c].  Returns true if a pivot was <= 0 (or NaN).
__device__ __forceinline__ bool potrf_tile(const double* sD, double* sL, double* sLT, double* rinv, double* col, int lane) {
  double a[kTile];
#pragma unroll
  for (int c = 0; c < kTile; c += 2) {
    const double2 v = *reinterpret_cast<const double2*>(sD + lane * kTS + c);
    a[c] = v.x;
    a[c + 1] = v.y;
  }
  bool bad = false;
  double my_d = 1.0;
#pragma unroll
  for (int j = 0; j < kTile; j++) {
    double* cb = col + (j & 1) * kTile;
    cb[lane] = a[j];
    __syncwarp();
    double d = cb[j];
    if (!(d > 0.0)) {
      bad = true;
      d = 1.0;
    }
    if (lane == j) my_d = d;
    const double tj = a[j] * (1.0 / d);
#pragma unroll
    for (int c = (j + 1) & ~1; c < kTile; c += 2) {
      const double2 v = *reinterpret_cast<const double2*>(cb + c);
      if (c > j) a[c] -= tj * v.x;
      a[c + 1] -= tj * v.y;
    }
  }
  const double rs = rsqrt(my_d);  // 1 / L[lane][lane]
  col[lane] = rs;                 // (lower half of `col`: last used by pivot 30, every lane is past pivot 31's barrier)
  rinv[lane] = rs;
  __syncwarp();
#pragma unroll
  for (int c = 0; c < kTile; c++) {
    const double l = (c <= lane) ? a[c] * col[c] : 0.0;
    sL[lane * kTS + c] = l;
    sLT[c * kTS + lane] = l;
  }
  return bad;
}

#ifdef M3R_SOLVE_PROF
__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define DFP(slot) do { if (tid == 0) { const unsigned long long n_ = gtime_ns(); prof[slot] += (float)(n_ - prof_t); prof_t = n_; if (s_raw_k >= 0 && s_raw_k < 58 && (slot) >= 2 && (slot) <= 6) s_raw[s_raw_k][(slot) - 2] = (unsigned)n_; if ((slot) == 11 && s_raw_k >= 0 && s_raw_k < 58) s_raw[s_raw_k][5] = (unsigned)n_; } } while (0)
#else
#define DFP(slot) do { } while (0)
#endif

__global__ void __launch_bounds__(kDfThreads, 2)
gn_solve_df_kernel(const SolveArgs a) {
  if (a.ctrl[CTRL_DONE]) return;  // uniform over the grid: set only by a previous launch
  extern __shared__ __align__(16) double s_dyn[];  // kDfTiles tiles (phase B) | s_x: dimP doubles (phase C)
  __shared__ __align__(16) double s_col[2 * kTile];
  __shared__ double s_rinv[kTile];
  __shared__ double s_red[kDfThreads / 32];
  __shared__ int s_fail;
  double* const s_x = s_dyn + kDfTiles * kTileSm;
  cg::grid_group grid = cg::this_grid();

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nblk = a.nblk, ld = a.ld, dim = a.dim, dimP = a.nblk * kTile;
  double* H = a.hd;
  const int nfree = a.N - 1;
  if (tid == 0) s_fail = 0;
#ifdef M3R_SOLVE_PROF
  float prof[12] = {};
  __shared__ unsigned s_raw[58][6];
  int s_raw_k = -1;
  unsigned long long prof_t = gtime_ns();
  const unsigned long long prof_t0 = prof_t;
#endif

  // ---------------------------------------------------------------- A. assemble (+ flag reset)
  for (int t = blockIdx.x * kDfThreads + tid; t < 3 * (nblk + 1) * nblk; t += gridDim.x * kDfThreads) a.done[t] = 0;  // done | rdy | rdyh contiguous
  for (int t = blockIdx.x * kDfThreads + tid; t < kLlKinds * kLlRing * 2 * kTile * kTile; t += gridDim.x * kDfThreads) a.ll[t] = 0ull;
  const int nstrips = nfree + 2;
  for (int sidx = blockIdx.x; sidx < nstrips; sidx += gridDim.x) {
    if (sidx < nfree) {
      const int p = sidx;
      const int r0 = 7 * p, ncol = 7 * p + 7;
      for (int t = tid; t < 7 * ncol; t += kDfThreads) H[(size_t)(r0 + t / ncol) * ld + (t % ncol)] = 0.0;
      __syncthreads();
      const int c = p + 1;  // compact pose index
      const int lo = a.inc_off[c], hi = a.inc_off[c + 1];
      for (int q = lo; q < hi; q++) {
        const int ent = a.inc_list[q];
        const int e = ent >> 1, role = ent & 1;
        const int other = (role == 0 ? a.cj[e] : a.ci[e]) - 1;
        const double* blk = a.eb + (size_t)e * kEdgeBlock;
        if (tid < 49) {
          const int r = tid / 7, cc = tid % 7;
          const double v = blk[tid];
          H[(size_t)(r0 + r) * ld + (r0 + cc)] += v;
          if (other >= 0 && other < p) H[(size_t)(r0 + r) * ld + (7 * other + cc)] -= v;
        }
      }
      __syncthreads();
    } else if (sidx == nfree) {
      for (int t = tid; t < (dimP - dim) * ld; t += kDfThreads) {
        const int r = dim + t / ld, cc = t % ld;
        H[(size_t)r * ld + cc] = (r == cc) ? 1.0 : 0.0;
      }
    } else {
      // rhs tile row: row dimP holds g, rows dimP+1.. are zero
      for (int t = tid; t < kTile * ld; t += kDfThreads) H[(size_t)dimP * ld + t] = 0.0;
      __syncthreads();
      for (int p = tid; p < nfree; p += kDfThreads) {
        const int c = p + 1;
        double g[7] = {0, 0, 0, 0, 0, 0, 0};
        for (int q = a.inc_off[c]; q < a.inc_off[c + 1]; q++) {
          const int ent = a.inc_list[q];
          const double* bv = a.eb + (size_t)(ent >> 1) * kEdgeBlock + 49;
          const double sgn = (ent & 1) ? 1.0 : -1.0;  // g_i = -b_e, g_j = +b_e
#pragma unroll
          for (int k = 0; k < 7; k++) g[k] += sgn * bv[k];
        }
#pragma unroll
        for (int k = 0; k < 7; k++) H[(size_t)dimP * ld + 7 * p + k] = g[k];
      }
    }
  }
  DFP(0);
  grid.sync();
  DFP(1);

  // ---------------------------------------------------------------- B. Cholesky, dataflow
  // Roles: CTA 0 = chain; CTAs 1..kHelpers = row helpers; the rest = P x Q block-cyclic workers.
  // Who applies column m to tile (i,j) (m < j <= i), and who finishes the tile:
  //   i - j >= 3 (far tiles)     : the cyclic owner applies every column and does the TRSM at step j
  //   i - j <= 2 (band of row i) : columns m <= i-4 the cyclic owner (hand-off flag rdyh[i][j]);
  //                                columns i-3 and i-2 helper(i) (tiles stay in its shared memory); it also does
  //                                the TRSM of (i,i-2) -- ONE hop after the chain publishes L_(i-2,i-2) -- and hands
  //                                (i,i-1) and (i,i) to the chain (rdy[i][j]);
  //                                column i-1: the chain (TRSM of (i,i-1), SYRK on (i,i), potrf).
  // Measured motivation (r02b): with every near-diagonal tile two L2 hops away (worker TRSM -> worker update -> chain,
  // ~2.2 us per hop: store, release fence, poll, load) the chain idled 10-15 us per step waiting for its inputs.
  const int P = a.P, Q = a.Q;
  const size_t nflag = (size_t)(nblk + 1) * nblk;
  int* const rdyh = a.rdy + nflag;
  double* const s_tiles = s_dyn;
  auto gtile = [&](int i, int j) -> double* { return &H[(size_t)(i * kTile) * ld + j * kTile]; };
  // LL slot of kind (0: L_kk, 1: L_(k+1,k), 2: tile (k+2,k+1), 3: tile (k+2,k+2)) written at step k; its tag is k+1
  auto llslot = [&](int kind, int k) -> unsigned long long* { return a.ll + (size_t)(kind * kLlRing + (k & (kLlRing - 1))) * 2 * kTile * kTile; };
  // dst (shared, 4 warps) -= A * B^T
  auto smem_gemm_sub = [&](double* dst, const double* A, const double* B) {
    const int g = lane >> 2, t4 = lane & 3, r0 = (warp >> 1) * 16, c0 = (warp & 1) * 16;
    double acc[2][2][2] = {};
    tile_mma(A, B, r0, c0, g, t4, acc);
#pragma unroll
    for (int mi = 0; mi < 2; mi++)
#pragma unroll
      for (int ni = 0; ni < 2; ni++) {
        double2* d = reinterpret_cast<double2*>(dst + (r0 + mi * 8 + g) * kTS + c0 + ni * 8 + 2 * t4);
        double2 v = *d;
        v.x -= acc[mi][ni][0];
        v.y -= acc[mi][ni][1];
        *d = v;
      }
  };
  // rows of the tile in `sT` <- rows * L^-T (warp 0 only; sLT / s_rinv describe L)
  auto trsm_tile_inplace = [&](double* sT, const double* sLT_) {
    double x[kTile];
#pragma unroll
    for (int c = 0; c < kTile; c += 2) {
      const double2 v = *reinterpret_cast<const double2*>(sT + lane * kTS + c);
      x[c] = v.x;
      x[c + 1] = v.y;
    }
    trsm_rows(x, sLT_, s_rinv);
#pragma unroll
    for (int c = 0; c < kTile; c += 2) *reinterpret_cast<double2*>(sT + lane * kTS + c) = make_double2(x[c], x[c + 1]);
  };
  if (blockIdx.x == 0) {
    // ------------------------------------------------ chain
    // warps 0-2 compute (warp 0: potrf and TRSM; all three: the SYRK), warp 3 is the I/O warp: it fetches the two tiles of
    // the next step and publishes L_kk / L_(k+1,k).  A release store blocks its thread for 1-1.5 us (the fence waits for
    // the L2 write acknowledgements), so the compute warps never wait for it: the hand-overs are split named barriers
    //   bar 1 (128)  sync, both sides        : fetch of step k done  <->  potrf(k) done (sL complete)
    //   bar 2 (96)   sync, compute warps     : TRSM done -> SYRK
    //   bar 5 (64)   warp 0 ARRIVE, I/O sync  : TRSM done, sP final (I/O may send it)
    //   bar 4 (64)   I/O ARRIVE, warp 0 sync  : I/O has read sL (potrf(k+1) may overwrite it)
    // The three transfers of the chain <-> helper loop (L_kk and L_(k+1,k) out, the next two tiles in) are LL, fence-free.
    double* sL = s_tiles;
    double* sLT = s_tiles + kTileSm;
    double* sPb[2] = {s_tiles + 2 * kTileSm, s_tiles + 3 * kTileSm};
    double* sDb[2] = {s_tiles + 4 * kTileSm, s_tiles + 5 * kTileSm};  // diagonal tile of step k: sDb[k & 1]
    tile_g2s<kDfThreads>(sDb[0], H, ld, tid);
    __syncthreads();
    // The two tiles of step k -- (k+1,k) and (k+1,k+1), each with columns <= k-1 applied -- fetched by warps 1-3 (thread
    // t of 96) while warp 0 factors: helper(k+1) sent them at step k-1 (LL, tag k); step 0 reads the assembled matrix.
    auto chain_fetch = [&](int k, double* sP_, double* sDn_, int t) {
      if (k >= 1) ll_load_tile<96, false>(sP_, llslot(2, k - 1), (unsigned)k, t);
      else tile_g2s<96>(sP_, gtile(1, 0), ld, t);
      if (k + 1 < nblk) {
        if (k >= 1) ll_load_tile<96, false>(sDn_, llslot(3, k - 1), (unsigned)k, t);
        else tile_g2s<96>(sDn_, gtile(1, 1), ld, t);
      }
    };
    if (warp == 3) {
#ifdef M3R_SOLVE_PROF
      float ip[10] = {};
      unsigned long long ip_t = gtime_ns();
#define IOP(slot) do { __syncwarp(); if (lane == 0) { const unsigned long long n_ = gtime_ns(); ip[slot] += (float)(n_ - ip_t); ip_t = n_; } } while (0)
#else
#define IOP(slot) do { } while (0)
#endif
      for (int k = 0; k < nblk; k++) {
        double* sP = sPb[k & 1];
        double* sDn = sDb[(k + 1) & 1];
        chain_fetch(k, sP, sDn, 64 + lane);
        IOP(1);
        asm volatile("bar.sync 1, 128;" ::: "memory");
        IOP(2);
        ll_store_tile<96>(llslot(0, k), sL, (unsigned)(k + 1), 64 + lane);  // L_kk -> helper(k+2), no fence
        IOP(3);
        tile_s2g<32>(gtile(k, k), sL, ld, lane);
        if (k + 1 < nblk) asm volatile("bar.arrive 4, 64;" ::: "memory");
        __syncwarp();
        if (lane == 0) st_release_gpu(&a.done[k * nblk + k], 1);  // everyone else reads L_kk from H behind this flag
        IOP(4);
        asm volatile("bar.sync 5, 64;" ::: "memory");                   // TRSM done
        IOP(5);
        ll_store_tile<32>(llslot(1, k), sP, (unsigned)(k + 1), lane);  // L_(k+1,k) -> helper(k+2)
        IOP(6);
        tile_s2g<32>(gtile(k + 1, k), sP, ld, lane);
        __syncwarp();
        if (lane == 0) st_release_gpu(&a.done[(k + 1) * nblk + k], 1);
        __syncwarp();
        IOP(7);
      }
#ifdef M3R_SOLVE_PROF
      if (lane == 0 && a.ctrl[CTRL_ITERS] == 0)
        printf("DFIO: LL-load S %.1f, LL-load D %.1f, wait potrf %.1f, LL-store L %.1f, store L + release %.1f, wait trsm %.1f, LL-store P %.1f, "
               "store P + release %.1f us\n", ip[0] * 1e-3f, ip[1] * 1e-3f, ip[2] * 1e-3f, ip[3] * 1e-3f, ip[4] * 1e-3f, ip[5] * 1e-3f, ip[6] * 1e-3f, ip[7] * 1e-3f);
#endif
    } else {
      for (int k = 0; k < nblk; k++) {
#ifdef M3R_SOLVE_PROF
        s_raw_k = k;
#endif
        double* sP = sPb[k & 1];
        if (warp == 0) {
          if (k >= 1) asm volatile("bar.sync 4, 64;" ::: "memory");
          DFP(3);
          if (potrf_tile(sDb[k & 1], sL, sLT, s_rinv, s_col, lane) && lane == 0) s_fail = 1;
          DFP(5);
        } else {
          chain_fetch(k, sP, sDb[(k + 1) & 1], (warp - 1) * 32 + lane);  // idle during the factorisation: help the I/O warp
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        DFP(2);
        if (warp == 0) {
          trsm_tile_inplace(sP, sLT);
          asm volatile("bar.arrive 5, 64;" ::: "memory");
        } else {
          ll_store_tile<96>(llslot(0, k), sL, (unsigned)(k + 1), (warp - 1) * 32 + lane);  // (idle during the TRSM)
        }
        asm volatile("bar.sync 2, 96;" ::: "memory");
        DFP(4);
        if (k + 1 < nblk) {
          // last update of the next diagonal tile: sDn -= L_(k+1,k) L_(k+1,k)^T; 16 8x8 output tiles over 3 warps
          double* sDn = sDb[(k + 1) & 1];
          const int g = lane >> 2, t4 = lane & 3;
          // (6 / 5 / 5 output tiles per warp, all accumulators live: the 8 dependent DMMAs of a tile interleave with the others')
          double acc[6][2];
#pragma unroll
          for (int q = 0; q < 6; q++) acc[q][0] = acc[q][1] = 0.0;
#pragma unroll
          for (int kk = 0; kk < 8; kk++) {
#pragma unroll
            for (int q = 0; q < 6; q++) {
              const int ot = warp + 3 * q;
              if (ot < 16) {
                const int r0 = (ot >> 2) * 8, c0 = (ot & 3) * 8;
                dmma_m8n8k4(acc[q][0], acc[q][1], sP[(r0 + g) * kTS + kk * 4 + t4], sP[(c0 + g) * kTS + kk * 4 + t4]);
              }
            }
          }
#pragma unroll
          for (int q = 0; q < 6; q++) {
            const int ot = warp + 3 * q;
            if (ot < 16) {
              const int r0 = (ot >> 2) * 8, c0 = (ot & 3) * 8;
              double2* d = reinterpret_cast<double2*>(sDn + (r0 + g) * kTS + c0 + 2 * t4);
              double2 v = *d;
              v.x -= acc[q][0];
              v.y -= acc[q][1];
              *d = v;
            }
          }
          asm volatile("bar.sync 2, 96;" ::: "memory");
          DFP(6);
        }
      }
    }
    __syncthreads();
  } else if ((int)blockIdx.x <= kHelpers) {
    // ------------------------------------------------ row helper h: rows r = h (mod kHelpers), r >= 2
    // Row r's band tiles (r, r-3 .. r) live in this CTA's shared memory while it applies the last three columns:
    //   stage 0, column r-4: far operand L_(r,r-4) from its cyclic owner (flag: two fenced hops behind the chain, which
    //            is why it gets two steps of slack), the other operands from H behind their flags;
    //   stage 1, column r-3: own TRSM of (r,r-3) with L_mm by LL; operands: chain's L_(m+1,m) (LL), helper(r-1)'s
    //            stage-2 output L_(r-1,m) (LL);
    //   stage 2, column r-2: own TRSM of (r,r-2) (LL), chain's L_(m+1,m) (LL); tiles (r,r-1), (r,r) go to the chain by LL.
    const int h = (int)blockIdx.x - 1;
    double* sBd[4] = {s_tiles, s_tiles + kTileSm, s_tiles + 2 * kTileSm, s_tiles + 3 * kTileSm};  // (r,r-3) .. (r,r)
    double* sLT = s_tiles + 4 * kTileSm;
    double* sX = s_tiles + 5 * kTileSm;
    double* sY = s_tiles + 6 * kTileSm;
#ifdef M3R_SOLVE_PROF
    float hp[10] = {};
    unsigned long long hp_t = gtime_ns();
#define HP(slot) do { if (tid == 0) { const unsigned long long n_ = gtime_ns(); hp[slot] += (float)(n_ - hp_t); hp_t = n_; } } while (0)
#else
#define HP(slot) do { } while (0)
#endif
    for (int r = h; r <= nblk; r += kHelpers) {
      if (r < 2) continue;
      const bool has_diag = r < nblk;
      HP(9);
      // band tiles with every column the cyclic owners apply (m <= r-5); tile q is (r, r-3+q)
      if (r >= 5) {
        if (tid == 0)
          for (int q = 0; q < 4; q++)
            if (r - 3 + q < nblk)
              while (ld_acquire_gpu(&rdyh[r * nblk + r - 3 + q]) == 0) {}
        __syncthreads();
      }
      HP(0);
      for (int q = 0; q < 4; q++)
        if (r - 3 + q >= 0 && r - 3 + q < nblk) tile_g2s<kDfThreads>(sBd[q], gtile(r, r - 3 + q), ld, tid);
      if (r >= 4) {
        // stage 0: column m = r-4
        const int m = r - 4;
        if (tid < 4) {  // the far operand and the three older ones: poll in parallel
          const int j = tid == 0 ? r : r - 4 + tid;  // r, r-3, r-2, r-1
          while (ld_acquire_gpu(&a.done[j * nblk + m]) == 0) {}
        }
        __syncthreads();
        HP(1);
        tile_g2s<kDfThreads>(sX, gtile(r, m), ld, tid);
        tile_g2s<kDfThreads>(sY, gtile(r - 3, m), ld, tid);
        tile_g2s<kDfThreads>(sLT, gtile(r - 2, m), ld, tid);  // (sLT is free until stage 1)
        __syncthreads();
        smem_gemm_sub(sBd[0], sX, sY);
        smem_gemm_sub(sBd[1], sX, sLT);
        if (has_diag) smem_gemm_sub(sBd[3], sX, sX);
        __syncthreads();
        tile_g2s<kDfThreads>(sY, gtile(r - 1, m), ld, tid);
        __syncthreads();
        smem_gemm_sub(sBd[2], sX, sY);
        __syncthreads();
        HP(2);
      }
      if (r >= 3) {
        // stage 1: column m = r-3
        const int m = r - 3;
        ll_load_tile<kDfThreads, true>(sLT, llslot(0, m), (unsigned)(m + 1), tid);  // L_mm^T
        __syncthreads();
        if (tid < kTile) s_rinv[tid] = 1.0 / sLT[tid * kTS + tid];
        __syncthreads();
        HP(3);
        if (warp == 0) trsm_tile_inplace(sBd[0], sLT);                                // L_(r,m)
        else ll_load_tile<96, false>(sY, llslot(1, m), (unsigned)(m + 1), tid - 32);  // meanwhile: the chain's L_(m+1,m)
        __syncthreads();
        smem_gemm_sub(sBd[1], sBd[0], sY);  // (r, r-2)
        __syncthreads();
        ll_load_tile<kDfThreads, false>(sY, llslot(4, m), (unsigned)(m + 1), tid);  // helper(r-1)'s L_(r-1,m)
        tile_s2g<kDfThreads>(gtile(r, m), sBd[0], ld, tid);                           // L_(r,m) for everyone else
        __syncthreads();
        smem_gemm_sub(sBd[2], sBd[0], sY);  // (r, r-1)
        if (has_diag) smem_gemm_sub(sBd[3], sBd[0], sBd[0]);
        __syncthreads();
        if (tid == 0) st_release_gpu(&a.done[r * nblk + m], 1);  // (last: the fence blocks its warp for > 1 us)
        HP(4);
      }
      {
        // stage 2: column m = r-2, one LL hop behind the chain
        const int m = r - 2;
        ll_load_tile<kDfThreads, true>(sLT, llslot(0, m), (unsigned)(m + 1), tid);  // L_mm^T
        __syncthreads();
        HP(5);
        if (tid < kTile) s_rinv[tid] = 1.0 / sLT[tid * kTS + tid];
        __syncthreads();
        if (warp == 0) trsm_tile_inplace(sBd[1], sLT);                                // L_(r,m)
        else ll_load_tile<96, false>(sY, llslot(1, m), (unsigned)(m + 1), tid - 32);  // meanwhile: the chain's L_(m+1,m)
        __syncthreads();
        HP(6);
        smem_gemm_sub(sBd[2], sBd[1], sY);
        __syncthreads();
        ll_store_tile<kDfThreads>(llslot(2, m), sBd[2], (unsigned)(m + 1), tid);  // tile (r,r-1) -> chain
        if (has_diag) {
          smem_gemm_sub(sBd[3], sBd[1], sBd[1]);
          __syncthreads();
          ll_store_tile<kDfThreads>(llslot(3, m), sBd[3], (unsigned)(m + 1), tid);  // tile (r,r) -> chain
        }
        ll_store_tile<kDfThreads>(llslot(4, m), sBd[1], (unsigned)(m + 1), tid);  // L_(r,m) -> helper(r+1), stage 1
        HP(7);
        tile_s2g<kDfThreads>(gtile(r, m), sBd[1], ld, tid);  // ... and for everyone else, behind a flag
        __syncthreads();
        if (tid == 0) st_release_gpu(&a.done[r * nblk + m], 1);
        __syncthreads();
        HP(8);
      }
    }
#ifdef M3R_SOLVE_PROF
    if (tid == 0 && a.ctrl[CTRL_ITERS] == 0 && h < 2)
      printf("DFHELP %d: wait rdyh %.1f | stage0: wait far L(r,m) %.1f, rest %.1f | stage1: wait LL L_mm %.1f, rest %.1f | stage2: wait LL L_mm %.1f, "
             "trsm + wait LL P %.1f, gemms + LL sends %.1f, store + release %.1f | other %.1f us\n", h, hp[0] * 1e-3f, hp[1] * 1e-3f, hp[2] * 1e-3f, hp[3] * 1e-3f,
             hp[4] * 1e-3f, hp[5] * 1e-3f, hp[6] * 1e-3f, hp[7] * 1e-3f, hp[8] * 1e-3f, hp[9] * 1e-3f);
#endif
  } else if ((int)blockIdx.x - 1 - kHelpers < P * Q) {
    // ------------------------------------------------ worker (wr, wc): owns tiles i % P == wr, j % Q == wc
    const int w = (int)blockIdx.x - 1 - kHelpers, wr = w / Q, wc = w % Q;
    double* sLT = s_tiles;
    double* sW = s_tiles + kTileSm;  // 4 per-warp tiles (TRSM phase) / operands A, B (update phase)
    // one update task: tile (i,j) -= L_im L_jm^T, destination read-modify-written in global memory (L2)
    auto update_task = [&](int i, int j, int m) {
      const bool diag = (i == j);
      double* sA = sW;
      double* sBm = sW + kTileSm;
      cta_wait2(&a.done[i * nblk + m], &a.done[j * nblk + m]);
      tile_g2s<kDfThreads>(sA, gtile(i, m), ld, tid);
      if (!diag) tile_g2s<kDfThreads>(sBm, gtile(j, m), ld, tid);
      const int g = lane >> 2, t4 = lane & 3, r0 = (warp >> 1) * 16, c0 = (warp & 1) * 16;
      double* gd = gtile(i, j);
      double2 dst[2][2];
#pragma unroll
      for (int mi = 0; mi < 2; mi++)
#pragma unroll
        for (int ni = 0; ni < 2; ni++)
          dst[mi][ni] = __ldcg(reinterpret_cast<const double2*>(gd + (size_t)(r0 + mi * 8 + g) * ld + c0 + ni * 8 + 2 * t4));
      __syncthreads();
      double acc[2][2][2] = {};
      tile_mma(sA, diag ? sA : sBm, r0, c0, g, t4, acc);
#pragma unroll
      for (int mi = 0; mi < 2; mi++)
#pragma unroll
        for (int ni = 0; ni < 2; ni++)
          __stcg(reinterpret_cast<double2*>(gd + (size_t)(r0 + mi * 8 + g) * ld + c0 + ni * 8 + 2 * t4),
                 make_double2(dst[mi][ni].x - acc[mi][ni][0], dst[mi][ni].y - acc[mi][ni][1]));
      __syncthreads();  // operands are overwritten by the next task; every store of this tile is issued
      if (i - j <= kBand && m == i - kBand - 2 && tid == 0) st_release_gpu(&rdyh[i * nblk + j], 1);  // hand-off to helper(i)
    };
    for (int m = 0; m < nblk; m++) {
      // ---- column m: TRSM of the owned far tiles (i,m), i >= m+3, and L_mm^-1 (for phase C) if (m,m) is owned
      if (m % Q == wc) {
        const int i0 = m + ((wr - m % P) + P) % P;  // first i >= m with i % P == wr
        bool any = false;
        for (int i = i0; i <= nblk; i += P) any = any || (i == m || i > m + kBand);
        if (any) {
          cta_wait(&a.done[m * nblk + m]);
          tile_g2s_transposed<kDfThreads>(sLT, gtile(m, m), ld, tid);
          __syncthreads();
          if (tid < kTile) s_rinv[tid] = 1.0 / sLT[tid * kTS + tid];
          __syncthreads();
          int q = 0;
          for (int i = i0; i <= nblk; i += P) {
            if (i > m && i <= m + kBand) continue;  // chain / helper
            if (((q++) & 3) != warp) continue;
            double* sT = sW + warp * kTileSm;
            double* gt = gtile(i, m);
            double x[kTile];
            if (i == m) {
#pragma unroll
              for (int c = 0; c < kTile; c++) x[c] = (c == lane) ? 1.0 : 0.0;
            } else {
              __syncwarp();
              tile_g2s<32>(sT, gt, ld, lane);
              __syncwarp();
#pragma unroll
              for (int c = 0; c < kTile; c += 2) {
                const double2 v = *reinterpret_cast<const double2*>(sT + lane * kTS + c);
                x[c] = v.x;
                x[c + 1] = v.y;
              }
            }
            trsm_rows(x, sLT, s_rinv);
            if (i == m) {
              // x = row `lane` of L^-T, i.e. column `lane` of L^-1: linv[c][lane] (row-major L^-1, as phase C reads it)
#pragma unroll
              for (int c = 0; c < kTile; c++) a.linv[(size_t)m * kTile * kTile + c * kTile + lane] = x[c];
            } else {
              __syncwarp();
#pragma unroll
              for (int c = 0; c < kTile; c += 2) *reinterpret_cast<double2*>(sT + lane * kTS + c) = make_double2(x[c], x[c + 1]);
              __syncwarp();
              tile_s2g<32>(gt, sT, ld, lane);
              __syncwarp();
              if (lane == 0) st_release_gpu(&a.done[i * nblk + m], 1);
            }
          }
          __syncthreads();
        }
      }
      // ---- updates.  A tile's column-c update is LAGGED one step (done in phase c+1) unless it is the last one before the
      // tile's TRSM / hand-off: the on-time updates (whose operands were published moments ago) then never queue behind
      // bulk work, and the bulk work only needs operands that are a whole step old (no stall).  Per tile the columns are
      // still applied in ascending order.
      {  // (1) far tiles of column j = m+1 (their TRSM is next phase): lagged column m-1, then column m
        const int j = m + 1;
        if (j < nblk && j % Q == wc)
          for (int i = j + ((wr - j % P) + P) % P; i <= nblk; i += P)
            if (i - j > kBand) {
              if (m >= 1) update_task(i, j, m - 1);
              update_task(i, j, m);
            }
      }
      {  // (2) hand-offs of this phase: band tiles of row m+kBand+2
        const int i = m + kBand + 2;
        if (i <= nblk && i % P == wr)
          for (int j = i - kBand; j <= i && j < nblk; j++)
            if (j % Q == wc) {
              if (m >= 1) update_task(i, j, m - 1);
              update_task(i, j, m);
            }
      }
      if (m >= 1) {  // (3) everything else: lagged column m-1
        const int j0 = (m + 2) + ((wc - (m + 2) % Q) + Q) % Q;
        for (int j = j0; j < nblk; j += Q) {
          const int ii0 = j + ((wr - j % P) + P) % P;
          for (int i = ii0; i <= nblk; i += P) {
            if (i - j <= kBand && i < m + kBand + 3) continue;  // band tile at or past its hand-off
            update_task(i, j, m - 1);
          }
        }
      }
    }
  }
  DFP(7);
  grid.sync();
  DFP(8);
  const bool failed = s_fail != 0;  // (only CTA 0, the chain, can have set it; only CTA 0 uses it below)

  // ---------------------------------------------------------------- C. back-substitution  L^T x = y
  constexpr int kRG = kDfThreads / 32;  // row groups per tile
  constexpr int kRP = kTile / kRG;      // rows per group
  const bool par_back = nblk >= 2 && nblk <= (int)gridDim.x;
  if (par_back) {
    // column-block dataflow (see gn_solve_kernel): CTA kb owns x_kb, the x_j travel as {flag | half} words
    const int kb = blockIdx.x;
    if (kb >= nblk) return;
    unsigned long long* xw = reinterpret_cast<unsigned long long*>(&H[(size_t)(dimP + 2) * ld]);
    const unsigned long long flag = 0x5A5A0001ull << 32;
    double* red = s_dyn;               // [kRG][32] partial sums
    double* sLi = s_dyn + kTileSm;     // L_kk^-1, row-major stride kTS
    for (int t = tid; t < kTile * kTile; t += kDfThreads)
      sLi[(t >> 5) * kTS + (t & 31)] = __ldcg(&a.linv[(size_t)kb * kTile * kTile + t]);
    if (tid < kTile) s_x[kb * kTile + tid] = __ldcg(&H[(size_t)dimP * ld + kb * kTile + tid]);  // y_kb
    const int c = tid & 31, rg = tid >> 5;
    double tile[kRP];
    if (kb + 1 < nblk) {
#pragma unroll
      for (int q = 0; q < kRP; q++) tile[q] = __ldcg(&H[(size_t)((nblk - 1) * kTile + rg * kRP + q) * ld + kb * kTile + c]);
    }
    __syncthreads();
    for (int j = nblk - 1; j > kb; j--) {
      double cur[kRP];
#pragma unroll
      for (int q = 0; q < kRP; q++) cur[q] = tile[q];
      if (j - 1 > kb) {  // next tile in flight while we wait for x_j
#pragma unroll
        for (int q = 0; q < kRP; q++) tile[q] = __ldcg(&H[(size_t)((j - 1) * kTile + rg * kRP + q) * ld + kb * kTile + c]);
      }
      if (tid < 2 * kTile) {
        unsigned long long wv;
        const unsigned long long* src = xw + (size_t)j * 2 * kTile + tid;
        while (true) {
          asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(wv) : "l"(src) : "memory");
          if ((wv >> 32) == (flag >> 32)) break;
        }
        reinterpret_cast<unsigned*>(&s_x[j * kTile])[tid] = (unsigned)wv;  // word 2r = low half, 2r+1 = high half
      }
      __syncthreads();
      double part = 0.0;
#pragma unroll
      for (int q = 0; q < kRP; q++) part += cur[q] * s_x[j * kTile + rg * kRP + q];
      red[rg * kTile + c] = part;
      __syncthreads();
      if (tid < kTile) {
        double sum = 0.0;
#pragma unroll
        for (int g = 0; g < kRG; g++) sum += red[g * kTile + tid];
        s_x[kb * kTile + tid] -= sum;
      }
      __syncthreads();
    }
    double xv = 0.0;
    if (tid < kTile)
      for (int m = tid; m < kTile; m++) xv += sLi[m * kTS + tid] * s_x[kb * kTile + m];
    __syncthreads();
    if (tid < kTile) {
      s_x[kb * kTile + tid] = xv;
      const unsigned long long bits = (unsigned long long)__double_as_longlong(xv);
      unsigned long long* dstw = xw + (size_t)kb * 2 * kTile + 2 * tid;
      asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(dstw), "l"(flag | (bits & 0xffffffffull)) : "memory");
      asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(dstw + 1), "l"(flag | (bits >> 32)) : "memory");
    }
    __syncthreads();
    if (blockIdx.x != 0) return;
  } else {
    if (blockIdx.x != 0) return;
    for (int t = tid; t < dimP; t += kDfThreads) s_x[t] = __ldcg(&H[(size_t)dimP * ld + t]);  // y = L^-1 g
    __syncthreads();
    for (int kb = nblk - 1; kb >= 0; kb--) {
      double xv = 0.0;
      if (tid < kTile) {
        const double* Li = a.linv + (size_t)kb * kTile * kTile;
        for (int m = tid; m < kTile; m++) xv += __ldcg(&Li[m * kTile + tid]) * s_x[kb * kTile + m];
      }
      __syncthreads();
      if (tid < kTile) s_x[kb * kTile + tid] = xv;
      __syncthreads();
      for (int col = tid; col < kb * kTile; col += kDfThreads) {
        double sacc = 0.0;
#pragma unroll 8
        for (int r = 0; r < kTile; r++) sacc += __ldcg(&H[(size_t)(kb * kTile + r) * ld + col]) * s_x[kb * kTile + r];
        s_x[col] -= sacc;
      }
      __syncthreads();
    }
  }

  DFP(9);
  // ---------------------------------------------------------------- D. dx, retraction, stop test
  double ss = 0.0;
  for (int t = tid; t < dim; t += kDfThreads) {
    const float d = failed ? 0.0f : (float)(-s_x[t]);
    a.dx_out[t] = d;
    ss += (double)d * (double)d;
  }
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  if ((tid & 31) == 0) s_red[tid >> 5] = ss;
  __syncthreads();
  for (int k = 1 + tid; k < a.N; k += kDfThreads) {
    float T[8], xi[7];
#pragma unroll
    for (int c = 0; c < 8; c++) T[c] = a.Twc[(size_t)k * 8 + c];
#pragma unroll
    for (int c = 0; c < 7; c++) xi[c] = failed ? 0.0f : (float)(-s_x[(k - 1) * 7 + c]);
    sim3_retract(xi, T);
#pragma unroll
    for (int c = 0; c < 8; c++) a.Twc[(size_t)k * 8 + c] = T[c];
  }
  if (tid == 0) {
    double tot = 0.0;
    for (int w = 0; w < kDfThreads / 32; w++) tot += s_red[w];
    const float nrm = (float)sqrt(tot);
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
#ifdef M3R_SOLVE_PROF
    prof[10] = (float)(gtime_ns() - prof_t);
    if (a.ctrl[CTRL_ITERS] == 1)
      printf("DFPROF nblk %d grid %d P %d Q %d | assemble %.1f barrier %.1f | chain: wait-fetch(bar1) %.1f wait-bar4 %.1f trsm %.1f potrf %.1f "
             "syrk %.1f wait-workers-end %.1f | barrier %.1f backsub %.1f tail %.1f | total %.1f us\n",
             nblk, (int)gridDim.x, P, Q, prof[0] * 1e-3f, prof[1] * 1e-3f, prof[2] * 1e-3f, prof[3] * 1e-3f, prof[4] * 1e-3f, prof[5] * 1e-3f,
             prof[6] * 1e-3f, prof[7] * 1e-3f, prof[8] * 1e-3f, prof[9] * 1e-3f, prof[10] * 1e-3f, (float)(gtime_ns() - prof_t0) * 1e-3f);
    if (a.ctrl[CTRL_ITERS] == 1) {
      printf("DFPROF signalL total %.1f us\n", prof[11] * 1e-3f);
      for (int k = 0; k < nblk && k < 58; k += 5)
        printf("DFRAW k %d: potrf-done +0, storeL %lld, signalL %lld, trsm %lld, pubP %lld, syrk %lld | next potrf-done %lld (ns)\n", k,
               (long long)(s_raw[k][1] - s_raw[k][0]), (long long)(s_raw[k][5] - s_raw[k][0]), (long long)(s_raw[k][2] - s_raw[k][0]),
               (long long)(s_raw[k][3] - s_raw[k][0]), (long long)(s_raw[k][4] - s_raw[k][0]),
               k + 1 < nblk && k + 1 < 58 ? (long long)(s_raw[k + 1][0] - s_raw[k][0]) : 0ll);
    }
#endif
  }
}

// one CTA: systems with dim <= kSmallDim when the solve cannot ride in the accumulate kernel
// (sharded runs: the all-reduce sits between the two; or the stand-alone m3r_gn_solve_retract call)
__global__ void __launch_bounds__(kSolveThreads)
gn_small_solve_kernel(const SmallSolveArgs a) {
  if (a.ctrl[CTRL_DONE]) return;
  extern __shared__ double s_small[];
  small_solve(a, s_small);
}

int g_coop_grid = -1;  // CTAs that can be co-resident for the cooperative variant (per process)

}  // namespace

unsigned gn_next_epoch() {
  static std::atomic<unsigned> g_epoch{1u << 20};  // generation (epoch >> 20) starts at 1: never the 0 of a zeroed header
  unsigned e = g_epoch.fetch_add(1);
  if (e == 0) e = g_epoch.fetch_add(1);  // (0 means "no tracking kernel")
  return e;
}

int launch_gn_prepare(const int64_t* ii, const int64_t* jj, const GnLayout& L, const GnPtrs& P, cudaStream_t s,
                      float* zero_dx, int n_dx, int32_t* zero_info, unsigned ll_epoch) {
  // tracking-sized graphs may run in the persistent kernel: clear its flag words (two gather buffers of <= kNumSMs x
  // kAcc x 16 B, 192 pose words, two buffers of <= kNumSMs x 16 B cost words: 172 032 B at the start of P.partials)
  int n_ll16 = 0;
  if (ll_epoch != 0 && L.num_edges <= kTrackMaxEdges && L.dim <= kSmallDim)
    n_ll16 = (2 * kNumSMs * kAcc * 16 + 192 * 8 + 2 * kNumSMs * 16) / 16;
  gn_prepare_kernel<<<1, 1024, 0, s>>>(ii, jj, L.num_edges, L.num_kf, P.ci, P.cj, P.bitmap, P.prefix,
                                       P.inc_off, P.inc_list, P.order, P.counters, P.ctrl, zero_dx, n_dx, zero_info,
                                       reinterpret_cast<uint4*>(P.partials), n_ll16, static_cast<int>(ll_epoch >> 20) + 1);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_gn_solve(float* Twc, const GnLayout& L, const GnPtrs& P_, float delta_thresh, float* dx_out,
                    int32_t* info_out, cudaStream_t s) {
  SolveArgs a{};
  a.Twc = Twc; a.eb = P_.eb; a.ci = P_.ci; a.cj = P_.cj; a.inc_off = P_.inc_off; a.inc_list = P_.inc_list;
  a.hd = P_.hd; a.linv = P_.linv; a.ctrl = P_.ctrl; a.dx_out = dx_out; a.info_out = info_out;
  a.N = L.num_kf; a.dim = L.dim; a.nblk = L.nblk; a.ld = L.ld; a.delta_thresh = delta_thresh;
  size_t dyn = sizeof(double) * (size_t)L.nblk * kTile;
  if (dyn > 128 * 1024) return M3R_ERR_UNSUPPORTED;  // > 16k unknowns: out of scope for a dense solve
  if (L.dim <= kSmallDim) {
    SmallSolveArgs ss{};
    ss.Twc = Twc; ss.eb = P_.eb; ss.ci = P_.ci; ss.cj = P_.cj; ss.inc_off = P_.inc_off; ss.inc_list = P_.inc_list;
    ss.ctrl = P_.ctrl; ss.dx_out = dx_out; ss.info_out = info_out; ss.N = L.num_kf; ss.dim = L.dim;
    ss.delta_thresh = delta_thresh;
    gn_small_solve_kernel<<<1, kSolveThreads, small_solve_smem_bytes(L.dim), s>>>(ss);
    M3R_LAUNCH_CHECK();
    return M3R_OK;
  }
  static const bool force_grid = [] {
    const char* e = getenv("M3R_SOLVE");
    return e && e[0] == 'v' && e[1] == '1';
  }();
  if (L.nblk <= kResidentMaxTiles && !force_grid) {
    // up to 33 keyframes: the whole factorisation in the shared memory of one CTA (M3R_SOLVE=v1 keeps the grid kernel: A/B)
    ResidentSolveArgs r{};
    r.Twc = Twc; r.eb = P_.eb; r.ci = P_.ci; r.cj = P_.cj; r.inc_off = P_.inc_off; r.inc_list = P_.inc_list;
    r.ctrl = P_.ctrl; r.dx_out = dx_out; r.info_out = info_out; r.N = L.num_kf; r.dim = L.dim; r.nblk = L.nblk;
    r.delta_thresh = delta_thresh;
    return launch_gn_solve_resident(r, s);
  }
  // M3R_SOLVE=df: the experimental dataflow factorisation (gn_solve_df_kernel).  Parity-green, but on B200 its chain <->
  // helper <-> worker hand-offs (0.4 us per L2 hop, 1-1.5 us per release fence) keep it at or behind the grid-step
  // kernel (profiles/r02c_*): opt-in only.
  static const bool use_df = [] {
    const char* e = getenv("M3R_SOLVE");
    return e && e[0] == 'd' && e[1] == 'f';
  }();
  if (use_df) {
    // dataflow factorisation: 1 chain CTA + kHelpers row helpers + P x Q workers, all co-resident
    const size_t dyn2 = sizeof(double) * ((size_t)kDfTiles * kTileSm + (size_t)L.nblk * kTile);
    if (dyn2 > 200 * 1024) return M3R_ERR_UNSUPPORTED;
    int dev = 0;
    M3R_CUDA_CHECK(cudaGetDevice(&dev));
    static int cap_dev = -1, cap_ctas = 0;
    static size_t cap_dyn = 0;
    if (cap_dev != dev || cap_dyn != dyn2) {  // (cheap; keyed by device and dynamic shared memory)
      int sms = 0, per_sm = 0;
      M3R_CUDA_CHECK(cudaFuncSetAttribute(gn_solve_df_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn2));
      M3R_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      M3R_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gn_solve_df_kernel, kDfThreads, dyn2));
      if (per_sm < 1) return M3R_ERR_UNSUPPORTED;
      if (const char* e = getenv("M3R_SOLVE_CTAS_PER_SM")) per_sm = per_sm < atoi(e) ? per_sm : atoi(e);  // experiments
      cap_ctas = sms * (per_sm > 3 ? 3 : per_sm);
      cap_dev = dev;
      cap_dyn = dyn2;
    }
    const int max_workers = cap_ctas - 1 - kHelpers;
    if (max_workers < 1) return M3R_ERR_UNSUPPORTED;
    int Q = 1;
    while ((Q + 1) * (Q + 1) <= max_workers && Q + 1 <= L.nblk) Q++;
    int P = max_workers / Q;
    if (P > L.nblk + 1) P = L.nblk + 1;
    a.done = P_.flags;
    a.rdy = P_.flags + (size_t)(L.nblk + 1) * L.nblk;
    a.ll = P_.ll;
    a.P = P;
    a.Q = Q;
    int grid = 1 + kHelpers + P * Q;
    if (grid < L.nblk && L.nblk <= cap_ctas) grid = L.nblk;  // phase C wants one CTA per tile column
    void* args[] = {(void*)&a};
    M3R_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)gn_solve_df_kernel, dim3(grid), dim3(kDfThreads), args, dyn2, s));
    return M3R_OK;
  }
  static const bool no_pipe = getenv("M3R_SOLVE_NOPIPE") != nullptr;  // A/B knob: the round-1 trailing update
  if (L.nblk >= kLookaheadMinTiles && !no_pipe) {
    const int off = (L.nblk * kTile + 1) & ~1;  // 16-byte aligned behind s_x
    // ring + (v4) L_kk^-1, L_Ik, L_Jk and the chain's potrf scratch
    const size_t dyn_pipe = sizeof(double) * ((size_t)off + (size_t)kUpdStages * kUpdStageDoubles + 3 * kTile * kTile + 128);
    if (dyn_pipe <= 200 * 1024) {  // (beyond ~7000 unknowns the round-1 path runs)
      a.pipe_off = off;
      dyn = dyn_pipe;
      static const bool no_v4 = getenv("M3R_SOLVE_NOV4") != nullptr;  // A/B knob: the two-barrier v2 step
      a.v4 = no_v4 ? 0 : 1;
      a.ybuf = P_.partials;  // (the accumulate's per-CTA partials are dead by now: >= 624 x 35 doubles)
    }
  }
  if (dyn > 48 * 1024 - 20 * 1024) {
    M3R_CUDA_CHECK(cudaFuncSetAttribute(gn_solve_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
  }
  static int coop_dev = -1;
  static size_t coop_dyn = 0;
  int dev_now = 0;
  M3R_CUDA_CHECK(cudaGetDevice(&dev_now));
  if (g_coop_grid < 0 || coop_dev != dev_now || coop_dyn != dyn) {
    int sms = 0, per_sm = 0;
    M3R_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev_now));
    M3R_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gn_solve_kernel<true>, kSolveThreads, dyn));
    if (per_sm < 1) return M3R_ERR_UNSUPPORTED;
    // two CTAs per SM: the trailing update is bound by the latency of its tile loads, a second CTA hides some
    // of it (measured at dim 1785: 1.98 -> 1.87 ms per solve); more only makes the grid barriers dearer
    g_coop_grid = sms * (per_sm >= 2 ? 2 : 1);
    coop_dev = dev_now;
    coop_dyn = dyn;
  }
  int grid = g_coop_grid;
  const int max_useful = L.nblk * (L.nblk + 1) / 2 + L.nblk;
  if (grid > max_useful) grid = max_useful;
  void* args[] = {(void*)&a};
  M3R_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)gn_solve_kernel<true>, dim3(grid), dim3(kSolveThreads), args, dyn, s));
  return M3R_OK;
}

int launch_pose_retr(float* Twc, const float* dx, int num_kf, int num_fix, cudaStream_t s) {
  if (num_kf <= num_fix) return M3R_OK;
  pose_retr_kernel<<<(num_kf - num_fix + 127) / 128, 128, 0, s>>>(Twc, dx, num_kf, num_fix);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r
