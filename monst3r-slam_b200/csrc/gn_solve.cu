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
//   B. Cholesky   right-looking, 32x32 tiles, fp64: every CTA factors the diagonal tile redundantly
//                 in shared memory (no broadcast needed), inverts it, and does TRSM on its share of
//                 the panel as a tile GEMM with L_kk^-1; grid.sync; trailing SYRK/GEMM update tiles
//                 spread over all CTAs; grid.sync.
//   C. back-substitution  L^T x = y by CTA 0 with the stored L_kk^-1 tiles, column-parallel.
//   D. dx = -x (0 if a pivot was <= 0: gn_kernels.cu:142-150), retraction T <- Exp(dx) T in fp32
//      following :299-413, ||dx|| stop test, next-iteration bookkeeping.
#include <cooperative_groups.h>

#include <cstdlib>

#include "gn_device.cuh"

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
                  float* __restrict__ zero_dx, int n_dx, int32_t* __restrict__ zero_info) {
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
  if (tid < CTRL_WORDS) ctrl[tid] = 0;
  // outputs of the solve that follows start from zero (max_iter == 0, or a solve that never runs): folded in
  // here instead of two cudaMemsetAsync launches in front of every Gauss-Newton call
  if (zero_dx)
    for (int k = tid; k < n_dx; k += nt) zero_dx[k] = 0.f;
  if (zero_info && tid < 4) zero_info[tid] = 0;
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
};

template <bool COOP>
__device__ __forceinline__ void grid_barrier() {
  if (COOP) {
    cg::this_grid().sync();
  } else {
    __syncthreads();
  }
}

// Factor the 32x32 tile held in Ls (lower, stride 33) and invert the factor, all by WARP 0 with the
// tile in registers (lane r owns row r of A / L, then lane c owns column c of L^-1): no block
// barriers inside, pivots broadcast by shuffle.  Ls <- L (upper part zeroed), Li <- L^-1.
// Returns (to every thread, through s_flag) false on a pivot <= 0.
__device__ __forceinline__ bool tile_potrf_invert(double (*Ls)[kTile + 1], double (*Li)[kTile + 1], double* s_dinv, int* s_flag) {
  const int lane = threadIdx.x;
  if (threadIdx.x < 32) {
    double a[kTile];
#pragma unroll
    for (int c = 0; c < kTile; c++) a[c] = Ls[lane][c];
    bool bad = false;
    double my_rinv = 1.0;
#pragma unroll
    for (int j = 0; j < kTile; j++) {
      double d = __shfl_sync(0xffffffffu, a[j], j);
      if (!(d > 0.0)) {
        bad = true;
        d = 1.0;
      }
      const double rinv = rsqrt(d);
      const double l = a[j] * rinv;  // lane j: sqrt(d); lanes r > j: L[r][j]
      a[j] = l;
      if (lane == j) my_rinv = rinv;
#pragma unroll
      for (int c = j + 1; c < kTile; c++) a[c] -= l * __shfl_sync(0xffffffffu, l, c);
    }
#pragma unroll
    for (int c = 0; c < kTile; c++) Ls[lane][c] = (c <= lane) ? a[c] : 0.0;
    s_dinv[lane] = my_rinv;
    if (lane == 0) *s_flag = bad ? 1 : 0;
    __syncwarp();
    // column `lane` of L^-1 by forward substitution; L[r][m] is a shared-memory broadcast
    double x[kTile];
#pragma unroll
    for (int r = 0; r < kTile; r++) {
      double s0 = (r == lane) ? 1.0 : 0.0, s1 = 0.0;
#pragma unroll
      for (int m = 0; m + 1 < r; m += 2) {
        s0 -= Ls[r][m] * x[m];
        s1 -= Ls[r][m + 1] * x[m + 1];
      }
      if (r & 1) s0 -= Ls[r][r - 1] * x[r - 1];
      x[r] = (r < lane) ? 0.0 : (s0 + s1) * s_dinv[r];
    }
#pragma unroll
    for (int r = 0; r < kTile; r++) Li[r][lane] = x[r];
  }
  __syncthreads();
  return *s_flag == 0;
}

// measured (tools/solve_prof.py): dim 1785 (56 tiles) 1.90 -> 1.66 ms with the look-ahead, dim 217 (7 tiles) 139 -> 160 us
constexpr int kLookaheadMinTiles = 12;

template <bool COOP>
__global__ void __launch_bounds__(kSolveThreads, 2)
gn_solve_kernel(const SolveArgs a) {
  if (a.ctrl[CTRL_DONE]) return;  // uniform over the grid: set only by a previous launch
  __shared__ double sA[kTile][kTile + 1];
  __shared__ double sB[kTile][kTile + 1];
  __shared__ int s_flag;
  __shared__ double s_dinv[kTile];
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
  for (int kb = 0; kb < nblk; kb++) {
    // B1: diagonal tile, inverse, TRSM on this CTA's panel tiles
    if (kb == 0 || !lookahead) {
      for (int t = tid; t < kTile * kTile; t += kSolveThreads) {
        const int r = t >> 5, c = t & 31;
        sA[r][c] = (c <= r) ? __ldcg(&H[(size_t)(kb * kTile + r) * ld + kb * kTile + c]) : 0.0;
      }
      __syncthreads();
      if (!tile_potrf_invert(sA, sB, s_dinv, &s_flag)) failed = true;  // sA = L_kk, sB = L_kk^-1 (redundant per CTA)
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
    // B2: trailing update  A_IJ -= L_Ik * L_Jk^T  for kb < J <= I <= nblk (J < nblk)
    const int rem = nblk - kb;          // tile rows below kb, including the rhs row
    const int ncols = rem - 1;          // tile columns kb+1 .. nblk-1
    // tiles enumerated row by row: row i (1..rem) has min(i, ncols) tiles
    const int ntiles = ncols * (ncols + 1) / 2 + ncols;  // triangle + rhs row
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
        const bool ok = tile_potrf_invert(sB, sA, s_dinv, &s_flag);  // sB = L, sA = L^-1
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
    if (tid < kTile) s_x[kb * kTile + tid] = __ldcg(&H[(size_t)dimP * ld + kb * kTile + tid]);  // y_kb
    const int c = tid & 31, rg = tid >> 5;   // column of the tile, group of 4 rows
    double tile[4];
    if (kb + 1 < nblk) {
#pragma unroll
      for (int q = 0; q < 4; q++) tile[q] = __ldcg(&H[(size_t)((nblk - 1) * kTile + rg * 4 + q) * ld + kb * kTile + c]);
    }
    __syncthreads();
    for (int j = nblk - 1; j > kb; j--) {
      double cur[4];
#pragma unroll
      for (int q = 0; q < 4; q++) cur[q] = tile[q];
      if (j - 1 > kb) {  // next tile in flight while we wait for x_j
#pragma unroll
        for (int q = 0; q < 4; q++) tile[q] = __ldcg(&H[(size_t)((j - 1) * kTile + rg * 4 + q) * ld + kb * kTile + c]);
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

int launch_gn_prepare(const int64_t* ii, const int64_t* jj, const GnLayout& L, const GnPtrs& P, cudaStream_t s,
                      float* zero_dx, int n_dx, int32_t* zero_info) {
  gn_prepare_kernel<<<1, 1024, 0, s>>>(ii, jj, L.num_edges, L.num_kf, P.ci, P.cj, P.bitmap, P.prefix,
                                       P.inc_off, P.inc_list, P.order, P.counters, P.ctrl, zero_dx, n_dx, zero_info);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int launch_gn_solve(float* Twc, const GnLayout& L, const GnPtrs& P, float delta_thresh, float* dx_out,
                    int32_t* info_out, cudaStream_t s) {
  SolveArgs a{};
  a.Twc = Twc; a.eb = P.eb; a.ci = P.ci; a.cj = P.cj; a.inc_off = P.inc_off; a.inc_list = P.inc_list;
  a.hd = P.hd; a.linv = P.linv; a.ctrl = P.ctrl; a.dx_out = dx_out; a.info_out = info_out;
  a.N = L.num_kf; a.dim = L.dim; a.nblk = L.nblk; a.ld = L.ld; a.delta_thresh = delta_thresh;
  const size_t dyn = sizeof(double) * (size_t)L.nblk * kTile;
  if (dyn > 160 * 1024) return M3R_ERR_UNSUPPORTED;  // > 20k unknowns: out of scope for a dense solve
  if (L.dim <= kSmallDim) {
    SmallSolveArgs ss{};
    ss.Twc = Twc; ss.eb = P.eb; ss.ci = P.ci; ss.cj = P.cj; ss.inc_off = P.inc_off; ss.inc_list = P.inc_list;
    ss.ctrl = P.ctrl; ss.dx_out = dx_out; ss.info_out = info_out; ss.N = L.num_kf; ss.dim = L.dim;
    ss.delta_thresh = delta_thresh;
    gn_small_solve_kernel<<<1, kSolveThreads, small_solve_smem_bytes(L.dim), s>>>(ss);
    M3R_LAUNCH_CHECK();
    return M3R_OK;
  }
  if (dyn > 48 * 1024 - 20 * 1024) {
    M3R_CUDA_CHECK(cudaFuncSetAttribute(gn_solve_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
  }
  if (g_coop_grid < 0) {
    int dev = 0, sms = 0, per_sm = 0;
    M3R_CUDA_CHECK(cudaGetDevice(&dev));
    M3R_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    M3R_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gn_solve_kernel<true>, kSolveThreads, dyn));
    if (per_sm < 1) return M3R_ERR_UNSUPPORTED;
    // two CTAs per SM: the trailing update is bound by the latency of its tile loads, a second CTA hides some
    // of it (measured at dim 1785: 1.98 -> 1.87 ms per solve); more only makes the grid barriers dearer
    g_coop_grid = sms * (per_sm >= 2 ? 2 : 1);
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
