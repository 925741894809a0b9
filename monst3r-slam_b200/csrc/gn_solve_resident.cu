// Shared-memory-resident solve for graphs of up to 33 keyframes (system dimension <= 224 = 7 tiles of 32): the whole
// lower triangle of H (28 tiles, 224 KB) lives in the shared memory of ONE CTA from the assemble to the last
// substitution, so the factorisation has no grid barrier, no flag and no L2 round trip on its critical path.
//
// Replaces, like gn_solve_kernel, the HOST code of the reference: SparseBlock::update_lhs / update_rhs / solve with
// Eigen::SimplicialLLT<double> (gn_kernels.cu:57-159), pose_retr_kernel (:415-453) and the norm test (:1219-1222).
//
// Why (B200 measurements, profiles/r02a_ubench_solve.log and r02c_*): at 7 tiles the multi-CTA kernel spends its time
// in hand-offs -- a grid barrier costs 1.24 us, a flag hop through L2 0.4 us, a release fence 1-1.5 us -- while the
// arithmetic is tiny (3.4 MFLOP).  One SM has 243 GFLOP/s of DMMA: the trailing updates of the whole factorisation
// are ~15 us of one SM, the seven dependent 32x32 tile factorisations ~28 us, and nothing else has to be waited for.
//
// Layout: tile (i,j), j <= i, at s_dyn + tri(i,j) * 1024 doubles, row-major with stride 32 and the column index
// XOR-swizzled by (row & 7) << 2, which makes the 8x4 fragment loads of mma.m8n8k4.f64 conflict-free without padding
// (padding does not fit: 28 tiles are 229 376 of the 232 448 bytes a CTA can have).  The right-hand side g is carried
// as a vector y (the row "below" the matrix), so the factorisation also yields y = L^-1 g.
#include "gn_tile.cuh"

namespace m3r {
namespace {

constexpr int kResThreads = 256;
constexpr int kResWarps = kResThreads / 32;

__global__ void __launch_bounds__(kResThreads, 1)
gn_solve_resident_kernel(const ResidentSolveArgs a) {
  if (a.ctrl[CTRL_DONE]) return;
  extern __shared__ __align__(16) double s_dyn[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nb = a.nblk, dim = a.dim, dimP = nb * kTile, nfree = a.N - 1;
  const int ntile = nb * (nb + 1) / 2;
  double* Tm = s_dyn;                   // ntile tiles
  double* y = Tm + ntile * kTileEl;     // dimP: g, then y = L^-1 g, then x
  double* col = y + dimP;               // 2 x 32 (potrf column buffer)
  double* rinv = col + 2 * kTile;       // 32
  double* red = rinv + kTile;           // kResWarps
  int* s_fail = reinterpret_cast<int*>(red + kResWarps);
  auto tile = [&](int i, int j) -> double* { return Tm + tri(i, j) * kTileEl; };
  auto elem = [&](int R, int C) -> double& { return Tm[tri(R >> 5, C >> 5) * kTileEl + sw(R & 31, C & 31)]; };

#ifdef M3R_SOLVE_PROF
  long long pt[8];
  int pn = 0;
#define RP() do { if (tid == 0) pt[pn++] = clock64(); } while (0)
#else
#define RP() do { } while (0)
#endif
  RP();
  // ---------------------------------------------------------------- A. assemble, in shared memory
  for (int t = tid; t < (ntile * kTileEl + dimP) / 2; t += kResThreads) reinterpret_cast<double2*>(s_dyn)[t] = make_double2(0.0, 0.0);
  if (tid == 0) *s_fail = 0;
  __syncthreads();
#ifdef M3R_SOLVE_PROF
  const long long pz = clock64();
#endif
  if (tid < 7 * nfree) {
    // thread (p, r) owns row r of every 7x7 block in block row p: all poses in flight at once (<= 224 threads), incident
    // edges in list order (deterministic, no atomics): H_pp += A_e, H_pq -= A_e (q < p).  The three dependent global
    // reads per entry (list -> endpoint -> block) are issued eight entries at a time.
    const int p = tid / 7, r = tid % 7;
    double diag[7] = {0, 0, 0, 0, 0, 0, 0};
    double gsum = 0.0;  // g[7p + r] = sum of -b_e (pose is the edge's i) / +b_e (its j)
    const int lo = a.inc_off[p + 1], hi = a.inc_off[p + 2];
    for (int q0 = lo; q0 < hi; q0 += 8) {
      int ent[8], oth[8];
      double v[8][7], bv[8];
#pragma unroll
      for (int u = 0; u < 8; u++) ent[u] = q0 + u < hi ? __ldg(a.inc_list + q0 + u) : -1;
#pragma unroll
      for (int u = 0; u < 8; u++) {
        oth[u] = -1;
        if (ent[u] >= 0) {
          const int e = ent[u] >> 1;
          oth[u] = ((ent[u] & 1) ? __ldg(a.ci + e) : __ldg(a.cj + e)) - 1;
          const double* blk = a.eb + (size_t)e * kEdgeBlock + 7 * r;
#pragma unroll
          for (int cc = 0; cc < 7; cc++) v[u][cc] = __ldcg(blk + cc);
          bv[u] = __ldcg(a.eb + (size_t)e * kEdgeBlock + 49 + r);
        }
      }
#pragma unroll
      for (int u = 0; u < 8; u++)
        if (ent[u] >= 0) {
          gsum += (ent[u] & 1) ? bv[u] : -bv[u];
#pragma unroll
          for (int cc = 0; cc < 7; cc++) diag[cc] += v[u][cc];
          if (oth[u] >= 0 && oth[u] < p) {
            // seven loads, then seven stores: written as `elem -= v` the compiler has to assume the seven shared-memory
            // addresses may alias and serialises load -> add -> store seven times per entry
            double* ep[7];
            double t[7];
#pragma unroll
            for (int cc = 0; cc < 7; cc++) {
              ep[cc] = &elem(7 * p + r, 7 * oth[u] + cc);
              t[cc] = *ep[cc];
            }
#pragma unroll
            for (int cc = 0; cc < 7; cc++) *ep[cc] = t[cc] - v[u][cc];
          }
        }
    }
#pragma unroll
    for (int cc = 0; cc < 7; cc++)
      if (cc <= r) elem(7 * p + r, 7 * p + cc) = diag[cc];
    y[tid] = gsum;
  }
  for (int R = dim + tid; R < dimP; R += kResThreads) elem(R, R) = 1.0;  // padding rows: identity
  __syncthreads();

  RP();
  // ---------------------------------------------------------------- B. Cholesky (right-looking, look-ahead of one tile)
  if (warp == 0 && potrf_inplace(tile(0, 0), rinv, col, lane) && lane == 0) *s_fail = 1;
  __syncthreads();
#ifdef M3R_SOLVE_PROF
  long long fp[4] = {0, 0, 0, 0}, ft = clock64();
#define FP(i) do { if (tid == 0) { const long long n_ = clock64(); fp[i] += n_ - ft; ft = n_; } } while (0)
#else
#define FP(i) do { } while (0)
#endif
  for (int k = 0; k < nb; k++) {
    // panel: X <- X L_kk^-T for the tiles below the diagonal and for the right-hand side, one warp each
    const double* Lkk = tile(k, k);
    for (int task = k + 1 + warp; task <= nb; task += kResWarps) {
      double x[kTile];
      double* X = task < nb ? tile(task, k) : nullptr;  // task == nb: the right-hand side (every lane computes the same row)
#pragma unroll
      for (int c = 0; c < kTile; c += 2) {
        const double2 v = X ? *reinterpret_cast<const double2*>(X + sw(lane, c)) : *reinterpret_cast<const double2*>(y + k * kTile + c);
        x[c] = v.x;
        x[c + 1] = v.y;
      }
      trsm_rows_sw(x, Lkk, rinv);
      if (X) {
#pragma unroll
        for (int c = 0; c < kTile; c += 2) *reinterpret_cast<double2*>(X + sw(lane, c)) = make_double2(x[c], x[c + 1]);
      } else if (lane == 0) {
#pragma unroll
        for (int c = 0; c < kTile; c += 2) *reinterpret_cast<double2*>(y + k * kTile + c) = make_double2(x[c], x[c + 1]);
      }
    }
    __syncthreads();
    FP(0);
    if (k + 1 == nb) break;
    // trailing update with column k (tiles (i,j), k < j <= i < nb, and y_j -= L_jk y_k) with a look-ahead of one tile:
    // warp 1 updates the next diagonal tile first and hands it to warp 0 over named barrier 1; warp 0 factors it while
    // warps 1-7 run the rest of the update.  Same box, tools/ab_solve.py 32,120: 140-141 us per iteration against
    // 149.6 us with the factorisation after the update (profiles/r02d).
    if (warp == 0) {
      asm volatile("bar.sync 1, 64;" ::: "memory");
      if (potrf_inplace(tile(k + 1, k + 1), rinv, col, lane) && lane == 0) *s_fail = 1;
    } else {
      if (warp == 1) {
        warp_gemm_sub(tile(k + 1, k + 1), tile(k + 1, k), tile(k + 1, k), lane);
        asm volatile("bar.sync 1, 64;" ::: "memory");
      }
      int idx = 1;  // (warp 1 starts one task late: it has done the diagonal tile)
      for (int i = k + 1; i < nb; i++)
        for (int j = k + 1; j <= i; j++) {
          if (i == k + 1) continue;  // (k+1, k+1): done above
          if ((idx++) % (kResWarps - 1) == warp - 1) warp_gemm_sub(tile(i, j), tile(i, k), tile(j, k), lane);
        }
      for (int j = k + 1; j < nb; j++)
        if ((idx++) % (kResWarps - 1) == warp - 1) {
          const double* Ljk = tile(j, k);
          double acc = 0.0;
#pragma unroll 8
          for (int c = 0; c < kTile; c++) acc += Ljk[sw(lane, c)] * y[k * kTile + c];
          y[j * kTile + lane] -= acc;
        }
    }
    FP(1);
    __syncthreads();
    FP(3);
  }
#ifdef M3R_SOLVE_PROF
  if (tid == 0 && a.ctrl[CTRL_ITERS] == 0)
    printf("RESFACT: panel trsm %.1f, look-ahead potrf (warp 0) %.1f, wait for the update %.1f us\n", fp[0] / 1965.0, fp[1] / 1965.0, fp[3] / 1965.0);
#endif

  RP();
  // ---------------------------------------------------------------- C. back-substitution  L^T x = y (in place in y)
  for (int j = nb - 1; j >= 0; j--) {
    if (warp == 0) {
      const double* Ljj = tile(j, j);
      const double ri = 1.0 / Ljj[sw(lane, lane)];
      double s = y[j * kTile + lane];
#pragma unroll
      for (int c = kTile - 1; c >= 0; c--) {
        const double xc = __shfl_sync(0xffffffffu, s * ri, c);
        if (lane == c) s = xc;
        if (lane < c) s -= Ljj[sw(c, lane)] * xc;
      }
      y[j * kTile + lane] = s;
    }
    __syncthreads();
    for (int kb = warp; kb < j; kb += kResWarps) {  // y_kb -= L_(j,kb)^T x_j
      const double* Ljk = tile(j, kb);
      double acc = 0.0;
#pragma unroll 8
      for (int r = 0; r < kTile; r++) acc += Ljk[sw(r, lane)] * y[j * kTile + r];
      y[kb * kTile + lane] -= acc;
    }
    __syncthreads();
  }

  RP();
  // ---------------------------------------------------------------- D. dx, retraction, stop test
  const bool failed = *s_fail != 0;  // non-SPD => dx = 0 (gn_kernels.cu:147-150)
  double ss = 0.0;
  for (int t = tid; t < dim; t += kResThreads) {
    const float d = failed ? 0.0f : (float)(-y[t]);
    a.dx_out[t] = d;
    ss += (double)d * (double)d;
  }
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  if (lane == 0) red[warp] = ss;
  __syncthreads();
  for (int k = 1 + tid; k < a.N; k += kResThreads) {
    float T[8], xi[7];
#pragma unroll
    for (int c = 0; c < 8; c++) T[c] = a.Twc[(size_t)k * 8 + c];
#pragma unroll
    for (int c = 0; c < 7; c++) xi[c] = failed ? 0.0f : (float)(-y[(k - 1) * 7 + c]);
    sim3_retract(xi, T);
#pragma unroll
    for (int c = 0; c < 8; c++) a.Twc[(size_t)k * 8 + c] = T[c];
  }
  if (tid == 0) {
    double tot = 0.0;
    for (int w = 0; w < kResWarps; w++) tot += red[w];
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
    if (a.ctrl[CTRL_ITERS] == 1)
      printf("RESPROF nb %d: (zero fill %.1f) assemble %.1f, factor %.1f, backsub %.1f, tail %.1f us (at 1.965 GHz)\n", nb, (pz - pt[0]) / 1965.0, (pt[1] - pt[0]) / 1965.0,
             (pt[2] - pt[1]) / 1965.0, (pt[3] - pt[2]) / 1965.0, (clock64() - pt[3]) / 1965.0);
#endif
  }
}

}  // namespace

size_t gn_solve_resident_smem_bytes(int nblk) {
  return sizeof(double) * ((size_t)(nblk * (nblk + 1) / 2) * kTileEl + (size_t)nblk * kTile + 2 * kTile + kTile + kResWarps + 2);
}

int launch_gn_solve_resident(const ResidentSolveArgs& a, cudaStream_t s) {
  if (a.nblk < 1 || a.nblk > kResidentMaxTiles) return M3R_ERR_UNSUPPORTED;
  const size_t dyn = gn_solve_resident_smem_bytes(a.nblk);
  static DynSmemOptIn opt_in;  // (raised when a larger system shows up)
  M3R_CUDA_CHECK(opt_in.ensure(gn_solve_resident_kernel, (int)dyn));
  gn_solve_resident_kernel<<<1, kResThreads, dyn, s>>>(a);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r
