// Device pieces shared by the shared-memory-resident solve (gn_solve_resident.cu) and the look-ahead chain of the grid
// solve (gn_solve.cu): 32 x 32 fp64 tiles in shared memory, row-major with stride 32 and the column index XOR-swizzled
// by (row & 7) << 2 (conflict-free 8 x 4 fragment loads for mma.m8n8k4.f64 without padding; 16-byte pairs stay together).
#pragma once
#include "gn_device.cuh"

namespace m3r {
namespace {

constexpr int kTileEl = kTile * kTile;

__device__ __forceinline__ int sw(int r, int c) { return r * kTile + (c ^ ((r & 7) << 2)); }
__device__ __forceinline__ int tri(int i, int j) { return i * (i + 1) / 2 + j; }

__device__ __forceinline__ void dmma884(double& c0, double& c1, const double a, const double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// Cholesky of one tile IN PLACE by one warp (lane = row; see tile_potrf_invert in gn_solve.cu for the scheme):
// T <- L (upper part zeroed), rinv[c] = 1 / L[c][c].  Returns true on a pivot <= 0 (or NaN).
__device__ __forceinline__ bool potrf_inplace(double* T, double* rinv, double* col, int lane) {
  double a[kTile];
#pragma unroll
  for (int c = 0; c < kTile; c += 2) {
    const double2 v = *reinterpret_cast<const double2*>(T + sw(lane, c));
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
  const double rs = rsqrt(my_d);
  col[lane] = rs;  // (lower half: last used by pivot 30; every lane is past pivot 31's barrier)
  rinv[lane] = rs;
  __syncwarp();
#pragma unroll
  for (int c = 0; c < kTile; c += 2) {
    const double2 s = *reinterpret_cast<const double2*>(col + c);
    *reinterpret_cast<double2*>(T + sw(lane, c)) = make_double2(c <= lane ? a[c] * s.x : 0.0, c + 1 <= lane ? a[c + 1] * s.y : 0.0);
  }
  return bad;
}

// rows of X <- rows of X * L^-T by forward substitution (one warp, lane = row); L is a swizzled tile, read by column.
// Two columns per step: the swizzle keeps an even / odd column pair adjacent, so one LDS.128 brings L[c2][C] and
// L[c2][C+1] -- half the shared-memory instructions of a column-at-a-time sweep, same operations in the same order.
template <int C>
__device__ __forceinline__ void trsm_step2_sw(double (&x)[kTile], const double* L, const double* rinv) {
  const double x0 = x[C] * rinv[C];
  const double x1 = (x[C + 1] - x0 * L[(C + 1) * kTile + (C ^ (((C + 1) & 7) << 2))]) * rinv[C + 1];
  x[C] = x0;
  x[C + 1] = x1;
#pragma unroll
  for (int c2 = 0; c2 < kTile; c2++)
    if (c2 > C + 1) {
      const double2 l = *reinterpret_cast<const double2*>(L + c2 * kTile + (C ^ ((c2 & 7) << 2)));
      x[c2] -= x0 * l.x;
      x[c2] -= x1 * l.y;
    }
  if constexpr (C + 2 < kTile) trsm_step2_sw<C + 2>(x, L, rinv);
}
// (template recursion instead of nested `#pragma unroll`: the compiler left the inner loop rolled, which put x[] in
// local memory -- 4x slower, profiles/r02d)
__device__ __forceinline__ void trsm_rows_sw(double (&x)[kTile], const double* L, const double* rinv) {
  trsm_step2_sw<0>(x, L, rinv);
}

// D -= A * B^T, all three swizzled tiles, by ONE warp (16 m8n8 accumulators)
__device__ __forceinline__ void warp_gemm_sub(double* D, const double* A, const double* B, int lane) {
  const int g = lane >> 2, t = lane & 3;
  double acc[4][4][2];
#pragma unroll
  for (int mi = 0; mi < 4; mi++)
#pragma unroll
    for (int ni = 0; ni < 4; ni++) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
#pragma unroll
  for (int kk = 0; kk < 8; kk++) {
    const int pc = (kk * 4 + t) ^ (g << 2);  // rows 8*mi + g: (row & 7) == g
    double af[4], bf[4];
#pragma unroll
    for (int mi = 0; mi < 4; mi++) {
      af[mi] = A[(mi * 8 + g) * kTile + pc];
      bf[mi] = B[(mi * 8 + g) * kTile + pc];
    }
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
      for (int ni = 0; ni < 4; ni++) dmma884(acc[mi][ni][0], acc[mi][ni][1], af[mi], bf[ni]);
  }
#pragma unroll
  for (int mi = 0; mi < 4; mi++)
#pragma unroll
    for (int ni = 0; ni < 4; ni++) {
      double2* d = reinterpret_cast<double2*>(D + (mi * 8 + g) * kTile + ((ni * 8 + 2 * t) ^ (g << 2)));
      double2 v = *d;
      v.x -= acc[mi][ni][0];
      v.y -= acc[mi][ni][1];
      *d = v;
    }
}



}  // namespace
}  // namespace m3r
