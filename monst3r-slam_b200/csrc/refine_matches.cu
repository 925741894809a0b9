// refine_matches — dilated-window descriptor argmax (sm_100a).
//
// Replaces mast3r_slam/backend/src/matching_kernels.cu:25-116.  Spec: SURVEY.md Appendix B.
// Parity rules that decide ties (all kept):
//   * scan order: dilation d = D..1; u-offset outer, v-offset inner (:54-57)
//   * strict `>` against a best score that is carried across levels (:47,:65), initial best =
//     cuda::std::numeric_limits<T>::min(): 0 for c10::Half, FLT_MIN / DBL_MIN for float / double
//   * fp16: score = fold_k( round16( score + round16(q_k * c_k) ) ), k ascending.  c10::Half does
//     this through float; IEEE half mul/add with a single RN rounding give the same bits (the
//     float product of two halves is exact; a float sum of two halves rounded again to half is
//     innocuous double rounding because 24 >= 2*11+2).  HMUL2/HADD2 `_rn` intrinsics are never
//     contracted to HFMA2.
//
// fp16 / fdim = 24 paths (the shape every caller uses).  The arithmetic is the same in both: the
// query descriptor stays packed in 12 registers (HMUL2's .H0_H0/.H1_H1 operand swizzles broadcast one
// element for free) and TWO window candidates are scored at a time, one per half2 lane (candidate
// words are lane-transposed with PRMT), so a pair of MACs for two candidates costs 2 PRMT + 2 HMUL2 +
// 2 HADD2.  Measured on B200 (profiles/r01b_ubench_pipes.log): HMUL2/HADD2 issue at 0.5 warp-inst/clk/
// SMSP, so 245 candidates x 24 (mul, add) pairs put a floor of ~63 us on a 512x384 pair; the
// 48 B per (query, candidate) that must reach the thread put ~62 us on the shared-memory pipe.
//
//   refine_staged_kernel (n == h*w, radius 3, dilation <= 8): a CTA owns a 16x16 tile of queries.
//     Per dilation level the CTA reduces the bounding box of its queries' current centres, TMA-bulk-
//     copies box + halo rows of D11 into shared memory (one cp.async.bulk per row, one mbarrier) and
//     every query scans its 7x7 lattice from there: 6 conflict-free LDS.128 per candidate pair
//     (neighbouring queries sit 48 B apart: 8 lanes x 16 B cover the 32 banks exactly once) instead
//     of 6 LDG.128 with a 12-line L1 footprint each.  If box x halo exceeds the 108 KB buffer the
//     level is split into passes over groups of u-offset columns (scan order is u-outer, so the
//     strict-'>' tie-breaking is unchanged).  Queries whose centre is far from the tile's mean
//     (outliers, wrong matches) are scored from global memory by the same code.
//   refine_gather_kernel: any n / radius / dilation; every candidate read from global (L1/L2).
#include <cuda_fp16.h>

#include <cfloat>
#include <climits>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "tma.cuh"

namespace m3r {
namespace {

constexpr int kRefineThreads = 128;
constexpr int kFD = 24;
constexpr int kPxBytes = kFD * 2;  // 48

__device__ __forceinline__ bool inside(int u, int v, int w, int h) {
  return v >= 0 && v < h && u >= 0 && u < w;
}

union Word4 {
  uint4 v;
  uint32_t w[4];
};
__device__ __forceinline__ __half2 as_h2(uint32_t x) { return *reinterpret_cast<const __half2*>(&x); }

struct Desc {  // one 24-d fp16 descriptor = 3 x 16 B
  Word4 t[3];
};

__device__ __forceinline__ Desc load_desc_global(const unsigned char* p) {
  Desc d;
#pragma unroll
  for (int t = 0; t < 3; t++) d.t[t].v = __ldg(reinterpret_cast<const uint4*>(p) + t);
  return d;
}
__device__ __forceinline__ Desc load_desc_shared(const unsigned char* p) {
  Desc d;
#pragma unroll
  for (int t = 0; t < 3; t++) d.t[t].v = *(reinterpret_cast<const uint4*>(p) + t);
  return d;
}

// scores of two candidates (low lane: a, high lane: b) against the packed query q
__device__ __forceinline__ __half2 score2(const Desc& q, const Desc& a, const Desc& b) {
  __half2 s = __float2half2_rn(0.0f);
#pragma unroll
  for (int t = 0; t < 3; t++) {
#pragma unroll
    for (int c = 0; c < 4; c++) {
      const __half2 lo = as_h2(__byte_perm(a.t[t].w[c], b.t[t].w[c], 0x5410));
      const __half2 hi = as_h2(__byte_perm(a.t[t].w[c], b.t[t].w[c], 0x7632));
      const __half2 qq = as_h2(q.t[t].w[c]);
      s = __hadd2_rn(s, __hmul2_rn(__low2half2(qq), lo));
      s = __hadd2_rn(s, __hmul2_rn(__high2half2(qq), hi));
    }
  }
  return s;
}

// One dilation level with every candidate gathered from global memory.  Scan order of the
// reference: u-offset outer, v-offset inner, strict '>' against the running best.
__device__ __forceinline__ void gather_level(const unsigned char* __restrict__ img, const Desc& q, int u0, int v0,
                                             int d, int radius, int w, int h, float& best, int& ub, int& vb) {
  const int rd = radius * d;
  const int cnt = 2 * radius + 1;
  const int total = cnt * cnt;
  int ia = 0, ja = 0;  // (u-offset index, v-offset index) of candidate A
  for (int c = 0; c < total; c += 2) {
    int ib = ia, jb = ja + 1;
    if (jb == cnt) {
      jb = 0;
      ib = ia + 1;
    }
    const int uA = u0 - rd + ia * d, vA = v0 - rd + ja * d;
    const int uB = u0 - rd + ib * d, vB = v0 - rd + jb * d;
    const bool inA = inside(uA, vA, w, h);
    const bool inB = (c + 1 < total) && inside(uB, vB, w, h);
    if (inA || inB) {
      const Desc a = load_desc_global(img + (inA ? (static_cast<size_t>(vA) * w + uA) * kPxBytes : 0));
      const Desc bq = load_desc_global(img + (inB ? (static_cast<size_t>(vB) * w + uB) * kPxBytes : 0));
      const __half2 s = score2(q, a, bq);
      const float sA = __low2float(s), sB = __high2float(s);
      if (inA && sA > best) {
        best = sA;
        ub = uA;
        vb = vA;
      }
      if (inB && sB > best) {
        best = sB;
        ub = uB;
        vb = vB;
      }
    }
    ja += 2;  // advance A by two candidates
    if (ja >= cnt) {
      ja -= cnt;
      ia += 1;
    }
  }
}

__global__ void __launch_bounds__(kRefineThreads)
refine_gather_kernel(const __half* __restrict__ D11, const __half* __restrict__ D21,
                     const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new, int h, int w,
                     int n, int radius, int dilation_max) {
  const int k = blockIdx.x * kRefineThreads + threadIdx.x;
  const int b = blockIdx.y;
  if (k >= n) return;
  const size_t qi = static_cast<size_t>(b) * n + k;
  const unsigned char* img = reinterpret_cast<const unsigned char*>(D11) + static_cast<size_t>(b) * h * w * kPxBytes;
  const Desc q = load_desc_global(reinterpret_cast<const unsigned char*>(D21) + qi * kPxBytes);
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + qi);
  int u0 = static_cast<int>(p.x), v0 = static_cast<int>(p.y);
  int ub = u0, vb = v0;
  float best = 0.0f;  // numeric_limits<c10::Half>::min() == Half() == 0
  for (int d = dilation_max; d > 0; d--) {
    gather_level(img, q, u0, v0, d, radius, w, h, best, ub, vb);
    u0 = ub;
    v0 = vb;
  }
  const bool same = ub == static_cast<int>(p.x) && vb == static_cast<int>(p.y);  // keep out-of-int-range inputs intact
  reinterpret_cast<longlong2*>(p1_new)[qi] = same ? p : make_longlong2(ub, vb);
}

// ---- staged path ---------------------------------------------------------------------------------
#ifndef M3R_TILE_X
#define M3R_TILE_X 16
#endif
// queries per CTA (x: a power of two).  32 x 16 with one 220 KB box per SM stages every level in one pass and was
// measured within 2 % of 16 x 16 with two CTAs per SM on the benchmark pair (and 5 % slower on perfectly coherent
// matches), so 16 x 16 stays the default; -DM3R_TILE_X=32 builds the other one.  Tile height, box capacity and
// CTAs per SM are build parameters too (same-box A/B at 512x384, 16 x 16 / 2304 px / 2 CTAs = 0.161 ms:
// 16 x 12 -> 0.162, 16 x 10 -> 0.164, 16 x 12 with 1536 px and 3 CTAs -> 0.182, 16 x 8 with 1152 px and 4 CTAs
// -> 0.191): the finer load balance of smaller tiles is paid back by their larger halo share.
#ifndef M3R_TILE_Y
#define M3R_TILE_Y 16
#endif
#ifndef M3R_CAP_PX
#define M3R_CAP_PX (M3R_TILE_X == 16 ? 2304 : 4608)
#endif
#ifndef M3R_REFINE_CTAS
#define M3R_REFINE_CTAS (M3R_TILE_X == 16 ? 2 : 1)
#endif
constexpr int kTileX = M3R_TILE_X, kTileY = M3R_TILE_Y;
constexpr int kStThreads = kTileX * kTileY;
static_assert(kStThreads % 32 == 0, "whole warps only: the box statistics use full-mask warp reductions");
constexpr int kStRadius = 3, kStCnt = 7;
constexpr int kStMaxLevels = 8;
constexpr int kCapPx = M3R_CAP_PX;  // staged pixels per pass: 110 592 B -> 2 CTAs / SM (3 CTAs x 72 KB measured slower: 177 vs 166 us)
constexpr int kStSmemBytes = kCapPx * kPxBytes;
constexpr int kSane = 1 << 16;                // |centre| beyond this never enters the box statistics
constexpr int kInlierRadX = kTileX / 2 + 4;   // robust pass: centres within +-(half tile + 4) px of the tile mean
constexpr int kInlierRadY = kTileY / 2 + 4;

// columns [i0, i0+c) of one level for one query, candidates read from the staged box.
// base/rowstep/colstep are byte offsets into the box; bc receives (i << 3 | j) of a new best.
template <bool CHECK>
__device__ __forceinline__ void staged_columns(const Desc& q, const unsigned char* __restrict__ sbox, int base,
                                               int rowstep, int colstep, int i0, int c, int u0, int v0, int rd,
                                               int d, int w, int h, float& best, int& bc) {
  unsigned vmask = 0x7f;
  if (CHECK) {
    vmask = 0;
#pragma unroll
    for (int j = 0; j < kStCnt; j++)
      vmask |= (static_cast<unsigned>(v0 - rd + j * d) < static_cast<unsigned>(h) ? 1u : 0u) << j;
  }
  int i = i0;
#pragma unroll 1
  for (; i + 1 < i0 + c; i += 2) {  // two columns at a time: lane lo = column i, lane hi = column i+1
    bool okA = true, okB = true;
    if (CHECK) {
      okA = static_cast<unsigned>(u0 - rd + i * d) < static_cast<unsigned>(w);
      okB = static_cast<unsigned>(u0 - rd + (i + 1) * d) < static_cast<unsigned>(w);
      if (!(okA || okB) || vmask == 0) continue;
    }
    const uint32_t colmask = (okA ? 0x0000ffffu : 0u) | (okB ? 0xffff0000u : 0u);
    __half2 s[kStCnt];
    int off = base + i * colstep;
#pragma unroll
    for (int j = 0; j < kStCnt; j++) {
      int offA = off, offB = off + colstep;
      if (CHECK) {
        const bool row = (vmask >> j) & 1u;
        offA = (row && okA) ? offA : 0;
        offB = (row && okB) ? offB : 0;
      }
      const Desc a = load_desc_shared(sbox + offA);
      const Desc bq = load_desc_shared(sbox + offB);
      s[j] = score2(q, a, bq);
      if (CHECK) {
        const uint32_t m = ((vmask >> j) & 1u) ? colmask : 0u;
        uint32_t bits = *reinterpret_cast<uint32_t*>(&s[j]) & m;  // out-of-image candidates score +0: never > best
        s[j] = *reinterpret_cast<__half2*>(&bits);
      }
      off += rowstep;
    }
    __half2 m2 = s[0];
#pragma unroll
    for (int j = 1; j < kStCnt; j++) m2 = __hmax2(m2, s[j]);
    if (fmaxf(__low2float(m2), __high2float(m2)) > best) {
#pragma unroll
      for (int j = 0; j < kStCnt; j++) {
        const float x = __low2float(s[j]);
        if (x > best) {
          best = x;
          bc = (i << 3) | j;
        }
      }
#pragma unroll
      for (int j = 0; j < kStCnt; j++) {
        const float x = __high2float(s[j]);
        if (x > best) {
          best = x;
          bc = ((i + 1) << 3) | j;
        }
      }
    }
  }
  if (i < i0 + c) {  // last single column: lane lo = row j, lane hi = row j+1
    bool okA = true;
    if (CHECK) okA = static_cast<unsigned>(u0 - rd + i * d) < static_cast<unsigned>(w);
    if (okA && vmask != 0) {
      __half2 s[4];
      int off = base + i * colstep;
#pragma unroll
      for (int jp = 0; jp < 4; jp++) {
        const int j = 2 * jp;
        int offA = off, offB = (j + 1 < kStCnt) ? off + rowstep : off;
        if (CHECK) {
          offA = ((vmask >> j) & 1u) ? offA : 0;
          offB = (j + 1 < kStCnt && ((vmask >> (j + 1)) & 1u)) ? offB : 0;
        }
        const Desc a = load_desc_shared(sbox + offA);
        const Desc bq = load_desc_shared(sbox + offB);
        s[jp] = score2(q, a, bq);
        uint32_t m = 0xffffffffu;
        if (j + 1 >= kStCnt) m = 0x0000ffffu;
        if (CHECK) m &= (((vmask >> j) & 1u) ? 0x0000ffffu : 0u) | (((vmask >> (j + 1)) & 1u) ? 0xffff0000u : 0u);
        uint32_t bits = *reinterpret_cast<uint32_t*>(&s[jp]) & m;
        s[jp] = *reinterpret_cast<__half2*>(&bits);
        off += 2 * rowstep;
      }
      const __half2 m2 = __hmax2(__hmax2(s[0], s[1]), __hmax2(s[2], s[3]));
      if (fmaxf(__low2float(m2), __high2float(m2)) > best) {
#pragma unroll
        for (int jp = 0; jp < 4; jp++) {
          const float x = __low2float(s[jp]);
          if (x > best) {
            best = x;
            bc = (i << 3) | (2 * jp);
          }
          const float y = __high2float(s[jp]);
          if (y > best) {  // the padded 8th lane holds +0 and never wins
            best = y;
            bc = (i << 3) | (2 * jp + 1);
          }
        }
      }
    }
  }
}

#ifdef M3R_REFINE_PROF  // per-CTA timeline (tools/refine_prof.py): [cta][0..3] = t0, t1 (globaltimer ns), smid, passes;
// [4+4*lvl ..] = cycles waiting for the box, cycles scoring, cycles in the global-gather fallback, outlier count
__device__ long long g_refine_prof[4096 * 40];
__device__ __forceinline__ long long gtime() {
  long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define RPROF(...) __VA_ARGS__
#else
#define RPROF(...)
#endif

__global__ void __launch_bounds__(kStThreads, M3R_REFINE_CTAS)
refine_staged_kernel(const __half* __restrict__ D11, const __half* __restrict__ D21,
                     const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new, int h, int w,
                     int tiles_x, int dilation_max, int reuse_ok) {
  extern __shared__ __align__(128) unsigned char sbox[];
  __shared__ int s_red[kStMaxLevels][8];   // umin, umax, vmin, vmax, sum u, sum v, count
  __shared__ int s_red2[kStMaxLevels][4];  // robust pass: umin, umax, vmin, vmax over inliers
  __shared__ __align__(8) uint64_t s_bar;

  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
  const int qx = tx * kTileX + (tid & (kTileX - 1)), qy = ty * kTileY + (tid / kTileX);
  const bool valid = qx < w && qy < h;
  const int n = h * w;
  const size_t qi = static_cast<size_t>(b) * n + (valid ? qy * w + qx : 0);
  const unsigned char* img = reinterpret_cast<const unsigned char*>(D11) + static_cast<size_t>(b) * n * kPxBytes;

  if (tid < kStMaxLevels) {
    s_red[tid][0] = INT_MAX; s_red[tid][1] = INT_MIN; s_red[tid][2] = INT_MAX; s_red[tid][3] = INT_MIN;
    s_red[tid][4] = 0; s_red[tid][5] = 0; s_red[tid][6] = 0; s_red[tid][7] = 0;
    s_red2[tid][0] = INT_MAX; s_red2[tid][1] = INT_MIN; s_red2[tid][2] = INT_MAX; s_red2[tid][3] = INT_MIN;
  }
  RPROF(long long* prof = g_refine_prof + (blockIdx.y * gridDim.x + blockIdx.x) * 40; long long pc0 = 0; int npass = 0;
        if (tid == 0) { unsigned sm; asm volatile("mov.u32 %0, %smid;" : "=r"(sm)); prof[0] = gtime(); prof[2] = sm; })
  const uint32_t bar = smem_u32(&s_bar);
  if (tid == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  const uint64_t pol = l2_policy_evict_last();  // D11 is re-read by the neighbouring tiles and levels

  const Desc q = load_desc_global(reinterpret_cast<const unsigned char*>(D21) + qi * kPxBytes);
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + qi);
  int u0 = static_cast<int>(p.x), v0 = static_cast<int>(p.y);
  int ub = u0, vb = v0;
  float best = 0.0f;  // numeric_limits<c10::Half>::min() == Half() == 0
  uint32_t phase = 0;
  // geometry of the box that is resident in shared memory when the previous level staged its whole window in
  // one pass (block-uniform): a finer level whose window lies inside it is scored without a refill
  bool res = false;
  int res_px0 = 0, res_px1 = -1, res_py0 = 0, res_py1 = -1, res_pitch = 0;
  __syncthreads();

  for (int d = dilation_max, lvl = 0; d > 0; d--, lvl++) {
    const int rd = kStRadius * d;
    // ---- bounding box of the tile's centres -----------------------------------------------------
    const bool sane = valid && u0 > -kSane && u0 < kSane && v0 > -kSane && v0 < kSane;
    {
      const int mnu = __reduce_min_sync(0xffffffffu, sane ? u0 : INT_MAX);
      const int mxu = __reduce_max_sync(0xffffffffu, sane ? u0 : INT_MIN);
      const int mnv = __reduce_min_sync(0xffffffffu, sane ? v0 : INT_MAX);
      const int mxv = __reduce_max_sync(0xffffffffu, sane ? v0 : INT_MIN);
      const int su = __reduce_add_sync(0xffffffffu, sane ? u0 : 0);
      const int sv = __reduce_add_sync(0xffffffffu, sane ? v0 : 0);
      const int sc = __popc(__ballot_sync(0xffffffffu, sane));
      if ((tid & 31) == 0 && sc > 0) {
        atomicMin(&s_red[lvl][0], mnu); atomicMax(&s_red[lvl][1], mxu);
        atomicMin(&s_red[lvl][2], mnv); atomicMax(&s_red[lvl][3], mxv);
        atomicAdd(&s_red[lvl][4], su); atomicAdd(&s_red[lvl][5], sv); atomicAdd(&s_red[lvl][6], sc);
      }
    }
    __syncthreads();
    int umin = s_red[lvl][0], umax = s_red[lvl][1], vmin = s_red[lvl][2], vmax = s_red[lvl][3];
    const int count = s_red[lvl][6];
    bool inl = sane;
    bool have_box = count > 0;
    int py0 = 0, py1 = -1;
    if (have_box) {
      py0 = max(vmin - rd, 0);
      py1 = min(vmax + rd, h - 1);
      if (py1 >= py0 && (umax - umin + 1) * (py1 - py0 + 1) > kCapPx) {
        // robust pass (block-uniform): keep the centres near the mean, the rest gathers from global
        const int mu = s_red[lvl][4] / count, mv = s_red[lvl][5] / count;
        inl = sane && abs(u0 - mu) <= kInlierRadX && abs(v0 - mv) <= kInlierRadY;
        const int mnu = __reduce_min_sync(0xffffffffu, inl ? u0 : INT_MAX);
        const int mxu = __reduce_max_sync(0xffffffffu, inl ? u0 : INT_MIN);
        const int mnv = __reduce_min_sync(0xffffffffu, inl ? v0 : INT_MAX);
        const int mxv = __reduce_max_sync(0xffffffffu, inl ? v0 : INT_MIN);
        if ((tid & 31) == 0 && mnu != INT_MAX) {
          atomicMin(&s_red2[lvl][0], mnu); atomicMax(&s_red2[lvl][1], mxu);
          atomicMin(&s_red2[lvl][2], mnv); atomicMax(&s_red2[lvl][3], mxv);
        }
        __syncthreads();
        umin = s_red2[lvl][0]; umax = s_red2[lvl][1]; vmin = s_red2[lvl][2]; vmax = s_red2[lvl][3];
        have_box = umin != INT_MAX;
        py0 = max(vmin - rd, 0);
        py1 = min(vmax + rd, h - 1);
      }
    }
    const int bh = py1 - py0 + 1;
    if (have_box && bh > 0 && (umax - umin + 1) * bh > kCapPx) {  // (large dilations: even the robust box is too big)
      have_box = false;
      inl = false;
    }
    // ---- the previous level's box still covers this level's window: no refill ---------------------
    const bool reuse = reuse_ok && res && have_box && bh > 0 && max(umin - rd, 0) >= res_px0 && min(umax + rd, w - 1) <= res_px1 &&
                       py0 >= res_py0 && py1 <= res_py1;
    if (reuse) {
      RPROF(pc0 = clock64();)
      const bool interior = u0 - rd >= 0 && u0 + rd < w && v0 - rd >= 0 && v0 + rd < h;
      const bool fast = __all_sync(0xffffffffu, !inl || interior);
      if (inl) {
        const int base = ((v0 - rd - res_py0) * res_pitch + (u0 - rd - res_px0)) * kPxBytes;
        const int rowstep = d * res_pitch * kPxBytes, colstep = d * kPxBytes;
        int bc = -1;
        if (fast)
          staged_columns<false>(q, sbox, base, rowstep, colstep, 0, kStCnt, u0, v0, rd, d, w, h, best, bc);
        else
          staged_columns<true>(q, sbox, base, rowstep, colstep, 0, kStCnt, u0, v0, rd, d, w, h, best, bc);
        if (bc >= 0) {
          ub = u0 - rd + (bc >> 3) * d;
          vb = v0 - rd + (bc & 7) * d;
        }
      }
      RPROF(__syncthreads(); if (tid == 0) prof[5 + 4 * lvl] += clock64() - pc0;)
    } else
    // ---- passes over groups of u-offset columns -------------------------------------------------
    if (have_box && bh > 0) {
      res = false;
      // Row pitch = width rounded up to 8 px (384 B): a 16-byte chunk of pixel (x, y) then lands in bank
      // group (3x + t) mod 8 whatever y is, so the 8 lanes of an LDS.128 wavefront stay conflict-free
      // when neighbouring queries' centres sit on different rows.
      const int spanu = umax - umin + 1;
      const int cmax = ((kCapPx / bh) / 8 * 8 - spanu) / d + 1;
      for (int i0 = 0; i0 < kStCnt;) {
        const int c = max(1, min(kStCnt - i0, cmax));
        const int px0 = max(umin - rd + i0 * d, 0);
        const int px1 = min(umax - rd + (i0 + c - 1) * d, w - 1);
        const int pw = px1 - px0 + 1;
        if (pw > 0) {  // block-uniform
          int pitch = (pw + 7) & ~7;
          if (pitch * bh > kCapPx) pitch = pw;  // (only when cmax < 1: a single column that just fits unpadded)
          const uint32_t rowbytes = static_cast<uint32_t>(pw) * kPxBytes;
          __syncthreads();  // every thread is done reading the previous box
          if (tid == 0) mbar_expect_tx(bar, rowbytes * static_cast<uint32_t>(bh));
          for (int r = tid; r < bh; r += kStThreads)
            bulk_g2s(smem_u32(sbox) + r * pitch * kPxBytes,
                     img + (static_cast<size_t>(py0 + r) * w + px0) * kPxBytes, rowbytes, bar, pol);
          RPROF(pc0 = clock64(); npass++;)
          mbar_wait(bar, phase);
          phase ^= 1u;
          RPROF(if (tid == 0) prof[4 + 4 * lvl] += clock64() - pc0; pc0 = clock64();)
          if (i0 == 0 && c == kStCnt) {  // the whole window of this level is resident now
            res = true;
            res_px0 = px0; res_px1 = px1; res_py0 = py0; res_py1 = py1; res_pitch = pitch;
          }
          const bool interior = u0 - rd >= 0 && u0 + rd < w && v0 - rd >= 0 && v0 + rd < h;
          const bool fast = __all_sync(0xffffffffu, !inl || interior);
          if (inl) {
            const int base = ((v0 - rd - py0) * pitch + (u0 - rd - px0)) * kPxBytes;
            const int rowstep = d * pitch * kPxBytes, colstep = d * kPxBytes;
            int bc = -1;
            if (fast)
              staged_columns<false>(q, sbox, base, rowstep, colstep, i0, c, u0, v0, rd, d, w, h, best, bc);
            else
              staged_columns<true>(q, sbox, base, rowstep, colstep, i0, c, u0, v0, rd, d, w, h, best, bc);
            if (bc >= 0) {
              ub = u0 - rd + (bc >> 3) * d;
              vb = v0 - rd + (bc & 7) * d;
            }
          }
          RPROF(__syncthreads(); if (tid == 0) prof[5 + 4 * lvl] += clock64() - pc0;)
        }
        i0 += c;
      }
    }
    RPROF(pc0 = clock64(); { const int no = __syncthreads_count(valid && !inl); if (tid == 0) prof[7 + 4 * lvl] = no; })
    if (valid && !inl) gather_level(img, q, u0, v0, d, kStRadius, w, h, best, ub, vb);
    RPROF(__syncthreads(); if (tid == 0) prof[6 + 4 * lvl] += clock64() - pc0;)
    u0 = ub;
    v0 = vb;
  }
  const bool same = ub == static_cast<int>(p.x) && vb == static_cast<int>(p.y);
  if (valid) reinterpret_cast<longlong2*>(p1_new)[qi] = same ? p : make_longlong2(ub, vb);
  RPROF(if (tid == 0) { prof[1] = gtime(); prof[3] = npass; })
}

// ---- generic path: any fdim, fp16 / fp32 / fp64, one candidate at a time ----------------------
template <typename T>
struct Acc;
template <>
struct Acc<__half> {
  using S = __half;
  static __device__ __forceinline__ S zero() { return __float2half_rn(0.0f); }
  static __device__ __forceinline__ S mac(S s, __half a, __half b) { return __hadd_rn(s, __hmul_rn(a, b)); }
  static __device__ __forceinline__ float cmp(S s) { return __half2float(s); }
  static __device__ __forceinline__ float init() { return 0.0f; }
};
template <>
struct Acc<float> {
  using S = float;
  static __device__ __forceinline__ S zero() { return 0.0f; }
  static __device__ __forceinline__ S mac(S s, float a, float b) { return s + a * b; }  // FFMA like the ref
  static __device__ __forceinline__ float cmp(S s) { return s; }
  static __device__ __forceinline__ float init() { return FLT_MIN; }
};
template <>
struct Acc<double> {
  using S = double;
  static __device__ __forceinline__ S zero() { return 0.0; }
  static __device__ __forceinline__ S mac(S s, double a, double b) { return s + a * b; }
  static __device__ __forceinline__ double cmp(S s) { return s; }
  static __device__ __forceinline__ double init() { return DBL_MIN; }
};

template <typename T>
__global__ void __launch_bounds__(kRefineThreads)
refine_generic_kernel(const T* __restrict__ D11, const T* __restrict__ D21,
                      const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new, int h, int w,
                      int n, int fdim, int radius, int dilation_max) {
  const int k = blockIdx.x * kRefineThreads + threadIdx.x;
  const int b = blockIdx.y;
  if (k >= n) return;
  const size_t qi = static_cast<size_t>(b) * n + k;
  const T* img = D11 + static_cast<size_t>(b) * h * w * fdim;
  const T* q = D21 + qi * fdim;
  int u0 = static_cast<int>(p1[qi * 2 + 0]), v0 = static_cast<int>(p1[qi * 2 + 1]);
  int ub = u0, vb = v0;
  auto best = Acc<T>::init();
  for (int d = dilation_max; d > 0; d--) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      for (int j = 0; j < diam; j += d) {
        const int u = u0 - rd + i, v = v0 - rd + j;
        if (inside(u, v, w, h)) {
          const T* c = img + (static_cast<size_t>(v) * w + u) * fdim;
          typename Acc<T>::S s = Acc<T>::zero();
          for (int t = 0; t < fdim; t++) s = Acc<T>::mac(s, q[t], c[t]);
          if (Acc<T>::cmp(s) > best) {
            best = Acc<T>::cmp(s);
            ub = u;
            vb = v;
          }
        }
      }
    }
    u0 = ub;
    v0 = vb;
  }
  p1_new[qi * 2 + 0] = ub;
  p1_new[qi * 2 + 1] = vb;
}

template <typename T>
int launch_generic(const T* D11, const T* D21, const int64_t* p1, int64_t* p1_new, int b, int h, int w,
                   int n, int fdim, int radius, int dilation_max, void* stream) {
  dim3 grid((n + kRefineThreads - 1) / kRefineThreads, b);
  refine_generic_kernel<T><<<grid, kRefineThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      D11, D21, p1, p1_new, h, w, n, fdim, radius, dilation_max);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

int check_args(const void* D11, const void* D21, const void* p1, const void* p1_new, int b, int h, int w,
               int n, int fdim, int radius, int dilation_max) {
  if (!D11 || !D21 || !p1 || !p1_new) return M3R_ERR_BAD_ARG;
  if (b < 0 || n < 0 || h < 1 || w < 1 || fdim < 1 || radius < 0 || dilation_max < 0) return M3R_ERR_BAD_ARG;
  if (b > 65535) return M3R_ERR_BAD_ARG;
  return M3R_OK;
}

}  // namespace
}  // namespace m3r

#ifdef M3R_REFINE_PROF  // debugging export of the -DM3R_REFINE_PROF build only (tools/refine_prof.py)
extern "C" int m3r_debug_refine_prof(long long* host_out, int clear) {
  if (clear) {
    void* p = nullptr;
    M3R_CUDA_CHECK(cudaGetSymbolAddress(&p, m3r::g_refine_prof));
    M3R_CUDA_CHECK(cudaMemset(p, 0, sizeof(m3r::g_refine_prof)));
    return M3R_OK;
  }
  M3R_CUDA_CHECK(cudaMemcpyFromSymbol(host_out, m3r::g_refine_prof, sizeof(m3r::g_refine_prof)));
  return M3R_OK;
}
#endif

extern "C" int m3r_refine_matches_f16(const void* D11, const void* D21, const int64_t* p1,
                                      int64_t* p1_new, int b, int h, int w, int n, int fdim,
                                      int radius, int dilation_max, void* stream) {
  int rc = m3r::check_args(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max);
  if (rc != M3R_OK) return rc;
  if (b == 0 || n == 0) return M3R_OK;
  const bool aligned = ((reinterpret_cast<uintptr_t>(D11) | reinterpret_cast<uintptr_t>(D21) |
                         reinterpret_cast<uintptr_t>(p1) | reinterpret_cast<uintptr_t>(p1_new)) & 15) == 0;
  if (fdim == m3r::kFD && aligned) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const __half* d11 = static_cast<const __half*>(D11);
    const __half* d21 = static_cast<const __half*>(D21);
    if (radius == m3r::kStRadius && dilation_max >= 1 && dilation_max <= m3r::kStMaxLevels &&
        static_cast<long long>(n) == static_cast<long long>(h) * w) {
      static m3r::DynSmemOptIn opt_in;
      M3R_CUDA_CHECK(opt_in.ensure(m3r::refine_staged_kernel, m3r::kStSmemBytes));
      const int tiles_x = (w + m3r::kTileX - 1) / m3r::kTileX, tiles_y = (h + m3r::kTileY - 1) / m3r::kTileY;
      dim3 grid(tiles_x * tiles_y, b);
      static const int reuse_ok = [] {  // debugging knob: M3R_REFINE_REUSE=0 refills the box at every level
        const char* e = getenv("M3R_REFINE_REUSE");
        return (e && e[0] == '0') ? 0 : 1;
      }();
      m3r::refine_staged_kernel<<<grid, m3r::kStThreads, m3r::kStSmemBytes, st>>>(d11, d21, p1, p1_new, h, w,
                                                                                 tiles_x, dilation_max, reuse_ok);
      M3R_LAUNCH_CHECK();
      return M3R_OK;
    }
    dim3 grid((n + m3r::kRefineThreads - 1) / m3r::kRefineThreads, b);
    m3r::refine_gather_kernel<<<grid, m3r::kRefineThreads, 0, st>>>(d11, d21, p1, p1_new, h, w, n, radius,
                                                                    dilation_max);
    M3R_LAUNCH_CHECK();
    return M3R_OK;
  }
  return m3r::launch_generic<__half>(static_cast<const __half*>(D11), static_cast<const __half*>(D21), p1,
                                     p1_new, b, h, w, n, fdim, radius, dilation_max, stream);
}

extern "C" int m3r_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1,
                                      int64_t* p1_new, int b, int h, int w, int n, int fdim,
                                      int radius, int dilation_max, void* stream) {
  int rc = m3r::check_args(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max);
  if (rc != M3R_OK) return rc;
  if (b == 0 || n == 0) return M3R_OK;
  return m3r::launch_generic<float>(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max, stream);
}

extern "C" int m3r_refine_matches_f64(const double* D11, const double* D21, const int64_t* p1,
                                      int64_t* p1_new, int b, int h, int w, int n, int fdim,
                                      int radius, int dilation_max, void* stream) {
  int rc = m3r::check_args(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max);
  if (rc != M3R_OK) return rc;
  if (b == 0 || n == 0) return M3R_OK;
  return m3r::launch_generic<double>(D11, D21, p1, p1_new, b, h, w, n, fdim, radius, dilation_max, stream);
}
