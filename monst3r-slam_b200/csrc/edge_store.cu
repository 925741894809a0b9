// Factor-graph staging kernels (sm_100a): SURVEY.md §8(f) ranks 2 and 4.
//
//  edge_admit_kernel — the arithmetic of FactorGraph.add_factors after the symmetric matcher
//    (mast3r_slam/global_opt.py:54-99): for every new keyframe pair and both directions
//        Q = sqrt(Q_src[b, idx[b,k]] * Q_dst[b,k])                         (:54-58)
//        count[b, dir] = #{k : valid[b,k] && Q > Q_conf}                   (:60-67, numerator of match_frac)
//    and writes {idx, valid, Q} straight into the edge store's slabs at directed slots
//    2*(slot0+b)+dir.  The reference materialises batch_inds, Qj, valid_Qj, valid_j and then
//    torch.cat's every slab of the WHOLE graph (:89-96, again in prep_two_way_edges :104-110); here a new
//    pair costs one pass over its own 30 B/pt and nothing else moves.
//    Streaming + one coherent gather: HBM-bound (30 B/pt algorithmic: idx 8 + valid 1 + Q_dst 4 +
//    Q_src 4 in, idx 8 + valid 1 + Q 4 out).
//    fp32 `__fmul_rn` then `__fsqrt_rn`: the same two roundings as torch's mul and sqrt kernels.
//
//  ray_constrain_kernel — constrain_points_to_ray (mast3r_slam/geometry.py:37-42 with backproject
//    :107-114 and get_pixel_coords :117-123):  X <- z * ((u-cx)/fx, (v-cy)/fy, 1), u = k mod w,
//    v = k div w.  Same operation order as the torch expressions (sub, div, mul; no contraction).
#include "common.cuh"

namespace m3r {
namespace {

constexpr int kAdmitThreads = 256;
constexpr int kAdmitPerThread = 4;
constexpr int kAdmitChunk = kAdmitThreads * kAdmitPerThread;  // 1024 points per CTA

struct AdmitArgs {
  const int64_t* idx[2];    // [batch, n]: i->j, j->i
  const uint8_t* valid[2];  // [batch, n]
  const float* q_src[2];    // gathered by idx: Qii (dir 0), Qjj (dir 1)
  const float* q_dst[2];    // Qji (dir 0), Qij (dir 1)
  int64_t* s_idx;           // store slabs [2*capacity, n]
  uint8_t* s_valid;
  float* s_q;
  int32_t* counts;          // [batch, 2]
  long long slot0;          // first undirected slot of this batch
  int n, batch;
  int iters;  // 1024-point sub-chunks per CTA (host: fewer, longer-lived CTAs once the grid fills the SMs)
  float q_conf;
};

__device__ __forceinline__ float admit_q(const float* __restrict__ src, long long m, int n, float qd) {
  // an index outside [0, n) would be an out-of-bounds read in the reference; here it reads entry 0
  const long long mm = (m >= 0 && m < n) ? m : 0;
  return __fsqrt_rn(__fmul_rn(__ldg(src + mm), qd));
}

__global__ void __launch_bounds__(kAdmitThreads) edge_admit_kernel(const AdmitArgs a) {
  const int dir = blockIdx.y & 1, b = blockIdx.y >> 1;
  const int n = a.n;
  const size_t in_off = static_cast<size_t>(b) * n;
  const size_t out_off = static_cast<size_t>(2 * (a.slot0 + b) + dir) * n;
  const int64_t* __restrict__ idx = a.idx[dir] + in_off;
  const uint8_t* __restrict__ valid = a.valid[dir] + in_off;
  const float* __restrict__ qd = a.q_dst[dir] + in_off;
  const float* __restrict__ qs = a.q_src[dir] + in_off;
  int64_t* __restrict__ o_idx = a.s_idx + out_off;
  uint8_t* __restrict__ o_valid = a.s_valid + out_off;
  float* __restrict__ o_q = a.s_q + out_off;

  // A warp owns 128 consecutive points as two halves of 64: lane l takes points {2l, 2l+1} of each half, so every
  // warp-wide load / store instruction covers whole, consecutive 32-byte sectors (16 B idx, 8 B Q, 2 B valid per lane).
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int cnt = 0;
  const bool vec = (n % kAdmitPerThread) == 0;  // rows start 16-byte aligned only then
  for (int sub = 0; sub < a.iters; sub++) {
  const int wbase = (blockIdx.x * a.iters + sub) * kAdmitChunk + warp * 128;
  if (wbase >= n) break;
  if (vec && wbase + 128 <= n) {
    longlong2 m[2];
    uchar2 vm[2];
    float2 qd2[2];
#pragma unroll
    for (int hlf = 0; hlf < 2; hlf++) {
      const int k = wbase + hlf * 64 + 2 * lane;
      m[hlf] = __ldcs(reinterpret_cast<const longlong2*>(idx + k));
      vm[hlf] = __ldcs(reinterpret_cast<const uchar2*>(valid + k));
      qd2[hlf] = __ldcs(reinterpret_cast<const float2*>(qd + k));
    }
#pragma unroll
    for (int hlf = 0; hlf < 2; hlf++) {
      const int k = wbase + hlf * 64 + 2 * lane;
      float2 q;
      q.x = admit_q(qs, m[hlf].x, n, qd2[hlf].x);
      q.y = admit_q(qs, m[hlf].y, n, qd2[hlf].y);
      cnt += (vm[hlf].x != 0 && q.x > a.q_conf) + (vm[hlf].y != 0 && q.y > a.q_conf);
      __stcs(reinterpret_cast<longlong2*>(o_idx + k), m[hlf]);
      __stcs(reinterpret_cast<uchar2*>(o_valid + k), make_uchar2(vm[hlf].x != 0, vm[hlf].y != 0));
      __stcs(reinterpret_cast<float2*>(o_q + k), q);
    }
  } else {
    for (int r = 0; r < kAdmitPerThread; r++) {
      const int k = wbase + r * 32 + lane;
      if (k >= n) break;
      const long long m = idx[k];
      const bool vm = valid[k] != 0;
      const float q = admit_q(qs, m, n, qd[k]);
      cnt += (vm && q > a.q_conf);
      o_idx[k] = m;
      o_valid[k] = vm ? 1 : 0;
      o_q[k] = q;
    }
  }
  }
  __shared__ int s_cnt[kAdmitThreads / 32];
  cnt = __reduce_add_sync(0xffffffffu, cnt);
  if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
#pragma unroll
    for (int i = 0; i < kAdmitThreads / 32; i++) t += s_cnt[i];
    if (t) atomicAdd(a.counts + blockIdx.y, t);  // integer: order-independent
  }
}

__global__ void __launch_bounds__(256) ray_constrain_kernel(const float* __restrict__ Xs, float* __restrict__ out,
                                                            const float* __restrict__ K, int n, int w) {
  const int k = blockIdx.x * 256 + threadIdx.x;
  if (k >= n) return;
  const size_t o = (static_cast<size_t>(blockIdx.y) * n + k) * 3;
  const float fx = __ldg(K + 0), cx = __ldg(K + 2), fy = __ldg(K + 4), cy = __ldg(K + 5);
  const float u = static_cast<float>(k % w), v = static_cast<float>(k / w);
  const float z = Xs[o + 2];
  const float tx = __fdiv_rn(__fsub_rn(u, cx), fx);
  const float ty = __fdiv_rn(__fsub_rn(v, cy), fy);
  out[o + 0] = __fmul_rn(z, tx);
  out[o + 1] = __fmul_rn(z, ty);
  out[o + 2] = z;
}

}  // namespace
}  // namespace m3r

using namespace m3r;

extern "C" int m3r_edge_admit(const int64_t* idx_i2j, const int64_t* idx_j2i, const uint8_t* valid_j,
                              const uint8_t* valid_i, const float* Qii, const float* Qjj, const float* Qji,
                              const float* Qij, int batch, int num_pts, float Q_conf, int64_t* store_idx,
                              uint8_t* store_valid, float* store_Q, int64_t slot0, int32_t* counts_out,
                              void* stream) {
  if (!idx_i2j || !idx_j2i || !valid_j || !valid_i || !Qii || !Qjj || !Qji || !Qij || !store_idx || !store_valid ||
      !store_Q || !counts_out)
    return M3R_ERR_BAD_ARG;
  if (batch < 0 || num_pts < 1 || slot0 < 0) return M3R_ERR_BAD_ARG;
  if (batch == 0) return M3R_OK;
  if (2 * (long long)batch > 65535) return M3R_ERR_BAD_ARG;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  M3R_CUDA_CHECK(cudaMemsetAsync(counts_out, 0, sizeof(int32_t) * 2 * (size_t)batch, s));
  AdmitArgs a;
  a.idx[0] = idx_i2j; a.idx[1] = idx_j2i;
  a.valid[0] = valid_j; a.valid[1] = valid_i;
  a.q_src[0] = Qii; a.q_src[1] = Qjj;
  a.q_dst[0] = Qji; a.q_dst[1] = Qij;
  a.s_idx = store_idx; a.s_valid = store_valid; a.s_q = store_Q;
  a.counts = counts_out;
  a.slot0 = slot0; a.n = num_pts; a.batch = batch; a.q_conf = Q_conf;
  const int chunks = (num_pts + kAdmitChunk - 1) / kAdmitChunk;
  a.iters = 1;
  while (a.iters < 4 && (long long)((chunks + 2 * a.iters - 1) / (2 * a.iters)) * 2 * batch >= 8LL * kNumSMs) a.iters *= 2;
  dim3 grid((chunks + a.iters - 1) / a.iters, 2 * batch);
  edge_admit_kernel<<<grid, kAdmitThreads, 0, s>>>(a);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

extern "C" int m3r_constrain_points_to_ray(const float* Xs, float* out, const float* K, int num_kf, int height,
                                           int width, void* stream) {
  if (!Xs || !out || !K) return M3R_ERR_BAD_ARG;
  if (num_kf < 0 || height < 1 || width < 1 || num_kf > 65535) return M3R_ERR_BAD_ARG;
  if (num_kf == 0) return M3R_OK;
  const int n = height * width;
  dim3 grid((n + 255) / 256, num_kf);
  ray_constrain_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(Xs, out, K, n, width);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}
