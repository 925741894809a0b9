// Microbenchmarks behind the design of gn_solve (fp64 Cholesky on sm_100a): what the fp64 pipes, the
// warp-level data exchange and the inter-CTA hand-offs actually cost on a B200.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/ubench_solve.bin tools/ubench_solve.cu
#include <cooperative_groups.h>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// ---------------------------------------------------------------- throughput
constexpr int T_ITERS = 2048;
__global__ void k_dfma(double* out, double a, double b) {
  double acc[8];
  for (int i = 0; i < 8; i++) acc[i] = threadIdx.x + i;
  for (int it = 0; it < T_ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
  for (int i = 0; i < 8; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmma(double* out, double a, double b) {
  double c[8][2];
  for (int i = 0; i < 8; i++) { c[i][0] = threadIdx.x + i; c[i][1] = i; }
  for (int it = 0; it < T_ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) dmma(c[i][0], c[i][1], a, b);
  }
  double s = 0;
  for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---------------------------------------------------------------- dependent-chain latencies (one warp)
constexpr int L_ITERS = 512;
__global__ void k_lat(long long* cyc, double* sink, double x0) {
  __shared__ double sm[64];
  const int lane = threadIdx.x;
  double x = x0 + lane * 1e-3;
  long long t0, t1;
  int slot = 0;
#define LAT(body)                                    \
  __syncwarp(); t0 = clock64();                      \
  for (int it = 0; it < L_ITERS; it++) { body; }     \
  t1 = clock64(); if (lane == 0) cyc[slot] = t1 - t0; slot++;
  LAT(x = fma(x, 1.0000001, 1e-9));                                   // 0 dfma
  LAT(x = rsqrt(x) + 1.5);                                            // 1 rsqrt + dadd
  LAT(x = 1.0 / x + 1.5);                                             // 2 div + dadd
  LAT(x = __drcp_rn(x) + 1.5);                                        // 3 drcp + dadd
  LAT(x = sqrt(x) + 1.5);                                             // 4 sqrt + dadd
  LAT(x = __shfl_sync(0xffffffffu, x, (lane + 1) & 31) + 1e-9);       // 5 shfl64 + dadd
  LAT(sm[lane] = x; __syncwarp(); x = sm[(lane + 1) & 31] + 1e-9; __syncwarp());  // 6 sts/lds + dadd
  { double c0 = x, c1 = x;
    LAT(dmma(c0, c1, 1.0000001, 1e-3));                               // 7 dmma dependent
    x = c0 + c1; }
  { float f = (float)x;
    LAT(f = rsqrtf(f) + 1.5f);                                        // 8 float rsqrt + fadd
    x += f; }
  LAT(x = (double)rsqrtf((float)x) + 1.5);                            // 9 d->f rsqrtf f->d + dadd
  LAT(x = x + 1.5);                                                   // 10 dadd
  sink[lane] = x;
}

// ---------------------------------------------------------------- 32x32 potrf in one warp: variants
// A: lane r owns row r in registers, pivot column broadcast by shuffles (the round-1 kernel)
__device__ __forceinline__ void potrf_shfl(double (&a)[32], int lane) {
#pragma unroll
  for (int j = 0; j < 32; j++) {
    double d = __shfl_sync(0xffffffffu, a[j], j);
    const double rinv = rsqrt(d);
    const double l = a[j] * rinv;
    a[j] = l;
#pragma unroll
    for (int c = j + 1; c < 32; c++) a[c] -= l * __shfl_sync(0xffffffffu, l, c);
  }
}
// B: same ownership, but the scaled pivot column goes through shared memory (1 STS + 16 broadcast LDS.128 per pivot)
__device__ __forceinline__ void potrf_smem(double (&a)[32], int lane, double* col /*[2][32]*/) {
#pragma unroll
  for (int j = 0; j < 32; j++) {
    double* cb = col + (j & 1) * 32;
    cb[lane] = a[j];  // unscaled column j (rows >= j valid)
    __syncwarp();
    const double d = cb[j];
    const double rinv = rsqrt(d);
    const double l = a[j] * rinv;
    a[j] = l;
    const double lr = l * rinv;  // a[r][j] / d
#pragma unroll
    for (int c = j + 1; c < 32; c++) a[c] -= lr * cb[c];
  }
}
// C: like B but with a 16-byte aligned sweep (start at the even column <= j+1, masked), reciprocal instead of rsqrt
// on the chain (LDL^T order: the sqrt scaling is applied off the chain at the end)
__device__ __forceinline__ void potrf_ldl(double (&a)[32], int lane, double* col /*[2][32]*/, double* dg) {
#pragma unroll
  for (int j = 0; j < 32; j++) {
    double* cb = col + (j & 1) * 32;
    cb[lane] = a[j];
    __syncwarp();
    const double d = cb[j];
    const double dinv = 1.0 / d;
    const double t = a[j] * dinv;
    if (lane == j) dg[j] = d;
    constexpr int dummy = 0; (void)dummy;
#pragma unroll
    for (int c = (j + 1) & ~1; c < 32; c += 2) {
      const double2 v = *reinterpret_cast<const double2*>(cb + c);
      if (c > j) a[c] -= t * v.x;
      a[c + 1] -= t * v.y;
    }
  }
}
template <int VAR>
__global__ void k_potrf(const double* A, double* out, long long* cyc, int reps) {
  __shared__ __align__(16) double col[64];
  __shared__ double dg[32];
  const int lane = threadIdx.x;
  double a[32];
  long long tot = 0;
  for (int r = 0; r < reps; r++) {
#pragma unroll
    for (int c = 0; c < 32; c++) a[c] = A[lane * 32 + c];
    __syncwarp();
    const long long t0 = clock64();
    if (VAR == 0) potrf_shfl(a, lane);
    if (VAR == 1) potrf_smem(a, lane, col);
    if (VAR == 2) potrf_ldl(a, lane, col, dg);
    __syncwarp();
    tot += clock64() - t0;
  }
#pragma unroll
  for (int c = 0; c < 32; c++) out[lane * 32 + c] = a[c];
  if (lane == 0) cyc[0] = tot / reps;
}

// ---------------------------------------------------------------- grid barrier / flag hand-off
__global__ void k_gridsync(long long* cyc, int n) {
  cg::grid_group g = cg::this_grid();
  g.sync();
  const long long t0 = clock64();
  for (int i = 0; i < n; i++) g.sync();
  if (blockIdx.x == 0 && threadIdx.x == 0) cyc[0] = (clock64() - t0) / n;
}
// CTA 0 and CTA `peer` bounce a counter through global memory (volatile / relaxed, like the LL words of the solver)
__global__ void k_pingpong(volatile unsigned long long* w, long long* cyc, int peer, int n) {
  if (threadIdx.x != 0) return;
  if (blockIdx.x == 0) {
    const long long t0 = clock64();
    for (int i = 1; i <= n; i++) {
      w[0] = (unsigned long long)i;
      while (w[1] != (unsigned long long)i) {}
    }
    cyc[0] = (clock64() - t0) / n;
  } else if ((int)blockIdx.x == peer) {
    for (int i = 1; i <= n; i++) {
      while (w[0] != (unsigned long long)i) {}
      w[1] = (unsigned long long)i;
    }
  }
}

int main() {
  int dev = 0;
  CK(cudaSetDevice(dev));
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, dev));
  int clk_khz = 0;
  CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev));
  printf("%s, %d SMs, clock %d MHz\n", p.name, p.multiProcessorCount, clk_khz / 1000);
  double* out;
  long long* cyc;
  CK(cudaMalloc(&out, 1 << 24));
  CK(cudaMallocManaged(&cyc, 64 * sizeof(long long)));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  // throughput at several warps/SM
  for (int wps : {4, 8, 16, 32}) {
    const int threads = 32 * (wps > 8 ? 8 : wps), ctas = p.multiProcessorCount * (wps > 8 ? wps / 8 : 1);
    for (int which = 0; which < 2; which++) {
      float best = 1e9f;
      for (int r = 0; r < 5; r++) {
        CK(cudaEventRecord(e0));
        if (which == 0) k_dfma<<<ctas, threads>>>(out, 1.0000001, 1e-9);
        else k_dmma<<<ctas, threads>>>(out, 1.0000001, 1e-9);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (r > 0 && ms < best) best = ms;
      }
      const double fma_per_thread = (double)T_ITERS * 8 * (which == 0 ? 1.0 : 8.0);  // m8n8k4 = 256 FMA / 32 lanes
      const double tf = 2.0 * fma_per_thread * threads * ctas / (best * 1e-3) / 1e12;
      printf("%s  %2d warps/SM: %.3f ms  %.2f TFLOP/s\n", which == 0 ? "DFMA" : "DMMA", wps, best, tf);
    }
  }
  // latencies
  k_lat<<<1, 32>>>(cyc, out, 1.7);
  CK(cudaDeviceSynchronize());
  const char* names[] = {"dfma", "rsqrt(double)+dadd", "1.0/x+dadd", "__drcp_rn+dadd", "sqrt(double)+dadd", "shfl64+dadd",
                         "sts+lds(+2 syncwarp)+dadd", "dmma m8n8k4 dependent", "rsqrtf+fadd", "d2f,rsqrtf,f2d+dadd", "dadd"};
  for (int i = 0; i < 11; i++) printf("latency %-28s %.1f cycles\n", names[i], (double)cyc[i] / L_ITERS);
  // potrf variants on an SPD tile
  {
    double hA[1024];
    srand(1);
    double B[1024];
    for (int i = 0; i < 1024; i++) B[i] = (rand() / (double)RAND_MAX) - 0.5;
    for (int r = 0; r < 32; r++)
      for (int c = 0; c < 32; c++) {
        double s = (r == c) ? 8.0 : 0.0;
        for (int k = 0; k < 32; k++) s += B[r * 32 + k] * B[c * 32 + k];
        hA[r * 32 + c] = s;
      }
    double* dA;
    double* dO;
    CK(cudaMalloc(&dA, sizeof(hA)));
    CK(cudaMalloc(&dO, sizeof(hA)));
    CK(cudaMemcpy(dA, hA, sizeof(hA), cudaMemcpyHostToDevice));
    double ref[1024], got[1024];
    for (int v = 0; v < 3; v++) {
      if (v == 0) k_potrf<0><<<1, 32>>>(dA, dO, cyc, 20);
      if (v == 1) k_potrf<1><<<1, 32>>>(dA, dO, cyc, 20);
      if (v == 2) k_potrf<2><<<1, 32>>>(dA, dO, cyc, 20);
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(got, dO, sizeof(got), cudaMemcpyDeviceToHost));
      if (v == 0) for (int i = 0; i < 1024; i++) ref[i] = got[i];
      double err = 0;
      if (v == 1)
        for (int r = 0; r < 32; r++) for (int c = 0; c <= r; c++) err = fmax(err, fabs(got[r * 32 + c] - ref[r * 32 + c]));
      printf("potrf32 variant %d (%s): %lld cycles  (max |L - L_A| %.2e)\n", v,
             v == 0 ? "shuffle broadcast" : v == 1 ? "smem column, rsqrt" : "smem column, LDL reciprocal", cyc[0], err);
    }
  }
  // grid barrier
  for (int per_sm : {1, 2}) {
    int n = 200, grid = p.multiProcessorCount * per_sm;
    void* args[] = {(void*)&cyc, (void*)&n};
    CK(cudaLaunchCooperativeKernel((const void*)k_gridsync, dim3(grid), dim3(256), args, 0, 0));
    CK(cudaDeviceSynchronize());
    printf("grid.sync() with %d CTAs x 256 thr: %lld cycles\n", grid, cyc[0]);
  }
  {
    unsigned long long* w;
    CK(cudaMalloc(&w, 256));
    for (int peer : {1, 2, 74, 147}) {
      CK(cudaMemset(w, 0, 256));
      int n = 2000;
      void* args[] = {(void*)&w, (void*)&cyc, (void*)&peer, (void*)&n};
      CK(cudaLaunchCooperativeKernel((const void*)k_pingpong, dim3(p.multiProcessorCount), dim3(32), args, 0, 0));
      CK(cudaDeviceSynchronize());
      printf("flag ping-pong CTA0 <-> CTA%d: %lld cycles per round trip\n", peer, cyc[0]);
    }
  }
  return 0;
}
