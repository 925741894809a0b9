// Microbenchmark: issue rate of scalar FFMA vs packed FFMA2 (fma.rn.f32x2) vs HFMA2/HADD2+HMUL2 on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_ffma2.bin tools/ubench_ffma2.cu
#include <cuda_fp16.h>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;
constexpr int ILP = 8;

__global__ void k_ffma(float* out, float a, float b) {
  float acc[ILP];
  for (int i = 0; i < ILP; i++) acc[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = fmaf(acc[i], a, b);
  }
  float s = 0;
  for (int i = 0; i < ILP; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma2(float* out, float a, float b) {
  float2 acc[ILP];
  for (int i = 0; i < ILP; i++) acc[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
  const float2 a2 = make_float2(a, a + 1), b2 = make_float2(b, b + 1);
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = __ffma2_rn(acc[i], a2, b2);
  }
  float s = 0;
  for (int i = 0; i < ILP; i++) s += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// mixed: one FFMA2 + one integer ALU op per slot pair — does the ALU op issue in the FFMA2's shadow?
__global__ void k_ffma2_alu(float* out, float a, float b) {
  float2 acc[ILP];
  unsigned x[ILP];
  for (int i = 0; i < ILP; i++) { acc[i] = make_float2(threadIdx.x + i, threadIdx.x - i); x[i] = threadIdx.x * 7 + i; }
  const float2 a2 = make_float2(a, a + 1), b2 = make_float2(b, b + 1);
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) { acc[i] = __ffma2_rn(acc[i], a2, b2); x[i] = (x[i] ^ (x[i] >> 3)) + 0x9e3779b9u; }
  }
  float s = 0;
  for (int i = 0; i < ILP; i++) s += acc[i].x + acc[i].y + (float)x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_hmul_hadd(float* out, float a, float b) {
  __half2 acc[ILP];
  for (int i = 0; i < ILP; i++) acc[i] = __floats2half2_rn(threadIdx.x + i, 1.0f);
  const __half2 a2 = __floats2half2_rn(a, a), b2 = __floats2half2_rn(b, b);
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = __hadd2_rn(__hmul2_rn(acc[i], a2), b2);
  }
  float s = 0;
  for (int i = 0; i < ILP; i++) s += __low2float(acc[i]) + __high2float(acc[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// the refine_matches inner mix: 1 PRMT (alu pipe) + 1 HMUL2 + 1 HADD2 (fma pipe) per step
__global__ void k_prmt_hmul_hadd(float* out, float a, float b) {
  __half2 acc[ILP];
  unsigned x[ILP];
  for (int i = 0; i < ILP; i++) { acc[i] = __floats2half2_rn(threadIdx.x + i, 1.0f); x[i] = threadIdx.x * 0x01010101u + i; }
  const __half2 a2 = __floats2half2_rn(a, a);
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) {
      x[i] = __byte_perm(x[i], x[(i + 1) % ILP], 0x5410 + (it & 1));
      acc[i] = __hadd2_rn(acc[i], __hmul2_rn(*reinterpret_cast<__half2*>(&x[i]), a2));
    }
  }
  float s = 0;
  for (int i = 0; i < ILP; i++) s += __low2float(acc[i]) + __high2float(acc[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// HADD2 only (dependent chain per acc, ILP independent chains)
__global__ void k_hadd(float* out, float a, float b) {
  __half2 acc[ILP];
  for (int i = 0; i < ILP; i++) acc[i] = __floats2half2_rn(threadIdx.x + i, 1.0f);
  const __half2 b2 = __floats2half2_rn(b, b);
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) acc[i] = __hadd2_rn(acc[i], b2);
  }
  float s = 0;
  for (int i = 0; i < ILP; i++) s += __low2float(acc[i]) + __high2float(acc[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
void run(const char* name, F launch, double inst_per_thread) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  launch();
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  for (int r = 0; r < 5; r++) launch();
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  ms /= 5;
  int dev, clk, sms;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const double warps = 148.0 * 8 * 1024 / 32;
  const double winst = warps * inst_per_thread;
  printf("%-14s %8.3f ms  %.3f warp-inst/clk/SMSP (at %d kHz nominal)  %.2f Tinst-lane/s\n", name, ms,
         winst / (ms * 1e-3) / (sms * 4.0) / (clk * 1e3), clk, winst * 32 / (ms * 1e-3) / 1e12);
}

int main() {
  float* out;
  cudaMalloc(&out, 148 * 8 * 1024 * sizeof(float));
  dim3 g(148 * 8), b(1024);
  run("FFMA", [&] { k_ffma<<<g, b>>>(out, 1.0001f, 0.5f); }, (double)ITERS * ILP);
  run("FFMA2", [&] { k_ffma2<<<g, b>>>(out, 1.0001f, 0.5f); }, (double)ITERS * ILP);
  run("FFMA2+2ALU", [&] { k_ffma2_alu<<<g, b>>>(out, 1.0001f, 0.5f); }, (double)ITERS * ILP * 3);
  run("HMUL2+HADD2", [&] { k_hmul_hadd<<<g, b>>>(out, 1.0001f, 0.5f); }, (double)ITERS * ILP * 2);
  run("HADD2", [&] { k_hadd<<<g, b>>>(out, 1.0001f, 0.5f); }, (double)ITERS * ILP);
  run("PRMT+HMUL2+HADD2", [&] { k_prmt_hmul_hadd<<<g, b>>>(out, 1.0001f, 0.5f); }, (double)ITERS * ILP * 3);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
