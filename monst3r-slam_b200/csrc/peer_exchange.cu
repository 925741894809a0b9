// All-gather of the Gauss-Newton edge blocks over NVLink peer memory, in ONE kernel of ours (sm_100a).
//
// The reference has no multi-GPU path; this replaces the ncclAllGather of the sharded global optimisation
// (gn_driver.cu: accumulate_step) where the exchange is latency-bound: 107 KB (config 3) to 0.9 MB (config 4) per
// iteration, ~20 us / ~100 us through NCCL on 8 GPUs against a few microseconds of NVLink time.
//
// Every rank owns an exchange buffer allocated by the library (cudaMalloc, so it can be shared by CUDA IPC whatever
// allocator the caller uses): two parity buffers of the full [E2, 56] fp64 edge-block array + two rows of flags.
// The accumulate kernel of iteration `it` writes this rank's blocks into parity it & 1 of its OWN buffer; then
// peer_allgather_kernel (one CTA per peer)
//   1. copies them into the same slots of the peer's buffer with 16-byte stores over NVLink,
//   2. fence.sys, then st.release.sys of this solve's epoch into flag[parity][rank] on the peer,
//   3. waits (ld.acquire.sys) until the peer's flag in its own buffer carries the epoch.
// When the kernel retires the local buffer holds all blocks and the solve that follows on the stream reads it.
// Two parities suffice: a rank can only push iteration it+1 after its own solve of iteration it, which needed every
// peer's blocks of iteration it, which a peer only pushes after ITS solve of iteration it-1 -- so nobody still reads
// parity (it+1) & 1 when it is overwritten.  Epochs (solve sequence number << 12 | iteration + 1) only grow, so flags
// are never reset.  The wait is bounded: a peer that never arrives sets CTRL_BAD instead of hanging the device.
#include <cstring>

#include "gn_common.cuh"

namespace m3r {
namespace {

constexpr int kMaxPeers = 16;
constexpr int kFlagStride = 64;  // ints per parity row of flags

struct PeerXchg {
  int rank = -1, world = 0;
  size_t slot_bytes = 0;  // one parity buffer
  char* local = nullptr;
  char* peer[kMaxPeers] = {};
  unsigned seq = 0;
  bool connected = false;
  int device = -1;
};

struct PushArgs {
  char* peer[kMaxPeers];
  int rank, world, parity, epoch;
  unsigned long long slot_bytes, off, nbytes;
  int* ctrl;
};

__device__ __forceinline__ void st_release_sys(int* p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(1024)
peer_allgather_kernel(const PushArgs a) {
  if (a.ctrl[CTRL_DONE]) return;  // (identical on every rank: the stop test runs on identical data)
  // one CTA per peer: push this rank's blocks into the peer's buffer, raise our flag there, wait for the peer's flag here
  const int tid = threadIdx.x;
  const int r = (int)blockIdx.x + ((int)blockIdx.x >= a.rank ? 1 : 0);
  const uint4* src = reinterpret_cast<const uint4*>(a.peer[a.rank] + (size_t)a.parity * a.slot_bytes + a.off);
  uint4* dst = reinterpret_cast<uint4*>(a.peer[r] + (size_t)a.parity * a.slot_bytes + a.off);
  const size_t n16 = a.nbytes / 16;
  for (size_t k = tid; k < n16; k += blockDim.x) dst[k] = src[k];
  // EVERY storing thread fences before the barrier.  Measured on 2 x B200: with only the flag-writing thread fencing
  // after the barrier (the usual "barrier; thread 0: fence, flag" idiom) the flag overtook other warps' NVLink stores
  // about every second time two ranks pushed at the same moment -- the receiver then solved with the peer's blocks of
  // two iterations earlier (3 of 3 runs of tests/test_gpu_multi.py failed, 8 of 8 pass with this line).
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<int*>(a.peer[r] + 2 * a.slot_bytes) + a.parity * kFlagStride + a.rank, a.epoch);
    const int* f = reinterpret_cast<const int*>(a.peer[a.rank] + 2 * a.slot_bytes) + a.parity * kFlagStride + r;
    long long spins = 0;
    while ((int)(ld_acquire_sys(f) - a.epoch) < 0) {
      if (++spins > (1ll << 27)) {  // seconds: a peer is gone
        a.ctrl[CTRL_BAD] = 3;
        break;
      }
    }
  }
  __syncthreads();
}

}  // namespace

// used by gn_driver.cu
int peer_xchg_check(const void* x, int num_edges, int lo, int hi) {
  const PeerXchg* p = static_cast<const PeerXchg*>(x);
  if (!p || !p->connected || p->world < 2) return M3R_ERR_BAD_ARG;
  if (num_edges % p->world != 0) return M3R_ERR_BAD_ARG;
  const int cnt = num_edges / p->world;
  if (lo != p->rank * cnt || hi != lo + cnt) return M3R_ERR_BAD_ARG;
  if ((size_t)num_edges * kEdgeBlock * sizeof(double) > p->slot_bytes) return M3R_ERR_WORKSPACE;
  return M3R_OK;
}
unsigned peer_xchg_next_seq(void* x) { return ++static_cast<PeerXchg*>(x)->seq; }
static int peer_epoch(unsigned seq, int it) { return (int)(((seq & 0x7ffffu) << 12) | (unsigned)((it + 1) & 0xfff)); }
double* peer_xchg_buffer(void* x, int parity) {
  PeerXchg* p = static_cast<PeerXchg*>(x);
  return reinterpret_cast<double*>(p->local + (size_t)(parity & 1) * p->slot_bytes);
}
int launch_peer_allgather(void* x, int lo, int hi, int it, unsigned seq, int* ctrl, cudaStream_t s) {
  PeerXchg* p = static_cast<PeerXchg*>(x);
  PushArgs a{};
  for (int r = 0; r < p->world; r++) a.peer[r] = p->peer[r];
  a.rank = p->rank;
  a.world = p->world;
  a.parity = it & 1;
  a.epoch = peer_epoch(seq, it);
  a.slot_bytes = p->slot_bytes;
  a.off = (unsigned long long)lo * kEdgeBlock * sizeof(double);
  a.nbytes = (unsigned long long)(hi - lo) * kEdgeBlock * sizeof(double);  // 448 B per edge: a multiple of 16
  a.ctrl = ctrl;
  peer_allgather_kernel<<<p->world - 1, 1024, 0, s>>>(a);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r

using namespace m3r;

extern "C" int m3r_xchg_create(size_t capacity_bytes, void** xchg_out, void* ipc_handle_out) {
  if (!xchg_out || !ipc_handle_out || capacity_bytes == 0) return M3R_ERR_BAD_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the handle travels as 64 opaque bytes");
  PeerXchg* p = new PeerXchg();
  p->slot_bytes = align_up(capacity_bytes, 256);
  const size_t total = 2 * p->slot_bytes + 2 * kFlagStride * sizeof(int);
  if (cudaGetDevice(&p->device) != cudaSuccess || cudaMalloc(reinterpret_cast<void**>(&p->local), total) != cudaSuccess ||
      cudaMemset(p->local, 0, total) != cudaSuccess) {
    set_last_error("m3r_xchg_create: cudaMalloc / cudaMemset", cudaGetLastError());
    delete p;
    return M3R_ERR_CUDA;
  }
  cudaIpcMemHandle_t h;
  const cudaError_t e = cudaIpcGetMemHandle(&h, p->local);
  if (e != cudaSuccess) {
    set_last_error("cudaIpcGetMemHandle", e);
    cudaFree(p->local);
    delete p;
    return M3R_ERR_CUDA;
  }
  memcpy(ipc_handle_out, &h, 64);
  M3R_CUDA_CHECK(cudaDeviceSynchronize());  // the zeroed flags are in memory before any peer can write them
  *xchg_out = p;
  return M3R_OK;
}

extern "C" int m3r_xchg_connect(void* xchg, int rank, int world, const void* all_handles) {
  PeerXchg* p = static_cast<PeerXchg*>(xchg);
  if (!p || !all_handles || world < 2 || world > kMaxPeers || rank < 0 || rank >= world || p->connected) return M3R_ERR_BAD_ARG;
  p->rank = rank;
  p->world = world;
  for (int r = 0; r < world; r++) {
    if (r == rank) {
      p->peer[r] = p->local;
      continue;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const char*>(all_handles) + (size_t)r * 64, 64);
    void* ptr = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      set_last_error("cudaIpcOpenMemHandle (peer exchange needs P2P access between the ranks' GPUs)", e);
      return M3R_ERR_CUDA;
    }
    p->peer[r] = static_cast<char*>(ptr);
  }
  p->connected = true;
  return M3R_OK;
}

extern "C" int m3r_xchg_destroy(void* xchg) {
  PeerXchg* p = static_cast<PeerXchg*>(xchg);
  if (!p) return M3R_OK;
  cudaDeviceSynchronize();
  for (int r = 0; r < p->world; r++)
    if (r != p->rank && p->peer[r]) cudaIpcCloseMemHandle(p->peer[r]);
  if (p->local) cudaFree(p->local);
  delete p;
  return M3R_OK;
}
