// dibr_peer.cu -- the path's only exchange as our own kernels over NVLink peer
// memory instead of a NCCL call: a one-shot SUM all-reduce of the shared mesh's vertex-level
// gradient buffer (V x (2 + D) floats, < 1 MB) across the GPUs of one NVSwitch box.
//
// Every rank owns one buffer of a symmetric allocation (same size on every rank, every peer's
// copy mapped into every process; the Python binding gets the mapping from
// torch.distributed._symmetric_memory, which is plumbing -- allocation + rendezvous) and a
// signal pad of 32-bit flags.  One exchange:
//   peer_wait_done   (before the local scatter refills the buffer) every peer has finished
//                    READING this rank's buffer of the previous exchange
//   ... memset + face -> vertex scatter into the own buffer (dibr_mesh.cu) ...
//   peer_allreduce   block 0: tell every peer "my buffer is ready", wait for every peer's
//                    "ready"; all blocks: out[i] = sum over ranks r = 0..N-1 of buf_r[i]
//                    (128-bit loads straight from the peers' HBM over NVLink, the same order
//                    on every rank, so all ranks hold bit-identical sums); last block: tell
//                    every peer "done reading".
// A message this small is latency bound: one kernel with two flag exchanges instead of a
// collective library's launch + protocol.  Flags follow the put / wait convention of
// torch's signal pads: put = CAS 0 -> 1 on the peer's pad (spins while the previous signal
// is unconsumed), wait = CAS 1 -> 0 on the own pad.
#include <cstring>

#include "../../include/dibr_b200.h"
#include "dibr_common.cuh"

namespace dibr {

int fail(int code, const char* fmt, ...);

namespace {

constexpr int kMaxPeers = 16;
constexpr int kChanReady = 8, kChanDone = 9;   // flag rows of the pad used here (torch's own
                                               // barriers use the low channels)

struct PeerArgs {
  const float* buf[kMaxPeers];     // every rank's buffer, as mapped in this process
  unsigned* pad[kMaxPeers];        // every rank's signal pad
  float* out;                      // local result (n floats)
  int* sync;                       // local int[2], zero: {go flag, finished-block counter}
  long long n;                     // floats, multiple of 4
  int rank, world;
};

__device__ __forceinline__ void put_signal(unsigned* peer_pad, int chan, int world, int me) {
  unsigned* slot = peer_pad + chan * world + me;
  while (atomicCAS_system(slot, 0u, 1u) != 0u) { }
}
__device__ __forceinline__ void wait_signal(unsigned* my_pad, int chan, int world, int src) {
  unsigned* slot = my_pad + chan * world + src;
  while (atomicCAS_system(slot, 1u, 0u) != 1u) { }
}

// mode 0: prime (once, at set-up): "done reading" to every peer so that the first wait passes
// mode 1: wait until every peer is done reading this rank's buffer
__global__ void peer_signal_kernel(const PeerArgs a, int mode) {
  const int t = threadIdx.x;
  if (t >= a.world) return;
  if (mode == 0) put_signal(a.pad[t], kChanDone, a.world, a.rank);
  else wait_signal(a.pad[a.rank], kChanDone, a.world, t);
}

__global__ void __launch_bounds__(256) peer_allreduce_kernel(const PeerArgs a) {
  volatile int* go = a.sync;
  if (blockIdx.x == 0) {
    if (threadIdx.x < a.world) {
      __threadfence_system();                       // the scatter's results, system wide
      put_signal(a.pad[threadIdx.x], kChanReady, a.world, a.rank);
      wait_signal(a.pad[a.rank], kChanReady, a.world, threadIdx.x);
      __threadfence_system();
    }
    __syncthreads();
    if (threadIdx.x == 0) { *go = 1; __threadfence(); }
  } else {
    if (threadIdx.x == 0) {
      while (*go == 0) { }
      __threadfence();
    }
    __syncthreads();
  }
  const long long n4 = a.n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < a.world; r++) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(a.buf[r]) + i);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    reinterpret_cast<float4*>(a.out)[i] = s;
  }
  __syncthreads();
  __shared__ int last;
  if (threadIdx.x == 0) {
    __threadfence();
    last = atomicAdd(a.sync + 1, 1) == (int)gridDim.x - 1;
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) { a.sync[0] = 0; a.sync[1] = 0; }   // every block is past its spin
    if (threadIdx.x < a.world) put_signal(a.pad[threadIdx.x], kChanDone, a.world, a.rank);
  }
}

int fill_args(PeerArgs& a, int rank, int world, const int64_t* bufs, const int64_t* pads) {
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world)
    return fail(DIBR_E_INVALID, "rank %d / world %d (at most %d peers)", rank, world, kMaxPeers);
  if (!pads) return fail(DIBR_E_INVALID, "null pointer");
  memset(&a, 0, sizeof(a));
  a.rank = rank; a.world = world;
  for (int r = 0; r < world; r++) {
    a.buf[r] = bufs ? (const float*)(uintptr_t)bufs[r] : nullptr;
    a.pad[r] = (unsigned*)(uintptr_t)pads[r];
    if (!a.pad[r]) return fail(DIBR_E_INVALID, "null signal pad of rank %d", r);
  }
  return DIBR_OK;
}

}  // namespace
}  // namespace dibr

using namespace dibr;

#define PEER_CUDA(expr)                                                              \
  do {                                                                               \
    cudaError_t e_ = (expr);                                                         \
    if (e_ != cudaSuccess)                                                           \
      return fail(DIBR_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_));      \
  } while (0)

extern "C" {

DIBR_API int dibr_peer_signal(int rank, int world, const int64_t* signal_pads, int mode,
                              void* stream) {
  PeerArgs a;
  int rc = fill_args(a, rank, world, nullptr, signal_pads);
  if (rc) return rc;
  if (mode != 0 && mode != 1) return fail(DIBR_E_INVALID, "mode must be 0 (prime) or 1 (wait)");
  peer_signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a, mode);
  PEER_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_peer_allreduce(int rank, int world, const int64_t* buffers,
                                 const int64_t* signal_pads, int64_t num_floats, float* out,
                                 int* sync, void* stream) {
  PeerArgs a;
  if (!buffers) return fail(DIBR_E_INVALID, "null pointer");
  int rc = fill_args(a, rank, world, buffers, signal_pads);
  if (rc) return rc;
  if (num_floats < 0 || (num_floats & 3)) return fail(DIBR_E_INVALID, "num_floats must be a multiple of 4");
  if (!out || !sync) return fail(DIBR_E_INVALID, "null pointer");
  for (int r = 0; r < world; r++)
    if (!a.buf[r] || ((uintptr_t)a.buf[r] & 15)) return fail(DIBR_E_INVALID, "buffer of rank %d null or unaligned", r);
  a.out = out; a.sync = sync; a.n = num_floats;
  // every block must be resident (block 0 releases the others): at most one per SM
  long long blocks = (num_floats / 4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 128) blocks = 128;
  peer_allreduce_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a);
  PEER_CUDA(cudaGetLastError());
  return DIBR_OK;
}

}  // extern "C"
