// the one exchange step of the path -- summing the packed per-Gaussian
// parameter gradients of a view-sharded batch over the GPUs of one NVSwitch box -- as ONE
// hand-written kernel over NVLink multicast memory (NVLS), replacing ncclAllReduce for this
// buffer (measured on 8 x B200 for the 6 MB block: NCCL 58-75 us, this kernel's scheme ~30 us).
//
// Every rank holds the gradient block in symmetric memory (same offset in a peer-mapped
// allocation, torch.distributed._symmetric_memory) that is also mapped at a MULTICAST address:
//   1. barrier over the ranks (every rank's per-Gaussian backward kernel has finished: it ran
//      earlier on the same stream as this kernel);
//   2. rank r owns the r-th 1/world slice: `multimem.ld_reduce.add.v4.f32` makes the switch
//      fetch that 16-byte word from all replicas and return the sum, `multimem.st.v4.f32`
//      broadcasts the sum back into all replicas -- each GPU moves 1/world of the block twice;
//   3. barrier (all slices have landed everywhere).
// The barriers are flag exchanges through the peers' signal pads: block b of rank r sets slot
// [b][r] of every peer's pad from 0 to 1 (release, system scope) and waits until its own pad's
// slots [b][*] are 1, resetting them to 0 (acquire) -- no epoch counters, reusable immediately.
#include <cstdlib>
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

constexpr int AR_THREADS = 256;
constexpr int AR_UNROLL = 4;
constexpr int AR_MAX_BLOCKS = 24;   // signal-pad slots used: [AR_PAD_OFFSET + b * world + peer]
constexpr int AR_PAD_OFFSET = 256;  // in uint32 words: the first KB of the pad is left to torch's own barriers

__device__ __forceinline__ void put_flag(uint32_t* addr) {
  uint32_t old;
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(addr) : "memory");
  } while (old != 0u);
}
__device__ __forceinline__ void take_flag(uint32_t* addr) {
  uint32_t old;
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(addr) : "memory");
  } while (old != 1u);
}

// all threads of the block arrive; thread t < world talks to peer t
__device__ __forceinline__ void rank_barrier(uint32_t* const* pads, int rank, int world) {
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const int peer = threadIdx.x;
    const size_t slot = AR_PAD_OFFSET + (size_t)blockIdx.x * world;
    put_flag(pads[peer] + slot + rank);   // "block b of `rank` is here", in peer's pad
    take_flag(pads[rank] + slot + peer);  // wait for block b of `peer`
  }
  __syncthreads();
}

__global__ void __launch_bounds__(AR_THREADS) nvls_allreduce_kernel(float* mc, uint32_t* const* pads, int rank, int world,
                                                                    size_t n_vec4) {
  rank_barrier(pads, rank, world);
  // this rank's slice of 16-byte words, spread over the blocks
  const size_t per = (n_vec4 + world - 1) / world;
  const size_t lo = min(n_vec4, per * rank), hi = min(n_vec4, lo + per);
  // AR_UNROLL independent switch round trips in flight per thread
  const size_t stride = (size_t)gridDim.x * AR_THREADS;
  for (size_t i = lo + (size_t)blockIdx.x * AR_THREADS + threadIdx.x; i < hi; i += stride * AR_UNROLL) {
    float4 v[AR_UNROLL];
#pragma unroll
    for (int u = 0; u < AR_UNROLL; u++)
      if (i + u * stride < hi)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(mc + 4 * (i + u * stride)) : "memory");
#pragma unroll
    for (int u = 0; u < AR_UNROLL; u++)
      if (i + u * stride < hi)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                     ::"l"(mc + 4 * (i + u * stride)), "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w) : "memory");
  }
  __threadfence_system();
  rank_barrier(pads, rank, world);
}

} // namespace igs

using namespace igs;

extern "C" {

int igs_allreduce_nvls(void* multicast_ptr, void* signal_pad_ptrs_dev, int32_t signal_pad_bytes, int32_t rank,
                       int32_t world, int64_t n_floats, igs_stream_t stream) {
  // blocks are limited by the flag slots the pad has behind AR_PAD_OFFSET: one per (block, peer)
  const int max_blocks_arg = signal_pad_bytes > 0 ? (signal_pad_bytes / 4 - AR_PAD_OFFSET) / (world > 0 ? world : 1) : 0;
  if (signal_pad_bytes > 0 && max_blocks_arg < 1) { set_error("allreduce_nvls: signal pad too small"); return IGS_ERR_INVALID; }
  if (!multicast_ptr || !signal_pad_ptrs_dev || world < 1 || world > 32 || rank < 0 || rank >= world || n_floats < 0 ||
      (n_floats & 3) || ((uintptr_t)multicast_ptr & 15)) {
    set_error("allreduce_nvls: bad arguments (n_floats must be a multiple of 4, the pointer 16-byte aligned)");
    return IGS_ERR_INVALID;
  }
  if (n_floats == 0 || world == 1) return IGS_OK;
  const size_t n4 = (size_t)n_floats / 4, per = (n4 + world - 1) / world;
  static const int vpt = [] { const char* e = getenv("IGS_AR_VPT"); const int v = e ? atoi(e) : 8; return v < 1 ? 1 : v; }();
  int blocks = (int)((per + AR_THREADS * vpt - 1) / (AR_THREADS * vpt)); // 16-byte words per thread (tuning: IGS_AR_VPT)
  const int max_blocks = max_blocks_arg > 0 ? max_blocks_arg : AR_MAX_BLOCKS;
  blocks = blocks < 1 ? 1 : blocks > max_blocks ? max_blocks : blocks;
  nvls_allreduce_kernel<<<blocks, AR_THREADS, 0, (cudaStream_t)stream>>>((float*)multicast_ptr,
                                                                          (uint32_t* const*)signal_pad_ptrs_dev, rank,
                                                                          world, n4);
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "allreduce_nvls");
}

} // extern "C"
