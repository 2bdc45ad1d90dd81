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
// The barriers are flag exchanges through the peers' signal pads: thread `peer` of CTA 0 of rank r
// sets slot [r] of the peer's pad from 0 to 1 (release, system scope) and waits until slot [peer] of
// its own pad is 1, resetting it to 0 (acquire) -- no epoch counters, reusable immediately.  Only CTA 0
// talks to the peers; the other CTAs of the grid (all co-resident: at most one per SM) are released by,
// and report to, CTA 0 through two words of the rank's own pad.  (r01: one flag per (CTA, peer), which
// capped the grid at 24 CTAs -- too few 16-byte round trips in flight to cover the NVLink latency.)
// The kernel is launched as a programmatic dependent of the per-Gaussian backward kernel that fills the block.
#include <cstdlib>
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

constexpr int AR_THREADS = 256;
#ifndef IGS_AR_UNROLL
#define IGS_AR_UNROLL 4
#endif
constexpr int AR_UNROLL = IGS_AR_UNROLL;
constexpr int AR_MAX_BLOCKS = 132;  // co-resident by construction (< 148 SMs)
constexpr int AR_PAD_OFFSET = 256;  // in uint32 words: the first KB of the pad is left to torch's own barriers
// words behind AR_PAD_OFFSET: [0, world) peer flags | [world] go | [world + 1] arrived

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
__device__ __forceinline__ uint32_t load_acquire(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.global.acquire.gpu.b32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void store_release(uint32_t* addr, uint32_t v) {
  asm volatile("st.global.release.gpu.b32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}

// CTA 0, all threads arrive; thread t < world talks to peer t
__device__ __forceinline__ void rank_barrier(uint32_t* const* pads, int rank, int world) {
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const int peer = threadIdx.x;
    put_flag(pads[peer] + AR_PAD_OFFSET + rank);   // "`rank` is here", in peer's pad
    take_flag(pads[rank] + AR_PAD_OFFSET + peer);  // wait for `peer`
  }
  __syncthreads();
}

__global__ void __launch_bounds__(AR_THREADS) nvls_allreduce_kernel(float* mc, uint32_t* const* pads, int rank, int world,
                                                                    size_t n_vec4) {
  pdl_wait(); // this rank's gradient block is complete
  uint32_t* const go = pads[rank] + AR_PAD_OFFSET + world;
  uint32_t* const arrived = go + 1;
  // 1. every rank's block is complete: CTA 0 meets the peers, then releases the local CTAs
  if (blockIdx.x == 0) {
    rank_barrier(pads, rank, world);
    if (threadIdx.x == 0) store_release(go, 1u);
  } else {
    if (threadIdx.x == 0)
      while (load_acquire(go) == 0u) {}
    __syncthreads();
  }
  // 2. this rank's slice of 16-byte words, spread over the CTAs; AR_UNROLL independent switch round
  //    trips in flight per thread
  const size_t per = (n_vec4 + world - 1) / world;
  const size_t lo = min(n_vec4, per * rank), hi = min(n_vec4, lo + per);
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
  // 3. all slices have landed everywhere: the local CTAs report to CTA 0, CTA 0 meets the peers again
  __threadfence_system();
  __syncthreads();
  if (blockIdx.x != 0) {
    if (threadIdx.x == 0) atomicAdd(arrived, 1u);
    return;
  }
  if (threadIdx.x == 0) {
    while (load_acquire(arrived) != gridDim.x - 1) {}
    *arrived = 0u; // the two words go back to zero for the next call (stream order: it starts after this one ends)
    *go = 0u;
    __threadfence();
  }
  rank_barrier(pads, rank, world);
}

} // namespace igs

using namespace igs;

extern "C" {

int igs_allreduce_nvls(void* multicast_ptr, void* signal_pad_ptrs_dev, int32_t signal_pad_bytes, int32_t rank,
                       int32_t world, int64_t n_floats, igs_stream_t stream) {
  if (!multicast_ptr || !signal_pad_ptrs_dev || world < 1 || world > 32 || rank < 0 || rank >= world || n_floats < 0 ||
      (n_floats & 3) || ((uintptr_t)multicast_ptr & 15)) {
    set_error("allreduce_nvls: bad arguments (n_floats must be a multiple of 4, the pointer 16-byte aligned)");
    return IGS_ERR_INVALID;
  }
  if (n_floats == 0 || world == 1) return IGS_OK;
  const size_t n4 = (size_t)n_floats / 4, per = (n4 + world - 1) / world;
  if (signal_pad_bytes > 0 && (size_t)signal_pad_bytes < (AR_PAD_OFFSET + (size_t)world + 2) * 4) {
    set_error("allreduce_nvls: signal pad too small"); return IGS_ERR_INVALID;
  }
  // enough CTAs that one sweep of AR_UNROLL 16-byte words per thread covers the slice, at most one per SM
  static const int max_blocks = [] {
    const char* e = getenv("IGS_AR_BLOCKS");
    int v = e ? atoi(e) : AR_MAX_BLOCKS, sms = 0, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms > 16 && v > sms - 16) v = sms - 16; // (co-residency is what the local release / report relies on)
    return v < 1 ? 1 : v;
  }();
  int blocks = (int)((per + AR_THREADS * AR_UNROLL - 1) / (AR_THREADS * AR_UNROLL));
  blocks = blocks < 1 ? 1 : blocks > max_blocks ? max_blocks : blocks;
  if (int r = check_cuda(launch_after(nvls_allreduce_kernel, dim3(blocks), dim3(AR_THREADS), 0, (cudaStream_t)stream,
                                      (float*)multicast_ptr, (uint32_t* const*)signal_pad_ptrs_dev, (int)rank, (int)world, n4),
                         "allreduce_nvls launch")) return r;
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "allreduce_nvls");
}

} // extern "C"
