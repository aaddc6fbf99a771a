// gs_allreduce.cu -- the one collective of data-parallel training: sum of the per-rank flat gradient slabs
// that pose_bwd_kernel writes, in place, as ONE hand-written kernel over NVLink / NVSwitch peer memory (the slab lives in
// symmetric memory: every rank holds the peers' addresses and, with NVLS, a multicast address of the same buffer).
// Two-shot over the NVLS multicast object: rank r owns the r-th slice; `multimem.ld_reduce` pulls the slice already summed by
// the switch, `multimem.st` writes the result into every replica: each GPU moves 2/N of the slab instead of NCCL's ring
// traffic, and there is no staging copy.  (A peer-to-peer variant without the multicast object existed in round 1; it never
// ran on hardware -- this pool always offers multicast -- and was removed; without multicast the caller uses NCCL.)
// Ranks meet at in-kernel barriers over the symmetric signal pads (one epoch word per peer, written with a single release store:
// no remote atomics, nothing to reset), so the launch sits inside the
// step's CUDA graph right behind pose_bwd_kernel and overlaps its tail through programmatic dependent launch.  Sums are
// identical on every rank (each slice is reduced once, then broadcast).
#include "gs_common.cuh"

namespace {

#define AR_THREADS 512
#define AR_SPIN_LIMIT (1u << 25)  // a stuck barrier gives up after tens of seconds (error word set) instead of hanging the GPU:
                                  // a rank may legitimately arrive seconds late (graph re-capture after an overflow, data stall)
#define AR_UNROLL 8

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Cross-rank barrier executed by ONE block per rank, EPOCH based: the k-th barrier of this slab's life writes the number k into
// word [AR_PAD_BASE + rank] of every peer's signal pad (one release store per peer: a one-way trip through the switch, no
// remote atomic and no reset traffic) and waits until its own words [AR_PAD_BASE + peer] have all reached k.  Every rank runs
// the same sequence of calls, so the counters agree; they only ever grow (compared with wrap-around arithmetic).  The words
// sit 1 KB into the pad, clear of the channels torch's own symmetric-memory barriers use.
// Returns false when a peer never arrived (error word set): the caller then leaves the slab untouched.
#define AR_PAD_BASE 256
__device__ __forceinline__ bool barrier_ranks(uint32_t* const* pads, int rank, int world, uint32_t epoch, uint32_t* error_flag) {
  __shared__ uint32_t s_fail;
  if (threadIdx.x == 0) s_fail = 0u;
  __syncthreads();
  if ((int)threadIdx.x < world) {
    __threadfence_system();  // this rank's earlier writes (its slab / its share of the result) are visible before the signal
    const int peer = (int)threadIdx.x;
    st_release_sys(pads[peer] + AR_PAD_BASE + rank, epoch);
    const uint32_t* wait = pads[rank] + AR_PAD_BASE + peer;
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(wait) - epoch) < 0)
      if (++spins > AR_SPIN_LIMIT) { *error_flag = 2u; s_fail = 1u; break; }
  }
  __syncthreads();
  return s_fail == 0u;
}

__device__ __forceinline__ float4 mm_ld_reduce(const float4* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mm_st(float4* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 ld_sys(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(AR_THREADS)
slab_allreduce_kernel(long long n4, int rank, int world, void* const* peers, uint32_t* const* pads, float4* mc, long long byte_off,
                      uint32_t* error_flag) {
  GS_PDL_PROLOGUE();  // the local slab is complete (pose_bwd_kernel has finished)
  // phase timestamps of block 0 (ns, %globaltimer) behind the error word: [2] start, [4] after barrier 1, [6] after the data
  // phase, [8] end -- read by profiles/allreduce_bench.py (a multi-rank barrier kernel cannot be replayed under ncu)
  unsigned long long* const stamps = reinterpret_cast<unsigned long long*>(error_flag + 2);
  auto stamp = [&](int k) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      stamps[k] = t;
    }
  };
  stamp(0);
  const long long per = (n4 + world - 1) / world;
  const long long begin = min(n4, per * rank), end = min(n4, begin + per);
  const long long stride = (long long)gridDim.x * AR_THREADS, t0 = (long long)blockIdx.x * AR_THREADS + threadIdx.x;
  // local control words behind the stamps: [12] go flag, [13] blocks done, [14] barriers completed so far (the epoch)
  uint32_t* const go = error_flag + 12;
  uint32_t* const done = error_flag + 13;
  uint32_t* const epoch = error_flag + 14;
  __shared__ bool s_last;
  auto open_gate = [&]() {   // every rank's slab is complete: block 0 meets the peers, the other blocks wait for block 0
    if (blockIdx.x == 0) {
      const bool ok = barrier_ranks(pads, rank, world, *epoch + 1u, error_flag);
      if (threadIdx.x == 0) st_release_gpu(go, ok ? 1u : 2u);  // 2 = a peer is missing: nobody touches the slab
      __syncthreads();  // the other threads of THIS block read the gate too
    } else {
      if (threadIdx.x == 0) {
        uint32_t spins = 0;
        while (ld_acquire_gpu(go) == 0u)
          if (++spins > AR_SPIN_LIMIT) { *error_flag = 3u; break; }
      }
      __syncthreads();
    }
  };
  auto last_block = [&]() -> bool {  // true in the block that finishes a phase last (its writes and everybody else's are out)
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      s_last = atomicAdd(done, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    return s_last;
  };
  open_gate();
  stamp(1);
  const bool healthy = ld_acquire_gpu(go) == 1u && *error_flag == 0u;
  {
    float4* const m = reinterpret_cast<float4*>(reinterpret_cast<char*>(mc) + byte_off);
    // AR_UNROLL independent switch reductions in flight per thread before the first store: the round trip through the
    // NVSwitch is microseconds, so the loop is latency-bound unless many are outstanding
    for (long long i0 = begin + t0; healthy && i0 < end; i0 += stride * AR_UNROLL) {
      float4 v[AR_UNROLL];
#pragma unroll
      for (int u = 0; u < AR_UNROLL; u++)
        if (i0 + u * stride < end) v[u] = mm_ld_reduce(m + i0 + u * stride);
#pragma unroll
      for (int u = 0; u < AR_UNROLL; u++)
        if (i0 + u * stride < end) mm_st(m + i0 + u * stride, v[u]);
    }
    stamp(2);
    if (last_block()) {  // every block's stores are out: meet the peers, then leave the control words clean
      if (healthy) barrier_ranks(pads, rank, world, *epoch + 2u, error_flag);  // every slice has reached every replica
      if (threadIdx.x == 0) { *go = 0u; *done = 0u; if (healthy) *epoch += 2u; }
      if (threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); stamps[3] = t; }
    }
    return;
  }
}

}  // namespace

extern "C" int slab_allreduce(const SlabAllReduceArgs* args, void* stream) {
  if (!args) return GS_ERR_BAD_ARG;
  const SlabAllReduceArgs& a = *args;
  if (a.n < 0 || (a.n & 3) || a.world < 1 || a.world > 64 || a.rank < 0 || a.rank >= a.world || !a.peer_ptrs || !a.signal_pads ||
      !a.error_flag || a.blocks < 1 || (a.byte_offset & 15))
    return GS_ERR_BAD_ARG;
  if (a.n == 0 || a.world == 1) return GS_OK;
  if (!a.multicast_ptr) return GS_ERR_BAD_ARG;  // NVLS multicast only (without it the caller uses NCCL)
  gs_launch(slab_allreduce_kernel, dim3((unsigned)a.blocks), dim3(AR_THREADS), 0, (cudaStream_t)stream, (long long)(a.n >> 2), (int)a.rank,
            (int)a.world, a.peer_ptrs, a.signal_pads, (float4*)a.multicast_ptr, (long long)a.byte_offset, a.error_flag);
  return gs_check(cudaGetLastError());
}
