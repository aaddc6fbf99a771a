// gs_umma.cuh -- the tcgen05 / TMEM building block shared by the per-Gaussian MLP kernels (gs_nonrigid.cu: non-rigid
// and skinning MLPs; gs_color.cu: colour MLP): canonical K-major no-swizzle operand layout, BF16 hi/lo split, PTX
// wrappers for tcgen05.mma / commit / ld and mbarriers, and the one-thread MMA issue of a whole layer.
#pragma once
#include <cuda_bf16.h>

#include "gs_common.cuh"

namespace {

constexpr int NR_TILE = 128;   // Gaussians per CTA = UMMA M

// element index of (row, k) in the canonical K-major no-swizzle layout of a [ROWS x K] BF16 operand:
// core matrix = 8 rows x 16 bytes; core matrices of one 8-wide K slab are contiguous over the rows (SBO = 128 B),
// consecutive K slabs are ROWS*16 bytes apart (LBO).
__host__ __device__ inline int core_idx(int row, int k, int rows) { return (k >> 3) * (rows * 8) + row * 8 + (k & 7); }

__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}

// ---- tcgen05 / TMEM / mbarrier wrappers (PTX) ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// shared-memory matrix descriptor, K-major, no swizzle (cute/arch/mma_sm100_desc.hpp: SmemDescriptor):
// [0,14) start>>4 | [16,30) LBO>>4 (K-slab stride) | [32,46) SBO>>4 (8-row group stride) | [46,48) version = 1
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
         (1ull << 46);
}
// instruction descriptor, kind::f16 (InstrDescriptor): D = F32 [4,6)=1, A = B = BF16 [7,10)=[10,13)=1, K-major both,
// N>>3 at [17,23), M>>4 at [24,29)
__device__ __forceinline__ uint32_t umma_idesc(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void nr_mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void nr_mbar_wait(unsigned long long* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "NRW_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra NRD_%=;\n"
      "bra NRW_%=;\n"
      "NRD_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// ---- bulk asynchronous copy global -> shared (TMA, 1-D): one thread moves a whole operand; the bytes land on an mbarrier ----
// (the per-thread LDG -> STS loops these replace kept ONE load in flight per thread: with 4 warps per SM the operand loads were
// a chain of L2 round trips and ~60 % of the MLP kernels' time, profiles/r2/r2_v2_mlp_before_tma.json)
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// One thread: the four operand halves of a layer (A hi | lo, W hi | lo; 16-byte aligned, sizes multiples of 16) on `bar`.
// The fence orders earlier generic-proxy accesses to the destination buffers (made visible to this thread by a barrier)
// before the async-proxy writes.
__device__ __forceinline__ void bulk_load_operands(unsigned long long* bar, void* a_hi, void* a_lo, const void* ga_hi, const void* ga_lo,
                                                   uint32_t a_bytes, void* w_hi, void* w_lo, const void* gw_hi, const void* gw_lo,
                                                   uint32_t w_bytes) {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  mbar_expect_tx(bar, 2 * a_bytes + 2 * w_bytes);
  if (a_bytes) { bulk_g2s(a_hi, ga_hi, a_bytes, bar); bulk_g2s(a_lo, ga_lo, a_bytes, bar); }
  if (w_bytes) { bulk_g2s(w_hi, gw_hi, w_bytes, bar); bulk_g2s(w_lo, gw_lo, w_bytes, bar); }
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {  // 32 lanes x 32 columns: lane = this thread's row
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}

// stores 8 consecutive-k activations of one row (one 16-byte core-matrix row) as BF16 hi and lo
__device__ __forceinline__ void store_act8(__nv_bfloat16* A_hi, __nv_bfloat16* A_lo, int row, int k0, const float* v) {
  __align__(16) __nv_bfloat16 h[8], l[8];
#pragma unroll
  for (int i = 0; i < 8; i++) split_bf16(v[i], h[i], l[i]);
  const int idx = core_idx(row, k0, NR_TILE);
  *reinterpret_cast<uint4*>(A_hi + idx) = *reinterpret_cast<const uint4*>(h);
  *reinterpret_cast<uint4*>(A_lo + idx) = *reinterpret_cast<const uint4*>(l);
}

__device__ __forceinline__ void copy_weights(__nv_bfloat16* dst, const __nv_bfloat16* __restrict__ src, int elems) {
  const uint4* s = reinterpret_cast<const uint4*>(src);
  uint4* d = reinterpret_cast<uint4*>(dst);
  for (int i = threadIdx.x; i < elems / 8; i += blockDim.x) d[i] = __ldg(s + i);
}

// D[128 x N] (TMEM) = A[128 x K] . W[N x K]^T with the 2-way BF16 split; issued by ONE thread
__device__ __forceinline__ void issue_layer(uint32_t tmem_d, const __nv_bfloat16* A_hi, const __nv_bfloat16* A_lo,
                                            const __nv_bfloat16* W_hi, const __nv_bfloat16* W_lo, int K, int N,
                                            unsigned long long* bar) {
  const uint32_t idesc = umma_idesc(NR_TILE, N);
  const uint32_t lbo_a = NR_TILE * 16, lbo_w = (uint32_t)N * 16, sbo = 128;
  const uint32_t a_hi = smem_u32(A_hi), a_lo = smem_u32(A_lo), w_hi = smem_u32(W_hi), w_lo = smem_u32(W_lo);
  for (int ks = 0; ks < K / 16; ks++) {  // one instruction consumes K = 16: two 8-wide K slabs
    const uint64_t dah = umma_desc(a_hi + ks * 2 * lbo_a, lbo_a, sbo), dal = umma_desc(a_lo + ks * 2 * lbo_a, lbo_a, sbo);
    const uint64_t dwh = umma_desc(w_hi + ks * 2 * lbo_w, lbo_w, sbo), dwl = umma_desc(w_lo + ks * 2 * lbo_w, lbo_w, sbo);
    umma_bf16(tmem_d, dah, dwh, idesc, ks > 0 ? 1u : 0u);
    umma_bf16(tmem_d, dal, dwh, idesc, 1u);
    umma_bf16(tmem_d, dah, dwl, idesc, 1u);
  }
  umma_commit(bar);  // implies tcgen05.fence::before_thread_sync
}

}  // namespace
