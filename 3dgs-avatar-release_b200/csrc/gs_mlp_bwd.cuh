// gs_mlp_bwd.cuh -- backward of the 128-wide per-Gaussian MLPs (non-rigid deformation, skinning field) on the tcgen05 tensor
// cores: replaces the autograd backward of VanillaCondMLP (models/network_utils.py:182-249) that the reference obtains from
// cuBLAS GEMMs.
//
// One CTA walks tiles of 128 Gaussians.  For a layer Z = A_in . W^T + b with upstream gradient G = dL/dZ ([128 points x n_out],
// in shared memory in the same BF16 hi/lo core-matrix layout the forward uses for activations):
//   dA_in = G . W           K-major operands exactly like the forward (A operand = G, B operand = W^T prepared once per call)
//   dW   += G^T . A_in      contraction over the tile's 128 POINTS.  G and A_in sit in shared memory as [point][unit] with 8 units
//                           (16 bytes) contiguous per point: read with the point index as K that IS the canonical MN-major
//                           no-swizzle layout (core matrix = 8 K-rows x 16 bytes of MN-contiguous data), so the same buffers feed
//                           this MMA through descriptors with the a_major / b_major bits set and LBO / SBO swapped -- no transposed
//                           copy.  The 128 x 128 partial leaves TMEM through red.global.add.v4.f32 once per tile.
//   db   += column sums of G (warp butterfly, shared-memory accumulators, flushed once per CTA)
// 2-way BF16 split on both operands (hi*hi + lo*hi + hi*lo, FP32 accumulate in TMEM) as in the forward.
#pragma once
#include "gs_umma.cuh"

namespace {

constexpr int BW_H = 128;  // hidden width = rows of a tile
// Two threads per row: warps w and w + 4 own the same TMEM lane quadrant (a warp reaches lanes 32 * (warp % 4) ...) and split
// the COLUMNS of every epilogue / flush between them.  With one thread per row (4 warps per SM, one per scheduler) every
// kernel of this family was bound by the latency of its own instruction chains.
constexpr int BW_THREADS = 256;

struct BwSmem {
  __nv_bfloat16 G_hi[BW_H * BW_H], G_lo[BW_H * BW_H];  // dL/dZ of the current layer, [point][unit]
  __nv_bfloat16 A_hi[BW_H * BW_H], A_lo[BW_H * BW_H];  // saved input activations of the current layer, [point][unit]
  __nv_bfloat16 W_hi[BW_H * BW_H], W_lo[BW_H * BW_H];  // W^T of the current layer as the B operand of dA = G . W
  float dbias[5][BW_H];                                // per-CTA bias gradients, flushed at the end
  unsigned long long bar;     // MMA completion (tcgen05.commit)
  unsigned long long ld_bar;  // operand loads (cp.async.bulk complete_tx)
  uint32_t tmem_base;
};

// Everyone: wait for the operand loads issued with bulk_load_operands on sm.ld_bar (`ld_phase` counts them, kept per thread).
__device__ __forceinline__ void bw_wait_loads(BwSmem& sm, uint32_t& ld_phase) {
  nr_mbar_wait(&sm.ld_bar, ld_phase & 1u);
  ld_phase++;
}

// Signs of this thread's row of the saved activations in sm.A (bit k of the 128-bit mask = activation (row, k) > 0), taken right
// after the load so that the buffer can be refilled while the epilogue that needs the signs still runs.
__device__ __forceinline__ void bw_row_signs(const BwSmem& sm, int row, int unit_begin, int unit_end, uint32_t* signs /*[4]*/) {
  signs[0] = signs[1] = signs[2] = signs[3] = 0u;
  for (int s8 = unit_begin / 8; s8 < unit_end / 8; s8++) {
    const uint4 h = *reinterpret_cast<const uint4*>(sm.A_hi + core_idx(row, s8 * 8, BW_H));
    const uint32_t w[4] = {h.x, h.y, h.z, h.w};
    uint32_t bits = 0u;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const uint32_t b = (w[i >> 1] >> ((i & 1) * 16)) & 0xFFFFu;  // bf16 bits of element i
      bits |= ((!(b & 0x8000u) && (b & 0x7FFFu) != 0u) ? 1u : 0u) << i;  // strictly positive
    }
    signs[s8 >> 2] |= bits << ((s8 & 3) * 8);
  }
}
// LeakyReLU'(z) from the sign mask: 1 for z > 0, 0.01 otherwise (a > 0 <=> z > 0)
__device__ __forceinline__ void lrelu_mask8_bits(const uint32_t* signs, int k0, float* g) {
  const uint32_t bits = (signs[k0 >> 5] >> (k0 & 31)) & 0xFFu;
#pragma unroll
  for (int i = 0; i < 8; i++) g[i] = ((bits >> i) & 1u) ? g[i] : 0.01f * g[i];
}

__device__ __forceinline__ void copy16(void* dst, const void* src, int bytes) {  // whole CTA, 16-byte granules
  const uint4* s = reinterpret_cast<const uint4*>(src);
  uint4* d = reinterpret_cast<uint4*>(dst);
  for (int i = threadIdx.x; i < bytes / 16; i += blockDim.x) d[i] = s[i];
}

// instruction descriptor with both operands MN-major (InstrDescriptor bits 15 / 16: a_major, b_major)
__device__ __forceinline__ uint32_t umma_idesc_mn(int M, int N) { return umma_idesc(M, N) | (1u << 15) | (1u << 16); }

// D[128 x N] (+)= Mop^T . Nop over the 128 points of the tile.  Mop: [point][128 units], Nop: [point][N units], both stored with
// core_idx(point, unit, 128).  MN-major descriptor of such a buffer: 8 consecutive points = the 8 K-rows of a core matrix (16 B
// apart), the next 8 points 128 B further (LBO), the next 8 units 128 * 16 B further (SBO); one MMA consumes 16 points = 256 B.
__device__ __forceinline__ void issue_dw(uint32_t tmem_d, const __nv_bfloat16* M_hi, const __nv_bfloat16* M_lo,
                                         const __nv_bfloat16* N_hi, const __nv_bfloat16* N_lo, int N) {
  const uint32_t idesc = umma_idesc_mn(BW_H, N);
  const uint32_t lbo = 128, sbo = BW_H * 16;
  const uint32_t mh = smem_u32(M_hi), ml = smem_u32(M_lo), nh = smem_u32(N_hi), nl = smem_u32(N_lo);
  for (int ks = 0; ks < BW_H / 16; ks++) {
    const uint64_t dmh = umma_desc(mh + ks * 256, lbo, sbo), dml = umma_desc(ml + ks * 256, lbo, sbo);
    const uint64_t dnh = umma_desc(nh + ks * 256, lbo, sbo), dnl = umma_desc(nl + ks * 256, lbo, sbo);
    umma_bf16(tmem_d, dmh, dnh, idesc, ks > 0 ? 1u : 0u);
    umma_bf16(tmem_d, dml, dnh, idesc, 1u);
    umma_bf16(tmem_d, dmh, dnl, idesc, 1u);
  }
}

// D[128 x N] = G[128 x K] . Wt[N x K]^T, K-major (the forward's issue_layer without the commit)
__device__ __forceinline__ void issue_da(uint32_t tmem_d, const __nv_bfloat16* G_hi, const __nv_bfloat16* G_lo,
                                         const __nv_bfloat16* W_hi, const __nv_bfloat16* W_lo, int K, int N) {
  const uint32_t idesc = umma_idesc(BW_H, N);
  const uint32_t lbo_a = BW_H * 16, lbo_w = (uint32_t)N * 16, sbo = 128;
  const uint32_t gh = smem_u32(G_hi), gl = smem_u32(G_lo), wh = smem_u32(W_hi), wl = smem_u32(W_lo);
  for (int ks = 0; ks < K / 16; ks++) {
    const uint64_t dgh = umma_desc(gh + ks * 2 * lbo_a, lbo_a, sbo), dgl = umma_desc(gl + ks * 2 * lbo_a, lbo_a, sbo);
    const uint64_t dwh = umma_desc(wh + ks * 2 * lbo_w, lbo_w, sbo), dwl = umma_desc(wl + ks * 2 * lbo_w, lbo_w, sbo);
    umma_bf16(tmem_d, dgh, dwh, idesc, ks > 0 ? 1u : 0u);
    umma_bf16(tmem_d, dgl, dwh, idesc, 1u);
    umma_bf16(tmem_d, dgh, dwl, idesc, 1u);
  }
}

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Warp total of 32 per-lane values per column: lane L ends with sum over the warp's 32 rows of v[L] (31 shuffles).
__device__ __forceinline__ float warp_colsum32(float* v, int lane) {
#define BW_STEP(N, X)                                                           \
  _Pragma("unroll") for (int k = 0; k < (N) / 2; k++) {                         \
    const bool up = lane & (X);                                                 \
    const float send = up ? v[k] : v[k + (N) / 2];                              \
    const float keepv = up ? v[k + (N) / 2] : v[k];                             \
    v[k] = keepv + __shfl_xor_sync(0xFFFFFFFFu, send, (X));                     \
  }
  BW_STEP(32, 16) BW_STEP(16, 8) BW_STEP(8, 4) BW_STEP(4, 2) BW_STEP(2, 1)
#undef BW_STEP
  return v[0];
}

// Row `row` of a [128 x C] gradient chunk (32 columns starting at c0, registers g[32]) -> G operand (hi/lo) + bias-gradient sums.
__device__ __forceinline__ void store_grad32(BwSmem& sm, int row, int c0, float* g, float* dbias, int lane) {
#pragma unroll
  for (int j = 0; j < 32; j += 8) store_act8(sm.G_hi, sm.G_lo, row, c0 + j, g + j);
  const float s = warp_colsum32(g, lane);  // destroys g
  atomicAdd(&dbias[c0 + lane], s);
}

// sign of the saved activation (row, col) -> LeakyReLU'(z): 1 for z > 0, 0.01 otherwise (a > 0 <=> z > 0)
__device__ __forceinline__ void lrelu_mask8(const BwSmem& sm, int row, int k0, float* g) {
  const uint4 h = *reinterpret_cast<const uint4*>(sm.A_hi + core_idx(row, k0, BW_H));
  const uint32_t w[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const uint32_t bits = (w[i >> 1] >> ((i & 1) * 16)) & 0xFFFFu;           // bf16 bits of element i
    const bool pos = !(bits & 0x8000u) && (bits & 0x7FFFu) != 0;             // strictly positive
    g[i] = pos ? g[i] : 0.01f * g[i];
  }
}

__device__ __forceinline__ void bw_sync_mma_inputs() {  // generic-proxy smem writes -> visible to the tensor core, CTA-wide
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// One thread: commit the MMAs issued so far; everyone: wait for them (`phase` counts the commits, kept per thread).
__device__ __forceinline__ void bw_commit_wait(BwSmem& sm, bool issuer, uint32_t& phase) {
  if (issuer) umma_commit(&sm.bar);
  nr_mbar_wait(&sm.bar, phase & 1u);
  phase++;
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// dW partial [128 rows (this thread = row) x ncols] in TMEM -> global, row-major destination with `ld` floats per row
__device__ __forceinline__ void flush_dw_rows(uint32_t tmem_row, int col_begin, int ncols, float* dst_row) {
  for (int c = col_begin; c < ncols; c += 32) {
    float v[32];
    tmem_ld32(tmem_row + c, v);
#pragma unroll
    for (int j = 0; j < 32; j += 4)
      if (c + j < ncols) red_add_v4(dst_row + c + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
  }
}

}  // namespace
