// gs_nonrigid.cu -- non-rigid deformation of the canonical Gaussians: multiresolution hash-grid encoding + conditional
// MLP + delta application (replaces HashGridwithMLP.forward, models/deformer/non_rigid.py:226-286; see
// include/gsplat_b200.h for the reference lines of every piece).
//
// The MLP is the only dense contraction of the hot path and the only place tensor cores are used: one CTA owns a tile
// of 128 Gaussians (= the M dimension of a 128xNx16 tcgen05.mma), activations and weights sit in shared memory in the
// canonical K-major no-swizzle core-matrix layout, accumulators live in TMEM, one elected thread issues the MMAs, and
// the epilogue (tcgen05.ld -> bias -> LeakyReLU -> BF16 hi/lo split) writes the next layer's A operand straight back
// into shared memory.  FP32 fidelity comes from a 2-way BF16 split of both operands (hi*hi + lo*hi + hi*lo).
#include "gs_umma.cuh"

namespace {

constexpr int NR_H = 128;      // hidden width
constexpr int NR_F = 32;       // hash-grid features (16 levels x 2), K of layer 0

struct NrWs {  // prepared once per call by nr_prep_kernel
  __nv_bfloat16 *w0_hi, *w0_lo;  // [128 x 32]
  __nv_bfloat16 *w1_hi, *w1_lo;  // [128 x 128]
  __nv_bfloat16 *w2_hi, *w2_lo;
  __nv_bfloat16 *w3_hi, *w3_lo;  // [NOUT_PAD x 128]
  float* bias0;                  // [128] b0 + W0[:, F:] . cond
};

__host__ inline int nout_pad(int n_out) { return n_out <= 16 ? 16 : 32; }

__host__ size_t carve_ws(char* base, int n_out, NrWs* w) {
  char* p = base;
  NrWs t;
  const size_t np = (size_t)nout_pad(n_out);
  gs_carve(p, t.w0_hi, (size_t)NR_H * NR_F); gs_carve(p, t.w0_lo, (size_t)NR_H * NR_F);
  gs_carve(p, t.w1_hi, (size_t)NR_H * NR_H); gs_carve(p, t.w1_lo, (size_t)NR_H * NR_H);
  gs_carve(p, t.w2_hi, (size_t)NR_H * NR_H); gs_carve(p, t.w2_lo, (size_t)NR_H * NR_H);
  gs_carve(p, t.w3_hi, np * NR_H); gs_carve(p, t.w3_lo, np * NR_H);
  gs_carve(p, t.bias0, (size_t)NR_H);
  if (w) *w = t;
  return (size_t)(p - base);
}

// Weights -> BF16 hi/lo in the operand layout; condition folded into the first bias.
__global__ void nr_prep_kernel(NonRigidArgs a, NrWs w, int npad) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  const int in0 = NR_F + a.n_cond;
  for (int e = tid; e < NR_H * NR_F; e += nth) {
    const int n = e / NR_F, k = e % NR_F;
    split_bf16(a.W0[(size_t)n * in0 + k], w.w0_hi[core_idx(n, k, NR_H)], w.w0_lo[core_idx(n, k, NR_H)]);
  }
  for (int e = tid; e < NR_H * NR_H; e += nth) {
    const int n = e / NR_H, k = e % NR_H;
    split_bf16(a.W1[e], w.w1_hi[core_idx(n, k, NR_H)], w.w1_lo[core_idx(n, k, NR_H)]);
    split_bf16(a.W2[e], w.w2_hi[core_idx(n, k, NR_H)], w.w2_lo[core_idx(n, k, NR_H)]);
  }
  for (int e = tid; e < npad * NR_H; e += nth) {
    const int n = e / NR_H, k = e % NR_H;
    const float v = n < a.n_out ? a.W3[(size_t)n * NR_H + k] : 0.f;
    split_bf16(v, w.w3_hi[core_idx(n, k, npad)], w.w3_lo[core_idx(n, k, npad)]);
  }
  for (int n = tid; n < NR_H; n += nth) {
    float acc = a.b0[n];
    for (int c = 0; c < a.n_cond; c++) acc = fmaf(a.W0[(size_t)n * in0 + NR_F + c], a.cond[c], acc);
    w.bias0[n] = acc;
  }
}

// ---- multiresolution hash encoding (Mueller et al. 2022, as implemented by tiny-cuda-nn's GridEncoding) ----
__device__ __forceinline__ uint32_t grid_index(uint32_t x, uint32_t y, uint32_t z, uint32_t res, uint32_t size) {
  uint32_t stride = 1, index = 0;
  const uint32_t pos[3] = {x, y, z};
#pragma unroll
  for (int d = 0; d < 3; d++) {
    if (stride <= size) { index += pos[d] * stride; stride *= res; }
  }
  if (size < stride) index = x ^ (y * 2654435761u) ^ (z * 805459861u);
  return index % size;
}

__device__ __forceinline__ void hash_encode(const NonRigidArgs& a, float x, float y, float z, float* feat /*[2L]*/) {
  for (int l = 0; l < a.n_levels; l++) {
    const float scale = a.level_scale[l];
    const uint32_t res = a.level_res[l];
    const uint32_t off = a.level_offset[l], size = a.level_offset[l + 1] - off;
    const float px = fmaf(x, scale, 0.5f), py = fmaf(y, scale, 0.5f), pz = fmaf(z, scale, 0.5f);
    const float fx0 = floorf(px), fy0 = floorf(py), fz0 = floorf(pz);
    const float wx = px - fx0, wy = py - fy0, wz = pz - fz0;
    const uint32_t gx = (uint32_t)(int)fx0, gy = (uint32_t)(int)fy0, gz = (uint32_t)(int)fz0;
    float f0 = 0.f, f1 = 0.f;
#pragma unroll
    for (int c = 0; c < 8; c++) {
      const uint32_t dx = c & 1, dy = (c >> 1) & 1, dz = (c >> 2) & 1;
      const float wgt = (dx ? wx : 1.f - wx) * (dy ? wy : 1.f - wy) * (dz ? wz : 1.f - wz);
      const uint32_t idx = grid_index(gx + dx, gy + dy, gz + dz, res, size);
      const float2 v = __ldg(reinterpret_cast<const float2*>(a.table) + off + idx);
      f0 = fmaf(wgt, v.x, f0);
      f1 = fmaf(wgt, v.y, f1);
    }
    feat[2 * l] = f0;
    feat[2 * l + 1] = f1;
  }
}

struct NrSmem {
  __nv_bfloat16 A_hi[NR_TILE * NR_H], A_lo[NR_TILE * NR_H];  // activations, operand layout (32 KB each)
  __nv_bfloat16 W_hi[NR_H * NR_H], W_lo[NR_H * NR_H];        // weights of the current layer
  float bias[NR_H];
  unsigned long long bar;
  uint32_t tmem_base;
};

// bias + LeakyReLU(0.01) epilogue of a hidden layer: TMEM accumulators of this thread's row -> next layer's A operand
__device__ __forceinline__ void hidden_epilogue(NrSmem& sm, uint32_t tmem_row, int row) {
#pragma unroll 1
  for (int c = 0; c < NR_H / 32; c++) {
    float v[32];
    tmem_ld32(tmem_row + c * 32, v);
#pragma unroll
    for (int j = 0; j < 32; j++) {
      const float t = v[j] + sm.bias[c * 32 + j];
      v[j] = t > 0.f ? t : 0.01f * t;  // nn.LeakyReLU() default slope
    }
#pragma unroll
    for (int j = 0; j < 32; j += 8) store_act8(sm.A_hi, sm.A_lo, row, c * 32 + j, v + j);
  }
}

__global__ void __launch_bounds__(NR_TILE, 1) nr_fwd_kernel(NonRigidArgs a, NrWs w, int npad) {
  extern __shared__ __align__(1024) unsigned char nr_smem_raw[];
  NrSmem& sm = *reinterpret_cast<NrSmem*>(nr_smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5;
  const int row = tid;                         // one Gaussian per thread = one row of the 128-row MMA tile
  const int i = blockIdx.x * NR_TILE + row;
  const bool live = i < a.P;

  if (warp == 0) {  // TMEM: 128 columns x 128 lanes of FP32 accumulators
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    nr_mbar_init(&sm.bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // ---- layer-0 operands: hash-grid features of this thread's Gaussian, W0, folded bias ----
  float x = 0.f, y = 0.f, z = 0.f;
  {
    float feat[NR_F];
#pragma unroll
    for (int k = 0; k < NR_F; k++) feat[k] = 0.f;
    if (live) {
      x = a.xyz[3 * (size_t)i]; y = a.xyz[3 * (size_t)i + 1]; z = a.xyz[3 * (size_t)i + 2];
      // AABB.normalize(sym=True) then HashGrid's (x+1)/2 (dataset_utils.py:66-71, network_utils.py:341-343)
      const float nx = 2.f * ((x - a.aabb_min[0]) / (a.aabb_max[0] - a.aabb_min[0])) - 1.f;
      const float ny = 2.f * ((y - a.aabb_min[1]) / (a.aabb_max[1] - a.aabb_min[1])) - 1.f;
      const float nz = 2.f * ((z - a.aabb_min[2]) / (a.aabb_max[2] - a.aabb_min[2])) - 1.f;
      hash_encode(a, (nx + 1.f) * 0.5f, (ny + 1.f) * 0.5f, (nz + 1.f) * 0.5f, feat);
      if (a.features)
        for (int k = 0; k < 2 * a.n_levels; k++) a.features[(size_t)i * 2 * a.n_levels + k] = feat[k];
    }
#pragma unroll
    for (int k = 0; k < NR_F; k += 8) store_act8(sm.A_hi, sm.A_lo, row, k, feat + k);
  }
  copy_weights(sm.W_hi, w.w0_hi, NR_H * NR_F);
  copy_weights(sm.W_lo, w.w0_lo, NR_H * NR_F);
  sm.bias[tid] = w.bias0[tid];
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> visible to the tensor core
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tmem_row = tmem + ((uint32_t)(warp * 32) << 16);  // this warp's 32-lane quadrant

  const __nv_bfloat16* next_hi[3] = {w.w1_hi, w.w2_hi, w.w3_hi};
  const __nv_bfloat16* next_lo[3] = {w.w1_lo, w.w2_lo, w.w3_lo};
  const float* next_b[3] = {a.b1, a.b2, a.b3};
#pragma unroll 1
  for (int layer = 0; layer < 3; layer++) {  // three hidden layers (N = 128); K = 32 for the first
    if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W_hi, sm.W_lo, layer == 0 ? NR_F : NR_H, NR_H, &sm.bar);
    nr_mbar_wait(&sm.bar, (uint32_t)layer & 1u);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    hidden_epilogue(sm, tmem_row, row);  // A and W are free: the MMAs that read them have completed
    const int nrows = layer == 2 ? npad : NR_H;
    copy_weights(sm.W_hi, next_hi[layer], nrows * NR_H);
    copy_weights(sm.W_lo, next_lo[layer], nrows * NR_H);
    __syncthreads();  // everyone has read the old bias
    if (layer < 2) sm.bias[tid] = next_b[layer][tid];
    else sm.bias[tid] = tid < a.n_out ? a.b3[tid] : 0.f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  // ---- output layer: N = npad (16 or 32) ----
  if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W_hi, sm.W_lo, NR_H, npad, &sm.bar);
  nr_mbar_wait(&sm.bar, 1u);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  float d[32];
  tmem_ld32(tmem_row, d);  // columns >= npad hold stale values of the previous layer: never used
  if (live) {
#pragma unroll
    for (int k = 0; k < 32; k++) d[k] += sm.bias[k];
    if (a.deltas)
      for (int k = 0; k < a.n_out; k++) a.deltas[(size_t)i * a.n_out + k] = d[k];
    a.xyz_out[3 * (size_t)i] = x + d[0]; a.xyz_out[3 * (size_t)i + 1] = y + d[1]; a.xyz_out[3 * (size_t)i + 2] = z + d[2];
#pragma unroll
    for (int k = 0; k < 3; k++) a.scaling_out[3 * (size_t)i + k] = a.scaling[3 * (size_t)i + k] + d[3 + k];  // scale_offset: logit
    // rot_offset: mult -- q1 = (1, d7, d8, d9) (non_rigid.py:276-277 overwrites the real part), Hamilton product q1 * q2
    const float4 q2 = *reinterpret_cast<const float4*>(a.rotation + 4 * (size_t)i);
    const float aw = 1.f, ax = d[7], ay = d[8], az = d[9];
    float4 o;
    o.x = aw * q2.x - ax * q2.y - ay * q2.z - az * q2.w;
    o.y = aw * q2.y + ax * q2.x + ay * q2.w - az * q2.z;
    o.z = aw * q2.z - ax * q2.w + ay * q2.x + az * q2.y;
    o.w = aw * q2.w + ax * q2.z - ay * q2.y + az * q2.x;
    *reinterpret_cast<float4*>(a.rotation_out + 4 * (size_t)i) = o;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
}

}  // namespace

static int nr_args_ok(const NonRigidArgs& a) {
  if (a.P < 0 || a.n_feat != 2 || a.hidden != NR_H || a.n_levels * a.n_feat != NR_F) return 0;
  if (a.n_out < 10 || a.n_out > 32 || a.n_cond < 0) return 0;
  if (!a.aabb_min || !a.aabb_max || !a.xyz || !a.scaling || !a.rotation || !a.table) return 0;
  if ((a.n_cond > 0 && !a.cond) || !a.W0 || !a.b0 || !a.W1 || !a.b1 || !a.W2 || !a.b2 || !a.W3 || !a.b3) return 0;
  if (!a.xyz_out || !a.scaling_out || !a.rotation_out) return 0;
  return 1;
}

extern "C" size_t nonrigid_workspace_bytes(const NonRigidArgs* args) {
  return carve_ws(nullptr, args ? args->n_out : 10, nullptr);
}

extern "C" int nonrigid_forward(const NonRigidArgs* args, void* workspace, size_t workspace_bytes, void* stream) {
  if (!args || !workspace || !nr_args_ok(*args)) return GS_ERR_BAD_ARG;
  const NonRigidArgs& a = *args;
  if (a.P == 0) return GS_OK;
  cudaStream_t st = (cudaStream_t)stream;
  NrWs w;
  if (carve_ws((char*)workspace, a.n_out, &w) > workspace_bytes) return GS_ERR_WORKSPACE_SMALL;
  const int npad = nout_pad(a.n_out);
  nr_prep_kernel<<<64, 256, 0, st>>>(a, w, npad);
  static bool configured = false;
  if (!configured) {
    if (gs_check(cudaFuncSetAttribute(nr_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NrSmem) + 1024)) != GS_OK)
      return GS_ERR_CUDA;
    configured = true;
  }
  nr_fwd_kernel<<<(a.P + NR_TILE - 1) / NR_TILE, NR_TILE, sizeof(NrSmem) + 1024, st>>>(a, w, npad);
  return gs_check(cudaGetLastError());
}

// ================================= skinning field (SkinningField, rigid.py:160-171) =================================
namespace {

constexpr int SK_K0 = 16;   // 3 input coordinates padded to one K = 16 MMA step
constexpr int SK_NOUT = 32; // 25 (or 24) logits padded to N = 32

struct SkWs {
  __nv_bfloat16 *w0_hi, *w0_lo;     // [128 x 16]
  __nv_bfloat16 *wh_hi[3], *wh_lo[3];  // lin1..lin3 [128 x 128]
  __nv_bfloat16 *w4_hi, *w4_lo;     // [32 x 128]
};

__host__ size_t carve_sk(char* base, SkWs* w) {
  char* p = base;
  SkWs t;
  gs_carve(p, t.w0_hi, (size_t)NR_H * SK_K0); gs_carve(p, t.w0_lo, (size_t)NR_H * SK_K0);
  for (int i = 0; i < 3; i++) { gs_carve(p, t.wh_hi[i], (size_t)NR_H * NR_H); gs_carve(p, t.wh_lo[i], (size_t)NR_H * NR_H); }
  gs_carve(p, t.w4_hi, (size_t)SK_NOUT * NR_H); gs_carve(p, t.w4_lo, (size_t)SK_NOUT * NR_H);
  if (w) *w = t;
  return (size_t)(p - base);
}


__global__ void sk_prep_kernel_impl(SkinArgs a, SkWs w) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  for (int e = tid; e < NR_H * SK_K0; e += nth) {
    const int n = e / SK_K0, k = e % SK_K0;
    const float v = k < 3 ? a.W[0][n * 3 + k] : 0.f;
    split_bf16(v, w.w0_hi[core_idx(n, k, NR_H)], w.w0_lo[core_idx(n, k, NR_H)]);
  }
  for (int l = 0; l < 3; l++)
    for (int e = tid; e < NR_H * NR_H; e += nth) {
      const int n = e / NR_H, k = e % NR_H;
      split_bf16(a.W[1 + l][e], w.wh_hi[l][core_idx(n, k, NR_H)], w.wh_lo[l][core_idx(n, k, NR_H)]);
    }
  for (int e = tid; e < SK_NOUT * NR_H; e += nth) {
    const int n = e / NR_H, k = e % NR_H;
    const float v = n < a.n_out ? a.W[4][(size_t)n * NR_H + k] : 0.f;
    split_bf16(v, w.w4_hi[core_idx(n, k, SK_NOUT)], w.w4_lo[core_idx(n, k, SK_NOUT)]);
  }
}

__device__ __forceinline__ float sigmoidf_(float v) { return 1.0f / (1.0f + expf(-v)); }

// rigid.py:85-129, statement by statement (x = 25 logits, p = 24 joint probabilities)
__device__ __forceinline__ void hierarchical_softmax(const float* x, float* p) {
  float s[25];
#pragma unroll
  for (int k = 0; k < 25; k++) s[k] = sigmoidf_(x[k]);
  auto softmax3 = [](float a, float b, float c, float* o) {
    const float m = fmaxf(a, fmaxf(b, c));
    const float ea = expf(a - m), eb = expf(b - m), ec = expf(c - m), inv = 1.0f / (ea + eb + ec);
    o[0] = ea * inv; o[1] = eb * inv; o[2] = ec * inv;
  };
  float sm[3];
  softmax3(x[1], x[2], x[3], sm);
  for (int k = 0; k < 3; k++) p[1 + k] = s[0] * sm[k];
  p[0] = 1.f - s[0];
  for (int k = 0; k < 3; k++) { p[4 + k] = p[1 + k] * s[4 + k]; p[1 + k] = p[1 + k] * (1.f - s[4 + k]); }
  for (int k = 0; k < 3; k++) { p[7 + k] = p[4 + k] * s[7 + k]; p[4 + k] = p[4 + k] * (1.f - s[7 + k]); }
  for (int k = 0; k < 2; k++) { p[10 + k] = p[7 + k] * s[10 + k]; p[7 + k] = p[7 + k] * (1.f - s[10 + k]); }
  softmax3(x[12], x[13], x[14], sm);
  for (int k = 0; k < 3; k++) p[12 + k] = p[9] * s[24] * sm[k];
  p[9] = p[9] * (1.f - s[24]);
  p[15] = p[12] * s[15]; p[12] = p[12] * (1.f - s[15]);
  for (int k = 0; k < 2; k++) { p[16 + k] = p[13 + k] * s[16 + k]; p[13 + k] = p[13 + k] * (1.f - s[16 + k]); }
  for (int k = 0; k < 2; k++) { p[18 + k] = p[16 + k] * s[18 + k]; p[16 + k] = p[16 + k] * (1.f - s[18 + k]); }
  for (int k = 0; k < 2; k++) { p[20 + k] = p[18 + k] * s[20 + k]; p[18 + k] = p[18 + k] * (1.f - s[20 + k]); }
  for (int k = 0; k < 2; k++) { p[22 + k] = p[20 + k] * s[22 + k]; p[20 + k] = p[20 + k] * (1.f - s[22 + k]); }
}

__global__ void __launch_bounds__(NR_TILE, 1) sk_fwd_kernel(SkinArgs a, SkWs w) {
  extern __shared__ __align__(1024) unsigned char nr_smem_raw[];
  NrSmem& sm = *reinterpret_cast<NrSmem*>(nr_smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, row = tid;
  const int i = blockIdx.x * NR_TILE + row;
  const bool live = i < a.P;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    nr_mbar_init(&sm.bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  {
    float in[SK_K0];
#pragma unroll
    for (int k = 0; k < SK_K0; k++) in[k] = 0.f;
    if (live) {
#pragma unroll
      for (int k = 0; k < 3; k++)  // AABB.normalize(sym=True), dataset_utils.py:66-71
        in[k] = 2.f * ((a.xyz[3 * (size_t)i + k] - a.aabb_min[k]) / (a.aabb_max[k] - a.aabb_min[k])) - 1.f;
    }
    store_act8(sm.A_hi, sm.A_lo, row, 0, in);
    store_act8(sm.A_hi, sm.A_lo, row, 8, in + 8);
  }
  copy_weights(sm.W_hi, w.w0_hi, NR_H * SK_K0);
  copy_weights(sm.W_lo, w.w0_lo, NR_H * SK_K0);
  sm.bias[tid] = a.b[0][tid];
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tmem_row = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
  for (int layer = 0; layer < 4; layer++) {  // four hidden layers
    if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W_hi, sm.W_lo, layer == 0 ? SK_K0 : NR_H, NR_H, &sm.bar);
    nr_mbar_wait(&sm.bar, (uint32_t)layer & 1u);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    hidden_epilogue(sm, tmem_row, row);
    if (layer < 3) {
      copy_weights(sm.W_hi, w.wh_hi[layer], NR_H * NR_H);
      copy_weights(sm.W_lo, w.wh_lo[layer], NR_H * NR_H);
    } else {
      copy_weights(sm.W_hi, w.w4_hi, SK_NOUT * NR_H);
      copy_weights(sm.W_lo, w.w4_lo, SK_NOUT * NR_H);
    }
    __syncthreads();
    if (layer < 3) sm.bias[tid] = a.b[layer + 1][tid];
    else sm.bias[tid] = tid < a.n_out ? a.b[4][tid] : 0.f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W_hi, sm.W_lo, NR_H, SK_NOUT, &sm.bar);
  nr_mbar_wait(&sm.bar, 0u);  // fifth commit
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  float x[32];
  tmem_ld32(tmem_row, x);
  if (live) {
#pragma unroll
    for (int k = 0; k < 32; k++) x[k] += sm.bias[k];
    if (a.logits)
      for (int k = 0; k < a.n_out; k++) a.logits[(size_t)i * a.n_out + k] = x[k];
    float p[24];
    if (a.n_out == 25) {
      hierarchical_softmax(x, p);
    } else {  // F.softmax over 24 logits
      float m = x[0];
#pragma unroll
      for (int k = 1; k < 24; k++) m = fmaxf(m, x[k]);
      float sum = 0.f;
#pragma unroll
      for (int k = 0; k < 24; k++) { p[k] = expf(x[k] - m); sum += p[k]; }
      const float inv = 1.0f / sum;
#pragma unroll
      for (int k = 0; k < 24; k++) p[k] *= inv;
    }
    float4* o = reinterpret_cast<float4*>(a.weights + (size_t)i * 24);
#pragma unroll
    for (int k = 0; k < 6; k++) o[k] = make_float4(p[4 * k], p[4 * k + 1], p[4 * k + 2], p[4 * k + 3]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
}

}  // namespace

extern "C" size_t skinning_workspace_bytes(const SkinArgs* args) { (void)args; return carve_sk(nullptr, nullptr); }

extern "C" int skinning_forward(const SkinArgs* args, void* workspace, size_t workspace_bytes, void* stream) {
  if (!args || !workspace) return GS_ERR_BAD_ARG;
  const SkinArgs& a = *args;
  if (a.P < 0 || a.hidden != NR_H || (a.n_out != 25 && a.n_out != 24) || !a.aabb_min || !a.aabb_max || !a.xyz || !a.weights)
    return GS_ERR_BAD_ARG;
  for (int l = 0; l < 5; l++) if (!a.W[l] || !a.b[l]) return GS_ERR_BAD_ARG;
  if (a.P == 0) return GS_OK;
  cudaStream_t st = (cudaStream_t)stream;
  SkWs w;
  if (carve_sk((char*)workspace, &w) > workspace_bytes) return GS_ERR_WORKSPACE_SMALL;
  sk_prep_kernel_impl<<<64, 256, 0, st>>>(a, w);
  static bool configured = false;
  if (!configured) {
    if (gs_check(cudaFuncSetAttribute(sk_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NrSmem) + 1024)) != GS_OK)
      return GS_ERR_CUDA;
    configured = true;
  }
  sk_fwd_kernel<<<(a.P + NR_TILE - 1) / NR_TILE, NR_TILE, sizeof(NrSmem) + 1024, st>>>(a, w);
  return gs_check(cudaGetLastError());
}
