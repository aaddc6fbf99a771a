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
#include "gs_mlp_bwd.cuh"

#define GS_TRY_NR(expr) do { if (gs_check(expr) != GS_OK) return GS_ERR_CUDA; } while (0)

namespace {

constexpr int NR_H = 128;      // hidden width
constexpr int NR_F = 32;       // hash-grid features (16 levels x 2), K of layer 0

struct NrWs {  // prepared once per call by nr_prep_kernel
  __nv_bfloat16 *w0_hi, *w0_lo;  // [128 x 32]
  __nv_bfloat16 *w1_hi, *w1_lo;  // [128 x 128]
  __nv_bfloat16 *w2_hi, *w2_lo;
  __nv_bfloat16 *w3_hi, *w3_lo;  // [NOUT_PAD x 128]
  float* bias0;                  // [128] b0 + W0[:, F:] . cond
  // backward: W^T as the B operand of dA = G . W, i.e. element (row = k_in, col = n_out) = W[n_out][k_in]
  __nv_bfloat16 *w0t_hi, *w0t_lo;  // [32 x 128]
  __nv_bfloat16 *w1t_hi, *w1t_lo;  // [128 x 128]
  __nv_bfloat16 *w2t_hi, *w2t_lo;
  __nv_bfloat16 *w3t_hi, *w3t_lo;  // [128 x NOUT_PAD]
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
  gs_carve(p, t.w0t_hi, (size_t)NR_F * NR_H); gs_carve(p, t.w0t_lo, (size_t)NR_F * NR_H);
  gs_carve(p, t.w1t_hi, (size_t)NR_H * NR_H); gs_carve(p, t.w1t_lo, (size_t)NR_H * NR_H);
  gs_carve(p, t.w2t_hi, (size_t)NR_H * NR_H); gs_carve(p, t.w2t_lo, (size_t)NR_H * NR_H);
  gs_carve(p, t.w3t_hi, (size_t)NR_H * np); gs_carve(p, t.w3t_lo, (size_t)NR_H * np);
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
  // bias0[n] = b0[n] + W0[n, F:] . cond -- one WARP per output (coalesced row reads, shuffle reduction); one thread per output
  // walked 144 strided loads in a row and was most of this kernel's 15 us
  const int lane = threadIdx.x & 31;
  for (int n = tid >> 5; n < NR_H; n += nth >> 5) {
    float acc = 0.f;
    for (int c = lane; c < a.n_cond; c += 32) acc = fmaf(a.W0[(size_t)n * in0 + NR_F + c], a.cond[c], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
    if (lane == 0) w.bias0[n] = a.b0[n] + acc;
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

// Levels are processed four at a time: the 32 table reads of a group are issued before the first one is used.  With one warp per
// scheduler nothing else hides the latency, and level by level the encode was 16 dependent L2 round trips per Gaussian.
constexpr int HG_GROUP = 4;
constexpr int HG_LEVELS = NR_F / 2;  // n_levels is fixed by the MLP's input width (nr_inputs_ok): loops unroll, features stay in registers
template <int L_BEGIN, int L_END>  // levels [L_BEGIN, L_END): the two threads of a Gaussian take half of the levels each
__device__ __forceinline__ void hash_encode(const NonRigidArgs& a, float x, float y, float z, float* feat /*[2L]*/) {
#pragma unroll
  for (int l0 = L_BEGIN; l0 < L_END; l0 += HG_GROUP) {
    float2 v[HG_GROUP][8];
    float wx[HG_GROUP], wy[HG_GROUP], wz[HG_GROUP];
#pragma unroll
    for (int q = 0; q < HG_GROUP; q++) {
      const int l = l0 + q;
      const float scale = a.level_scale[l];
      const uint32_t res = a.level_res[l];
      const uint32_t off = a.level_offset[l], size = a.level_offset[l + 1] - off;
      const float px = fmaf(x, scale, 0.5f), py = fmaf(y, scale, 0.5f), pz = fmaf(z, scale, 0.5f);
      const float fx0 = floorf(px), fy0 = floorf(py), fz0 = floorf(pz);
      wx[q] = px - fx0; wy[q] = py - fy0; wz[q] = pz - fz0;
      const uint32_t gx = (uint32_t)(int)fx0, gy = (uint32_t)(int)fy0, gz = (uint32_t)(int)fz0;
#pragma unroll
      for (int c = 0; c < 8; c++) {
        const uint32_t idx = grid_index(gx + (c & 1), gy + ((c >> 1) & 1), gz + ((c >> 2) & 1), res, size);
        v[q][c] = __ldg(reinterpret_cast<const float2*>(a.table) + off + idx);
      }
    }
#pragma unroll
    for (int q = 0; q < HG_GROUP; q++) {
      float f0 = 0.f, f1 = 0.f;
#pragma unroll
      for (int c = 0; c < 8; c++) {  // same corner order and FMA chain as before: bit-identical features
        const uint32_t dx = c & 1, dy = (c >> 1) & 1, dz = (c >> 2) & 1;
        const float wgt = (dx ? wx[q] : 1.f - wx[q]) * (dy ? wy[q] : 1.f - wy[q]) * (dz ? wz[q] : 1.f - wz[q]);
        f0 = fmaf(wgt, v[q][c].x, f0);
        f1 = fmaf(wgt, v[q][c].y, f1);
      }
      feat[2 * (l0 + q)] = f0;
      feat[2 * (l0 + q) + 1] = f1;
    }
  }
}

// ---- activations kept for the backward (training): per tile of 128 Gaussians, already in the BF16 hi/lo operand layout the
// tensor cores read, so that the backward kernels load them with plain 16-byte copies.  Non-rigid: input features
// [128 x 32] hi|lo, then the three hidden activations [128 x 128] hi|lo.  Skinning: input [128 x 16] hi|lo, four hidden. ----
constexpr size_t NR_SAVE_IN = 2 * (size_t)128 * 32, NR_SAVE_LAYER = 2 * (size_t)128 * 128;
constexpr size_t NR_SAVE_TILE = NR_SAVE_IN + 3 * NR_SAVE_LAYER;
constexpr size_t SK_SAVE_IN = 2 * (size_t)128 * 16;
constexpr size_t SK_SAVE_TILE = SK_SAVE_IN + 4 * NR_SAVE_LAYER;

struct NrSmem {
  __nv_bfloat16 A_hi[NR_TILE * NR_H], A_lo[NR_TILE * NR_H];  // activations, operand layout (32 KB each)
  __nv_bfloat16 W_hi[NR_H * NR_H], W_lo[NR_H * NR_H];        // weights of the current layer
  float bias[NR_H];
  unsigned long long bar;     // MMA completion
  unsigned long long ld_bar;  // weight loads (cp.async.bulk complete_tx)
  uint32_t tmem_base;
};

// bias + LeakyReLU(0.01) epilogue of a hidden layer: TMEM accumulators of this thread's row -> next layer's A operand
__device__ __forceinline__ void hidden_epilogue(NrSmem& sm, uint32_t tmem_row, int row, int half) {
#pragma unroll 1
  for (int c = 2 * half; c < 2 * half + 2; c++) {  // this thread's 64 of the 128 columns (two threads per row)
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

__global__ void __launch_bounds__(BW_THREADS, 1) nr_fwd_kernel(NonRigidArgs a, NrWs w, int npad, __nv_bfloat16* __restrict__ saved) {
  extern __shared__ __align__(1024) unsigned char nr_smem_raw[];
  NrSmem& sm = *reinterpret_cast<NrSmem*>(nr_smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5;
  // one Gaussian = one row of the 128-row MMA tile, TWO threads per row (warps w and w + 4 share a TMEM lane quadrant): they
  // split the hash-grid levels and the columns of every epilogue (gs_mlp_bwd.cuh: BW_THREADS)
  const int row = tid & (NR_TILE - 1), half = tid >> 7;
  const int i = blockIdx.x * NR_TILE + row;
  const bool live = i < a.P;

  if (warp == 0) {  // TMEM: 128 columns x 128 lanes of FP32 accumulators
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    nr_mbar_init(&sm.bar, 1);
    nr_mbar_init(&sm.ld_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // W0 streams in (one bulk copy per half) while every thread encodes its Gaussian
    bulk_load_operands(&sm.ld_bar, nullptr, nullptr, nullptr, nullptr, 0u, sm.W_hi, sm.W_lo, w.w0_hi, w.w0_lo, NR_H * NR_F * 2);
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
      if (half == 0) hash_encode<0, HG_LEVELS / 2>(a, (nx + 1.f) * 0.5f, (ny + 1.f) * 0.5f, (nz + 1.f) * 0.5f, feat);
      else hash_encode<HG_LEVELS / 2, HG_LEVELS>(a, (nx + 1.f) * 0.5f, (ny + 1.f) * 0.5f, (nz + 1.f) * 0.5f, feat);
      if (a.features) {
#pragma unroll
        for (int k = 0; k < NR_F; k++)
          if ((k >= NR_F / 2) == (half == 1)) a.features[(size_t)i * NR_F + k] = feat[k];
      }
    }
#pragma unroll
    for (int k = 0; k < NR_F; k += 8)
      if ((k >= NR_F / 2) == (half == 1)) store_act8(sm.A_hi, sm.A_lo, row, k, feat + k);
  }
  if (tid < NR_H) sm.bias[tid] = w.bias0[tid];
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> visible to the tensor core
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  nr_mbar_wait(&sm.ld_bar, 0u);  // W0 has landed (after the barrier: the mbarrier was initialised by thread 0)
  __nv_bfloat16* const save_tile = saved ? saved + (size_t)blockIdx.x * NR_SAVE_TILE : nullptr;
  if (save_tile) {  // the first 32 K columns (4 K-slabs of 128 rows x 16 B) of A_hi / A_lo are the layer-0 operand
    copy16(save_tile, sm.A_hi, 128 * 32 * 2);
    copy16(save_tile + 128 * 32, sm.A_lo, 128 * 32 * 2);
  }
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tmem_row = tmem + ((uint32_t)((warp & 3) * 32) << 16);  // this warp's 32-lane quadrant

  const __nv_bfloat16* next_hi[3] = {w.w1_hi, w.w2_hi, w.w3_hi};
  const __nv_bfloat16* next_lo[3] = {w.w1_lo, w.w2_lo, w.w3_lo};
  const float* next_b[3] = {a.b1, a.b2, a.b3};
#pragma unroll 1
  for (int layer = 0; layer < 3; layer++) {  // three hidden layers (N = 128); K = 32 for the first
    if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W_hi, sm.W_lo, layer == 0 ? NR_F : NR_H, NR_H, &sm.bar);
    nr_mbar_wait(&sm.bar, (uint32_t)layer & 1u);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // A and W are free: the MMAs that read them have completed.  The next layer's weights stream in during the epilogue.
    if (tid == 0)
      bulk_load_operands(&sm.ld_bar, nullptr, nullptr, nullptr, nullptr, 0u, sm.W_hi, sm.W_lo, next_hi[layer], next_lo[layer],
                         (uint32_t)((layer == 2 ? npad : NR_H) * NR_H * 2));
    hidden_epilogue(sm, tmem_row, row, half);
    __syncthreads();  // everyone has read the old bias (and written its part of the new activations)
    if (save_tile) {
      copy16(save_tile + NR_SAVE_IN + (size_t)layer * NR_SAVE_LAYER, sm.A_hi, 128 * 128 * 2);
      copy16(save_tile + NR_SAVE_IN + (size_t)layer * NR_SAVE_LAYER + 128 * 128, sm.A_lo, 128 * 128 * 2);
    }
    if (tid < NR_H) {
      if (layer < 2) sm.bias[tid] = next_b[layer][tid];
      else sm.bias[tid] = tid < a.n_out ? a.b3[tid] : 0.f;
    }
    nr_mbar_wait(&sm.ld_bar, (uint32_t)(layer + 1) & 1u);
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
  if (half == 0) tmem_ld32(tmem_row, d);  // (warp-uniform) columns >= npad hold stale values of the previous layer: never used
  if (live && half == 0) {
#pragma unroll
    for (int k = 0; k < 32; k++) d[k] += sm.bias[k];
    if (a.deltas) {
#pragma unroll
      for (int k = 0; k < 32; k++)
        if (k < a.n_out) a.deltas[(size_t)i * a.n_out + k] = d[k];
    }
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

static int nr_inputs_ok(const NonRigidArgs& a) {
  if (a.P < 0 || a.n_feat != 2 || a.hidden != NR_H || a.n_levels * a.n_feat != NR_F) return 0;
  if (a.n_out < 10 || a.n_out > 32 || a.n_cond < 0) return 0;
  if (!a.aabb_min || !a.aabb_max || !a.xyz || !a.scaling || !a.rotation || !a.table) return 0;
  if ((a.n_cond > 0 && !a.cond) || !a.W0 || !a.b0 || !a.W1 || !a.b1 || !a.W2 || !a.b2 || !a.W3 || !a.b3) return 0;
  return 1;
}
static int nr_args_ok(const NonRigidArgs& a) { return nr_inputs_ok(a) && a.xyz_out && a.scaling_out && a.rotation_out; }

extern "C" size_t nonrigid_workspace_bytes(const NonRigidArgs* args) {
  return carve_ws(nullptr, args ? args->n_out : 10, nullptr);
}

extern "C" size_t nonrigid_saved_bytes(int32_t P) { return (size_t)((P + NR_TILE - 1) / NR_TILE) * NR_SAVE_TILE * sizeof(__nv_bfloat16); }

static int nonrigid_forward_impl(const NonRigidArgs* args, void* workspace, size_t workspace_bytes, void* saved, void* stream);
extern "C" int nonrigid_forward(const NonRigidArgs* args, void* workspace, size_t workspace_bytes, void* stream) {
  return nonrigid_forward_impl(args, workspace, workspace_bytes, nullptr, stream);
}
extern "C" int nonrigid_forward_train(const NonRigidArgs* args, void* workspace, size_t workspace_bytes, void* saved,
                                      size_t saved_bytes, void* stream) {
  if (!args || !saved || saved_bytes < nonrigid_saved_bytes(args->P)) return GS_ERR_WORKSPACE_SMALL;
  return nonrigid_forward_impl(args, workspace, workspace_bytes, saved, stream);
}
static int nonrigid_forward_impl(const NonRigidArgs* args, void* workspace, size_t workspace_bytes, void* saved, void* stream) {
  if (!args || !workspace || !nr_args_ok(*args)) return GS_ERR_BAD_ARG;
  if (((uintptr_t)workspace | (uintptr_t)saved) & 15) return GS_ERR_BAD_ARG;  // bulk copies: 16-byte aligned operands
  const NonRigidArgs& a = *args;
  if (a.P == 0) return GS_OK;
  cudaStream_t st = (cudaStream_t)stream;
  NrWs w;
  if (carve_ws((char*)workspace, a.n_out, &w) > workspace_bytes) return GS_ERR_WORKSPACE_SMALL;
  const int npad = nout_pad(a.n_out);
  nr_prep_kernel<<<64, 256, 0, st>>>(a, w, npad);
  // per call: function attributes are per DEVICE, and one process may drive several
  if (gs_check(cudaFuncSetAttribute(nr_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NrSmem) + 1024)) != GS_OK)
    return GS_ERR_CUDA;
  nr_fwd_kernel<<<(a.P + NR_TILE - 1) / NR_TILE, BW_THREADS, sizeof(NrSmem) + 1024, st>>>(a, w, npad, (__nv_bfloat16*)saved);
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
  // backward: W^T as the B operand of dA = G . W (element (row = k_in, col = n_out) = W[n_out][k_in])
  __nv_bfloat16 *w0t_hi, *w0t_lo;      // [16 x 128]
  __nv_bfloat16 *wht_hi[3], *wht_lo[3];  // lin1..lin3 [128 x 128]
  __nv_bfloat16 *w4t_hi, *w4t_lo;      // [128 x 32]
};

__host__ size_t carve_sk(char* base, SkWs* w) {
  char* p = base;
  SkWs t;
  gs_carve(p, t.w0_hi, (size_t)NR_H * SK_K0); gs_carve(p, t.w0_lo, (size_t)NR_H * SK_K0);
  for (int i = 0; i < 3; i++) { gs_carve(p, t.wh_hi[i], (size_t)NR_H * NR_H); gs_carve(p, t.wh_lo[i], (size_t)NR_H * NR_H); }
  gs_carve(p, t.w4_hi, (size_t)SK_NOUT * NR_H); gs_carve(p, t.w4_lo, (size_t)SK_NOUT * NR_H);
  gs_carve(p, t.w0t_hi, (size_t)SK_K0 * NR_H); gs_carve(p, t.w0t_lo, (size_t)SK_K0 * NR_H);
  for (int i = 0; i < 3; i++) { gs_carve(p, t.wht_hi[i], (size_t)NR_H * NR_H); gs_carve(p, t.wht_lo[i], (size_t)NR_H * NR_H); }
  gs_carve(p, t.w4t_hi, (size_t)NR_H * SK_NOUT); gs_carve(p, t.w4t_lo, (size_t)NR_H * SK_NOUT);
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

__global__ void __launch_bounds__(BW_THREADS, 1) sk_fwd_kernel(SkinArgs a, SkWs w, __nv_bfloat16* __restrict__ saved) {
  extern __shared__ __align__(1024) unsigned char nr_smem_raw[];
  NrSmem& sm = *reinterpret_cast<NrSmem*>(nr_smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5;
  const int row = tid & (NR_TILE - 1), half = tid >> 7;  // two threads per Gaussian (see nr_fwd_kernel)
  const int i = blockIdx.x * NR_TILE + row;
  const bool live = i < a.P;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    nr_mbar_init(&sm.bar, 1);
    nr_mbar_init(&sm.ld_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    bulk_load_operands(&sm.ld_bar, nullptr, nullptr, nullptr, nullptr, 0u, sm.W_hi, sm.W_lo, w.w0_hi, w.w0_lo, NR_H * SK_K0 * 2);
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
    if (half == 0) store_act8(sm.A_hi, sm.A_lo, row, 0, in);
    else store_act8(sm.A_hi, sm.A_lo, row, 8, in + 8);
  }
  if (tid < NR_H) sm.bias[tid] = a.b[0][tid];
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  nr_mbar_wait(&sm.ld_bar, 0u);  // W0 has landed (after the barrier: the mbarrier was initialised by thread 0)
  __nv_bfloat16* const save_tile = saved ? saved + (size_t)blockIdx.x * SK_SAVE_TILE : nullptr;
  if (save_tile) {
    copy16(save_tile, sm.A_hi, 128 * SK_K0 * 2);
    copy16(save_tile + 128 * SK_K0, sm.A_lo, 128 * SK_K0 * 2);
  }
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tmem_row = tmem + ((uint32_t)((warp & 3) * 32) << 16);
#pragma unroll 1
  for (int layer = 0; layer < 4; layer++) {  // four hidden layers
    if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W_hi, sm.W_lo, layer == 0 ? SK_K0 : NR_H, NR_H, &sm.bar);
    nr_mbar_wait(&sm.bar, (uint32_t)layer & 1u);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (tid == 0) {  // W is free (the MMAs have completed): the next layer's weights stream in during the epilogue
      if (layer < 3)
        bulk_load_operands(&sm.ld_bar, nullptr, nullptr, nullptr, nullptr, 0u, sm.W_hi, sm.W_lo, w.wh_hi[layer], w.wh_lo[layer], NR_H * NR_H * 2);
      else
        bulk_load_operands(&sm.ld_bar, nullptr, nullptr, nullptr, nullptr, 0u, sm.W_hi, sm.W_lo, w.w4_hi, w.w4_lo, SK_NOUT * NR_H * 2);
    }
    hidden_epilogue(sm, tmem_row, row, half);
    __syncthreads();
    if (save_tile) {
      copy16(save_tile + SK_SAVE_IN + (size_t)layer * NR_SAVE_LAYER, sm.A_hi, 128 * 128 * 2);
      copy16(save_tile + SK_SAVE_IN + (size_t)layer * NR_SAVE_LAYER + 128 * 128, sm.A_lo, 128 * 128 * 2);
    }
    if (tid < NR_H) {
      if (layer < 3) sm.bias[tid] = a.b[layer + 1][tid];
      else sm.bias[tid] = tid < a.n_out ? a.b[4][tid] : 0.f;
    }
    nr_mbar_wait(&sm.ld_bar, (uint32_t)(layer + 1) & 1u);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W_hi, sm.W_lo, NR_H, SK_NOUT, &sm.bar);
  nr_mbar_wait(&sm.bar, 0u);  // fifth commit
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  float x[32];
  if (half == 0) tmem_ld32(tmem_row, x);  // warp-uniform
  if (live && half == 0) {
#pragma unroll
    for (int k = 0; k < 32; k++) x[k] += sm.bias[k];
    if (a.logits) {
#pragma unroll
      for (int k = 0; k < 25; k++)
        if (k < a.n_out) a.logits[(size_t)i * a.n_out + k] = x[k];
    }
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

extern "C" size_t skinning_saved_bytes(int32_t P) { return (size_t)((P + NR_TILE - 1) / NR_TILE) * SK_SAVE_TILE * sizeof(__nv_bfloat16); }

static int skinning_forward_impl(const SkinArgs* args, void* workspace, size_t workspace_bytes, void* saved, void* stream);
extern "C" int skinning_forward(const SkinArgs* args, void* workspace, size_t workspace_bytes, void* stream) {
  return skinning_forward_impl(args, workspace, workspace_bytes, nullptr, stream);
}
extern "C" int skinning_forward_train(const SkinArgs* args, void* workspace, size_t workspace_bytes, void* saved, size_t saved_bytes,
                                      void* stream) {
  if (!args || !saved || saved_bytes < skinning_saved_bytes(args->P)) return GS_ERR_WORKSPACE_SMALL;
  return skinning_forward_impl(args, workspace, workspace_bytes, saved, stream);
}
static int skinning_forward_impl(const SkinArgs* args, void* workspace, size_t workspace_bytes, void* saved, void* stream) {
  if (!args || !workspace) return GS_ERR_BAD_ARG;
  if (((uintptr_t)workspace | (uintptr_t)saved) & 15) return GS_ERR_BAD_ARG;  // bulk copies: 16-byte aligned operands
  const SkinArgs& a = *args;
  if (a.P < 0 || a.hidden != NR_H || (a.n_out != 25 && a.n_out != 24) || !a.aabb_min || !a.aabb_max || !a.xyz || !a.weights)
    return GS_ERR_BAD_ARG;
  for (int l = 0; l < 5; l++) if (!a.W[l] || !a.b[l]) return GS_ERR_BAD_ARG;
  if (a.P == 0) return GS_OK;
  cudaStream_t st = (cudaStream_t)stream;
  SkWs w;
  if (carve_sk((char*)workspace, &w) > workspace_bytes) return GS_ERR_WORKSPACE_SMALL;
  sk_prep_kernel_impl<<<64, 256, 0, st>>>(a, w);
  // per call: function attributes are per DEVICE, and one process may drive several
  if (gs_check(cudaFuncSetAttribute(sk_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NrSmem) + 1024)) != GS_OK)
    return GS_ERR_CUDA;
  sk_fwd_kernel<<<(a.P + NR_TILE - 1) / NR_TILE, BW_THREADS, sizeof(NrSmem) + 1024, st>>>(a, w, (__nv_bfloat16*)saved);
  return gs_check(cudaGetLastError());
}

// ================================= non-rigid deformation: backward =================================
namespace {

// W^T operands of dA = G . W (see NrWs): element (row = k_in, col = n_out) = W[n_out][k_in], hi/lo split
__global__ void nr_prep_bwd_kernel(NonRigidArgs a, NrWs w, int npad) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  const int in0 = NR_F + a.n_cond;
  for (int e = tid; e < NR_F * NR_H; e += nth) {
    const int k = e / NR_H, n = e % NR_H;
    split_bf16(a.W0[(size_t)n * in0 + k], w.w0t_hi[core_idx(k, n, NR_F)], w.w0t_lo[core_idx(k, n, NR_F)]);
  }
  for (int e = tid; e < NR_H * NR_H; e += nth) {
    const int k = e / NR_H, n = e % NR_H;
    split_bf16(a.W1[(size_t)n * NR_H + k], w.w1t_hi[core_idx(k, n, NR_H)], w.w1t_lo[core_idx(k, n, NR_H)]);
    split_bf16(a.W2[(size_t)n * NR_H + k], w.w2t_hi[core_idx(k, n, NR_H)], w.w2t_lo[core_idx(k, n, NR_H)]);
  }
  for (int e = tid; e < NR_H * npad; e += nth) {
    const int k = e / npad, n = e % npad;
    const float v = n < a.n_out ? a.W3[(size_t)n * NR_H + k] : 0.f;
    split_bf16(v, w.w3t_hi[core_idx(k, n, NR_H)], w.w3t_lo[core_idx(k, n, NR_H)]);
  }
}

// Backward of hash_encode for one point: scatters dL/dtable (8 corners x levels) and returns dL/d(x01) through the trilinear
// weights (tiny-cuda-nn's grid backward + input gradient; same indexing as hash_encode above).
template <int L_BEGIN, int L_END>  // levels [L_BEGIN, L_END): the two threads of a Gaussian take half of the levels each
__device__ __forceinline__ void hash_backward(const NonRigidArgs& a, float x, float y, float z, const float* dfeat /*[2L]*/,
                                              float* __restrict__ dtable, float* dpos /*[3]*/) {
  dpos[0] = dpos[1] = dpos[2] = 0.f;
#pragma unroll
  for (int l0 = L_BEGIN; l0 < L_END; l0 += HG_GROUP) {  // four levels at a time: 32 table reads in flight (see hash_encode)
    float2 v[HG_GROUP][8];
    uint32_t ix[HG_GROUP][8];
    float wx[HG_GROUP], wy[HG_GROUP], wz[HG_GROUP];
#pragma unroll
    for (int q = 0; q < HG_GROUP; q++) {
      const int l = l0 + q;
      const float scale = a.level_scale[l];
      const uint32_t res = a.level_res[l];
      const uint32_t off = a.level_offset[l], size = a.level_offset[l + 1] - off;
      const float px = fmaf(x, scale, 0.5f), py = fmaf(y, scale, 0.5f), pz = fmaf(z, scale, 0.5f);
      const float fx0 = floorf(px), fy0 = floorf(py), fz0 = floorf(pz);
      wx[q] = px - fx0; wy[q] = py - fy0; wz[q] = pz - fz0;
      const uint32_t gx = (uint32_t)(int)fx0, gy = (uint32_t)(int)fy0, gz = (uint32_t)(int)fz0;
#pragma unroll
      for (int c = 0; c < 8; c++) {
        ix[q][c] = off + grid_index(gx + (c & 1), gy + ((c >> 1) & 1), gz + ((c >> 2) & 1), res, size);
        v[q][c] = __ldg(reinterpret_cast<const float2*>(a.table) + ix[q][c]);
      }
    }
#pragma unroll
    for (int q = 0; q < HG_GROUP; q++) {
      const int l = l0 + q;
      const float d0 = dfeat[2 * l], d1 = dfeat[2 * l + 1];
      float gwx = 0.f, gwy = 0.f, gwz = 0.f;
#pragma unroll
      for (int c = 0; c < 8; c++) {
        const uint32_t dx = c & 1, dy = (c >> 1) & 1, dz = (c >> 2) & 1;
        const float ax = dx ? wx[q] : 1.f - wx[q], ay = dy ? wy[q] : 1.f - wy[q], az = dz ? wz[q] : 1.f - wz[q];
        const float wgt = ax * ay * az;
        float* dst = dtable + 2 * (size_t)ix[q][c];
        asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst), "f"(wgt * d0), "f"(wgt * d1) : "memory");
        const float dv = v[q][c].x * d0 + v[q][c].y * d1;  // dL/dwgt of this corner
        gwx += (dx ? dv : -dv) * ay * az;
        gwy += (dy ? dv : -dv) * ax * az;
        gwz += (dz ? dv : -dv) * ax * ay;
      }
      const float scale = a.level_scale[l];
      dpos[0] = fmaf(gwx, scale, dpos[0]);
      dpos[1] = fmaf(gwy, scale, dpos[1]);
      dpos[2] = fmaf(gwz, scale, dpos[2]);
    }
  }
}

__global__ void __launch_bounds__(BW_THREADS, 1)
nr_bwd_kernel(NonRigidBackwardArgs b, NrWs w, int npad, const __nv_bfloat16* __restrict__ saved, int ntiles) {
  extern __shared__ __align__(1024) unsigned char bw_smem_raw[];
  BwSmem& sm = *reinterpret_cast<BwSmem*>(bw_smem_raw);
  const NonRigidArgs& a = b.fwd;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row = tid & (NR_TILE - 1), half = tid >> 7;  // two threads per Gaussian: half 0 / 1 split the columns (BW_THREADS)
  if (warp == 0) {  // TMEM: columns [0,128) dA of the current layer, [128,256) the tile's dW partial
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    nr_mbar_init(&sm.bar, 1);
    nr_mbar_init(&sm.ld_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int k = tid; k < 5 * BW_H; k += BW_THREADS) (&sm.dbias[0][0])[k] = 0.f;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tmem_row = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const int in0 = NR_F + a.n_cond;
  uint32_t phase = 0, ld_phase = 0;
  uint32_t signs[4];  // LeakyReLU branches of this thread's row in the activations loaded last

  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int i = tile * NR_TILE + row;
    const bool live = i < a.P;
    const __nv_bfloat16* const st = saved + (size_t)tile * NR_SAVE_TILE;
    // operands of the output layer (A2, W3^T) stream in while the upstream gradient is read: A / W are free (the previous tile's
    // MMAs have completed and the barrier at its end has been passed)
    if (tid == 0)
      bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st + NR_SAVE_IN + 2 * NR_SAVE_LAYER, st + NR_SAVE_IN + 2 * NR_SAVE_LAYER + 128 * 128,
                         128 * 128 * 2, sm.W_hi, sm.W_lo, w.w3t_hi, w.w3t_lo, (uint32_t)(NR_H * npad * 2));
    // ---- gradient of the raw MLP output (delta application, non_rigid.py:262-282) ----
    float g[32];
#pragma unroll
    for (int k = 0; k < 32; k++) g[k] = 0.f;
    float gx = 0.f, gy = 0.f, gz = 0.f;
    if (live && half == 0) {
      if (b.dL_ddeltas) {
#pragma unroll
        for (int k = 0; k < 32; k++)
          if (k < a.n_out) g[k] = b.dL_ddeltas[(size_t)i * a.n_out + k];
      }
      if (b.dL_dxyz_out) { gx = b.dL_dxyz_out[3 * (size_t)i]; gy = b.dL_dxyz_out[3 * (size_t)i + 1]; gz = b.dL_dxyz_out[3 * (size_t)i + 2]; }
      g[0] += gx; g[1] += gy; g[2] += gz;
      float gs[3] = {0.f, 0.f, 0.f};
      if (b.dL_dscaling_out)
        for (int k = 0; k < 3; k++) gs[k] = b.dL_dscaling_out[3 * (size_t)i + k];
      g[3] += gs[0]; g[4] += gs[1]; g[5] += gs[2];
      for (int k = 0; k < 3; k++) b.dL_dscaling[3 * (size_t)i + k] = gs[k];
      float4 gr = make_float4(0.f, 0.f, 0.f, 0.f);
      if (b.dL_drotation_out) gr = *reinterpret_cast<const float4*>(b.dL_drotation_out + 4 * (size_t)i);
      const float4 q2 = *reinterpret_cast<const float4*>(a.rotation + 4 * (size_t)i);
      const float ax = a.deltas[(size_t)i * a.n_out + 7], ay = a.deltas[(size_t)i * a.n_out + 8], az = a.deltas[(size_t)i * a.n_out + 9];
      // rotation' = (1, ax, ay, az) * q2 (Hamilton product): adjoints of both factors
      g[7] += -q2.y * gr.x + q2.x * gr.y - q2.w * gr.z + q2.z * gr.w;
      g[8] += -q2.z * gr.x + q2.w * gr.y + q2.x * gr.z - q2.y * gr.w;
      g[9] += -q2.w * gr.x - q2.z * gr.y + q2.y * gr.z + q2.x * gr.w;
      float4 dq;
      dq.x = gr.x + ax * gr.y + ay * gr.z + az * gr.w;
      dq.y = -ax * gr.x + gr.y + az * gr.z - ay * gr.w;
      dq.z = -ay * gr.x - az * gr.y + gr.z + ax * gr.w;
      dq.w = -az * gr.x + ay * gr.y - ax * gr.z + gr.w;
      *reinterpret_cast<float4*>(b.dL_drotation + 4 * (size_t)i) = dq;
    }
    if (half == 0) store_grad32(sm, row, 0, g, sm.dbias[3], lane);
    // ---- layer 3: Z3 = A2 . W3^T ----
    bw_wait_loads(sm, ld_phase);
    bw_row_signs(sm, row, 64 * half, 64 * half + 64, signs);  // of A2 (this thread's columns): used by the epilogue that forms G2
    bw_sync_mma_inputs();
    if (tid == 0) {
      issue_dw(tmem + 128, sm.A_hi, sm.A_lo, sm.G_hi, sm.G_lo, npad);   // dW3^T [k_in][n_out]
      issue_da(tmem, sm.G_hi, sm.G_lo, sm.W_hi, sm.W_lo, npad, NR_H);   // dA2
    }
    bw_commit_wait(sm, tid == 0, phase);
    if (tid == 0)  // A / W are free again (MMAs done, signs taken before the barrier above): next layer's operands, A1 and W2^T
      bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st + NR_SAVE_IN + NR_SAVE_LAYER, st + NR_SAVE_IN + NR_SAVE_LAYER + 128 * 128,
                         128 * 128 * 2, sm.W_hi, sm.W_lo, w.w2t_hi, w.w2t_lo, NR_H * NR_H * 2);
    if (half == 0) {  // warp-uniform
      float v[32];
      tmem_ld32(tmem_row + 128, v);
#pragma unroll
      for (int n = 0; n < 32; n++)
        if (n < a.n_out) atomicAdd(b.dL_dW3 + (size_t)n * NR_H + row, v[n]);
    }
    const float* const dWh[2] = {b.dL_dW2, b.dL_dW1};
#pragma unroll 1
    for (int hl = 0; hl < 3; hl++) {  // hl = 0: layer 2 (input A1), 1: layer 1 (input A0), 2: layer 0 (input = hash features)
      // G of this layer from dA (TMEM) and the sign of the saved activation still in sm.A
#pragma unroll 1
      for (int c = 2 * half; c < 2 * half + 2; c++) {  // this thread's 64 columns
        float v[32];
        tmem_ld32(tmem_row + c * 32, v);
#pragma unroll
        for (int j = 0; j < 32; j += 8) lrelu_mask8_bits(signs, c * 32 + j, v + j);
        store_grad32(sm, row, c * 32, v, sm.dbias[2 - hl], lane);
      }
      const int kin = hl < 2 ? NR_H : NR_F;
      bw_wait_loads(sm, ld_phase);                             // this layer's A_in and W^T (issued after the previous MMAs)
      if (hl < 2) bw_row_signs(sm, row, 64 * half, 64 * half + 64, signs);  // of A1 / A0, for the next epilogue
      bw_sync_mma_inputs();                                    // G complete; every row has read its dA from TMEM and its signs
      if (tid == 0) {
        issue_dw(tmem + 128, sm.G_hi, sm.G_lo, sm.A_hi, sm.A_lo, kin);      // dW [n_out][k_in]
        issue_da(tmem, sm.G_hi, sm.G_lo, sm.W_hi, sm.W_lo, NR_H, kin);      // dA_in
      }
      bw_commit_wait(sm, tid == 0, phase);
      if (tid == 0 && hl < 2) {  // the next layer's operands stream in during the dW flush and the epilogue
        if (hl == 0)
          bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st + NR_SAVE_IN, st + NR_SAVE_IN + 128 * 128, 128 * 128 * 2, sm.W_hi, sm.W_lo,
                             w.w1t_hi, w.w1t_lo, NR_H * NR_H * 2);
        else
          bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st, st + 128 * NR_F, 128 * NR_F * 2, sm.W_hi, sm.W_lo, w.w0t_hi, w.w0t_lo,
                             NR_F * NR_H * 2);
      }
      if (hl < 2) flush_dw_rows(tmem_row + 128, 64 * half, 64 * half + 64, const_cast<float*>(dWh[hl]) + (size_t)row * NR_H);
      else if (half == 0) flush_dw_rows(tmem_row + 128, 0, NR_F, b.dL_dW0 + (size_t)row * in0);
    }
    // ---- hash grid: dL/dfeatures of this Gaussian -> table scatter + dL/dxyz; the two threads of a row take eight levels each,
    // the second one passes its share of dL/dx through shared memory (sm.G is free: its last readers, the MMAs, are done) ----
    {
      float df[32];
      tmem_ld32(tmem_row, df);
      float dpos[3] = {0.f, 0.f, 0.f};
      float ex = 1.f, ey = 1.f, ez = 1.f;
      if (live) {
        const float x = a.xyz[3 * (size_t)i], y = a.xyz[3 * (size_t)i + 1], z = a.xyz[3 * (size_t)i + 2];
        ex = a.aabb_max[0] - a.aabb_min[0]; ey = a.aabb_max[1] - a.aabb_min[1]; ez = a.aabb_max[2] - a.aabb_min[2];
        const float nx = 2.f * ((x - a.aabb_min[0]) / ex) - 1.f, ny = 2.f * ((y - a.aabb_min[1]) / ey) - 1.f;
        const float nz = 2.f * ((z - a.aabb_min[2]) / ez) - 1.f;
        if (half == 0) hash_backward<0, HG_LEVELS / 2>(a, (nx + 1.f) * 0.5f, (ny + 1.f) * 0.5f, (nz + 1.f) * 0.5f, df, b.dL_dtable, dpos);
        else hash_backward<HG_LEVELS / 2, HG_LEVELS>(a, (nx + 1.f) * 0.5f, (ny + 1.f) * 0.5f, (nz + 1.f) * 0.5f, df, b.dL_dtable, dpos);
      }
      float* const xch = reinterpret_cast<float*>(sm.G_hi);
      if (half == 1) { xch[3 * row] = dpos[0]; xch[3 * row + 1] = dpos[1]; xch[3 * row + 2] = dpos[2]; }
      __syncthreads();
      if (live && half == 0) {  // x01 = (x - min) / extent
        b.dL_dxyz[3 * (size_t)i] = gx + (dpos[0] + xch[3 * row]) / ex;
        b.dL_dxyz[3 * (size_t)i + 1] = gy + (dpos[1] + xch[3 * row + 1]) / ey;
        b.dL_dxyz[3 * (size_t)i + 2] = gz + (dpos[2] + xch[3 * row + 2]) / ez;
      }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();  // sm.G / sm.A / TMEM are reused by the next tile
  }
  // ---- bias gradients of this CTA ----
  __syncthreads();
  {
    float* const dbg[4] = {b.dL_db0, b.dL_db1, b.dL_db2, b.dL_db3};
    for (int l = 0; l < 4 && tid < BW_H; l++) {
      const float v = sm.dbias[l][tid];
      if (v != 0.f && (l < 3 || tid < a.n_out)) atomicAdd(dbg[l] + tid, v);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

// The condition enters layer 0 through the bias (forward: bias0 = b0 + W0[:, F:] . cond), so its gradients are rank one:
// dL/dW0[n][F + c] = db0[n] * cond[c], dL/dcond[c] = sum_n W0[n][F + c] * db0[n].
__global__ void nr_bwd_finish_kernel(NonRigidBackwardArgs b) {
  const NonRigidArgs& a = b.fwd;
  const int in0 = NR_F + a.n_cond;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < NR_H * a.n_cond; e += gridDim.x * blockDim.x) {
    const int n = e / a.n_cond, c = e % a.n_cond;
    b.dL_dW0[(size_t)n * in0 + NR_F + c] = b.dL_db0[n] * a.cond[c];
  }
  if (b.dL_dcond) {  // dcond[c] = sum_n W0[n][F + c] db0[n]: a warp per output, lanes over n
    const int lane = threadIdx.x & 31, gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (int c = gw; c < a.n_cond; c += nw) {
      float acc = 0.f;
      for (int n = lane; n < NR_H; n += 32) acc = fmaf(a.W0[(size_t)n * in0 + NR_F + c], b.dL_db0[n], acc);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
      if (lane == 0) b.dL_dcond[c] = acc;
    }
  }
}

}  // namespace

extern "C" int nonrigid_backward(const NonRigidBackwardArgs* args, void* workspace, size_t workspace_bytes, const void* saved,
                                 size_t saved_bytes, void* stream) {
  if (!args || !workspace || !saved || !nr_inputs_ok(args->fwd)) return GS_ERR_BAD_ARG;
  if (((uintptr_t)workspace | (uintptr_t)saved) & 15) return GS_ERR_BAD_ARG;  // bulk copies: 16-byte aligned operands
  const NonRigidBackwardArgs& b = *args;
  const NonRigidArgs& a = b.fwd;
  if (!a.deltas || !b.dL_dxyz || !b.dL_dscaling || !b.dL_drotation || !b.dL_dtable) return GS_ERR_BAD_ARG;
  if (!b.dL_dW0 || !b.dL_db0 || !b.dL_dW1 || !b.dL_db1 || !b.dL_dW2 || !b.dL_db2 || !b.dL_dW3 || !b.dL_db3) return GS_ERR_BAD_ARG;
  if (saved_bytes < nonrigid_saved_bytes(a.P)) return GS_ERR_WORKSPACE_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  NrWs w;
  if (carve_ws((char*)workspace, a.n_out, &w) > workspace_bytes) return GS_ERR_WORKSPACE_SMALL;
  const int npad = nout_pad(a.n_out), in0 = NR_F + a.n_cond;
  const size_t f = sizeof(float);
  GS_TRY_NR(cudaMemsetAsync(b.dL_dtable, 0, (size_t)a.level_offset[a.n_levels] * 2 * f, st));
  GS_TRY_NR(cudaMemsetAsync(b.dL_dW0, 0, (size_t)NR_H * in0 * f, st));
  GS_TRY_NR(cudaMemsetAsync(b.dL_dW1, 0, (size_t)NR_H * NR_H * f, st));
  GS_TRY_NR(cudaMemsetAsync(b.dL_dW2, 0, (size_t)NR_H * NR_H * f, st));
  GS_TRY_NR(cudaMemsetAsync(b.dL_dW3, 0, (size_t)a.n_out * NR_H * f, st));
  GS_TRY_NR(cudaMemsetAsync(b.dL_db0, 0, NR_H * f, st));
  GS_TRY_NR(cudaMemsetAsync(b.dL_db1, 0, NR_H * f, st));
  GS_TRY_NR(cudaMemsetAsync(b.dL_db2, 0, NR_H * f, st));
  GS_TRY_NR(cudaMemsetAsync(b.dL_db3, 0, (size_t)a.n_out * f, st));
  if (b.dL_dcond) GS_TRY_NR(cudaMemsetAsync(b.dL_dcond, 0, (size_t)a.n_cond * f, st));
  if (a.P == 0) return GS_OK;
  nr_prep_bwd_kernel<<<64, 256, 0, st>>>(a, w, npad);
  const int smem = (int)sizeof(BwSmem) + 1024;
  if (gs_check(cudaFuncSetAttribute(nr_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != GS_OK) return GS_ERR_CUDA;
  const int ntiles = (a.P + NR_TILE - 1) / NR_TILE;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  nr_bwd_kernel<<<ntiles < sms ? ntiles : sms, BW_THREADS, smem, st>>>(b, w, npad, (const __nv_bfloat16*)saved, ntiles);
  nr_bwd_finish_kernel<<<32, 256, 0, st>>>(b);
  return gs_check(cudaGetLastError());
}

// ================================= skinning field: backward =================================
namespace {

__global__ void sk_prep_bwd_kernel(SkinArgs a, SkWs w) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  for (int e = tid; e < SK_K0 * NR_H; e += nth) {
    const int k = e / NR_H, n = e % NR_H;
    const float v = k < 3 ? a.W[0][n * 3 + k] : 0.f;
    split_bf16(v, w.w0t_hi[core_idx(k, n, SK_K0)], w.w0t_lo[core_idx(k, n, SK_K0)]);
  }
  for (int l = 0; l < 3; l++)
    for (int e = tid; e < NR_H * NR_H; e += nth) {
      const int k = e / NR_H, n = e % NR_H;
      split_bf16(a.W[1 + l][(size_t)n * NR_H + k], w.wht_hi[l][core_idx(k, n, NR_H)], w.wht_lo[l][core_idx(k, n, NR_H)]);
    }
  for (int e = tid; e < NR_H * SK_NOUT; e += nth) {
    const int k = e / SK_NOUT, n = e % SK_NOUT;
    const float v = n < a.n_out ? a.W[4][(size_t)n * NR_H + k] : 0.f;
    split_bf16(v, w.w4t_hi[core_idx(k, n, NR_H)], w.w4t_lo[core_idx(k, n, NR_H)]);
  }
}

// Adjoint of hierarchical_softmax (rigid.py:85-129): x = 25 logits, gp = dL/dp (24) -> gx = dL/dx (25).
__device__ __forceinline__ void hierarchical_softmax_bwd(const float* x, const float* gp, float* gx) {
  float s[25];
#pragma unroll
  for (int k = 0; k < 25; k++) s[k] = sigmoidf_(x[k]);
  auto softmax3 = [](float a, float b, float c, float* o) {
    const float m = fmaxf(a, fmaxf(b, c));
    const float ea = expf(a - m), eb = expf(b - m), ec = expf(c - m), inv = 1.0f / (ea + eb + ec);
    o[0] = ea * inv; o[1] = eb * inv; o[2] = ec * inv;
  };
  float sm1[3], sm2[3];
  softmax3(x[1], x[2], x[3], sm1);
  softmax3(x[12], x[13], x[14], sm2);
  // forward intermediates
  float l1[3], l2[3], l3[3];
  for (int k = 0; k < 3; k++) { l1[k] = s[0] * sm1[k]; l2[k] = l1[k] * s[4 + k]; l3[k] = l2[k] * s[7 + k]; }
  const float n12 = l3[2] * s[24] * sm2[0], n13 = l3[2] * s[24] * sm2[1], n14 = l3[2] * s[24] * sm2[2];
  const float c16[2] = {n13 * s[16], n14 * s[17]};
  const float c18[2] = {c16[0] * s[18], c16[1] * s[19]};
  const float c20[2] = {c18[0] * s[20], c18[1] * s[21]};
  float gs[25];
#pragma unroll
  for (int k = 0; k < 25; k++) gs[k] = 0.f;
  // arms: 13 -> 16 -> 18 -> 20 -> 22 and 14 -> 17 -> 19 -> 21 -> 23
  float g_c16[2];
  for (int k = 0; k < 2; k++) {
    const float g_c22 = gp[22 + k];
    const float g_c20 = g_c22 * s[22 + k] + gp[20 + k] * (1.f - s[22 + k]);
    gs[22 + k] = (g_c22 - gp[20 + k]) * c20[k];
    const float g_c18 = g_c20 * s[20 + k] + gp[18 + k] * (1.f - s[20 + k]);
    gs[20 + k] = (g_c20 - gp[18 + k]) * c18[k];
    g_c16[k] = g_c18 * s[18 + k] + gp[16 + k] * (1.f - s[18 + k]);
    gs[18 + k] = (g_c18 - gp[16 + k]) * c16[k];
  }
  const float g_n13 = g_c16[0] * s[16] + gp[13] * (1.f - s[16]);
  gs[16] = (g_c16[0] - gp[13]) * n13;
  const float g_n14 = g_c16[1] * s[17] + gp[14] * (1.f - s[17]);
  gs[17] = (g_c16[1] - gp[14]) * n14;
  const float g_n12 = gp[15] * s[15] + gp[12] * (1.f - s[15]);
  gs[15] = (gp[15] - gp[12]) * n12;
  // (n12, n13, n14) = l3[2] * s24 * softmax(x[12:15]); p9 = l3[2] * (1 - s24)
  const float gb[3] = {g_n12, g_n13, g_n14};
  const float gb_dot = gb[0] * sm2[0] + gb[1] * sm2[1] + gb[2] * sm2[2];
  float g_l3[3];
  g_l3[2] = gb_dot * s[24] + gp[9] * (1.f - s[24]);
  gs[24] = (gb_dot - gp[9]) * l3[2];
  float g_sm2[3];
  for (int k = 0; k < 3; k++) g_sm2[k] = gb[k] * l3[2] * s[24];
  // l4_k = l3_k * s[10+k] -> p[10+k]; p[7+k] = l3_k * (1 - s[10+k])   (k = 0, 1)
  for (int k = 0; k < 2; k++) {
    g_l3[k] = gp[10 + k] * s[10 + k] + gp[7 + k] * (1.f - s[10 + k]);
    gs[10 + k] = (gp[10 + k] - gp[7 + k]) * l3[k];
  }
  float g_sm1[3];
  float gs0 = -gp[0];
  for (int k = 0; k < 3; k++) {
    const float g_l2 = g_l3[k] * s[7 + k] + gp[4 + k] * (1.f - s[7 + k]);
    gs[7 + k] = (g_l3[k] - gp[4 + k]) * l2[k];
    const float g_l1 = g_l2 * s[4 + k] + gp[1 + k] * (1.f - s[4 + k]);
    gs[4 + k] = (g_l2 - gp[1 + k]) * l1[k];
    gs0 += g_l1 * sm1[k];
    g_sm1[k] = g_l1 * s[0];
  }
  gs[0] = gs0;
#pragma unroll
  for (int k = 0; k < 25; k++) gx[k] = gs[k] * s[k] * (1.f - s[k]);
  const float d1 = g_sm1[0] * sm1[0] + g_sm1[1] * sm1[1] + g_sm1[2] * sm1[2];
  const float d2 = g_sm2[0] * sm2[0] + g_sm2[1] * sm2[1] + g_sm2[2] * sm2[2];
  for (int k = 0; k < 3; k++) {  // s[1..3] and s[12..14] are not used by the forward: softmax terms only
    gx[1 + k] = sm1[k] * (g_sm1[k] - d1);
    gx[12 + k] = sm2[k] * (g_sm2[k] - d2);
  }
}

__global__ void __launch_bounds__(BW_THREADS, 1)
sk_bwd_kernel(SkinBackwardArgs b, SkWs w, const __nv_bfloat16* __restrict__ saved, int ntiles) {
  extern __shared__ __align__(1024) unsigned char bw_smem_raw[];
  BwSmem& sm = *reinterpret_cast<BwSmem*>(bw_smem_raw);
  const SkinArgs& a = b.fwd;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row = tid & (NR_TILE - 1), half = tid >> 7;  // two threads per Gaussian: half 0 / 1 split the columns (BW_THREADS)
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    nr_mbar_init(&sm.bar, 1);
    nr_mbar_init(&sm.ld_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int k = tid; k < 5 * BW_H; k += BW_THREADS) (&sm.dbias[0][0])[k] = 0.f;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tmem_row = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t phase = 0, ld_phase = 0;
  uint32_t signs[4];  // LeakyReLU branches of this thread's row in the activations loaded last

  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int i = tile * NR_TILE + row;
    const bool live = i < a.P;
    const __nv_bfloat16* const st = saved + (size_t)tile * SK_SAVE_TILE;
    if (tid == 0)  // operands of the output layer (A3, W4^T) stream in while the softmax adjoint runs
      bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st + SK_SAVE_IN + 3 * NR_SAVE_LAYER, st + SK_SAVE_IN + 3 * NR_SAVE_LAYER + 128 * 128,
                         128 * 128 * 2, sm.W_hi, sm.W_lo, w.w4t_hi, w.w4t_lo, NR_H * SK_NOUT * 2);
    // ---- dL/dlogits through the (hierarchical) softmax ----
    float g[32];
#pragma unroll
    for (int k = 0; k < 32; k++) g[k] = 0.f;
    if (live && half == 0) {
      float x[25], gp[24];
#pragma unroll
      for (int k = 0; k < 25; k++) x[k] = k < a.n_out ? a.logits[(size_t)i * a.n_out + k] : 0.f;
      const float4* gw = reinterpret_cast<const float4*>(b.dL_dweights + (size_t)i * 24);
#pragma unroll
      for (int k = 0; k < 6; k++) { const float4 t = gw[k]; gp[4 * k] = t.x; gp[4 * k + 1] = t.y; gp[4 * k + 2] = t.z; gp[4 * k + 3] = t.w; }
      if (a.n_out == 25) {
        hierarchical_softmax_bwd(x, gp, g);
      } else {
        float m = x[0], p[24];
        for (int k = 1; k < 24; k++) m = fmaxf(m, x[k]);
        float sum = 0.f, dot = 0.f;
        for (int k = 0; k < 24; k++) { p[k] = expf(x[k] - m); sum += p[k]; }
        const float inv = 1.f / sum;
        for (int k = 0; k < 24; k++) { p[k] *= inv; dot += p[k] * gp[k]; }
        for (int k = 0; k < 24; k++) g[k] = p[k] * (gp[k] - dot);
      }
    }
    if (half == 0) store_grad32(sm, row, 0, g, sm.dbias[4], lane);
    // ---- output layer: Z4 = A3 . W4^T ----
    bw_wait_loads(sm, ld_phase);
    bw_row_signs(sm, row, 64 * half, 64 * half + 64, signs);  // of A3 (this thread's columns)
    bw_sync_mma_inputs();
    if (tid == 0) {
      issue_dw(tmem + 128, sm.A_hi, sm.A_lo, sm.G_hi, sm.G_lo, SK_NOUT);   // dW4^T [k_in][n_out]
      issue_da(tmem, sm.G_hi, sm.G_lo, sm.W_hi, sm.W_lo, SK_NOUT, NR_H);   // dA3
    }
    bw_commit_wait(sm, tid == 0, phase);
    if (tid == 0)  // A / W are free again: A2 and W3^T
      bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st + SK_SAVE_IN + 2 * NR_SAVE_LAYER, st + SK_SAVE_IN + 2 * NR_SAVE_LAYER + 128 * 128,
                         128 * 128 * 2, sm.W_hi, sm.W_lo, w.wht_hi[2], w.wht_lo[2], NR_H * NR_H * 2);
    if (half == 0) {  // warp-uniform
      float v[32];
      tmem_ld32(tmem_row + 128, v);
#pragma unroll
      for (int n = 0; n < 25; n++)
        if (n < a.n_out) atomicAdd(b.dL_dW[4] + (size_t)n * NR_H + row, v[n]);
    }
#pragma unroll 1
    for (int hl = 0; hl < 4; hl++) {  // hl = 0..2: layers 3..1 (inputs A2..A0), hl = 3: layer 0 (input = normalised xyz)
#pragma unroll 1
      for (int c = 2 * half; c < 2 * half + 2; c++) {  // this thread's 64 columns
        float v[32];
        tmem_ld32(tmem_row + c * 32, v);
#pragma unroll
        for (int j = 0; j < 32; j += 8) lrelu_mask8_bits(signs, c * 32 + j, v + j);
        store_grad32(sm, row, c * 32, v, sm.dbias[3 - hl], lane);
      }
      const int kin = hl < 3 ? NR_H : SK_K0;
      bw_wait_loads(sm, ld_phase);                             // this layer's A_in and W^T
      if (hl < 3) bw_row_signs(sm, row, 64 * half, 64 * half + 64, signs);  // for the next epilogue
      bw_sync_mma_inputs();                                    // G complete; every row has read its dA from TMEM and its signs
      if (tid == 0) {
        issue_dw(tmem + 128, sm.G_hi, sm.G_lo, sm.A_hi, sm.A_lo, kin);
        issue_da(tmem, sm.G_hi, sm.G_lo, sm.W_hi, sm.W_lo, NR_H, kin);
      }
      bw_commit_wait(sm, tid == 0, phase);
      if (tid == 0 && hl < 3) {  // the next layer's operands stream in during the dW flush and the epilogue
        if (hl < 2)
          bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st + SK_SAVE_IN + (size_t)(1 - hl) * NR_SAVE_LAYER,
                             st + SK_SAVE_IN + (size_t)(1 - hl) * NR_SAVE_LAYER + 128 * 128, 128 * 128 * 2, sm.W_hi, sm.W_lo,
                             w.wht_hi[1 - hl], w.wht_lo[1 - hl], NR_H * NR_H * 2);
        else
          bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st, st + 128 * SK_K0, 128 * SK_K0 * 2, sm.W_hi, sm.W_lo, w.w0t_hi, w.w0t_lo,
                             SK_K0 * NR_H * 2);
      }
      if (hl < 3) {
        flush_dw_rows(tmem_row + 128, 64 * half, 64 * half + 64, b.dL_dW[3 - hl] + (size_t)row * NR_H);
      } else if (half == 0) {
        float v[32];
        tmem_ld32(tmem_row + 128, v);
        for (int k = 0; k < 3; k++) atomicAdd(b.dL_dW[0] + (size_t)row * 3 + k, v[k]);
      }
    }
    if (half == 0) {
      float dx[32];
      tmem_ld32(tmem_row, dx);  // dL/d(normalised xyz) in columns 0..2
      if (live)
        for (int k = 0; k < 3; k++) b.dL_dxyz[3 * (size_t)i + k] = dx[k] * 2.f / (a.aabb_max[k] - a.aabb_min[k]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
  }
  __syncthreads();
  for (int l = 0; l < 5 && tid < BW_H; l++) {
    const float v = sm.dbias[l][tid];
    if (v != 0.f && (l < 4 || tid < a.n_out)) atomicAdd(b.dL_db[l] + tid, v);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

}  // namespace

extern "C" int skinning_backward(const SkinBackwardArgs* args, void* workspace, size_t workspace_bytes, const void* saved,
                                 size_t saved_bytes, void* stream) {
  if (!args || !workspace || !saved) return GS_ERR_BAD_ARG;
  if (((uintptr_t)workspace | (uintptr_t)saved) & 15) return GS_ERR_BAD_ARG;  // bulk copies: 16-byte aligned operands
  const SkinBackwardArgs& b = *args;
  const SkinArgs& a = b.fwd;
  if (a.P < 0 || a.hidden != NR_H || (a.n_out != 25 && a.n_out != 24) || !a.aabb_min || !a.aabb_max || !a.xyz || !a.logits)
    return GS_ERR_BAD_ARG;
  if (!b.dL_dweights || !b.dL_dxyz) return GS_ERR_BAD_ARG;
  for (int l = 0; l < 5; l++) if (!a.W[l] || !a.b[l] || !b.dL_dW[l] || !b.dL_db[l]) return GS_ERR_BAD_ARG;
  if (saved_bytes < skinning_saved_bytes(a.P)) return GS_ERR_WORKSPACE_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  SkWs w;
  if (carve_sk((char*)workspace, &w) > workspace_bytes) return GS_ERR_WORKSPACE_SMALL;
  const size_t f = sizeof(float);
  const size_t wn[5] = {(size_t)NR_H * 3, (size_t)NR_H * NR_H, (size_t)NR_H * NR_H, (size_t)NR_H * NR_H, (size_t)a.n_out * NR_H};
  for (int l = 0; l < 5; l++) {
    GS_TRY_NR(cudaMemsetAsync(b.dL_dW[l], 0, wn[l] * f, st));
    GS_TRY_NR(cudaMemsetAsync(b.dL_db[l], 0, (size_t)(l < 4 ? NR_H : a.n_out) * f, st));
  }
  if (a.P == 0) return GS_OK;
  sk_prep_bwd_kernel<<<64, 256, 0, st>>>(a, w);
  const int smem = (int)sizeof(BwSmem) + 1024;
  if (gs_check(cudaFuncSetAttribute(sk_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != GS_OK) return GS_ERR_CUDA;
  const int ntiles = (a.P + NR_TILE - 1) / NR_TILE;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  sk_bwd_kernel<<<ntiles < sms ? ntiles : sms, BW_THREADS, smem, st>>>(b, w, (const __nv_bfloat16*)saved, ntiles);
  return gs_check(cudaGetLastError());
}
