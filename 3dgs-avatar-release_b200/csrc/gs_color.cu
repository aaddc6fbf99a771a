// gs_color.cu -- colour MLP texture, forward (replaces ColorMLP.forward, models/texture/texture.py:40-125; see
// include/gsplat_b200.h for the reference lines of every piece).
//
// Three 64-wide linear layers per Gaussian: the tensor-core workload of the default configuration.  Persistent kernel,
// B200-first: every CTA keeps ALL weights of the three layers resident in shared memory (BF16 hi/lo, operand layout,
// 36 KB) and walks tiles of 128 Gaussians (= the M of a 128xNx16 tcgen05.mma).  Per tile each thread composes its
// Gaussian's input row (feature | SH basis of the canonical view direction | non-rigid feature) straight into the A
// operand, one elected thread issues the MMAs of a layer (accumulators in TMEM), and the epilogue
// (tcgen05.ld -> bias -> LeakyReLU -> BF16 split) writes the next layer's A operand back into shared memory.  The
// per-frame latent code is folded into the first bias by the prep kernel.
#include "gs_umma.cuh"
#include "gs_mlp_bwd.cuh"

namespace {

constexpr int CL_H = 64;     // hidden width (configs/texture/shallow_mlp.yaml: n_neurons 64)
constexpr int CL_K0 = 64;    // per-Gaussian inputs padded to 64 (32 + 15 + 16 = 63 in the default configuration)
constexpr int CL_NOUT = 16;  // 3 colour channels padded to the smallest UMMA N

struct ClWs {
  __nv_bfloat16 *w0_hi, *w0_lo;  // [64 x 64]
  __nv_bfloat16 *w1_hi, *w1_lo;  // [64 x 64]
  __nv_bfloat16 *w2_hi, *w2_lo;  // [16 x 64]
  float* bias;                   // [64] b0 + W0[:, per-Gaussian inputs:] . latent | [64] b1 | [16] b2
  // backward: W^T as the B operand of dA = G . W (element (row = k_in, col = n_out) = W[n_out][k_in])
  __nv_bfloat16 *w0t_hi, *w0t_lo;  // [64 x 64]
  __nv_bfloat16 *w1t_hi, *w1t_lo;  // [64 x 64]
  __nv_bfloat16 *w2t_hi, *w2t_lo;  // [64 x 16]
};
// activations kept for the backward, per tile of 128 Gaussians: composed input, lin0 and lin1 activations, each [128 x 64] hi|lo
constexpr size_t CL_SAVE_LAYER = 2 * (size_t)128 * 64, CL_SAVE_TILE = 3 * CL_SAVE_LAYER;

__host__ size_t carve_cl(char* base, ClWs* w) {
  char* p = base;
  ClWs t;
  gs_carve(p, t.w0_hi, (size_t)CL_H * CL_K0); gs_carve(p, t.w0_lo, (size_t)CL_H * CL_K0);
  gs_carve(p, t.w1_hi, (size_t)CL_H * CL_H); gs_carve(p, t.w1_lo, (size_t)CL_H * CL_H);
  gs_carve(p, t.w2_hi, (size_t)CL_NOUT * CL_H); gs_carve(p, t.w2_lo, (size_t)CL_NOUT * CL_H);
  gs_carve(p, t.bias, (size_t)(2 * CL_H + CL_NOUT));
  gs_carve(p, t.w0t_hi, (size_t)CL_K0 * CL_H); gs_carve(p, t.w0t_lo, (size_t)CL_K0 * CL_H);
  gs_carve(p, t.w1t_hi, (size_t)CL_H * CL_H); gs_carve(p, t.w1t_lo, (size_t)CL_H * CL_H);
  gs_carve(p, t.w2t_hi, (size_t)CL_H * CL_NOUT); gs_carve(p, t.w2t_lo, (size_t)CL_H * CL_NOUT);
  if (w) *w = t;
  return (size_t)(p - base);
}

__host__ __device__ inline int cl_n_sh(int deg) { return (deg + 1) * (deg + 1) - 1; }

__global__ void color_prep_kernel(ColorArgs a, ClWs w) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  const int n_in = a.feature_dim + cl_n_sh(a.sh_degree) + a.non_rigid_dim;  // per-Gaussian inputs
  const int d_in = n_in + a.latent_dim;
  for (int e = tid; e < CL_H * CL_K0; e += nth) {
    const int n = e / CL_K0, k = e % CL_K0;
    const float v = k < n_in ? a.W0[(size_t)n * d_in + k] : 0.f;
    split_bf16(v, w.w0_hi[core_idx(n, k, CL_H)], w.w0_lo[core_idx(n, k, CL_H)]);
  }
  for (int e = tid; e < CL_H * CL_H; e += nth) {
    const int n = e / CL_H, k = e % CL_H;
    split_bf16(a.W1[e], w.w1_hi[core_idx(n, k, CL_H)], w.w1_lo[core_idx(n, k, CL_H)]);
  }
  for (int e = tid; e < CL_NOUT * CL_H; e += nth) {
    const int n = e / CL_H, k = e % CL_H;
    const float v = n < 3 ? a.W2[(size_t)n * CL_H + k] : 0.f;
    split_bf16(v, w.w2_hi[core_idx(n, k, CL_NOUT)], w.w2_lo[core_idx(n, k, CL_NOUT)]);
  }
  for (int n = tid; n < CL_H; n += nth) {
    float acc = a.b0[n];
    for (int c = 0; c < a.latent_dim; c++) acc = fmaf(a.W0[(size_t)n * d_in + n_in + c], a.latent[c], acc);
    w.bias[n] = acc;
    w.bias[CL_H + n] = a.b1[n];
    if (n < CL_NOUT) w.bias[2 * CL_H + n] = n < 3 ? a.b2[n] : 0.f;
  }
}

// eval_sh_bases(deg, dir)[1:], utils/sh_utils.py:115-161 (same constants as eval_sh)
__device__ __forceinline__ void sh_bases_1(int deg, float x, float y, float z, float* o) {
  const float C1 = 0.4886025119029199f;
  const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f};
  const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                       -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
  if (deg > 0) {
    o[0] = -C1 * y; o[1] = C1 * z; o[2] = -C1 * x;
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      o[3] = C2[0] * xy; o[4] = C2[1] * yz; o[5] = C2[2] * (2.0f * zz - xx - yy); o[6] = C2[3] * xz; o[7] = C2[4] * (xx - yy);
      if (deg > 2) {
        o[8] = C3[0] * y * (3 * xx - yy);
        o[9] = C3[1] * xy * z;
        o[10] = C3[2] * y * (4 * zz - xx - yy);
        o[11] = C3[3] * z * (2 * zz - 3 * xx - 3 * yy);
        o[12] = C3[4] * x * (4 * zz - xx - yy);
        o[13] = C3[5] * z * (xx - yy);
        o[14] = C3[6] * x * (xx - 3 * yy);
      }
    }
  }
}

struct ClSmem {
  __nv_bfloat16 A_hi[NR_TILE * CL_K0], A_lo[NR_TILE * CL_K0];  // activations of the current tile (16 KB each)
  __nv_bfloat16 W0_hi[CL_H * CL_K0], W0_lo[CL_H * CL_K0];      // resident weights (8 KB each)
  __nv_bfloat16 W1_hi[CL_H * CL_H], W1_lo[CL_H * CL_H];
  __nv_bfloat16 W2_hi[CL_NOUT * CL_H], W2_lo[CL_NOUT * CL_H];  // 2 KB each
  float bias[2 * CL_H + CL_NOUT];
  unsigned long long bar;
  uint32_t tmem_base;
};

// bias + LeakyReLU(0.01) of a 64-wide hidden layer: this thread's TMEM row -> next layer's A operand
__device__ __forceinline__ void cl_hidden_epilogue(ClSmem& sm, uint32_t tmem_row, int row, const float* bias) {
#pragma unroll 1
  for (int c = 0; c < CL_H / 32; c++) {
    float v[32];
    tmem_ld32(tmem_row + c * 32, v);
#pragma unroll
    for (int j = 0; j < 32; j++) {
      const float t = v[j] + bias[c * 32 + j];
      v[j] = t > 0.f ? t : 0.01f * t;  // nn.LeakyReLU() default slope
    }
#pragma unroll
    for (int j = 0; j < 32; j += 8) store_act8(sm.A_hi, sm.A_lo, row, c * 32 + j, v + j);
  }
}

// generic-proxy writes of the A operand -> visible to the tensor core, all threads past the hand-over point
__device__ __forceinline__ void cl_publish() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// FAST: the default configuration (feature 32 | SH deg 3 -> 15 | non-rigid 16): every index of the input row is a
// compile-time constant, so the row lives in registers (the generic variant keeps it in local memory).
template <bool FAST>
__global__ void __launch_bounds__(NR_TILE, 3) color_fwd_kernel(ColorArgs a, ClWs w, int n_tiles, __nv_bfloat16* __restrict__ saved) {
  extern __shared__ __align__(1024) unsigned char cl_smem_raw[];
  ClSmem& sm = *reinterpret_cast<ClSmem*>(cl_smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, row = tid;
  if (warp == 0) {  // TMEM: 64 FP32 accumulator columns x 128 lanes
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "r"(64u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    nr_mbar_init(&sm.bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // weights of all three layers: loaded once per CTA, resident for every tile
  copy_weights(sm.W0_hi, w.w0_hi, CL_H * CL_K0); copy_weights(sm.W0_lo, w.w0_lo, CL_H * CL_K0);
  copy_weights(sm.W1_hi, w.w1_hi, CL_H * CL_H); copy_weights(sm.W1_lo, w.w1_lo, CL_H * CL_H);
  copy_weights(sm.W2_hi, w.w2_hi, CL_NOUT * CL_H); copy_weights(sm.W2_lo, w.w2_lo, CL_NOUT * CL_H);
  for (int k = tid; k < 2 * CL_H + CL_NOUT; k += NR_TILE) sm.bias[k] = w.bias[k];
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tmem_row = tmem + ((uint32_t)(warp * 32) << 16);  // this warp's 32-lane quadrant
  const int F = FAST ? 32 : a.feature_dim, S = FAST ? 15 : cl_n_sh(a.sh_degree), R = FAST ? 16 : a.non_rigid_dim, n_in = F + S + R;
  const int deg = FAST ? 3 : a.sh_degree;
  float cam[3] = {0.f, 0.f, 0.f};
  if (S > 0) { cam[0] = a.camera_center[0]; cam[1] = a.camera_center[1]; cam[2] = a.camera_center[2]; }
  uint32_t phase = 0;

#pragma unroll 1
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int i = tile * NR_TILE + row;
    const bool live = i < a.P;
    // ---- compose_input (texture.py:74-118) straight into the A operand ----
    {
      float in[CL_K0];
#pragma unroll
      for (int k = 0; k < CL_K0; k++) in[k] = 0.f;
      if (live) {
        const float* f = a.features + (size_t)i * F;
        if (FAST) {  // 128-byte rows: 8 x LDG.128
#pragma unroll
          for (int k = 0; k < 32; k += 4) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(f + k));
            in[k] = v.x; in[k + 1] = v.y; in[k + 2] = v.z; in[k + 3] = v.w;
          }
        } else {
          for (int k = 0; k < F; k++) in[k] = f[k];
        }
        if (S > 0) {
          float dx = a.xyz[3 * (size_t)i] - cam[0], dy = a.xyz[3 * (size_t)i + 1] - cam[1], dz = a.xyz[3 * (size_t)i + 2] - cam[2];
          if (a.cano_view_dir) {  // R_bwd = T_fwd[:3,:3]^T; dir <- R_bwd . dir
            const float* T = a.T_fwd + 16 * (size_t)i;
            const float cx = T[0] * dx + T[4] * dy + T[8] * dz;
            const float cy = T[1] * dx + T[5] * dy + T[9] * dz;
            const float cz = T[2] * dx + T[6] * dy + T[10] * dz;
            dx = cx; dy = cy; dz = cz;
            if (a.view_noise) {  // dir <- dir @ view_noise (row vector times matrix)
              const float* N = a.view_noise;
              const float nx = dx * N[0] + dy * N[3] + dz * N[6];
              const float ny = dx * N[1] + dy * N[4] + dz * N[7];
              const float nz = dx * N[2] + dy * N[5] + dz * N[8];
              dx = nx; dy = ny; dz = nz;
            }
          }
          const float inv = 1.0f / (sqrtf(dx * dx + dy * dy + dz * dz) + 1e-12f);
          float sh[15];
          sh_bases_1(deg, dx * inv, dy * inv, dz * inv, sh);
          if (FAST) {
#pragma unroll
            for (int k = 0; k < 15; k++) in[32 + k] = sh[k];
          } else {
            for (int k = 0; k < S; k++) in[F + k] = sh[k];
          }
        }
        if (R > 0) {
          const float* nr = a.non_rigid_feature + (size_t)i * R;
          if (FAST) {
#pragma unroll
            for (int k = 0; k < 16; k += 4) {
              const float4 v = __ldg(reinterpret_cast<const float4*>(nr + k));
              in[47 + k] = v.x; in[48 + k] = v.y; in[49 + k] = v.z; in[50 + k] = v.w;
            }
          } else {
            for (int k = 0; k < R; k++) in[F + S + k] = nr[k];
          }
        }
        if (a.inputs) {
          if (FAST) {
#pragma unroll
            for (int k = 0; k < 63; k++) a.inputs[(size_t)i * 63 + k] = in[k];
          } else {
            for (int k = 0; k < n_in; k++) a.inputs[(size_t)i * n_in + k] = in[k];
          }
        }
      }
#pragma unroll
      for (int k = 0; k < CL_K0; k += 8) store_act8(sm.A_hi, sm.A_lo, row, k, in + k);
    }
    cl_publish();
    __nv_bfloat16* const save_tile = saved ? saved + (size_t)tile * CL_SAVE_TILE : nullptr;
    auto save_act = [&](int which) {  // whole CTA, after a cl_publish (every row written, no MMA reading A yet matters: reads only)
      if (!save_tile) return;
      const uint4* sh = reinterpret_cast<const uint4*>(sm.A_hi);
      const uint4* sl = reinterpret_cast<const uint4*>(sm.A_lo);
      uint4* d = reinterpret_cast<uint4*>(save_tile + (size_t)which * CL_SAVE_LAYER);
      for (int q = tid; q < 128 * 64 * 2 / 16; q += NR_TILE) { d[q] = sh[q]; d[q + 128 * 64 * 2 / 16] = sl[q]; }
    };
    save_act(0);
    // ---- lin0 ----
    if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W0_hi, sm.W0_lo, CL_K0, CL_H, &sm.bar);
    nr_mbar_wait(&sm.bar, phase); phase ^= 1u;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    cl_hidden_epilogue(sm, tmem_row, row, sm.bias);
    cl_publish();
    save_act(1);
    // ---- lin1 ----
    if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W1_hi, sm.W1_lo, CL_H, CL_H, &sm.bar);
    nr_mbar_wait(&sm.bar, phase); phase ^= 1u;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    cl_hidden_epilogue(sm, tmem_row, row, sm.bias + CL_H);
    cl_publish();
    save_act(2);
    // ---- lin2 (N = 16) + sigmoid ----
    if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W2_hi, sm.W2_lo, CL_H, CL_NOUT, &sm.bar);
    nr_mbar_wait(&sm.bar, phase); phase ^= 1u;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    float d[32];
    tmem_ld32(tmem_row, d);  // columns >= 16 hold stale values of the previous layer: never used
    if (live) {
#pragma unroll
      for (int k = 0; k < 3; k++) a.colors[3 * (size_t)i + k] = 1.0f / (1.0f + expf(-(d[k] + sm.bias[2 * CL_H + k])));
    }
    // the next tile's input rows overwrite A, and its first MMA overwrites the accumulators this thread just read
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64u) : "memory");
}

}  // namespace

extern "C" size_t color_workspace_bytes(const ColorArgs* args) { (void)args; return carve_cl(nullptr, nullptr); }

extern "C" size_t color_saved_bytes(int32_t P) { return (size_t)((P + NR_TILE - 1) / NR_TILE) * CL_SAVE_TILE * sizeof(__nv_bfloat16); }

static int color_inputs_ok(const ColorArgs& a) {
  if (a.P < 0 || a.hidden != CL_H || a.feature_dim < 0 || a.sh_degree < 0 || a.sh_degree > 3 || a.non_rigid_dim < 0 || a.latent_dim < 0)
    return 0;
  const int n_in = a.feature_dim + cl_n_sh(a.sh_degree) + a.non_rigid_dim;
  if (n_in <= 0 || n_in > CL_K0) return 0;
  if ((a.feature_dim > 0 && !a.features) || !a.W0 || !a.b0 || !a.W1 || !a.b1 || !a.W2 || !a.b2) return 0;
  if (a.sh_degree > 0 && (!a.xyz || !a.camera_center || (a.cano_view_dir && !a.T_fwd))) return 0;
  if ((a.non_rigid_dim > 0 && !a.non_rigid_feature) || (a.latent_dim > 0 && !a.latent)) return 0;
  return 1;
}

static int color_forward_impl(const ColorArgs* args, void* workspace, size_t workspace_bytes, void* saved, void* stream);
extern "C" int color_forward(const ColorArgs* args, void* workspace, size_t workspace_bytes, void* stream) {
  return color_forward_impl(args, workspace, workspace_bytes, nullptr, stream);
}
extern "C" int color_forward_train(const ColorArgs* args, void* workspace, size_t workspace_bytes, void* saved, size_t saved_bytes,
                                   void* stream) {
  if (!args || !saved || saved_bytes < color_saved_bytes(args->P)) return GS_ERR_WORKSPACE_SMALL;
  return color_forward_impl(args, workspace, workspace_bytes, saved, stream);
}
static int color_forward_impl(const ColorArgs* args, void* workspace, size_t workspace_bytes, void* saved, void* stream) {
  if (!args || !workspace) return GS_ERR_BAD_ARG;
  const ColorArgs& a = *args;
  if (a.P < 0 || a.hidden != CL_H || a.feature_dim < 0 || a.sh_degree < 0 || a.sh_degree > 3 || a.non_rigid_dim < 0 || a.latent_dim < 0)
    return GS_ERR_BAD_ARG;
  const int n_in = a.feature_dim + cl_n_sh(a.sh_degree) + a.non_rigid_dim;
  if (n_in <= 0 || n_in > CL_K0) return GS_ERR_BAD_ARG;
  if ((a.feature_dim > 0 && !a.features) || !a.colors || !a.W0 || !a.b0 || !a.W1 || !a.b1 || !a.W2 || !a.b2) return GS_ERR_BAD_ARG;
  if (a.sh_degree > 0 && (!a.xyz || !a.camera_center || (a.cano_view_dir && !a.T_fwd))) return GS_ERR_BAD_ARG;
  if ((a.non_rigid_dim > 0 && !a.non_rigid_feature) || (a.latent_dim > 0 && !a.latent)) return GS_ERR_BAD_ARG;
  if (a.P == 0) return GS_OK;
  cudaStream_t st = (cudaStream_t)stream;
  ClWs w;
  if (carve_cl((char*)workspace, &w) > workspace_bytes) return GS_ERR_WORKSPACE_SMALL;
  color_prep_kernel<<<16, 256, 0, st>>>(a, w);
  int sms = 0;
  {  // per call: function attributes are per DEVICE, and one process may drive several
    if (gs_check(cudaFuncSetAttribute(color_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ClSmem) + 1024)) != GS_OK ||
        gs_check(cudaFuncSetAttribute(color_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ClSmem) + 1024)) != GS_OK)
      return GS_ERR_CUDA;
    int dev = 0;
    if (gs_check(cudaGetDevice(&dev)) != GS_OK || gs_check(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != GS_OK)
      return GS_ERR_CUDA;
  }
  const int n_tiles = (a.P + NR_TILE - 1) / NR_TILE;
  const int grid = n_tiles < 3 * sms ? n_tiles : 3 * sms;  // persistent: three CTAs per SM (70 KB of shared memory, 64 TMEM columns each)
  const bool aligned = ((uintptr_t)a.features % 16 == 0) && (!a.non_rigid_feature || (uintptr_t)a.non_rigid_feature % 16 == 0);
  if (a.feature_dim == 32 && a.sh_degree == 3 && a.non_rigid_dim == 16 && aligned)
    color_fwd_kernel<true><<<grid, NR_TILE, sizeof(ClSmem) + 1024, st>>>(a, w, n_tiles, (__nv_bfloat16*)saved);
  else
    color_fwd_kernel<false><<<grid, NR_TILE, sizeof(ClSmem) + 1024, st>>>(a, w, n_tiles, (__nv_bfloat16*)saved);
  return gs_check(cudaGetLastError());
}

// ================================= colour MLP: backward =================================
// Same tensor-core engine as the 128-wide MLPs (gs_mlp_bwd.cuh): the 64-wide layers use the first 64 units of the 128-unit
// operand buffers; units 64..127 of the buffers that serve as the M operand of dW = G^T . A stay zero.
namespace {

__global__ void color_prep_bwd_kernel(ColorArgs a, ClWs w) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  const int n_in = a.feature_dim + cl_n_sh(a.sh_degree) + a.non_rigid_dim;
  const int d_in = n_in + a.latent_dim;
  for (int e = tid; e < CL_K0 * CL_H; e += nth) {
    const int k = e / CL_H, n = e % CL_H;
    const float v = k < n_in ? a.W0[(size_t)n * d_in + k] : 0.f;
    split_bf16(v, w.w0t_hi[core_idx(k, n, CL_K0)], w.w0t_lo[core_idx(k, n, CL_K0)]);
    split_bf16(a.W1[(size_t)n * CL_H + k], w.w1t_hi[core_idx(k, n, CL_H)], w.w1t_lo[core_idx(k, n, CL_H)]);
  }
  for (int e = tid; e < CL_H * CL_NOUT; e += nth) {
    const int k = e / CL_NOUT, n = e % CL_NOUT;
    const float v = n < 3 ? a.W2[(size_t)n * CL_H + k] : 0.f;
    split_bf16(v, w.w2t_hi[core_idx(k, n, CL_H)], w.w2t_lo[core_idx(k, n, CL_H)]);
  }
}

// adjoint of sh_bases_1: g = dL/d(basis values) -> dL/d(unit direction)
__device__ __forceinline__ void sh_bases_1_bwd(int deg, float x, float y, float z, const float* g, float* d) {
  const float C1 = 0.4886025119029199f;
  const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f};
  const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                       -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
  float dx = 0.f, dy = 0.f, dz = 0.f;
  if (deg > 0) {
    dy += -C1 * g[0]; dz += C1 * g[1]; dx += -C1 * g[2];
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z;
      dx += C2[0] * y * g[3]; dy += C2[0] * x * g[3];
      dy += C2[1] * z * g[4]; dz += C2[1] * y * g[4];
      dx += C2[2] * -2.f * x * g[5]; dy += C2[2] * -2.f * y * g[5]; dz += C2[2] * 4.f * z * g[5];
      dx += C2[3] * z * g[6]; dz += C2[3] * x * g[6];
      dx += C2[4] * 2.f * x * g[7]; dy += C2[4] * -2.f * y * g[7];
      if (deg > 2) {
        // o8 = C3[0] y (3xx - yy)
        dx += C3[0] * 6.f * x * y * g[8]; dy += C3[0] * (3.f * xx - 3.f * yy) * g[8];
        // o9 = C3[1] x y z
        dx += C3[1] * y * z * g[9]; dy += C3[1] * x * z * g[9]; dz += C3[1] * x * y * g[9];
        // o10 = C3[2] y (4zz - xx - yy)
        dx += C3[2] * -2.f * x * y * g[10]; dy += C3[2] * (4.f * zz - xx - 3.f * yy) * g[10]; dz += C3[2] * 8.f * y * z * g[10];
        // o11 = C3[3] z (2zz - 3xx - 3yy)
        dx += C3[3] * -6.f * x * z * g[11]; dy += C3[3] * -6.f * y * z * g[11]; dz += C3[3] * (6.f * zz - 3.f * xx - 3.f * yy) * g[11];
        // o12 = C3[4] x (4zz - xx - yy)
        dx += C3[4] * (4.f * zz - 3.f * xx - yy) * g[12]; dy += C3[4] * -2.f * x * y * g[12]; dz += C3[4] * 8.f * x * z * g[12];
        // o13 = C3[5] z (xx - yy)
        dx += C3[5] * 2.f * x * z * g[13]; dy += C3[5] * -2.f * y * z * g[13]; dz += C3[5] * (xx - yy) * g[13];
        // o14 = C3[6] x (xx - 3yy)
        dx += C3[6] * (3.f * xx - 3.f * yy) * g[14]; dy += C3[6] * -6.f * x * y * g[14];
      }
    }
  }
  d[0] = dx; d[1] = dy; d[2] = dz;
}

__global__ void __launch_bounds__(BW_THREADS, 1)
color_bwd_kernel(ColorBackwardArgs b, ClWs w, const __nv_bfloat16* __restrict__ saved, int ntiles) {
  extern __shared__ __align__(1024) unsigned char bw_smem_raw[];
  BwSmem& sm = *reinterpret_cast<BwSmem*>(bw_smem_raw);
  const ColorArgs& a = b.fwd;
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
  {  // units 64..127 of G and A are never written afterwards: zero the buffers once
    uint4* z = reinterpret_cast<uint4*>(sm.G_hi);  // G_hi, G_lo, A_hi, A_lo are contiguous
    for (int q = tid; q < 4 * BW_H * BW_H * 2 / 16; q += BW_THREADS) z[q] = make_uint4(0u, 0u, 0u, 0u);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = sm.tmem_base;
  const uint32_t tmem_row = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const int F = a.feature_dim, S = cl_n_sh(a.sh_degree), R = a.non_rigid_dim, n_in = F + S + R, d_in = n_in + a.latent_dim;
  uint32_t phase = 0, ld_phase = 0;
  uint32_t signs[4];  // LeakyReLU branches of this thread's row in the activations loaded last
  const int HALF = 128 * 64;  // elements of one saved half (hi or lo)

  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int i = tile * NR_TILE + row;
    const bool live = i < a.P;
    const __nv_bfloat16* const st = saved + (size_t)tile * CL_SAVE_TILE;
    if (tid == 0)  // operands of lin2 (A1, W2^T) stream in while the sigmoid adjoint runs (A / W are free: see the tile's last barrier)
      bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st + 2 * CL_SAVE_LAYER, st + 2 * CL_SAVE_LAYER + HALF, HALF * 2, sm.W_hi, sm.W_lo,
                         w.w2t_hi, w.w2t_lo, CL_H * CL_NOUT * 2);
    float g[32];
#pragma unroll
    for (int k = 0; k < 32; k++) g[k] = 0.f;
    if (live && half == 0) {
#pragma unroll
      for (int k = 0; k < 3; k++) {  // sigmoid'
        const float c = a.colors[3 * (size_t)i + k];
        g[k] = b.dL_dcolors[3 * (size_t)i + k] * c * (1.f - c);
      }
    }
    if (half == 0) store_grad32(sm, row, 0, g, sm.dbias[2], lane);
    // ---- lin2: Z2 = A1 . W2^T ----
    bw_wait_loads(sm, ld_phase);
    bw_row_signs(sm, row, 32 * half, 32 * half + 32, signs);  // of A1 (this thread's 32 columns)
    bw_sync_mma_inputs();
    if (tid == 0) {
      issue_dw(tmem + 128, sm.A_hi, sm.A_lo, sm.G_hi, sm.G_lo, CL_NOUT);    // dW2^T [k_in][n_out] (rows >= 64 are zero)
      issue_da(tmem, sm.G_hi, sm.G_lo, sm.W_hi, sm.W_lo, CL_NOUT, CL_H);    // dA1
    }
    bw_commit_wait(sm, tid == 0, phase);
    if (tid == 0)  // A / W are free again: A0 and W1^T
      bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st + CL_SAVE_LAYER, st + CL_SAVE_LAYER + HALF, HALF * 2, sm.W_hi, sm.W_lo, w.w1t_hi,
                         w.w1t_lo, CL_H * CL_H * 2);
    if (half == 0) {  // warp-uniform
      float v[32];
      tmem_ld32(tmem_row + 128, v);
      if (row < CL_H)
        for (int n = 0; n < 3; n++) atomicAdd(b.dL_dW2 + (size_t)n * CL_H + row, v[n]);
    }
#pragma unroll 1
    for (int hl = 0; hl < 2; hl++) {  // hl = 0: lin1 (input A0), hl = 1: lin0 (input = composed per-Gaussian row)
      {
        const int c = half;  // this thread's 32 of the 64 columns
        float v[32];
        tmem_ld32(tmem_row + c * 32, v);
#pragma unroll
        for (int j = 0; j < 32; j += 8) lrelu_mask8_bits(signs, c * 32 + j, v + j);
        store_grad32(sm, row, c * 32, v, sm.dbias[1 - hl], lane);
      }
      bw_wait_loads(sm, ld_phase);                          // this layer's A_in and W^T
      if (hl == 0) bw_row_signs(sm, row, 32 * half, 32 * half + 32, signs);  // of A0, for the next epilogue
      bw_sync_mma_inputs();                                 // G complete; every row has read its dA from TMEM and its signs
      if (tid == 0) {
        issue_dw(tmem + 128, sm.G_hi, sm.G_lo, sm.A_hi, sm.A_lo, CL_H);     // dW [n_out][k_in] (rows >= 64 are zero)
        issue_da(tmem, sm.G_hi, sm.G_lo, sm.W_hi, sm.W_lo, CL_H, CL_H);     // dA_in
      }
      bw_commit_wait(sm, tid == 0, phase);
      if (tid == 0 && hl == 0)  // lin0's operands (composed input, W0^T) stream in during the dW flush and the epilogue
        bulk_load_operands(&sm.ld_bar, sm.A_hi, sm.A_lo, st, st + HALF, HALF * 2, sm.W_hi, sm.W_lo, w.w0t_hi, w.w0t_lo, CL_H * CL_H * 2);
      if (row < CL_H) {
        if (hl == 0) {
          flush_dw_rows(tmem_row + 128, 32 * half, 32 * half + 32, b.dL_dW1 + (size_t)row * CL_H);
        } else {  // rows of W0 are d_in floats apart (not 16-byte aligned in general): scalar reductions
          {
            const int c = 32 * half;
            float v[32];
            tmem_ld32(tmem_row + 128 + c, v);
#pragma unroll
            for (int j = 0; j < 32; j++)
              if (c + j < n_in) atomicAdd(b.dL_dW0 + (size_t)row * d_in + c + j, v[j]);
          }
        }
      }
    }
    // ---- dL/d(composed input) of this thread's Gaussian: features | SH basis of the view direction | non-rigid feature ----
    if (half == 0) {  // warp-uniform: one thread per Gaussian composes the input adjoint
      float din[CL_K0];
      tmem_ld32(tmem_row, din);
      tmem_ld32(tmem_row + 32, din + 32);
      if (live) {
        if (b.dL_dfeatures)
          for (int k = 0; k < F; k++) b.dL_dfeatures[(size_t)i * F + k] = din[k];
        if (b.dL_dnon_rigid_feature)
          for (int k = 0; k < R; k++) b.dL_dnon_rigid_feature[(size_t)i * R + k] = din[F + S + k];
        float gd[3] = {0.f, 0.f, 0.f};   // dL/d(xyz - camera_center)
        float gT[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // dL/dT_fwd[r][c], r, c < 3
        if (S > 0) {
          const float ox = a.xyz[3 * (size_t)i] - a.camera_center[0], oy = a.xyz[3 * (size_t)i + 1] - a.camera_center[1];
          const float oz = a.xyz[3 * (size_t)i + 2] - a.camera_center[2];
          float dx = ox, dy = oy, dz = oz;
          const float* T = a.cano_view_dir ? a.T_fwd + 16 * (size_t)i : nullptr;
          float cx = 0.f, cy = 0.f, cz = 0.f;
          if (a.cano_view_dir) {
            cx = T[0] * dx + T[4] * dy + T[8] * dz; cy = T[1] * dx + T[5] * dy + T[9] * dz; cz = T[2] * dx + T[6] * dy + T[10] * dz;
            dx = cx; dy = cy; dz = cz;
            if (a.view_noise) {
              const float* N = a.view_noise;
              const float nx = dx * N[0] + dy * N[3] + dz * N[6], ny = dx * N[1] + dy * N[4] + dz * N[7], nz = dx * N[2] + dy * N[5] + dz * N[8];
              dx = nx; dy = ny; dz = nz;
            }
          }
          const float r = sqrtf(dx * dx + dy * dy + dz * dz), inv = 1.0f / (r + 1e-12f);
          float gn[3];
          float gsh[15];
          for (int k = 0; k < 15; k++) gsh[k] = k < S ? din[F + k] : 0.f;
          sh_bases_1_bwd(a.sh_degree, dx * inv, dy * inv, dz * inv, gsh, gn);
          // n = d / (r + eps): dL/dd = gn / (r + eps) - d (d . gn) / (r (r + eps)^2)
          const float dot = dx * gn[0] + dy * gn[1] + dz * gn[2];
          const float k2 = r > 0.f ? dot * inv * inv / r : 0.f;
          float g3[3] = {gn[0] * inv - dx * k2, gn[1] * inv - dy * k2, gn[2] * inv - dz * k2};
          if (a.cano_view_dir) {
            if (a.view_noise) {  // d' = d @ N  ->  dL/dd_j = sum_k N[j][k] g'_k
              const float* N = a.view_noise;
              const float t0 = N[0] * g3[0] + N[1] * g3[1] + N[2] * g3[2], t1 = N[3] * g3[0] + N[4] * g3[1] + N[5] * g3[2];
              const float t2 = N[6] * g3[0] + N[7] * g3[1] + N[8] * g3[2];
              g3[0] = t0; g3[1] = t1; g3[2] = t2;
            }
            // c_j = sum_i T[4 i + j] o_i
            const float o[3] = {ox, oy, oz};
            for (int ii = 0; ii < 3; ii++) {
              gd[ii] = T[4 * ii] * g3[0] + T[4 * ii + 1] * g3[1] + T[4 * ii + 2] * g3[2];
              for (int j = 0; j < 3; j++) gT[3 * ii + j] = o[ii] * g3[j];
            }
          } else {
            gd[0] = g3[0]; gd[1] = g3[1]; gd[2] = g3[2];
          }
        }
        if (b.dL_dxyz)
          for (int k = 0; k < 3; k++) b.dL_dxyz[3 * (size_t)i + k] = gd[k];
        if (b.dL_dT_fwd) {
          float* o = b.dL_dT_fwd + 16 * (size_t)i;
          for (int rr = 0; rr < 4; rr++)
            for (int cc = 0; cc < 4; cc++) o[4 * rr + cc] = (rr < 3 && cc < 3) ? gT[3 * rr + cc] : 0.f;
        }
      }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
  }
  __syncthreads();
  if (tid < CL_H) {
    if (sm.dbias[0][tid] != 0.f) atomicAdd(b.dL_db0 + tid, sm.dbias[0][tid]);
    if (sm.dbias[1][tid] != 0.f) atomicAdd(b.dL_db1 + tid, sm.dbias[1][tid]);
    if (tid < 3 && sm.dbias[2][tid] != 0.f) atomicAdd(b.dL_db2 + tid, sm.dbias[2][tid]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

// the latent code enters lin0 through the bias: dL/dW0[n][n_in + c] = db0[n] * latent[c], dL/dlatent[c] = sum_n W0[n][n_in + c] db0[n]
__global__ void color_bwd_finish_kernel(ColorBackwardArgs b) {
  const ColorArgs& a = b.fwd;
  const int n_in = a.feature_dim + cl_n_sh(a.sh_degree) + a.non_rigid_dim, d_in = n_in + a.latent_dim;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < CL_H * a.latent_dim; e += gridDim.x * blockDim.x) {
    const int n = e / a.latent_dim, c = e % a.latent_dim;
    b.dL_dW0[(size_t)n * d_in + n_in + c] = b.dL_db0[n] * a.latent[c];
  }
  if (b.dL_dlatent && blockIdx.x == 0)
    for (int c = threadIdx.x; c < a.latent_dim; c += blockDim.x) {
      float acc = 0.f;
      for (int n = 0; n < CL_H; n++) acc = fmaf(a.W0[(size_t)n * d_in + n_in + c], b.dL_db0[n], acc);
      b.dL_dlatent[c] = acc;
    }
}

}  // namespace

extern "C" int color_backward(const ColorBackwardArgs* args, void* workspace, size_t workspace_bytes, const void* saved,
                              size_t saved_bytes, void* stream) {
  if (!args || !workspace || !saved || !color_inputs_ok(args->fwd)) return GS_ERR_BAD_ARG;
  if (((uintptr_t)workspace | (uintptr_t)saved) & 15) return GS_ERR_BAD_ARG;  // bulk copies: 16-byte aligned operands
  const ColorBackwardArgs& b = *args;
  const ColorArgs& a = b.fwd;
  if (!a.colors || !b.dL_dcolors || !b.dL_dW0 || !b.dL_db0 || !b.dL_dW1 || !b.dL_db1 || !b.dL_dW2 || !b.dL_db2) return GS_ERR_BAD_ARG;
  if (saved_bytes < color_saved_bytes(a.P)) return GS_ERR_WORKSPACE_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  ClWs w;
  if (carve_cl((char*)workspace, &w) > workspace_bytes) return GS_ERR_WORKSPACE_SMALL;
  const int n_in = a.feature_dim + cl_n_sh(a.sh_degree) + a.non_rigid_dim, d_in = n_in + a.latent_dim;
  const size_t f = sizeof(float);
#define CL_TRY(expr) do { if (gs_check(expr) != GS_OK) return GS_ERR_CUDA; } while (0)
  CL_TRY(cudaMemsetAsync(b.dL_dW0, 0, (size_t)CL_H * d_in * f, st));
  CL_TRY(cudaMemsetAsync(b.dL_dW1, 0, (size_t)CL_H * CL_H * f, st));
  CL_TRY(cudaMemsetAsync(b.dL_dW2, 0, (size_t)3 * CL_H * f, st));
  CL_TRY(cudaMemsetAsync(b.dL_db0, 0, CL_H * f, st));
  CL_TRY(cudaMemsetAsync(b.dL_db1, 0, CL_H * f, st));
  CL_TRY(cudaMemsetAsync(b.dL_db2, 0, 3 * f, st));
  if (b.dL_dlatent && a.latent_dim) CL_TRY(cudaMemsetAsync(b.dL_dlatent, 0, (size_t)a.latent_dim * f, st));
#undef CL_TRY
  if (a.P == 0) return GS_OK;
  color_prep_bwd_kernel<<<16, 256, 0, st>>>(a, w);
  const int smem = (int)sizeof(BwSmem) + 1024;
  if (gs_check(cudaFuncSetAttribute(color_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != GS_OK) return GS_ERR_CUDA;
  const int ntiles = (a.P + NR_TILE - 1) / NR_TILE;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  color_bwd_kernel<<<ntiles < sms ? ntiles : sms, BW_THREADS, smem, st>>>(b, w, (const __nv_bfloat16*)saved, ntiles);
  if (a.latent_dim > 0) color_bwd_finish_kernel<<<4, 256, 0, st>>>(b);
  return gs_check(cudaGetLastError());
}
