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

namespace {

constexpr int CL_H = 64;     // hidden width (configs/texture/shallow_mlp.yaml: n_neurons 64)
constexpr int CL_K0 = 64;    // per-Gaussian inputs padded to 64 (32 + 15 + 16 = 63 in the default configuration)
constexpr int CL_NOUT = 16;  // 3 colour channels padded to the smallest UMMA N

struct ClWs {
  __nv_bfloat16 *w0_hi, *w0_lo;  // [64 x 64]
  __nv_bfloat16 *w1_hi, *w1_lo;  // [64 x 64]
  __nv_bfloat16 *w2_hi, *w2_lo;  // [16 x 64]
  float* bias;                   // [64] b0 + W0[:, per-Gaussian inputs:] . latent | [64] b1 | [16] b2
};

__host__ size_t carve_cl(char* base, ClWs* w) {
  char* p = base;
  ClWs t;
  gs_carve(p, t.w0_hi, (size_t)CL_H * CL_K0); gs_carve(p, t.w0_lo, (size_t)CL_H * CL_K0);
  gs_carve(p, t.w1_hi, (size_t)CL_H * CL_H); gs_carve(p, t.w1_lo, (size_t)CL_H * CL_H);
  gs_carve(p, t.w2_hi, (size_t)CL_NOUT * CL_H); gs_carve(p, t.w2_lo, (size_t)CL_NOUT * CL_H);
  gs_carve(p, t.bias, (size_t)(2 * CL_H + CL_NOUT));
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
__global__ void __launch_bounds__(NR_TILE, 3) color_fwd_kernel(ColorArgs a, ClWs w, int n_tiles) {
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
    // ---- lin0 ----
    if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W0_hi, sm.W0_lo, CL_K0, CL_H, &sm.bar);
    nr_mbar_wait(&sm.bar, phase); phase ^= 1u;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    cl_hidden_epilogue(sm, tmem_row, row, sm.bias);
    cl_publish();
    // ---- lin1 ----
    if (tid == 0) issue_layer(tmem, sm.A_hi, sm.A_lo, sm.W1_hi, sm.W1_lo, CL_H, CL_H, &sm.bar);
    nr_mbar_wait(&sm.bar, phase); phase ^= 1u;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    cl_hidden_epilogue(sm, tmem_row, row, sm.bias + CL_H);
    cl_publish();
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

extern "C" int color_forward(const ColorArgs* args, void* workspace, size_t workspace_bytes, void* stream) {
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
  static int sms = 0;
  if (!sms) {
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
    color_fwd_kernel<true><<<grid, NR_TILE, sizeof(ClSmem) + 1024, st>>>(a, w, n_tiles);
  else
    color_fwd_kernel<false><<<grid, NR_TILE, sizeof(ClSmem) + 1024, st>>>(a, w, n_tiles);
  return gs_check(cudaGetLastError());
}
