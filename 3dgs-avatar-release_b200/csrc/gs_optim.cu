// gs_optim.cu -- the per-Gaussian work either side of the hot path in a training iteration:
//   * adam_step: torch.optim.Adam over the six per-Gaussian parameter groups (scene/gaussian_model.py:201-221,
//     lr per group, eps 1e-15, no weight decay) as ONE launch over the flat parameter slab, whose layout is the flat
//     gradient slab pose_backward writes and the data-parallel all-reduce runs over (28 B of HBM traffic per scalar).
//   * densify_stats: max_radii2D / xyz_gradient_accum / denom update of train.py:219-220 and
//     scene/gaussian_model.py:464-466 in one pass.
// Both are HBM-bound streaming kernels.  The step counter lives on the device so that the launch can sit inside the
// step's CUDA graph; the last block to finish advances it.
#include "gs_common.cuh"

namespace {

#define OPT_THREADS 256

struct AdamState {  // 16 bytes on the device, zero-initialised by the caller
  long long step;
  unsigned int ticket;
  unsigned int pad;
};

struct AdamSeg {
  long long end[ADAM_MAX_GROUPS];  // exclusive end offset of every group in the slab
  int n;
};

__device__ __forceinline__ float adam_one(float& p, float g, float& m, float& v, float lr_over_bc1, float beta2, float w1,
                                          float w2, float bc2_sqrt, float eps) {
  // torch/optim/adam.py _single_tensor_adam: exp_avg.lerp_(grad, 1 - beta1); exp_avg_sq.mul_(beta2).addcmul_(grad, grad,
  // value = 1 - beta2); denom = (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps); param.addcdiv_(exp_avg, denom,
  // value = -step_size)
  m = __fadd_rn(m, __fmul_rn(w1, __fsub_rn(g, m)));
  v = __fadd_rn(__fmul_rn(v, beta2), __fmul_rn(__fmul_rn(w2, g), g));
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), eps);
  p = __fadd_rn(p, __fmul_rn(-lr_over_bc1, __fdiv_rn(m, denom)));
  return p;
}

__global__ void __launch_bounds__(OPT_THREADS)
adam_step_kernel(long long n, AdamSeg seg, const float* __restrict__ lrs, float beta1, float beta2, float eps,
                 float* __restrict__ params, const float* __restrict__ grads, float* __restrict__ exp_avg,
                 float* __restrict__ exp_avg_sq, AdamState* __restrict__ state) {
  GS_PDL_PROLOGUE();
  __shared__ float s_step_size[ADAM_MAX_GROUPS];
  __shared__ float s_bc2_sqrt;
  const long long t = state->step + 1;  // only the last block to finish writes it back
  if (threadIdx.x < ADAM_MAX_GROUPS) {
    const double bc1 = 1.0 - pow((double)beta1, (double)t);
    s_step_size[threadIdx.x] = threadIdx.x < seg.n ? (float)((double)lrs[threadIdx.x] / bc1) : 0.f;
    if (threadIdx.x == 0) s_bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)t));
  }
  __syncthreads();
  const float w1 = (float)(1.0 - (double)beta1), w2 = (float)(1.0 - (double)beta2), bc2s = s_bc2_sqrt;
  auto group_of = [&](long long i) {
    int gidx = 0;
#pragma unroll
    for (int k = 0; k < ADAM_MAX_GROUPS - 1; k++) gidx += (k < seg.n - 1 && i >= seg.end[k]) ? 1 : 0;
    return gidx;
  };
  const long long n4 = n >> 2, stride = (long long)gridDim.x * OPT_THREADS;
  for (long long i = (long long)blockIdx.x * OPT_THREADS + threadIdx.x; i < n4; i += stride) {
    float4 p = reinterpret_cast<float4*>(params)[i];
    const float4 g = __ldg(reinterpret_cast<const float4*>(grads) + i);
    float4 m = reinterpret_cast<float4*>(exp_avg)[i], v = reinterpret_cast<float4*>(exp_avg_sq)[i];
    const long long e = i << 2;
    const int g0 = group_of(e), g3 = group_of(e + 3);
    const float s0 = s_step_size[g0], s3 = s_step_size[g3];
    const float s1 = g0 == g3 ? s0 : s_step_size[group_of(e + 1)], s2 = g0 == g3 ? s0 : s_step_size[group_of(e + 2)];
    adam_one(p.x, g.x, m.x, v.x, s0, beta2, w1, w2, bc2s, eps);
    adam_one(p.y, g.y, m.y, v.y, s1, beta2, w1, w2, bc2s, eps);
    adam_one(p.z, g.z, m.z, v.z, s2, beta2, w1, w2, bc2s, eps);
    adam_one(p.w, g.w, m.w, v.w, s3, beta2, w1, w2, bc2s, eps);
    reinterpret_cast<float4*>(params)[i] = p;
    reinterpret_cast<float4*>(exp_avg)[i] = m;
    reinterpret_cast<float4*>(exp_avg_sq)[i] = v;
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * OPT_THREADS + threadIdx.x; i < n; i += stride) {  // tail
    float p = params[i], m = exp_avg[i], v = exp_avg_sq[i];
    adam_one(p, grads[i], m, v, s_step_size[group_of(i)], beta2, w1, w2, bc2s, eps);
    params[i] = p; exp_avg[i] = m; exp_avg_sq[i] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (atomicAdd(&state->ticket, 1u) == gridDim.x - 1) {  // every block has read the old step
      state->step = t;
      state->ticket = 0u;
    }
  }
}

__global__ void __launch_bounds__(OPT_THREADS)
densify_stats_kernel(int P, const int* __restrict__ radii, const float* __restrict__ viewspace_grad,
                     float* __restrict__ max_radii2D, float* __restrict__ xyz_gradient_accum, float* __restrict__ denom) {
  GS_PDL_PROLOGUE();
  const int i = blockIdx.x * OPT_THREADS + threadIdx.x;
  if (i >= P) return;
  const int r = radii[i];
  if (r <= 0) return;  // visibility_filter = radii > 0 (gaussian_renderer/__init__.py:149)
  max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);                         // train.py:219
  const float gx = viewspace_grad[3 * (size_t)i], gy = viewspace_grad[3 * (size_t)i + 1];
  xyz_gradient_accum[i] += __fsqrt_rn(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy)));  // gaussian_model.py:465
  denom[i] += 1.0f;                                                          // :466
}

}  // namespace

extern "C" {

size_t adam_state_bytes(void) { return sizeof(AdamState); }

int adam_step(const AdamArgs* args, void* stream) {
  if (!args) return GS_ERR_BAD_ARG;
  const AdamArgs& a = *args;
  if (a.n < 0 || a.n_groups < 1 || a.n_groups > ADAM_MAX_GROUPS || !a.lrs || !a.params || !a.grads || !a.exp_avg ||
      !a.exp_avg_sq || !a.state)
    return GS_ERR_BAD_ARG;
  if (((uintptr_t)a.params | (uintptr_t)a.grads | (uintptr_t)a.exp_avg | (uintptr_t)a.exp_avg_sq) & 15) return GS_ERR_BAD_ARG;
  AdamSeg seg;
  long long prev = 0;
  for (int k = 0; k < ADAM_MAX_GROUPS; k++) {
    seg.end[k] = k < a.n_groups ? a.group_end[k] : a.n;
    if (k < a.n_groups && (a.group_end[k] < prev || a.group_end[k] > a.n)) return GS_ERR_BAD_ARG;
    if (k < a.n_groups) prev = a.group_end[k];
  }
  if (a.group_end[a.n_groups - 1] != a.n) return GS_ERR_BAD_ARG;
  seg.n = a.n_groups;
  if (a.n == 0) return GS_OK;
  long long blocks = (a.n / 4 + OPT_THREADS - 1) / OPT_THREADS;
  if (blocks > 148 * 8) blocks = 148 * 8;  // grid-stride: 8 CTAs of 256 threads per SM
  if (blocks < 1) blocks = 1;
  gs_launch(adam_step_kernel, dim3((unsigned)blocks), dim3(OPT_THREADS), 0, (cudaStream_t)stream, (long long)a.n, seg, a.lrs,
            a.beta1, a.beta2, a.eps, a.params, a.grads, a.exp_avg, a.exp_avg_sq, (AdamState*)a.state);
  return gs_check(cudaGetLastError());
}

int densify_stats(int32_t P, const int32_t* radii, const float* viewspace_grad, float* max_radii2D,
                  float* xyz_gradient_accum, float* denom, void* stream) {
  if (P < 0 || (P > 0 && (!radii || !viewspace_grad || !max_radii2D || !xyz_gradient_accum || !denom))) return GS_ERR_BAD_ARG;
  if (P == 0) return GS_OK;
  gs_launch(densify_stats_kernel, dim3((P + OPT_THREADS - 1) / OPT_THREADS), dim3(OPT_THREADS), 0, (cudaStream_t)stream,
            (int)P, (const int*)radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom);
  return gs_check(cudaGetLastError());
}

}  // extern "C"
