// gs_encoder.cu -- hierarchical pose encoder (replaces HierarchicalPoseEncoder.forward,
// models/network_utils.py:152-178): the per-frame 144-float condition of the non-rigid MLP.  The reference walks the
// kinematic tree joint by joint with two tiny nn.Linear per joint: ~150 launches per frame for ~11k multiply-adds.  Here:
// ONE single-CTA launch, one warp per joint; a joint's warp runs when its parent's feature is in shared memory (joints are
// processed level by level of the kinematic tree, nine levels, one CTA barrier each).
#include "gs_common.cuh"

namespace {

constexpr int PE_J = 24;          // joints
constexpr int PE_D = 6;           // dim_per_joint (configs/non_rigid/*.yaml)
constexpr int PE_F = 13 + PE_D;   // rot 9 + Jtr 3 + bone length 1 + parent/global feature

__global__ void __launch_bounds__(PE_J * 32) pose_encoder_kernel(PoseEncoderArgs a) {
  GS_PDL_PROLOGUE();
  __shared__ float s_rots[PE_J * 9], s_J[PE_J * 3], s_glob[PE_D], s_out[PE_J * PE_D];
  __shared__ int s_depth[PE_J];
  const int tid = threadIdx.x, lane = tid & 31, j = tid >> 5;
  for (int i = tid; i < PE_J * 9; i += blockDim.x) s_rots[i] = a.rots[i];
  if (tid < PE_J) {
    int d = 0;
    for (int p = a.parents[tid]; p >= 0; p = a.parents[p]) d++;
    s_depth[tid] = d;
  }
  if (tid < PE_J * 3) {
    const int jj = tid / 3, c = tid % 3, p = a.parents[jj];
    float v = a.Jtrs[tid];
    if (a.rel_joints && jj > 0) v -= a.Jtrs[p * 3 + c];  // network_utils.py:155-159
    s_J[tid] = v;
  }
  __syncthreads();
  // layer_0: global feature = W0 [6, 288] . [rots | Jtrs] + b0 (:161-162); warp w < 6 computes output w
  if (j < PE_D) {
    float acc = 0.f;
    const float* w = a.W0 + (size_t)j * (PE_J * 12);
    for (int k = lane; k < PE_J * 12; k += 32) acc = fmaf(w[k], k < PE_J * 9 ? s_rots[k] : s_J[k - PE_J * 9], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
    if (lane == 0) s_glob[j] = acc + a.b0[j];
  }
  __syncthreads();
  const int parent = a.parents[j], depth = s_depth[j];
  // this warp's weights go to registers up front: every load is in flight at once, instead of one L2 round trip per
  // tree level on the serial path
  float wa[PE_F], wb[PE_F];
  {
    const float* pa = a.Wa + ((size_t)j * PE_F + (lane < PE_F ? lane : 0)) * PE_F;
    const float* pb = a.Wb + ((size_t)j * PE_D + (lane < PE_D ? lane : 0)) * PE_F;
#pragma unroll
    for (int k = 0; k < PE_F; k++) { wa[k] = __ldg(pa + k); wb[k] = __ldg(pb + k); }
  }
  const float bias_a = lane < PE_F ? a.ba[j * PE_F + lane] : 0.f, bias_b = lane < PE_D ? a.bb[j * PE_D + lane] : 0.f;
  for (int level = 0; level < PE_J; level++) {  // at most the tree height + 1 rounds do any work
    if (depth == level) {
      // in_feat = [rot 9 | Jtr 3 | bone length 1 | parent feature (or the global feature for the root) 6]  (:166-176)
      float in = 0.f;
      if (lane < 9) in = s_rots[j * 9 + lane];
      else if (lane < 12) in = s_J[j * 3 + lane - 9];
      else if (lane == 12) {
        float dx = s_J[j * 3], dy = s_J[j * 3 + 1], dz = s_J[j * 3 + 2];
        if (parent >= 0 && !a.rel_joints) { dx -= s_J[parent * 3]; dy -= s_J[parent * 3 + 1]; dz -= s_J[parent * 3 + 2]; }
        in = sqrtf(dx * dx + dy * dy + dz * dz);
      } else if (lane < PE_F) in = parent < 0 ? s_glob[lane - 13] : s_out[parent * PE_D + lane - 13];
      // hidden = ReLU(Wa[j] [19,19] . in + ba[j]): lane i < 19 owns output i, inputs broadcast by shuffle
      float h = bias_a;
#pragma unroll
      for (int k = 0; k < PE_F; k++) {
        const float x = __shfl_sync(0xFFFFFFFFu, in, k);
        if (lane < PE_F) h = fmaf(wa[k], x, h);
      }
      h = fmaxf(h, 0.f);
      // out = Wb[j] [6,19] . hidden + bb[j]: lane i < 6 owns output i
      float o = bias_b;
#pragma unroll
      for (int k = 0; k < PE_F; k++) {
        const float x = __shfl_sync(0xFFFFFFFFu, h, k);
        if (lane < PE_D) o = fmaf(wb[k], x, o);
      }
      if (lane < PE_D) { s_out[j * PE_D + lane] = o; a.out[j * PE_D + lane] = o; }
    }
    if (__syncthreads_and(depth <= level)) break;  // every joint has run
  }
}

}  // namespace

extern "C" int pose_encoder_forward(const PoseEncoderArgs* args, void* stream) {
  if (!args) return GS_ERR_BAD_ARG;
  const PoseEncoderArgs& a = *args;
  if (a.num_joints != PE_J || a.dim_per_joint != PE_D) return GS_ERR_BAD_ARG;
  if (!a.rots || !a.Jtrs || !a.W0 || !a.b0 || !a.Wa || !a.ba || !a.Wb || !a.bb || !a.out) return GS_ERR_BAD_ARG;
  for (int j = 0; j < PE_J; j++)
    if (a.parents[j] >= j || a.parents[j] < -1) return GS_ERR_BAD_ARG;  // topological order, like ktree_parents
  gs_launch(pose_encoder_kernel, dim3(1), dim3(PE_J * 32), 0, (cudaStream_t)stream, a);
  return gs_check(cudaGetLastError());
}
