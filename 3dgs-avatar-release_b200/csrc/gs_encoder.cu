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

// Backward of the encoder in ONE single-CTA launch (the reference gets it from autograd over ~150 tiny nn.Linear nodes: several
// hundred launches per step for ~22k multiply-adds).  One warp per joint again.  The forward is re-evaluated level by level
// (inputs and hidden activations stay in shared memory), then the tree is walked from the leaves to the root: a joint adds the
// gradients its children left for their parent feature to its own upstream gradient, and runs
//   dWb = go (x) h, dbb = go, dh = Wb^T go, dz = dh * [h > 0], dWa = dz (x) in, dba = dz, din = Wa^T dz.
// din splits into the joint's rotation / position / bone-length inputs and the parent feature (the global feature for the root,
// whose gradient ends in layer_0).  Children are summed in index order by the parent (no atomics): the result is deterministic.
// Every output element is written, so the caller need not clear anything.
__global__ void __launch_bounds__(PE_J * 32) pose_encoder_bwd_kernel(PoseEncoderBackwardArgs b) {
  GS_PDL_PROLOGUE();
  const PoseEncoderArgs& a = b.fwd;
  __shared__ float s_rots[PE_J * 9], s_J[PE_J * 3], s_glob[PE_D], s_out[PE_J * PE_D];
  __shared__ float s_in[PE_J][PE_F + 1], s_h[PE_J][PE_F + 1];
  __shared__ float s_dpar[PE_J][PE_D], s_dlen[PE_J][3], s_dJown[PE_J][3], s_drot[PE_J * 9], s_dglob[PE_D];
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
    if (a.rel_joints && jj > 0) v -= a.Jtrs[p * 3 + c];
    s_J[tid] = v;
  }
  __syncthreads();
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
  int max_depth = 0;
  for (int k = 0; k < PE_J; k++) max_depth = max(max_depth, s_depth[k]);
  // ---- forward re-evaluation: the same operations in the same order as pose_encoder_kernel (same ReLU branches) ----
  {
    float wa[PE_F], wb[PE_F];
    const float* pa = a.Wa + ((size_t)j * PE_F + (lane < PE_F ? lane : 0)) * PE_F;
    const float* pb = a.Wb + ((size_t)j * PE_D + (lane < PE_D ? lane : 0)) * PE_F;
#pragma unroll
    for (int k = 0; k < PE_F; k++) { wa[k] = __ldg(pa + k); wb[k] = __ldg(pb + k); }
    const float bias_a = lane < PE_F ? a.ba[j * PE_F + lane] : 0.f, bias_b = lane < PE_D ? a.bb[j * PE_D + lane] : 0.f;
    for (int level = 0; level <= max_depth; level++) {
      if (depth == level) {
        float in = 0.f;
        if (lane < 9) in = s_rots[j * 9 + lane];
        else if (lane < 12) in = s_J[j * 3 + lane - 9];
        else if (lane == 12) {
          float dx = s_J[j * 3], dy = s_J[j * 3 + 1], dz = s_J[j * 3 + 2];
          if (parent >= 0 && !a.rel_joints) { dx -= s_J[parent * 3]; dy -= s_J[parent * 3 + 1]; dz -= s_J[parent * 3 + 2]; }
          in = sqrtf(dx * dx + dy * dy + dz * dz);
        } else if (lane < PE_F) in = parent < 0 ? s_glob[lane - 13] : s_out[parent * PE_D + lane - 13];
        float h = bias_a;
#pragma unroll
        for (int k = 0; k < PE_F; k++) {
          const float x = __shfl_sync(0xFFFFFFFFu, in, k);
          if (lane < PE_F) h = fmaf(wa[k], x, h);
        }
        h = fmaxf(h, 0.f);
        float o = bias_b;
#pragma unroll
        for (int k = 0; k < PE_F; k++) {
          const float x = __shfl_sync(0xFFFFFFFFu, h, k);
          if (lane < PE_D) o = fmaf(wb[k], x, o);
        }
        if (lane < PE_D) s_out[j * PE_D + lane] = o;
        if (lane < PE_F) { s_in[j][lane] = in; s_h[j][lane] = h; }
      }
      __syncthreads();
    }
  }
  // ---- backward: leaves first ----
  // the transposed weight columns this lane needs (lane k < 19: Wb[j][i][k], Wa[j][i][k]) go to registers up front: every
  // load is in flight at once instead of two L2 round trips per tree level on the serial path
  float wbT[PE_D], waT[PE_F];
  {
    const int kk = lane < PE_F ? lane : 0;
#pragma unroll
    for (int i = 0; i < PE_D; i++) wbT[i] = __ldg(a.Wb + ((size_t)j * PE_D + i) * PE_F + kk);
#pragma unroll
    for (int i = 0; i < PE_F; i++) waT[i] = __ldg(a.Wa + ((size_t)j * PE_F + i) * PE_F + kk);
  }
  for (int level = max_depth; level >= 0; level--) {
    if (depth == level) {
      float go = 0.f;  // lane i < 6: dL/dout[j][i], own upstream + what the children left for their parent feature
      if (lane < PE_D) {
        go = b.dL_dout[j * PE_D + lane];
        for (int c = j + 1; c < PE_J; c++)
          if (a.parents[c] == j) go += s_dpar[c][lane];
        b.dL_dbb[j * PE_D + lane] = go;
      }
      for (int e0 = 0; e0 < PE_D * PE_F; e0 += 32) {  // dWb[j][i][k] = go_i * h_k (all lanes take part in the shuffle)
        const int e = e0 + lane, ec = e < PE_D * PE_F ? e : 0, i = ec / PE_F, k = ec - i * PE_F;
        const float v = __shfl_sync(0xFFFFFFFFu, go, i) * s_h[j][k];
        if (e < PE_D * PE_F) b.dL_dWb[(size_t)j * PE_D * PE_F + e] = v;
      }
      float dz = 0.f;  // lane k < 19
      {
        float dh = 0.f;
#pragma unroll
        for (int i = 0; i < PE_D; i++) {
          const float g = __shfl_sync(0xFFFFFFFFu, go, i);
          if (lane < PE_F) dh = fmaf(wbT[i], g, dh);
        }
        if (lane < PE_F) {
          dz = s_h[j][lane] > 0.f ? dh : 0.f;
          b.dL_dba[j * PE_F + lane] = dz;
        }
      }
      float din = 0.f;  // lane k < 19: dL/din[k] = sum_i Wa[j][i][k] dz_i
#pragma unroll
      for (int i = 0; i < PE_F; i++) {
        const float g = __shfl_sync(0xFFFFFFFFu, dz, i);
        if (lane < PE_F) din = fmaf(waT[i], g, din);
      }
      for (int e0 = 0; e0 < PE_F * PE_F; e0 += 32) {  // dWa[j][i][k] = dz_i * in_k (all lanes take part in the shuffle)
        const int e = e0 + lane, ec = e < PE_F * PE_F ? e : 0, i = ec / PE_F, k = ec - i * PE_F;
        const float v = __shfl_sync(0xFFFFFFFFu, dz, i) * s_in[j][k];
        if (e < PE_F * PE_F) b.dL_dWa[(size_t)j * PE_F * PE_F + e] = v;
      }
      if (lane < 9) s_drot[j * 9 + lane] = din;
      else if (lane < 12) s_dJown[j][lane - 9] = din;
      else if (lane == 12) {  // bone length |d|: d(len)/dd = d / len (0 at len = 0, like torch.norm's backward)
        float dx = s_J[j * 3], dy = s_J[j * 3 + 1], dzz = s_J[j * 3 + 2];
        if (parent >= 0 && !a.rel_joints) { dx -= s_J[parent * 3]; dy -= s_J[parent * 3 + 1]; dzz -= s_J[parent * 3 + 2]; }
        const float len = s_in[j][12], sc = len > 0.f ? din / len : 0.f;
        s_dlen[j][0] = sc * dx; s_dlen[j][1] = sc * dy; s_dlen[j][2] = sc * dzz;
      } else if (lane < PE_F) {
        if (parent >= 0) s_dpar[j][lane - 13] = din; else s_dglob[lane - 13] = din;
      }
    }
    __syncthreads();
  }
  // ---- layer_0 and the inputs ----
  for (int e = tid; e < PE_D * PE_J * 12; e += blockDim.x) {
    const int i = e / (PE_J * 12), k = e - i * (PE_J * 12);
    b.dL_dW0[e] = s_dglob[i] * (k < PE_J * 9 ? s_rots[k] : s_J[k - PE_J * 9]);
  }
  if (tid < PE_D) b.dL_db0[tid] = s_dglob[tid];
  for (int k = tid; k < PE_J * 12; k += blockDim.x) {
    float dx = 0.f;
#pragma unroll
    for (int i = 0; i < PE_D; i++) dx = fmaf(__ldg(a.W0 + (size_t)i * (PE_J * 12) + k), s_dglob[i], dx);
    if (k < PE_J * 9) {
      if (b.dL_drots) b.dL_drots[k] = s_drot[k] + dx;
    } else if (b.dL_dJtrs) {
      const int q = k - PE_J * 9, jj = q / 3, c = q - jj * 3;
      float v = 0.f;  // rel_joints: the reference forms the relative joints under no_grad (network_utils.py:155-159): no gradient
      if (!a.rel_joints) {
        v = s_dJown[jj][c] + dx + s_dlen[jj][c];
        for (int ch = jj + 1; ch < PE_J; ch++)
          if (a.parents[ch] == jj) v -= s_dlen[ch][c];  // d = J[child] - J[parent]
      }
      b.dL_dJtrs[q] = v;
    }
  }
}

}  // namespace

extern "C" int pose_encoder_backward(const PoseEncoderBackwardArgs* args, void* stream) {
  if (!args) return GS_ERR_BAD_ARG;
  const PoseEncoderBackwardArgs& b = *args;
  const PoseEncoderArgs& a = b.fwd;
  if (a.num_joints != PE_J || a.dim_per_joint != PE_D) return GS_ERR_BAD_ARG;
  if (!a.rots || !a.Jtrs || !a.W0 || !a.b0 || !a.Wa || !a.ba || !a.Wb || !a.bb || !b.dL_dout) return GS_ERR_BAD_ARG;
  if (!b.dL_dW0 || !b.dL_db0 || !b.dL_dWa || !b.dL_dba || !b.dL_dWb || !b.dL_dbb) return GS_ERR_BAD_ARG;
  for (int j = 0; j < PE_J; j++)
    if (a.parents[j] >= j || a.parents[j] < -1) return GS_ERR_BAD_ARG;
  gs_launch(pose_encoder_bwd_kernel, dim3(1), dim3(PE_J * 32), 0, (cudaStream_t)stream, b);
  return gs_check(cudaGetLastError());
}

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
