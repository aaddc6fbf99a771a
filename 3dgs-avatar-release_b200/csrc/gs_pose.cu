// gs_pose.cu -- fused posing of canonical Gaussians: one kernel forward, one kernel backward, replacing the ~60-kernel
// PyTorch chain of the reference per frame:
//   rigid LBS of mean and rotation          models/deformer/rigid.py:49-71 (SMPLNN) / :215-236 (SkinningField, after
//                                           get_forward_transform produced the weights)
//   build_rotation (normalise + matrix)     utils/general_utils.py:87-108
//   covariance R S S^T R^T packed to 6      scene/gaussian_model.py:28-32,154-157; utils/general_utils.py:194-207,73-85
//   activations sigmoid / exp               scene/gaussian_model.py:132-152
//   SH2RGB: canonical view dir + eval_sh    models/texture/texture.py:21-38; utils/sh_utils.py:58-113
// and the two GaussianModel.clone() copies (scene/gaussian_model.py:68-87) disappear.
// HBM-bound: 384 B/Gaussian forward, 716 B/Gaussian backward; one thread per Gaussian, the
// 24 bone matrices staged in shared memory.
#include "gs_common.cuh"

namespace {

#define C0f 0.28209479177387814f
#define C1f 0.4886025119029199f
__constant__ float pC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f,
                             0.5462742152960396f};
__constant__ float pC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                             -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

__device__ __forceinline__ void sh_basis(int deg, float x, float y, float z, float* b) {
  b[0] = C0f;
  if (deg > 0) {
    b[1] = -C1f * y; b[2] = C1f * z; b[3] = -C1f * x;
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      b[4] = pC2[0] * xy; b[5] = pC2[1] * yz; b[6] = pC2[2] * (2.0f * zz - xx - yy); b[7] = pC2[3] * xz;
      b[8] = pC2[4] * (xx - yy);
      if (deg > 2) {
        b[9] = pC3[0] * y * (3 * xx - yy); b[10] = pC3[1] * xy * z; b[11] = pC3[2] * y * (4 * zz - xx - yy);
        b[12] = pC3[3] * z * (2 * zz - 3 * xx - 3 * yy); b[13] = pC3[4] * x * (4 * zz - xx - yy);
        b[14] = pC3[5] * z * (xx - yy); b[15] = pC3[6] * x * (xx - 3 * yy);
      }
    }
  }
}

// d(basis_k)/d(x,y,z)
__device__ __forceinline__ void sh_basis_grad(int deg, float x, float y, float z, float* gx, float* gy, float* gz) {
#pragma unroll
  for (int k = 0; k < 16; k++) { gx[k] = 0.f; gy[k] = 0.f; gz[k] = 0.f; }
  if (deg > 0) {
    gy[1] = -C1f; gz[2] = C1f; gx[3] = -C1f;
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      gx[4] = pC2[0] * y; gy[4] = pC2[0] * x;
      gy[5] = pC2[1] * z; gz[5] = pC2[1] * y;
      gx[6] = -2.f * pC2[2] * x; gy[6] = -2.f * pC2[2] * y; gz[6] = 4.f * pC2[2] * z;
      gx[7] = pC2[3] * z; gz[7] = pC2[3] * x;
      gx[8] = 2.f * pC2[4] * x; gy[8] = -2.f * pC2[4] * y;
      if (deg > 2) {
        gx[9] = pC3[0] * 6.f * xy; gy[9] = pC3[0] * 3.f * (xx - yy);
        gx[10] = pC3[1] * yz; gy[10] = pC3[1] * xz; gz[10] = pC3[1] * xy;
        gx[11] = pC3[2] * -2.f * xy; gy[11] = pC3[2] * (4.f * zz - xx - 3.f * yy); gz[11] = pC3[2] * 8.f * yz;
        gx[12] = pC3[3] * -6.f * xz; gy[12] = pC3[3] * -6.f * yz; gz[12] = pC3[3] * (6.f * zz - 3.f * xx - 3.f * yy);
        gx[13] = pC3[4] * (4.f * zz - 3.f * xx - yy); gy[13] = pC3[4] * -2.f * xy; gz[13] = pC3[4] * 8.f * xz;
        gx[14] = pC3[5] * 2.f * xz; gy[14] = pC3[5] * -2.f * yz; gz[14] = pC3[5] * (xx - yy);
        gx[15] = pC3[6] * 3.f * (xx - yy); gy[15] = pC3[6] * -6.f * xy;
      }
    }
  }
}

// All 16 basis values / gradients, unconditionally: the callers' loops over k are fully unrolled and predicated with k < nb, so
// the arrays live in registers (with run-time bounds they sat in local memory and every use was a long-scoreboard LDL).
__device__ __forceinline__ void sh_basis16(float x, float y, float z, float (&b)[16]) { sh_basis(3, x, y, z, b); }
__device__ __forceinline__ void sh_basis_grad16(float x, float y, float z, float (&gx)[16], float (&gy)[16], float (&gz)[16]) {
  sh_basis_grad(3, x, y, z, gx, gy, gz);
}

struct PoseCore {
  float A[3][3], t[3];     // blended transform (rows 0..2 of T_fwd)
  float qn[4], inv_norm;   // normalised quaternion
  float R[3][3];           // rotation of qn
  float Rb[3][3];          // A * R
  float s[3];              // mod * exp(scaling)
  float xb[3];             // posed mean
};

// Per-warp coalesced staging of the SH rows of the warp's 32 Gaussians (a contiguous chunk of up to 32*45 floats): 45
// independent coalesced loads in flight per lane instead of 45 dependent-latency strided ones; the rows are then read
// back lane-private with an odd stride (conflict-free).
#define POSE_MAXREST 45
// Two halves, so that the caller can issue every other global load of the Gaussian between them: one memory round trip per
// warp instead of two.
__device__ __forceinline__ void warp_stage_load(float (&v)[POSE_MAXREST], const float* __restrict__ g, int n, int lane) {
#pragma unroll
  for (int k = 0; k < POSE_MAXREST; k++) { const int e = k * 32 + lane; v[k] = e < n ? __ldg(g + e) : 0.f; }
}
__device__ __forceinline__ void warp_stage_commit(float* __restrict__ s, const float (&v)[POSE_MAXREST], int n, int lane) {
#pragma unroll
  for (int k = 0; k < POSE_MAXREST; k++) { const int e = k * 32 + lane; if (e < n) s[e] = v[k]; }
  __syncwarp();
}
__device__ __forceinline__ void warp_stage_out(float* __restrict__ g, const float* __restrict__ s, int n, int lane) {
  __syncwarp();
#pragma unroll
  for (int k = 0; k < POSE_MAXREST; k++) { const int e = k * 32 + lane; if (e < n) g[e] = s[e]; }
  __syncwarp();
}

// The Gaussian's own inputs, loaded before anything is computed (and before the SH rows are committed to shared memory).
struct PoseRaw {
  float w[24];
  float x, y, z;
  float4 q;
  float sc[3];
  float op;
  float dc[3];
};
__device__ __forceinline__ void pose_load(const PoseArgs& a, int i, PoseRaw& r) {
  const float4* wp = reinterpret_cast<const float4*>(a.weights + (size_t)i * 24);  // 96-byte rows: 16-byte aligned
#pragma unroll
  for (int k = 0; k < 6; k++) { const float4 t = __ldg(wp + k); r.w[4 * k] = t.x; r.w[4 * k + 1] = t.y; r.w[4 * k + 2] = t.z; r.w[4 * k + 3] = t.w; }
  r.x = a.xyz[3 * (size_t)i]; r.y = a.xyz[3 * (size_t)i + 1]; r.z = a.xyz[3 * (size_t)i + 2];
  if ((reinterpret_cast<uintptr_t>(a.rotation) & 15) == 0) {
    r.q = *reinterpret_cast<const float4*>(a.rotation + 4 * (size_t)i);
  } else {  // a view into a flat parameter slab (avatar_pose.optim) starts at float 3P: aligned only when P % 4 == 0
    const float* qp = a.rotation + 4 * (size_t)i;
    r.q = make_float4(qp[0], qp[1], qp[2], qp[3]);
  }
#pragma unroll
  for (int k = 0; k < 3; k++) { r.sc[k] = a.scaling[3 * (size_t)i + k]; r.dc[k] = a.features_dc[3 * (size_t)i + k]; }
  r.op = a.opacity[i];
}

__device__ __forceinline__ void pose_core(const PoseArgs& a, const PoseRaw& raw, const float* __restrict__ s_bones, PoseCore& c) {
  float T[12];
#pragma unroll
  for (int k = 0; k < 12; k++) T[k] = 0.f;
#pragma unroll
  for (int j = 0; j < 24; j++) {
    const float wj = raw.w[j];
#pragma unroll
    for (int k = 0; k < 12; k++) T[k] = fmaf(wj, s_bones[j * 16 + k], T[k]);
  }
#pragma unroll
  for (int r = 0; r < 3; r++) {
    c.A[r][0] = T[4 * r]; c.A[r][1] = T[4 * r + 1]; c.A[r][2] = T[4 * r + 2]; c.t[r] = T[4 * r + 3];
  }
  const float x = raw.x, y = raw.y, z = raw.z;
#pragma unroll
  for (int r = 0; r < 3; r++) c.xb[r] = c.A[r][0] * x + c.A[r][1] * y + c.A[r][2] * z + c.t[r];
  const float4 q = raw.q;
  const float norm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  c.inv_norm = 1.0f / norm;
  const float qr = q.x * c.inv_norm, qx = q.y * c.inv_norm, qy = q.z * c.inv_norm, qz = q.w * c.inv_norm;
  c.qn[0] = qr; c.qn[1] = qx; c.qn[2] = qy; c.qn[3] = qz;
  c.R[0][0] = 1 - 2 * (qy * qy + qz * qz); c.R[0][1] = 2 * (qx * qy - qr * qz); c.R[0][2] = 2 * (qx * qz + qr * qy);
  c.R[1][0] = 2 * (qx * qy + qr * qz); c.R[1][1] = 1 - 2 * (qx * qx + qz * qz); c.R[1][2] = 2 * (qy * qz - qr * qx);
  c.R[2][0] = 2 * (qx * qz - qr * qy); c.R[2][1] = 2 * (qy * qz + qr * qx); c.R[2][2] = 1 - 2 * (qx * qx + qy * qy);
#pragma unroll
  for (int r = 0; r < 3; r++)
#pragma unroll
    for (int k = 0; k < 3; k++) c.Rb[r][k] = c.A[r][0] * c.R[0][k] + c.A[r][1] * c.R[1][k] + c.A[r][2] * c.R[2][k];
#pragma unroll
  for (int k = 0; k < 3; k++) c.s[k] = a.scaling_modifier * expf(raw.sc[k]);
}

// view direction fed to eval_sh (texture.py:23-35): returns d (un-normalised), n = |d|
__device__ __forceinline__ void view_dir(const PoseArgs& a, const PoseCore& c, const float* cam, float* d, float& n) {
  const float v0 = c.xb[0] - cam[0], v1 = c.xb[1] - cam[1], v2 = c.xb[2] - cam[2];
  if (a.cano_view_dir) {  // R_bwd = T_fwd[:3,:3]^T (transpose, not inverse)
    d[0] = c.A[0][0] * v0 + c.A[1][0] * v1 + c.A[2][0] * v2;
    d[1] = c.A[0][1] * v0 + c.A[1][1] * v1 + c.A[2][1] * v2;
    d[2] = c.A[0][2] * v0 + c.A[1][2] * v1 + c.A[2][2] * v2;
    if (a.view_noise) {  // training augmentation, texture.py:28-33: dir <- dir @ view_noise (row vector times matrix)
      const float* N = a.view_noise;
      const float n0 = d[0] * N[0] + d[1] * N[3] + d[2] * N[6], n1 = d[0] * N[1] + d[1] * N[4] + d[2] * N[7];
      const float n2 = d[0] * N[2] + d[1] * N[5] + d[2] * N[8];
      d[0] = n0; d[1] = n1; d[2] = n2;
    }
  } else { d[0] = v0; d[1] = v1; d[2] = v2; }
  n = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
}

// SH coefficient k of channel ch: dc from global (12 B per thread), the rest from this lane's staged row
__device__ __forceinline__ float sh_coef(const PoseRaw& raw, const float* __restrict__ rest_row, int k, int ch) {
  return k == 0 ? raw.dc[ch] : rest_row[(k - 1) * 3 + ch];
}

__global__ void __launch_bounds__(128) pose_fwd_kernel(PoseArgs a) {
  GS_PDL_PROLOGUE();
  __shared__ float s_bones[24 * 16];
  __shared__ float s_cam[3];
  for (int k = threadIdx.x; k < 24 * 16; k += blockDim.x) s_bones[k] = a.bone_transforms[k];
  if (threadIdx.x < 3) s_cam[threadIdx.x] = a.camera_center[threadIdx.x];
  __shared__ float s_rest_all[4][32 * POSE_MAXREST];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  float* const s_rest = s_rest_all[threadIdx.x >> 5];
  const int nrest = (a.M - 1) * 3;  // floats per Gaussian in features_rest (<= 45)
  for (int i0 = blockIdx.x * blockDim.x + threadIdx.x - lane; i0 < a.P; i0 += gridDim.x * blockDim.x) {  // warp-uniform
    const int i = i0 + lane;
    const int nstaged = min(32, a.P - i0) * nrest;
    float v[POSE_MAXREST];
    PoseRaw raw;
    if (nrest) warp_stage_load(v, a.features_rest + (size_t)i0 * nrest, nstaged, lane);
    if (i < a.P) pose_load(a, i, raw);
    if (nrest) warp_stage_commit(s_rest, v, nstaged, lane);
    const float* rest_row = s_rest + lane * nrest;
    if (i < a.P) {
    PoseCore c;
    pose_core(a, raw, s_bones, c);
    a.means3D[3 * (size_t)i] = c.xb[0]; a.means3D[3 * (size_t)i + 1] = c.xb[1]; a.means3D[3 * (size_t)i + 2] = c.xb[2];
    float L[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
      for (int k = 0; k < 3; k++) L[r][k] = c.Rb[r][k] * c.s[k];
    float* cov = a.cov3D + 6 * (size_t)i;
    cov[0] = L[0][0] * L[0][0] + L[0][1] * L[0][1] + L[0][2] * L[0][2];
    cov[1] = L[0][0] * L[1][0] + L[0][1] * L[1][1] + L[0][2] * L[1][2];
    cov[2] = L[0][0] * L[2][0] + L[0][1] * L[2][1] + L[0][2] * L[2][2];
    cov[3] = L[1][0] * L[1][0] + L[1][1] * L[1][1] + L[1][2] * L[1][2];
    cov[4] = L[1][0] * L[2][0] + L[1][1] * L[2][1] + L[1][2] * L[2][2];
    cov[5] = L[2][0] * L[2][0] + L[2][1] * L[2][1] + L[2][2] * L[2][2];
    a.opacity_out[i] = 1.0f / (1.0f + expf(-raw.op));
    float d[3], n;
    view_dir(a, c, s_cam, d, n);
    const float inv = 1.0f / (n + 1e-12f);
    float b[16];
    sh_basis16(d[0] * inv, d[1] * inv, d[2] * inv, b);
    const int nb = (a.sh_degree + 1) * (a.sh_degree + 1);
#pragma unroll
    for (int ch = 0; ch < 3; ch++) {
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < 16; k++)
        if (k < nb) acc += b[k] * sh_coef(raw, rest_row, k, ch);
      a.colors[3 * (size_t)i + ch] = fmaxf(acc + 0.5f, 0.0f);
    }
    if (a.T_fwd) {
      float* T = a.T_fwd + 16 * (size_t)i;
#pragma unroll
      for (int r = 0; r < 3; r++) { T[4 * r] = c.A[r][0]; T[4 * r + 1] = c.A[r][1]; T[4 * r + 2] = c.A[r][2]; T[4 * r + 3] = c.t[r]; }
      float t3[4] = {0.f, 0.f, 0.f, 0.f};
      const float* w = a.weights + (size_t)i * 24;
      for (int j = 0; j < 24; j++)
#pragma unroll
        for (int k = 0; k < 4; k++) t3[k] = fmaf(w[j], s_bones[j * 16 + 12 + k], t3[k]);
#pragma unroll
      for (int k = 0; k < 4; k++) T[12 + k] = t3[k];
    }
    if (a.rotation_bar) {
#pragma unroll
      for (int r = 0; r < 3; r++)
#pragma unroll
        for (int k = 0; k < 3; k++) a.rotation_bar[9 * (size_t)i + 3 * r + k] = c.Rb[r][k];
    }
    }  // i < P
    __syncwarp();  // rows are re-staged in the next iteration
  }
}

__global__ void __launch_bounds__(128, 3) pose_bwd_kernel(PoseBackwardArgs ba) {
  GS_PDL_PROLOGUE();
  const PoseArgs& a = ba.fwd;
  __shared__ float s_bones[24 * 16];
  __shared__ float s_cam[3];
  for (int k = threadIdx.x; k < 24 * 16; k += blockDim.x) s_bones[k] = a.bone_transforms[k];
  if (threadIdx.x < 3) s_cam[threadIdx.x] = a.camera_center[threadIdx.x];
  __shared__ float s_rest_all[4][32 * POSE_MAXREST];  // SH coefficients in, their gradients out
  __shared__ float s_gbones[288];                     // dL/dbone_transforms[:, :3, :4] of this block
  for (int k = threadIdx.x; k < 288; k += blockDim.x) s_gbones[k] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  float* const s_rest = s_rest_all[threadIdx.x >> 5];
  const int nrest = (a.M - 1) * 3;
  for (int i0 = blockIdx.x * blockDim.x + threadIdx.x - lane; i0 < a.P; i0 += gridDim.x * blockDim.x) {  // warp-uniform
    const int i = i0 + lane;
    const int nstaged = min(32, a.P - i0) * nrest;
    float v[POSE_MAXREST];
    PoseRaw raw;
    float gxb[3] = {0.f, 0.f, 0.f}, gcol[3] = {0.f, 0.f, 0.f}, g6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, gop = 0.f;
    if (nrest) warp_stage_load(v, a.features_rest + (size_t)i0 * nrest, nstaged, lane);
    if (i < a.P) {  // every global load of the Gaussian, upstream gradients included, in the same round trip
      pose_load(a, i, raw);
#pragma unroll
      for (int k = 0; k < 3; k++) { gxb[k] = ba.dL_dmeans3D[3 * (size_t)i + k]; gcol[k] = ba.dL_dcolors[3 * (size_t)i + k]; }
#pragma unroll
      for (int k = 0; k < 6; k++) g6[k] = ba.dL_dcov3D[6 * (size_t)i + k];
      gop = ba.dL_dopacity[i];
    }
    if (nrest) warp_stage_commit(s_rest, v, nstaged, lane);
    float* rest_row = s_rest + lane * nrest;
    float gT[12];  // dL/dT_fwd[:3, :4] of this Gaussian (zero for the lanes past P)
#pragma unroll
    for (int k = 0; k < 12; k++) gT[k] = 0.f;
    if (i < a.P) {
    PoseCore c;
    pose_core(a, raw, s_bones, c);
    // ---- colour -> SH coefficients and view direction ----
    {
      float d[3], n;
      view_dir(a, c, s_cam, d, n);
      const float inv = 1.0f / (n + 1e-12f);
      const float ux = d[0] * inv, uy = d[1] * inv, uz = d[2] * inv;
      float b[16], bx[16], by[16], bz[16];
      sh_basis16(ux, uy, uz, b);
      sh_basis_grad16(ux, uy, uz, bx, by, bz);
      const int nb = (a.sh_degree + 1) * (a.sh_degree + 1);
      float gu[3] = {0.f, 0.f, 0.f};
#pragma unroll
      for (int ch = 0; ch < 3; ch++) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 16; k++)
          if (k < nb) acc += b[k] * sh_coef(raw, rest_row, k, ch);
        const float gc = (acc + 0.5f < 0.0f) ? 0.f : gcol[ch];
        ba.dL_dfeatures_dc[3 * (size_t)i + ch] = b[0] * gc;
#pragma unroll
        for (int k = 1; k < 16; k++) {
          if (k < a.M) {
            const float coef = k < nb ? rest_row[(k - 1) * 3 + ch] : 0.f;
            rest_row[(k - 1) * 3 + ch] = k < nb ? b[k] * gc : 0.f;  // the staged coefficient is consumed: its slot takes the gradient
            if (k < nb) { gu[0] += bx[k] * coef * gc; gu[1] += by[k] * coef * gc; gu[2] += bz[k] * coef * gc; }
          }
        }
      }
      // u = d / (n + eps):  dL/dd = g/(n+eps) - d (d.g) / (n (n+eps)^2)
      const float dg = d[0] * gu[0] + d[1] * gu[1] + d[2] * gu[2];
      const float k2 = n > 0.f ? dg * inv * inv / n : 0.f;
      float gd[3] = {gu[0] * inv - d[0] * k2, gu[1] * inv - d[1] * k2, gu[2] * inv - d[2] * k2};
      if (a.cano_view_dir && a.view_noise) {  // d' = d @ N  ->  dL/dd_j = sum_k N[j][k] g'_k
        const float* N = a.view_noise;
        const float t0 = N[0] * gd[0] + N[1] * gd[1] + N[2] * gd[2], t1 = N[3] * gd[0] + N[4] * gd[1] + N[5] * gd[2];
        const float t2 = N[6] * gd[0] + N[7] * gd[1] + N[8] * gd[2];
        gd[0] = t0; gd[1] = t1; gd[2] = t2;
      }
      if (a.cano_view_dir) {  // d = A^T v with A detached -> dL/dv = A gd
#pragma unroll
        for (int r = 0; r < 3; r++) gxb[r] += c.A[r][0] * gd[0] + c.A[r][1] * gd[1] + c.A[r][2] * gd[2];
      } else {
#pragma unroll
        for (int r = 0; r < 3; r++) gxb[r] += gd[r];
      }
    }
    // ---- covariance -> R_bar, scaling ----
    const float G[3][3] = {{g6[0], 0.5f * g6[1], 0.5f * g6[2]}, {0.5f * g6[1], g6[3], 0.5f * g6[4]}, {0.5f * g6[2], 0.5f * g6[4], g6[5]}};
    float L[3][3], dLm[3][3], gRb[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
      for (int k = 0; k < 3; k++) L[r][k] = c.Rb[r][k] * c.s[k];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
      for (int k = 0; k < 3; k++) dLm[r][k] = 2.f * (G[r][0] * L[0][k] + G[r][1] * L[1][k] + G[r][2] * L[2][k]);
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const float gs = dLm[0][k] * c.Rb[0][k] + dLm[1][k] * c.Rb[1][k] + dLm[2][k] * c.Rb[2][k];
      ba.dL_dscaling[3 * (size_t)i + k] = gs * c.s[k];  // s = mod * exp(sigma)
#pragma unroll
      for (int r = 0; r < 3; r++) gRb[r][k] = dLm[r][k] * c.s[k];
    }
    // R_bar = A R:  dL/dR = A^T gRb ; dL/dA = gRb R^T (+ gxb x^T below)
    float gR[3][3], gA[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
      for (int k = 0; k < 3; k++) {
        gR[r][k] = c.A[0][r] * gRb[0][k] + c.A[1][r] * gRb[1][k] + c.A[2][r] * gRb[2][k];
        gA[r][k] = gRb[r][0] * c.R[k][0] + gRb[r][1] * c.R[k][1] + gRb[r][2] * c.R[k][2];
      }
    // rotation matrix -> normalised quaternion (w,x,y,z)
    {
      const float qr = c.qn[0], qx = c.qn[1], qy = c.qn[2], qz = c.qn[3];
      float gq[4];
      gq[0] = 2.f * (-qz * gR[0][1] + qy * gR[0][2] + qz * gR[1][0] - qx * gR[1][2] - qy * gR[2][0] + qx * gR[2][1]);
      gq[1] = 2.f * (qy * gR[0][1] + qz * gR[0][2] + qy * gR[1][0] - 2.f * qx * gR[1][1] - qr * gR[1][2] + qz * gR[2][0] + qr * gR[2][1] - 2.f * qx * gR[2][2]);
      gq[2] = 2.f * (-2.f * qy * gR[0][0] + qx * gR[0][1] + qr * gR[0][2] + qx * gR[1][0] + qz * gR[1][2] - qr * gR[2][0] + qz * gR[2][1] - 2.f * qy * gR[2][2]);
      gq[3] = 2.f * (-2.f * qz * gR[0][0] - qr * gR[0][1] + qx * gR[0][2] + qr * gR[1][0] - 2.f * qz * gR[1][1] + qy * gR[1][2] + qx * gR[2][0] + qy * gR[2][1]);
      const float dotq = gq[0] * qr + gq[1] * qx + gq[2] * qy + gq[3] * qz;
      float4 out;
      out.x = (gq[0] - qr * dotq) * c.inv_norm; out.y = (gq[1] - qx * dotq) * c.inv_norm;
      out.z = (gq[2] - qy * dotq) * c.inv_norm; out.w = (gq[3] - qz * dotq) * c.inv_norm;
      float* const gq_out = ba.dL_drotation + 4 * (size_t)i;
      if ((reinterpret_cast<uintptr_t>(ba.dL_drotation) & 15) == 0) {
        *reinterpret_cast<float4*>(gq_out) = out;
      } else {  // the rotation part of the flat gradient slab starts at float 3P: 16-byte aligned only when P % 4 == 0
        gq_out[0] = out.x; gq_out[1] = out.y; gq_out[2] = out.z; gq_out[3] = out.w;
      }
    }
    // mean: x_bar = A x + t
    const float x = raw.x, y = raw.y, z = raw.z;
#pragma unroll
    for (int k = 0; k < 3; k++) ba.dL_dxyz[3 * (size_t)i + k] = c.A[0][k] * gxb[0] + c.A[1][k] * gxb[1] + c.A[2][k] * gxb[2];
    const float so = 1.0f / (1.0f + expf(-raw.op));
    ba.dL_dopacity_logit[i] = gop * so * (1.0f - so);
    if (ba.dL_dweights || ba.dL_dbone_transforms) {
#pragma unroll
      for (int r = 0; r < 3; r++) {
        gT[4 * r] = gA[r][0] + gxb[r] * x; gT[4 * r + 1] = gA[r][1] + gxb[r] * y; gT[4 * r + 2] = gA[r][2] + gxb[r] * z;
        gT[4 * r + 3] = gxb[r];
      }
    }
    if (ba.dL_dweights) {
      float* gw = ba.dL_dweights + (size_t)i * 24;
      for (int j = 0; j < 24; j++) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 12; k++) acc = fmaf(gT[k], s_bones[j * 16 + k], acc);
        gw[j] = acc;
      }
    }
    }  // i < P
    if (ba.dL_dbone_transforms) {
      // T_fwd = sum_j w_j B_j (rigid.py:215-236)  ->  dL/dB_j[:3, :4] = sum over Gaussians of w_j dL/dT_fwd: the route by which the image
      // loss reaches the per-frame pose correction.  288 sums over the warp's 32 Gaussians, 32 at a time (transposing butterfly:
      // lane L ends with the total of value L), then shared-memory and finally global accumulation.
      const float* wrow = a.weights + (size_t)min(i, a.P - 1) * 24;
#pragma unroll 1
      for (int c0 = 0; c0 < 288; c0 += 32) {
        float v[32];
#pragma unroll
        for (int q = 0; q < 32; q++) {
          const int e = c0 + q;  // e = j * 12 + k
          v[q] = (i < a.P) ? wrow[e / 12] * gT[e % 12] : 0.f;
        }
#define POSE_BSTEP(N, X)                                                        \
  _Pragma("unroll") for (int k = 0; k < (N) / 2; k++) {                         \
    const bool up = lane & (X);                                                 \
    const float send = up ? v[k] : v[k + (N) / 2];                              \
    const float keepv = up ? v[k + (N) / 2] : v[k];                             \
    v[k] = keepv + __shfl_xor_sync(0xFFFFFFFFu, send, (X));                     \
  }
        POSE_BSTEP(32, 16) POSE_BSTEP(16, 8) POSE_BSTEP(8, 4) POSE_BSTEP(4, 2) POSE_BSTEP(2, 1)
#undef POSE_BSTEP
        if (v[0] != 0.f) atomicAdd(&s_gbones[c0 + lane], v[0]);
      }
    }
    if (nrest) warp_stage_out(ba.dL_dfeatures_rest + (size_t)i0 * nrest, s_rest, nstaged, lane);
  }
  if (ba.dL_dbone_transforms) {
    __syncthreads();
    for (int e = threadIdx.x; e < 288; e += blockDim.x)
      if (s_gbones[e] != 0.f) atomicAdd(ba.dL_dbone_transforms + (e / 12) * 16 + (e % 12), s_gbones[e]);
  }
}

}  // namespace

static int pose_args_ok(const PoseArgs& a) {
  if (a.P < 0 || a.sh_degree < 0 || a.sh_degree > 3) return 0;
  if (a.M < (a.sh_degree + 1) * (a.sh_degree + 1) || a.M > 16) return 0;
  if (!a.xyz || !a.rotation || !a.scaling || !a.opacity || !a.features_dc || !a.weights || !a.bone_transforms || !a.camera_center) return 0;
  if (a.M > 1 && !a.features_rest) return 0;
  return 1;
}

extern "C" int pose_forward(const PoseArgs* args, void* stream) {
  if (!args || !pose_args_ok(*args)) return GS_ERR_BAD_ARG;
  if (!args->means3D || !args->cov3D || !args->colors || !args->opacity_out) return GS_ERR_BAD_ARG;
  if (args->P == 0) return GS_OK;
  int blocks = (args->P + 127) / 128;
  if (blocks > 148 * 16) blocks = 148 * 16;
  gs_prof_mark(5, 0, (cudaStream_t)stream);
  gs_launch(pose_fwd_kernel, dim3(blocks), dim3(128), 0, (cudaStream_t)stream, *args);
  gs_prof_mark(5, 1, (cudaStream_t)stream);
  return gs_check(cudaGetLastError());
}

extern "C" int pose_backward(const PoseBackwardArgs* args, void* stream) {
  if (!args || !pose_args_ok(args->fwd)) return GS_ERR_BAD_ARG;
  if (!args->dL_dmeans3D || !args->dL_dcov3D || !args->dL_dcolors || !args->dL_dopacity) return GS_ERR_BAD_ARG;
  if (!args->dL_dxyz || !args->dL_drotation || !args->dL_dscaling || !args->dL_dopacity_logit || !args->dL_dfeatures_dc) return GS_ERR_BAD_ARG;
  if (args->fwd.M > 1 && !args->dL_dfeatures_rest) return GS_ERR_BAD_ARG;
  if (args->dL_dbone_transforms &&
      gs_check(cudaMemsetAsync(args->dL_dbone_transforms, 0, 24 * 16 * sizeof(float), (cudaStream_t)stream)) != GS_OK) return GS_ERR_CUDA;
  if (args->fwd.P == 0) return GS_OK;
  int blocks = (args->fwd.P + 127) / 128;
  if (blocks > 148 * 16) blocks = 148 * 16;
  gs_prof_mark(6, 0, (cudaStream_t)stream);
  gs_launch(pose_bwd_kernel, dim3(blocks), dim3(128), 0, (cudaStream_t)stream, *args);
  gs_prof_mark(6, 1, (cudaStream_t)stream);
  return gs_check(cudaGetLastError());
}

// ================================= smpl_nn: nearest SMPL vertex per Gaussian =================================
// Replaces SMPLNN.query_weights (models/deformer/rigid.py:41-47: pytorch3d knn_points(xyz, smpl_verts, K = 1) + gather of the
// vertex skinning weights).  Exact brute force: a thread owns a Gaussian, the 6890 vertices stream through shared memory in
// tiles; squared distance dx*dx + dy*dy + dz*dz in FP32, ties resolved to the lowest vertex index (first minimum).
namespace {
constexpr int NN_TILE = 1024;
__global__ void __launch_bounds__(256)
smpl_nn_kernel(int P, const float* __restrict__ xyz, int V, const float* __restrict__ verts, const float* __restrict__ vert_weights,
               int32_t* __restrict__ idx_out, float* __restrict__ weights_out) {
  __shared__ float s_v[NN_TILE * 3];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < P;
  const float x = live ? xyz[3 * (size_t)i] : 0.f, y = live ? xyz[3 * (size_t)i + 1] : 0.f, z = live ? xyz[3 * (size_t)i + 2] : 0.f;
  float best = 3.0e38f;
  int best_j = 0;
  for (int v0 = 0; v0 < V; v0 += NN_TILE) {
    const int n = min(NN_TILE, V - v0);
    __syncthreads();
    for (int k = threadIdx.x; k < n * 3; k += blockDim.x) s_v[k] = verts[(size_t)v0 * 3 + k];
    __syncthreads();
#pragma unroll 4
    for (int j = 0; j < n; j++) {
      const float dx = x - s_v[3 * j], dy = y - s_v[3 * j + 1], dz = z - s_v[3 * j + 2];
      const float d = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
      if (d < best) { best = d; best_j = v0 + j; }  // strict: the first minimum wins
    }
  }
  if (!live) return;
  if (idx_out) idx_out[i] = best_j;
  if (weights_out) {
    const float4* src = reinterpret_cast<const float4*>(vert_weights + (size_t)best_j * 24);
    float4* dst = reinterpret_cast<float4*>(weights_out + (size_t)i * 24);
#pragma unroll
    for (int k = 0; k < 6; k++) dst[k] = __ldg(src + k);
  }
}
}  // namespace

extern "C" int smpl_nn_forward(int32_t P, const float* xyz, int32_t V, const float* verts, const float* vert_weights,
                               int32_t* idx_out, float* weights_out, void* stream) {
  if (P < 0 || V <= 0 || !xyz || !verts || (weights_out && !vert_weights) || (!idx_out && !weights_out)) return GS_ERR_BAD_ARG;
  if (P == 0) return GS_OK;
  smpl_nn_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, xyz, V, verts, vert_weights, idx_out, weights_out);
  return gs_check(cudaGetLastError());
}
