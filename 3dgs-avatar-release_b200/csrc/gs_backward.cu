// gs_backward.cu -- per-Gaussian backward: computeCov2DCUDA (cuda_rasterizer/backward.cu:144-274) and preprocessCUDA
// backward (:346-396) with the SH (:20-139) and scale/rotation (:278-341) backward, fused into ONE streaming pass
// (the reference runs two kernels and round-trips dL_dmean3D / dL_dcov3D through memory between them).
#include "gs_common.cuh"

namespace {

__constant__ float cSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f};
__constant__ float cSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                                -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

struct F3 { float x, y, z; };
__device__ __forceinline__ F3 operator*(float s, F3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ float dot(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// backward.cu:20-139: writes dL_dsh[M][3]; returns the mean gradient caused by the view direction.
__device__ F3 sh_backward(int deg, F3 pos, F3 cam, const float* __restrict__ sh, const uint8_t* __restrict__ clamped,
                          const float* __restrict__ dL_dcolor, float* __restrict__ dL_dsh) {
  const float SH_C0 = 0.28209479177387814f, SH_C1 = 0.4886025119029199f;
  F3 dir_orig = {pos.x - cam.x, pos.y - cam.y, pos.z - cam.z};
  float inv_len = 1.0f / sqrtf(dot(dir_orig, dir_orig));
  float x = dir_orig.x * inv_len, y = dir_orig.y * inv_len, z = dir_orig.z * inv_len;
  F3 dRGB = {clamped[0] ? 0.f : dL_dcolor[0], clamped[1] ? 0.f : dL_dcolor[1], clamped[2] ? 0.f : dL_dcolor[2]};
  F3 dx = {0, 0, 0}, dy = {0, 0, 0}, dz = {0, 0, 0};
  auto SH = [&](int k) -> F3 { return {sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]}; };
  auto DSH = [&](int k, float wgt) { dL_dsh[3 * k] = wgt * dRGB.x; dL_dsh[3 * k + 1] = wgt * dRGB.y; dL_dsh[3 * k + 2] = wgt * dRGB.z; };
  DSH(0, SH_C0);
  if (deg > 0) {
    DSH(1, -SH_C1 * y); DSH(2, SH_C1 * z); DSH(3, -SH_C1 * x);
    dx = -SH_C1 * SH(3); dy = -SH_C1 * SH(1); dz = SH_C1 * SH(2);
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      DSH(4, cSH_C2[0] * xy); DSH(5, cSH_C2[1] * yz); DSH(6, cSH_C2[2] * (2.f * zz - xx - yy));
      DSH(7, cSH_C2[3] * xz); DSH(8, cSH_C2[4] * (xx - yy));
      dx = dx + (cSH_C2[0] * y) * SH(4) + (cSH_C2[2] * 2.f * -x) * SH(6) + (cSH_C2[3] * z) * SH(7) + (cSH_C2[4] * 2.f * x) * SH(8);
      dy = dy + (cSH_C2[0] * x) * SH(4) + (cSH_C2[1] * z) * SH(5) + (cSH_C2[2] * 2.f * -y) * SH(6) + (cSH_C2[4] * 2.f * -y) * SH(8);
      dz = dz + (cSH_C2[1] * y) * SH(5) + (cSH_C2[2] * 2.f * 2.f * z) * SH(6) + (cSH_C2[3] * x) * SH(7);
      if (deg > 2) {
        DSH(9, cSH_C3[0] * y * (3.f * xx - yy)); DSH(10, cSH_C3[1] * xy * z);
        DSH(11, cSH_C3[2] * y * (4.f * zz - xx - yy)); DSH(12, cSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
        DSH(13, cSH_C3[4] * x * (4.f * zz - xx - yy)); DSH(14, cSH_C3[5] * z * (xx - yy));
        DSH(15, cSH_C3[6] * x * (xx - 3.f * yy));
        dx = dx + (cSH_C3[0] * 3.f * 2.f * xy) * SH(9) + (cSH_C3[1] * yz) * SH(10) + (cSH_C3[2] * -2.f * xy) * SH(11) +
             (cSH_C3[3] * -3.f * 2.f * xz) * SH(12) + (cSH_C3[4] * (-3.f * xx + 4.f * zz - yy)) * SH(13) +
             (cSH_C3[5] * 2.f * xz) * SH(14) + (cSH_C3[6] * 3.f * (xx - yy)) * SH(15);
        dy = dy + (cSH_C3[0] * 3.f * (xx - yy)) * SH(9) + (cSH_C3[1] * xz) * SH(10) +
             (cSH_C3[2] * (-3.f * yy + 4.f * zz - xx)) * SH(11) + (cSH_C3[3] * -3.f * 2.f * yz) * SH(12) +
             (cSH_C3[4] * -2.f * xy) * SH(13) + (cSH_C3[5] * -2.f * yz) * SH(14) + (cSH_C3[6] * -3.f * 2.f * xy) * SH(15);
        dz = dz + (cSH_C3[1] * xy) * SH(10) + (cSH_C3[2] * 4.f * 2.f * yz) * SH(11) +
             (cSH_C3[3] * 3.f * (2.f * zz - xx - yy)) * SH(12) + (cSH_C3[4] * 4.f * 2.f * xz) * SH(13) +
             (cSH_C3[5] * (xx - yy)) * SH(14);
      }
    }
  }
  F3 ddir = {dot(dx, dRGB), dot(dy, dRGB), dot(dz, dRGB)};
  // dnormvdv, auxiliary.h:107-117
  F3 v = dir_orig;
  float sum2 = dot(v, v);
  float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  F3 r;
  r.x = ((+sum2 - v.x * v.x) * ddir.x - v.y * v.x * ddir.y - v.z * v.x * ddir.z) * invsum32;
  r.y = (-v.x * v.y * ddir.x + (sum2 - v.y * v.y) * ddir.y - v.z * v.y * ddir.z) * invsum32;
  r.z = (-v.x * v.z * ddir.x - v.y * v.z * ddir.y + (sum2 - v.z * v.z) * ddir.z) * invsum32;
  return r;
}

// backward.cu:278-341
__device__ void cov3d_backward(const float* __restrict__ s3, float mod, const float* __restrict__ q4, const float* g6,
                               float* __restrict__ dL_dscale, float* __restrict__ dL_drot) {
  float r = q4[0], x = q4[1], y = q4[2], z = q4[3];
  float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                   {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                   {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};  // [col][row]
  float s[3] = {mod * s3[0], mod * s3[1], mod * s3[2]};
  float dSig[3][3];
  dSig[0][0] = g6[0]; dSig[1][1] = g6[3]; dSig[2][2] = g6[5];
  dSig[0][1] = dSig[1][0] = 0.5f * g6[1];
  dSig[0][2] = dSig[2][0] = 0.5f * g6[2];
  dSig[1][2] = dSig[2][1] = 0.5f * g6[4];
  float dMt[3][3];  // dMt[c][rr] = dL_dM[rr][c],  dL_dM = 2 * M * dSig,  M[c][rr] = s[rr] * R[c][rr]
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int rr = 0; rr < 3; rr++) {
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < 3; k++) acc += (2.0f * s[c] * R[k][c]) * dSig[rr][k];
      dMt[c][rr] = acc;
    }
#pragma unroll
  for (int c = 0; c < 3; c++) {
    dL_dscale[c] = R[0][c] * dMt[c][0] + R[1][c] * dMt[c][1] + R[2][c] * dMt[c][2];
#pragma unroll
    for (int rr = 0; rr < 3; rr++) dMt[c][rr] *= s[c];
  }
  dL_drot[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
  dL_drot[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
  dL_drot[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
  dL_drot[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
}

template <bool PRE>
__global__ void __launch_bounds__(256, PRE ? 3 : 4) preprocess_bwd_kernel(GsBackwardArgs a, const float* __restrict__ cov3Ds,
                                                             const uint8_t* __restrict__ clamped, const GsRec* __restrict__ recs,
                                                             float* __restrict__ accs) {
  if (!PRE) GS_PDL_PROLOGUE();
  // The direct predecessor of this grid is always blend_bwd_kernel of the same call, which waited for everything before it:
  // whatever blend_bwd does not write (camera, the first Gaussian's forward-time state) is fetched BEFORE the wait, under the
  // tail of the blend; only the accumulators come after it.
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  if (threadIdx.x < 16) { s_view[threadIdx.x] = a.viewmatrix[threadIdx.x]; s_proj[threadIdx.x] = a.projmatrix[threadIdx.x]; }
  if (threadIdx.x < 3 && a.campos) s_cam[threadIdx.x] = a.campos[threadIdx.x];
  const float* view = s_view;
  const float* proj = s_proj;
  const float h_y = a.height / (2.0f * a.tan_fovy), h_x = a.width / (2.0f * a.tan_fovx);
  const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
  const int idx0 = blockIdx.x * blockDim.x + threadIdx.x;
  int radius = 0;
  float mx = 0.f, my = 0.f, mz = 0.f, c6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float4 rb = make_float4(0.f, 0.f, 0.f, 0.f);
  // every load of a Gaussian is issued BEFORE the visibility branch (all of them are in bounds for a culled Gaussian too):
  // one memory round trip per thread instead of two -- at 1M Gaussians the kernel sat on long-scoreboard stalls
  auto load_state = [&](int idx) {
    radius = a.radii[idx];
    mx = a.means3D[3 * (size_t)idx]; my = a.means3D[3 * (size_t)idx + 1]; mz = a.means3D[3 * (size_t)idx + 2];
#pragma unroll
    for (int k = 0; k < 6; k++) c6[k] = cov3Ds[6 * (size_t)idx + k];
    rb = reinterpret_cast<const float4*>(recs + idx)[1];  // conic a, b, c, opacity
  };
  if (idx0 < a.P) load_state(idx0);
  __syncthreads();
  if (PRE) GS_PDL_PROLOGUE();
  for (int idx = idx0; idx < a.P; idx += gridDim.x * blockDim.x) {
    // the blend's accumulators: consumed here and zeroed again (a second backward over the same forward starts clean)
    float4* const ac = reinterpret_cast<float4*>(accs + (size_t)GS_ACC * idx);
    const float4 ac0 = ac[0], ac1 = ac[1], ac2 = ac[2];
    if (idx != idx0) load_state(idx);
    ac[0] = ac[1] = ac[2] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!(radius > 0)) {  // culled: every gradient is zero (the outputs need no initialisation by the caller)
      a.dL_dmean2D[3 * (size_t)idx] = a.dL_dmean2D[3 * (size_t)idx + 1] = a.dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
      a.dL_dopacity[idx] = 0.f;
      if (a.dL_dconic) *reinterpret_cast<float4*>(a.dL_dconic + 4 * (size_t)idx) = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int k = 0; k < 3; k++) { a.dL_dcolor[3 * (size_t)idx + k] = 0.f; a.dL_dmean3D[3 * (size_t)idx + k] = 0.f; }
#pragma unroll
      for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * (size_t)idx + k] = 0.f;
      if (a.shs) for (int k = 0; k < 3 * a.M; k++) a.dL_dsh[(size_t)idx * a.M * 3 + k] = 0.f;
      if (a.scales) {
#pragma unroll
        for (int k = 0; k < 3; k++) a.dL_dscale[3 * (size_t)idx + k] = 0.f;
#pragma unroll
        for (int k = 0; k < 4; k++) a.dL_drot[4 * (size_t)idx + k] = 0.f;
      }
      continue;
    }
    // blend_bwd_kernel left RAW MOMENTS of g = G * dL/dalpha over the blended (pixel, Gaussian) pairs: sum g dx, g dy, g dx^2,
    // g dx dy, g dy^2, g and the colour gradient.  With the Gaussian's conic and opacity they become the sums of
    // backward.cu:536-549: dL_dmean2D = -ddel * o * (conic * m1), dL_dconic = -0.5 * o * m2, dL_dopacity = sum g.
    const float m1x = ac0.x, m1y = ac0.y;
    const float g2x = -(0.5f * a.width) * rb.w * (rb.x * m1x + rb.y * m1y);
    const float g2y = -(0.5f * a.height) * rb.w * (rb.z * m1y + rb.y * m1x);
    a.dL_dmean2D[3 * (size_t)idx] = g2x;
    a.dL_dmean2D[3 * (size_t)idx + 1] = g2y;
    a.dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
    const float dcx = -0.5f * rb.w * ac0.z, dcy = -0.5f * rb.w * ac0.w, dcz = -0.5f * rb.w * ac1.x;
    if (a.dL_dconic) *reinterpret_cast<float4*>(a.dL_dconic + 4 * (size_t)idx) = make_float4(dcx, dcy, 0.f, dcz);  // optional output
    a.dL_dopacity[idx] = ac1.y;
    const float dcol[3] = {ac1.z, ac1.w, ac2.x};
#pragma unroll
    for (int k = 0; k < 3; k++) a.dL_dcolor[3 * (size_t)idx + k] = dcol[k];
    // ---- computeCov2DCUDA ----
    float tx = view[0] * mx + view[4] * my + view[8] * mz + view[12];
    float ty = view[1] * mx + view[5] * my + view[9] * mz + view[13];
    float tz = view[2] * mx + view[6] * my + view[10] * mz + view[14];
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
    ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    const float J00 = h_x / tz, J11 = h_y / tz, J02 = -(h_x * tx) / (tz * tz), J12 = -(h_y * ty) / (tz * tz);
    float T0[3], T1[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
      T0[r] = view[4 * r] * J00 + view[2 + 4 * r] * J02;
      T1[r] = view[1 + 4 * r] * J11 + view[2 + 4 * r] * J12;
    }
    const float V[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
    float vt0[3], vt1[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      vt0[k] = T0[0] * V[k][0] + T0[1] * V[k][1] + T0[2] * V[k][2];
      vt1[k] = T1[0] * V[k][0] + T1[1] * V[k][1] + T1[2] * V[k][2];
    }
    const float ca = T0[0] * vt0[0] + T0[1] * vt0[1] + T0[2] * vt0[2] + 0.3f;
    const float cb = T0[0] * vt1[0] + T0[1] * vt1[1] + T0[2] * vt1[2];
    const float cc = T1[0] * vt1[0] + T1[1] * vt1[1] + T1[2] * vt1[2] + 0.3f;
    const float denom = ca * cc - cb * cb;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float g6[6] = {0, 0, 0, 0, 0, 0};
    if (denom2inv != 0) {
      dL_da = denom2inv * (-cc * cc * dcx + 2 * cb * cc * dcy + (denom - ca * cc) * dcz);
      dL_dc = denom2inv * (-ca * ca * dcz + 2 * ca * cb * dcy + (denom - ca * cc) * dcx);
      dL_db = denom2inv * 2 * (cb * cc * dcx - (denom + 2 * cb * cb) * dcy + ca * cb * dcz);
      g6[0] = T0[0] * T0[0] * dL_da + T0[0] * T1[0] * dL_db + T1[0] * T1[0] * dL_dc;
      g6[3] = T0[1] * T0[1] * dL_da + T0[1] * T1[1] * dL_db + T1[1] * T1[1] * dL_dc;
      g6[5] = T0[2] * T0[2] * dL_da + T0[2] * T1[2] * dL_db + T1[2] * T1[2] * dL_dc;
      g6[1] = 2 * T0[0] * T0[1] * dL_da + (T0[0] * T1[1] + T0[1] * T1[0]) * dL_db + 2 * T1[0] * T1[1] * dL_dc;
      g6[2] = 2 * T0[0] * T0[2] * dL_da + (T0[0] * T1[2] + T0[2] * T1[0]) * dL_db + 2 * T1[0] * T1[2] * dL_dc;
      g6[4] = 2 * T0[2] * T0[1] * dL_da + (T0[1] * T1[2] + T0[2] * T1[1]) * dL_db + 2 * T1[1] * T1[2] * dL_dc;
    }
#pragma unroll
    for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * (size_t)idx + k] = g6[k];
    float dT0[3], dT1[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      dT0[k] = 2 * vt0[k] * dL_da + vt1[k] * dL_db;
      dT1[k] = 2 * vt1[k] * dL_dc + vt0[k] * dL_db;
    }
    // W[c][r] = view[c + 4 r]
    const float dJ00 = view[0] * dT0[0] + view[4] * dT0[1] + view[8] * dT0[2];
    const float dJ02 = view[2] * dT0[0] + view[6] * dT0[1] + view[10] * dT0[2];
    const float dJ11 = view[1] * dT1[0] + view[5] * dT1[1] + view[9] * dT1[2];
    const float dJ12 = view[2] * dT1[0] + view[6] * dT1[1] + view[10] * dT1[2];
    const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
    const float dtx = x_grad_mul * -h_x * itz2 * dJ02;
    const float dty = y_grad_mul * -h_y * itz2 * dJ12;
    const float dtz = -h_x * itz2 * dJ00 - h_y * itz2 * dJ11 + (2 * h_x * tx) * itz3 * dJ02 + (2 * h_y * ty) * itz3 * dJ12;
    float dmx = view[0] * dtx + view[1] * dty + view[2] * dtz;
    float dmy = view[4] * dtx + view[5] * dty + view[6] * dtz;
    float dmz = view[8] * dtx + view[9] * dty + view[10] * dtz;
    // ---- preprocessCUDA backward ----
    const float m_hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
    const float m_w = 1.0f / (m_hw + 0.0000001f);
    const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
    dmx += (proj[0] * m_w - proj[3] * mul1) * g2x + (proj[1] * m_w - proj[3] * mul2) * g2y;
    dmy += (proj[4] * m_w - proj[7] * mul1) * g2x + (proj[5] * m_w - proj[7] * mul2) * g2y;
    dmz += (proj[8] * m_w - proj[11] * mul1) * g2x + (proj[9] * m_w - proj[11] * mul2) * g2y;
    if (a.shs) {
      F3 r = sh_backward(a.D, {mx, my, mz}, {s_cam[0], s_cam[1], s_cam[2]}, a.shs + (size_t)idx * a.M * 3,
                         clamped + 3 * (size_t)idx, dcol, a.dL_dsh + (size_t)idx * a.M * 3);
      dmx += r.x; dmy += r.y; dmz += r.z;
    }
    a.dL_dmean3D[3 * (size_t)idx] = dmx;
    a.dL_dmean3D[3 * (size_t)idx + 1] = dmy;
    a.dL_dmean3D[3 * (size_t)idx + 2] = dmz;
    if (a.scales)
      cov3d_backward(a.scales + 3 * (size_t)idx, a.scale_modifier, a.rotations + 4 * (size_t)idx, g6,
                     a.dL_dscale + 3 * (size_t)idx, a.dL_drot + 4 * (size_t)idx);
  }
}

}  // namespace

int gs_launch_preprocess_bwd(const GsBackwardArgs& a, const GsGeom& g, cudaStream_t st) {
  int blocks = (a.P + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const float* cov = a.cov3D_precomp ? a.cov3D_precomp : g.cov3D;
  gs_launch((gs_prewait_mask() & 4) ? preprocess_bwd_kernel<true> : preprocess_bwd_kernel<false>, dim3(blocks), dim3(256), 0, st, a, cov,
            g.clamped, (const GsRec*)g.rec, g.acc);
  return gs_check(cudaGetLastError());
}
