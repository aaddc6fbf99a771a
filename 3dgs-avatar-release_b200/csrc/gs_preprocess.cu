// gs_preprocess.cu -- per-Gaussian forward preprocessing (replaces preprocessCUDA, cuda_rasterizer/forward.cu:155-256,
// and checkFrustum, rasterizer_impl.cu:54-66).
//
// One thread per Gaussian, grid-stride.  Besides the reference's outputs it (a) packs everything the blend
// kernels gather per list entry into one 48-byte record, (b) stores the tile rectangle so the key-emission
// kernel does not recompute it, (c) writes the 32-bit depth key + index for the depth sort and (d) accumulates
// the four 8-bit digit histograms of the depth keys (shared-memory atomics, one flush per CTA), which removes
// the separate histogram launch of the onesweep sort.
#include "gs_common.cuh"

namespace {

__device__ __forceinline__ int f2i_rz(float v) { return __float2int_rz(v); }  // cvt.rzi.s32.f32 (saturating)

// auxiliary.h:41-44 -- evaluated in FP64 exactly as nvcc compiles the reference
__device__ __forceinline__ float ndc2pix(float v, int S) {
  return (float)(__dmul_rn(__fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0), 0.5));
}

__device__ __forceinline__ void cov3d_from_scale_rot(const float* __restrict__ s3, float mod,
                                                     const float* __restrict__ q4, float* c6) {
  // forward.cu:118-152; contraction pattern of the reference PTX (see oracle/gs_oracle.c)
  float sx = __fmul_rn(mod, s3[0]), sy = __fmul_rn(mod, s3[1]), sz = __fmul_rn(mod, s3[2]);
  float r = q4[0], x = q4[1], y = q4[2], z = q4[3];
  // fusion pattern of the reference SASS (sm_100): xy-rz = fma(x,y,-rz), ry+xz = fma(r,y,xz), xz-ry = fma(-r,y,xz) ...
  float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z), xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
  float a = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(yy, zz)));
  float b = __fmul_rn(2.f, __fmaf_rn(x, y, -rz));
  float c = __fmul_rn(2.f, __fmaf_rn(r, y, xz));
  float d = __fmul_rn(2.f, __fmaf_rn(x, y, rz));
  float e = __fsub_rn(1.f, __fmul_rn(2.f, __fmaf_rn(x, x, zz)));
  float f = __fmul_rn(2.f, __fmaf_rn(y, z, -rx));
  float g = __fmul_rn(2.f, __fmaf_rn(-r, y, xz));
  float h = __fmul_rn(2.f, __fmaf_rn(y, z, rx));
  float i = __fsub_rn(1.f, __fmul_rn(2.f, __fmaf_rn(x, x, yy)));
  float M00 = __fmul_rn(sx, a), M01 = __fmul_rn(sy, b), M02 = __fmul_rn(sz, c);
  float M10 = __fmul_rn(sx, d), M11 = __fmul_rn(sy, e), M12 = __fmul_rn(sz, f);
  float M20 = __fmul_rn(sx, g), M21 = __fmul_rn(sy, h), M22 = __fmul_rn(sz, i);
  c6[0] = gs_dot3(M00, M00, M01, M01, M02, M02);
  c6[1] = gs_dot3(M10, M00, M11, M01, M12, M02);
  c6[2] = gs_dot3(M20, M00, M21, M01, M22, M02);
  c6[3] = gs_dot3(M10, M10, M11, M11, M12, M12);
  c6[4] = gs_dot3(M20, M10, M21, M11, M22, M12);
  c6[5] = gs_dot3(M20, M20, M21, M21, M22, M22);
}

__constant__ float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f};
__constant__ float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                                -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

// forward.cu:20-71.  sh points at this Gaussian's [M][3] block.
__device__ __forceinline__ void color_from_sh(int deg, float px, float py, float pz, const float* __restrict__ cam,
                                              const float* __restrict__ sh, float* rgb, uint8_t* clamped) {
  const float SH_C0 = 0.28209479177387814f, SH_C1 = 0.4886025119029199f;
  float dx = __fsub_rn(px, cam[0]), dy = __fsub_rn(py, cam[1]), dz = __fsub_rn(pz, cam[2]);
  float len = __fsqrt_rn(__fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy))));
  float x = __fdiv_rn(dx, len), y = __fdiv_rn(dy, len), z = __fdiv_rn(dz, len);
  float res[3];
#pragma unroll
  for (int c = 0; c < 3; c++) res[c] = __fmul_rn(SH_C0, sh[c]);
  if (deg > 0) {
    float b1 = __fmul_rn(SH_C1, y), b2 = __fmul_rn(SH_C1, z), b3 = __fmul_rn(SH_C1, x);
#pragma unroll
    for (int c = 0; c < 3; c++) {
      float r = __fsub_rn(res[c], __fmul_rn(b1, sh[3 + c]));
      r = __fmaf_rn(b2, sh[6 + c], r);
      res[c] = __fsub_rn(r, __fmul_rn(b3, sh[9 + c]));
    }
    if (deg > 1) {
      float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
      float xy = __fmul_rn(x, y), yz = __fmul_rn(y, z), xz = __fmul_rn(x, z);
      float b4 = __fmul_rn(kSH_C2[0], xy), b5 = __fmul_rn(kSH_C2[1], yz);
      float b6 = __fmul_rn(kSH_C2[2], __fsub_rn(__fsub_rn(__fmul_rn(2.0f, zz), xx), yy));
      float b7 = __fmul_rn(kSH_C2[3], xz), b8 = __fmul_rn(kSH_C2[4], __fsub_rn(xx, yy));
#pragma unroll
      for (int c = 0; c < 3; c++) {
        float r = res[c];
        r = __fmaf_rn(b4, sh[12 + c], r);
        r = __fmaf_rn(b5, sh[15 + c], r);
        r = __fmaf_rn(b6, sh[18 + c], r);
        r = __fmaf_rn(b7, sh[21 + c], r);
        res[c] = __fmaf_rn(b8, sh[24 + c], r);
      }
      if (deg > 2) {
        float b9 = __fmul_rn(__fmul_rn(kSH_C3[0], y), __fsub_rn(__fmul_rn(3.0f, xx), yy));
        float b10 = __fmul_rn(__fmul_rn(kSH_C3[1], xy), z);
        float zz4 = __fsub_rn(__fsub_rn(__fmul_rn(4.0f, zz), xx), yy);
        float b11 = __fmul_rn(__fmul_rn(kSH_C3[2], y), zz4);
        float b12 = __fmul_rn(__fmul_rn(kSH_C3[3], z),
                              __fsub_rn(__fsub_rn(__fmul_rn(2.0f, zz), __fmul_rn(3.0f, xx)), __fmul_rn(3.0f, yy)));
        float b13 = __fmul_rn(__fmul_rn(kSH_C3[4], x), zz4);
        float b14 = __fmul_rn(__fmul_rn(kSH_C3[5], z), __fsub_rn(xx, yy));
        float b15 = __fmul_rn(__fmul_rn(kSH_C3[6], x), __fsub_rn(xx, __fmul_rn(3.0f, yy)));
#pragma unroll
        for (int c = 0; c < 3; c++) {
          float r = res[c];
          r = __fmaf_rn(b9, sh[27 + c], r);
          r = __fmaf_rn(b10, sh[30 + c], r);
          r = __fmaf_rn(b11, sh[33 + c], r);
          r = __fmaf_rn(b12, sh[36 + c], r);
          r = __fmaf_rn(b13, sh[39 + c], r);
          r = __fmaf_rn(b14, sh[42 + c], r);
          res[c] = __fmaf_rn(b15, sh[45 + c], r);
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 3; c++) {
    float v = __fadd_rn(res[c], 0.5f);
    clamped[c] = (v < 0.0f);
    rgb[c] = (v < 0.0f) ? 0.0f : v;
  }
}

__global__ void __launch_bounds__(256) preprocess_kernel(GsForwardArgs a, GsGeom g, uint32_t* __restrict__ dkey,
                                                         uint32_t* __restrict__ dval, uint32_t* __restrict__ hist) {
  GS_PDL_PROLOGUE();
  __shared__ uint32_t s_hist[4 * 256];
  for (int i = threadIdx.x; i < 4 * 256; i += blockDim.x) s_hist[i] = 0;
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  if (threadIdx.x < 16) { s_view[threadIdx.x] = a.viewmatrix[threadIdx.x]; s_proj[threadIdx.x] = a.projmatrix[threadIdx.x]; }
  if (threadIdx.x < 3 && a.campos) s_cam[threadIdx.x] = a.campos[threadIdx.x];
  // The Gaussian's own inputs (all in bounds for a culled Gaussian too) are requested BEFORE the barrier behind the camera
  // loads and before the visibility / empty-rectangle branches: one memory round trip per thread instead of four.
  const int idx0 = blockIdx.x * blockDim.x + threadIdx.x;
  float mx = 0.f, my = 0.f, mz = 0.f, c6in[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, sc[3] = {0.f, 0.f, 0.f}, rq[4] = {0.f, 0.f, 0.f, 0.f};
  float colin[3] = {0.f, 0.f, 0.f}, oin = 0.f;
  auto load_inputs = [&](int idx) {
    mx = a.means3D[3 * idx]; my = a.means3D[3 * idx + 1]; mz = a.means3D[3 * idx + 2];
    if (a.cov3D_precomp) {
#pragma unroll
      for (int k = 0; k < 6; k++) c6in[k] = a.cov3D_precomp[6 * (size_t)idx + k];
    } else {
#pragma unroll
      for (int k = 0; k < 3; k++) sc[k] = a.scales[3 * (size_t)idx + k];
#pragma unroll
      for (int k = 0; k < 4; k++) rq[k] = a.rotations[4 * (size_t)idx + k];
    }
    if (a.colors_precomp) {
#pragma unroll
      for (int k = 0; k < 3; k++) colin[k] = a.colors_precomp[3 * (size_t)idx + k];
    }
    oin = a.opacities[idx];
  };
  if (idx0 < a.P) load_inputs(idx0);
  __syncthreads();
  const float* v = s_view;
  const float* pm = s_proj;
  const int W = a.width, H = a.height;
  const int gx = (W + GS_TILE - 1) / GS_TILE, gy = (H + GS_TILE - 1) / GS_TILE;
  const float focal_y = H / (2.0f * a.tan_fovy), focal_x = W / (2.0f * a.tan_fovx);  // rasterizer_impl.cu:222-223
  const float limx = __fmul_rn(a.tan_fovx, 1.3f), limy = __fmul_rn(a.tan_fovy, 1.3f);

  for (int idx = idx0; idx < a.P; idx += gridDim.x * blockDim.x) {
    uint32_t key = 0xFFFFFFFFu;
    int radius = 0;
    uint32_t touched = 0;
    if (idx != idx0) load_inputs(idx);
    {  // accumulators of the backward blend start at zero (no separate fill launch)
      float4* const ac = reinterpret_cast<float4*>(g.acc + (size_t)GS_ACC * idx);
      ac[0] = ac[1] = ac[2] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const float pvz = gs_dot3p(mx, v[2], my, v[6], mz, v[10], v[14]);
    if (pvz > 0.2f) {  // auxiliary.h:154
      float hx = gs_dot3p(mx, pm[0], my, pm[4], mz, pm[8], pm[12]);
      float hy = gs_dot3p(mx, pm[1], my, pm[5], mz, pm[9], pm[13]);
      float hw = gs_dot3p(mx, pm[3], my, pm[7], mz, pm[11], pm[15]);
      float p_w = __frcp_rn(__fadd_rn(hw, 0.0000001f));
      float projx = __fmul_rn(hx, p_w), projy = __fmul_rn(hy, p_w);
      float c6[6];
      if (a.cov3D_precomp) {
#pragma unroll
        for (int k = 0; k < 6; k++) c6[k] = c6in[k];
      } else {
        cov3d_from_scale_rot(sc, a.scale_modifier, rq, c6);
#pragma unroll
        for (int k = 0; k < 6; k++) g.cov3D[6 * (size_t)idx + k] = c6[k];
      }
      // computeCov2D, forward.cu:74-113
      float tx = gs_dot3p(mx, v[0], my, v[4], mz, v[8], v[12]);
      float ty = gs_dot3p(mx, v[1], my, v[5], mz, v[9], v[13]);
      float tz = pvz;
      float cx = fminf(limx, fmaxf(-limx, __fdiv_rn(tx, tz)));
      float cy = fminf(limy, fmaxf(-limy, __fdiv_rn(ty, tz)));
      float J00 = __fdiv_rn(focal_x, tz), J11 = __fdiv_rn(focal_y, tz);
      float tz2 = __fmul_rn(tz, tz);
      float J02 = __fdiv_rn(__fmul_rn(focal_x, __fmul_rn(cx, -tz)), tz2);
      float J12 = __fdiv_rn(__fmul_rn(focal_y, __fmul_rn(cy, -tz)), tz2);
      float T0[3], T1[3];
#pragma unroll
      for (int r = 0; r < 3; r++) {
        T0[r] = __fmaf_rn(v[2 + 4 * r], J02, __fmul_rn(v[4 * r], J00));
        T1[r] = __fmaf_rn(v[2 + 4 * r], J12, __fmul_rn(J11, v[1 + 4 * r]));
      }
      float a0 = gs_dot3(T0[0], c6[0], T0[1], c6[1], T0[2], c6[2]);
      float a1 = gs_dot3(T0[0], c6[1], T0[1], c6[3], T0[2], c6[4]);
      float a2 = gs_dot3(T0[0], c6[2], T0[1], c6[4], T0[2], c6[5]);
      float b0 = gs_dot3(T1[0], c6[0], T1[1], c6[1], T1[2], c6[2]);
      float b1 = gs_dot3(T1[0], c6[1], T1[1], c6[3], T1[2], c6[4]);
      float b2 = gs_dot3(T1[0], c6[2], T1[1], c6[4], T1[2], c6[5]);
      float cov_a = __fadd_rn(gs_dot3(T0[0], a0, T0[1], a1, T0[2], a2), 0.3f);
      float cov_b = gs_dot3(T0[0], b0, T0[1], b1, T0[2], b2);
      float cov_c = __fadd_rn(gs_dot3(T1[0], b0, T1[1], b1, T1[2], b2), 0.3f);
      float det = __fmaf_rn(cov_a, cov_c, -__fmul_rn(cov_b, cov_b));  // SASS: FFMA a*c - (b*b)
      if (det != 0.0f) {
        float det_inv = __frcp_rn(det);
        float con_a = __fmul_rn(cov_c, det_inv), con_b = __fmul_rn(-cov_b, det_inv), con_c = __fmul_rn(cov_a, det_inv);
        float mid = __fmul_rn(__fadd_rn(cov_a, cov_c), 0.5f);
        float sq = __fsqrt_rn(fmaxf(__fmaf_rn(mid, mid, -det), 0.1f));
        float lam = fmaxf(__fadd_rn(mid, sq), __fsub_rn(mid, sq));
        float my_radius = ceilf(__fmul_rn(__fsqrt_rn(lam), 3.f));
        float px = ndc2pix(projx, W), py = ndc2pix(projy, H);
        int irad = f2i_rz(my_radius);
        float rf = (float)irad;
        // getRect, auxiliary.h:46-56
        int x0 = min(gx, max(0, f2i_rz(__fmul_rn(__fsub_rn(px, rf), 0.0625f))));
        int y0 = min(gy, max(0, f2i_rz(__fmul_rn(__fsub_rn(py, rf), 0.0625f))));
        int x1 = min(gx, max(0, f2i_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(px, rf), 16.0f), -1.0f), 0.0625f))));
        int y1 = min(gy, max(0, f2i_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(py, rf), 16.0f), -1.0f), 0.0625f))));
        int area = (x1 - x0) * (y1 - y0);
        if (area != 0) {
          GsRec rec;
          if (a.colors_precomp) {
            rec.r = colin[0]; rec.g = colin[1]; rec.b = colin[2];
          } else {
            float rgb[3];
            uint8_t cl[3];
            color_from_sh(a.D, mx, my, mz, s_cam, a.shs + (size_t)idx * a.M * 3, rgb, cl);
            rec.r = rgb[0]; rec.g = rgb[1]; rec.b = rgb[2];
            g.clamped[3 * (size_t)idx] = cl[0]; g.clamped[3 * (size_t)idx + 1] = cl[1]; g.clamped[3 * (size_t)idx + 2] = cl[2];
          }
          float o = oin;
          rec.x = px; rec.y = py; rec.depth = pvz;
          rec.tau = __logf(255.0f * o);  // cull helper only (never decides a blend; see gs_blend.cu)
          rec.ca = con_a; rec.cb = con_b; rec.cc = con_c; rec.o = o;
          rec.pad = 0.f;
          float4* dst = reinterpret_cast<float4*>(g.rec + idx);
          dst[0] = make_float4(rec.x, rec.y, rec.tau, rec.depth);
          dst[1] = make_float4(rec.ca, rec.cb, rec.cc, rec.o);
          dst[2] = make_float4(rec.r, rec.g, rec.b, 0.f);
          g.rect[2 * (size_t)idx] = (uint32_t)x0 | ((uint32_t)y0 << 16);
          g.rect[2 * (size_t)idx + 1] = (uint32_t)x1 | ((uint32_t)y1 << 16);
          radius = irad;
          touched = (uint32_t)area;
          key = __float_as_uint(pvz);
        }
      }
    }
    a.radii[idx] = radius;
    g.tiles_touched[idx] = touched;
    dkey[idx] = key;
    dval[idx] = (uint32_t)idx;
#pragma unroll
    for (int p = 0; p < 4; p++) atomicAdd(&s_hist[p * 256 + ((key >> (8 * p)) & 255u)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 4 * 256; i += blockDim.x) {
    uint32_t c = s_hist[i];
    if (c) atomicAdd(&hist[i], c);
  }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  float z = gs_dot3p(means3D[3 * idx], view[2], means3D[3 * idx + 1], view[6], means3D[3 * idx + 2], view[10], view[14]);
  present[idx] = z > 0.2f;
}

}  // namespace

int gs_launch_preprocess(const GsForwardArgs& a, const GsGeom& g, const GsScratch& s, cudaStream_t st) {
  int blocks = (a.P + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  gs_launch(preprocess_kernel, dim3(blocks), dim3(256), 0, st, a, g, s.dkey[0], s.dval[0], s.hist);
  return gs_check(cudaGetLastError());
}

int gs_launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st) {
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, view, present);
  return gs_check(cudaGetLastError());
}
