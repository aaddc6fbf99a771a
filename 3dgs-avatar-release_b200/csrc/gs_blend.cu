// gs_blend.cu -- forward and backward alpha blending (replaces renderCUDA, cuda_rasterizer/forward.cu:261-374 and
// backward.cu:399-557).
//
// Work decomposition (B200-first, not the reference's): one CTA per 16x16 tile (the tile size defines the lists,
// ranges and n_contrib, so it is kept), but each of the 8 warps owns an 8x4 pixel block and walks the tile's
// depth-sorted list independently -- no CTA-wide barriers, per-warp early termination.  Each step a warp
//   1. loads 32 list entries lane-parallel (coalesced index load + one 48-byte record gather per lane),
//   2. culls them lane-parallel against its 8x4 block with a conservative bound of the Gaussian's maximum exponent
//      over the block (an entry is dropped only if no pixel of the block could pass the reference's alpha >= 1/255
//      test), stages the records in a per-warp shared-memory slab,
//   3. ballots the survivors and blends only those, all 32 lanes (= pixels) evaluating one entry per iteration
//      with shared-memory broadcast reads.
// For avatar frames ~90% of the (pixel, entry) evaluations of the reference are rejections; the cull removes most
// of them at 1/32 of the cost.  The surviving evaluations run the reference's exact test sequence in the
// reference's operation order, so final_T / n_contrib / colour are bit-identical to the oracle.
//
// Backward: same traversal back-to-front; the nine per-Gaussian gradient terms are reduced over the warp's 32 pixels
// with a transposing butterfly (14 shuffles instead of 45) BEFORE any atomic: 9 atomics per (warp, entry) instead
// of 9 per (pixel, entry).
#include "gs_common.cuh"

namespace {

constexpr int kWarps = 8;

struct Slab {  // per-warp staging of 32 list entries (SoA -> conflict-free lane-parallel stores, broadcast loads)
  float x[32], y[32], ca[32], cb[32], cc[32], o[32], r[32], g[32], b[32];
  uint32_t id[32];
};

// Conservative test: can Gaussian (x,y,conic,tau) reach alpha >= 1/255 anywhere in pixel block [X0,X1]x[Y0,Y1]?
__device__ __forceinline__ bool may_contribute(float x, float y, float ca, float cb, float cc, float tau, float X0,
                                               float X1, float Y0, float Y1) {
  float dxl = x - X1, dxh = x - X0, dyl = y - Y1, dyh = y - Y0;  // d = mean - pixel
  bool pd = (ca > 0.f) && (cc > 0.f) && (ca * cc - cb * cb > 0.f);
  bool centre_in = (dxl <= 0.f) && (dxh >= 0.f) && (dyl <= 0.f) && (dyh >= 0.f);
  if (!pd || centre_in) return true;
  float m = 3.0e38f;
  {
    float ica = 1.0f / ca, icc = 1.0f / cc;
#pragma unroll
    for (int k = 0; k < 2; k++) {
      float e = k ? dxh : dxl;  // edge dx = e
      float t = fminf(dyh, fmaxf(dyl, -cb * e * icc));
      float s = ca * e * e + cc * t * t;
      m = fminf(m, s + 2.f * cb * e * t - 1e-5f * s);
      e = k ? dyh : dyl;  // edge dy = e
      t = fminf(dxh, fmaxf(dxl, -cb * e * ica));
      s = ca * t * t + cc * e * e;
      m = fminf(m, s + 2.f * cb * e * t - 1e-5f * s);
    }
  }
  // power_max = -m/2 ; contributes only if power >= -tau.  NaN tau (opacity < 0) -> keep.
  return !(0.5f * m > tau + 0.01f);
}

__global__ void __launch_bounds__(kWarps * 32)
blend_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int W, int H,
                 const GsRec* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ final_T,
                 uint32_t* __restrict__ n_contrib, float* __restrict__ out_color) {
  __shared__ Slab s_slab[kWarps];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Slab& sl = s_slab[w];
  const int tile = blockIdx.y * gridDim.x + blockIdx.x;
  const uint2 range = ranges[tile];
  const int bx = blockIdx.x * GS_TILE + (w & 1) * 8, by = blockIdx.y * GS_TILE + (w >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pfx = (float)px, pfy = (float)py;
  const float X0 = (float)bx, X1 = (float)(bx + 7), Y0 = (float)by, Y1 = (float)(by + 3);
  bool done = !inside;
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
  uint32_t last = 0;

  for (uint32_t base = range.x; base < range.y; base += 32) {
    if (__all_sync(0xFFFFFFFFu, done)) break;
    const uint32_t e = base + lane;
    bool keep = false;
    if (e < range.y) {
      const uint32_t id = list[e];
      const float4* rp = reinterpret_cast<const float4*>(rec + id);
      const float4 r0 = __ldg(rp), r1 = __ldg(rp + 1), r2 = __ldg(rp + 2);
      keep = may_contribute(r0.x, r0.y, r1.x, r1.y, r1.z, r0.z, X0, X1, Y0, Y1);
      sl.x[lane] = r0.x; sl.y[lane] = r0.y;
      sl.ca[lane] = r1.x; sl.cb[lane] = r1.y; sl.cc[lane] = r1.z; sl.o[lane] = r1.w;
      sl.r[lane] = r2.x; sl.g[lane] = r2.y; sl.b[lane] = r2.z;
    }
    __syncwarp();
    uint32_t mask = __ballot_sync(0xFFFFFFFFu, keep);
    while (mask) {
      const int j = __ffs(mask) - 1;
      mask &= mask - 1;
      // reference test sequence, forward.cu:325-361, in the reference's FP32 operation order
      const float dx = __fsub_rn(sl.x[j], pfx), dy = __fsub_rn(sl.y[j], pfy);
      const float power = __fmaf_rn(__fmaf_rn(dx, __fmul_rn(dx, sl.ca[j]), __fmul_rn(dy, __fmul_rn(dy, sl.cc[j]))), -0.5f,
                                    -__fmul_rn(dy, __fmul_rn(dx, sl.cb[j])));  // SASS: FFMA q*-0.5 - cross
      const float alpha = fminf(0.99f, __fmul_rn(sl.o[j], gs_exp(power)));
      bool ok = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
      const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
      if (ok && test_T < 0.0001f) { done = true; ok = false; }
      if (ok) {
        C0 = __fmaf_rn(T, __fmul_rn(alpha, sl.r[j]), C0);
        C1 = __fmaf_rn(T, __fmul_rn(alpha, sl.g[j]), C1);
        C2 = __fmaf_rn(T, __fmul_rn(alpha, sl.b[j]), C2);
        T = test_T;
        last = base + j - range.x + 1;  // 1-based position in the tile list (contributor, forward.cu:328)
      }
      if (__all_sync(0xFFFFFFFFu, done)) { mask = 0; }
    }
    __syncwarp();
  }
  if (inside) {
    const size_t pix = (size_t)py * W + px, hw = (size_t)H * W;
    final_T[pix] = T;
    n_contrib[pix] = last;
    out_color[pix] = __fmaf_rn(T, bg[0], C0);
    out_color[hw + pix] = __fmaf_rn(T, bg[1], C1);
    out_color[2 * hw + pix] = __fmaf_rn(T, bg[2], C2);
  }
}

// Sum v[0..7] over the 32 lanes with a transposing butterfly: afterwards lane 4*k (k = 0..7) holds the warp total of
// v[k].  7 + 2 = 9 shuffles instead of 40.
__device__ __forceinline__ float warp_reduce8(float (&v)[8], int lane) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
  float a[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    float send = b4 ? v[k] : v[4 + k];
    float keepv = b4 ? v[4 + k] : v[k];
    a[k] = keepv + __shfl_xor_sync(0xFFFFFFFFu, send, 16);
  }
  float c[2];
#pragma unroll
  for (int k = 0; k < 2; k++) {
    float send = b3 ? a[k] : a[2 + k];
    float keepv = b3 ? a[2 + k] : a[k];
    c[k] = keepv + __shfl_xor_sync(0xFFFFFFFFu, send, 8);
  }
  float send = b2 ? c[0] : c[1];
  float keepv = b2 ? c[1] : c[0];
  float r = keepv + __shfl_xor_sync(0xFFFFFFFFu, send, 4);
  r += __shfl_xor_sync(0xFFFFFFFFu, r, 2);
  r += __shfl_xor_sync(0xFFFFFFFFu, r, 1);
  return r;  // valid in lanes with (lane & 3) == 0: value index = (b4?4:0) + (b3?2:0) + (b2?1:0)
}

__global__ void __launch_bounds__(kWarps * 32)
blend_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int W, int H,
                 const float* __restrict__ bg, const GsRec* __restrict__ rec, const float* __restrict__ final_Ts,
                 const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                 float* __restrict__ dL_dmean2D /*[P,3]*/, float* __restrict__ dL_dconic /*[P,4]*/,
                 float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolors) {
  __shared__ Slab s_slab[kWarps];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Slab& sl = s_slab[w];
  const int tile = blockIdx.y * gridDim.x + blockIdx.x;
  const uint2 range = ranges[tile];
  if (range.y <= range.x) return;
  const int bx = blockIdx.x * GS_TILE + (w & 1) * 8, by = blockIdx.y * GS_TILE + (w >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pfx = (float)px, pfy = (float)py;
  const float X0 = (float)bx, X1 = (float)(bx + 7), Y0 = (float)by, Y1 = (float)(by + 3);
  const size_t pix = (size_t)py * W + px, hw = (size_t)H * W;
  const float T_final = inside ? final_Ts[pix] : 0.f;
  float T = T_final;
  const int last_contributor = inside ? (int)n_contrib[pix] : 0;
  float dp0 = 0.f, dp1 = 0.f, dp2 = 0.f;
  if (inside) { dp0 = dL_dpixels[pix]; dp1 = dL_dpixels[hw + pix]; dp2 = dL_dpixels[2 * hw + pix]; }
  const float bg_dot_dpixel = bg[0] * dp0 + bg[1] * dp1 + bg[2] * dp2;
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, last_alpha = 0.f;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;  // backward.cu:460-461

  // entries at 0-based positions >= max(last_contributor) over the warp are never touched
  int wmax = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xFFFFFFFFu, wmax, o));
  if (wmax == 0) return;
  // batches aligned to the start of the range so that positions map to lanes: pos = b*32 + lane
  for (int b = (wmax - 1) >> 5; b >= 0; b--) {
    const int pos = b * 32 + lane;
    bool keep = false;
    if (pos < wmax) {
      const uint32_t id = list[range.x + pos];
      const float4* rp = reinterpret_cast<const float4*>(rec + id);
      const float4 r0 = __ldg(rp), r1 = __ldg(rp + 1), r2 = __ldg(rp + 2);
      keep = may_contribute(r0.x, r0.y, r1.x, r1.y, r1.z, r0.z, X0, X1, Y0, Y1);
      sl.x[lane] = r0.x; sl.y[lane] = r0.y;
      sl.ca[lane] = r1.x; sl.cb[lane] = r1.y; sl.cc[lane] = r1.z; sl.o[lane] = r1.w;
      sl.r[lane] = r2.x; sl.g[lane] = r2.y; sl.b[lane] = r2.z;
      sl.id[lane] = id;
    }
    __syncwarp();
    uint32_t mask = __ballot_sync(0xFFFFFFFFu, keep);
    while (mask) {
      const int j = 31 - __clz(mask);  // back to front
      mask &= ~(1u << j);
      const int epos = b * 32 + j;
      const float cx = sl.ca[j], cy = sl.cb[j], cz = sl.cc[j], o = sl.o[j];
      const float dx = __fsub_rn(sl.x[j], pfx), dy = __fsub_rn(sl.y[j], pfy);
      const float power = __fmaf_rn(__fmaf_rn(dx, __fmul_rn(dx, cx), __fmul_rn(dy, __fmul_rn(dy, cz))), -0.5f,
                                    -__fmul_rn(dy, __fmul_rn(dx, cy)));
      const float G = gs_exp(power);
      const float alpha = fminf(0.99f, __fmul_rn(o, G));
      const bool ok = (epos < last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
      if (!__any_sync(0xFFFFFFFFu, ok)) continue;
      float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, v8 = 0.f;
      if (ok) {
        T = T / (1.f - alpha);
        const float dchannel_dcolor = alpha * T;
        const float c0 = sl.r[j], c1 = sl.g[j], c2 = sl.b[j];
        acc0 = last_alpha * lc0 + (1.f - last_alpha) * acc0; lc0 = c0;
        acc1 = last_alpha * lc1 + (1.f - last_alpha) * acc1; lc1 = c1;
        acc2 = last_alpha * lc2 + (1.f - last_alpha) * acc2; lc2 = c2;
        float dL_dalpha = (c0 - acc0) * dp0 + (c1 - acc1) * dp1 + (c2 - acc2) * dp2;
        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
        const float dL_dG = o * dL_dalpha;
        const float gdx = G * dx, gdy = G * dy;
        const float dG_ddelx = -gdx * cx - gdy * cy;
        const float dG_ddely = -gdy * cz - gdx * cy;
        v[0] = dL_dG * dG_ddelx * ddelx_dx;
        v[1] = dL_dG * dG_ddely * ddely_dy;
        v[2] = -0.5f * gdx * dx * dL_dG;
        v[3] = -0.5f * gdx * dy * dL_dG;
        v[4] = -0.5f * gdy * dy * dL_dG;
        v[5] = G * dL_dalpha;
        v[6] = dchannel_dcolor * dp0;
        v[7] = dchannel_dcolor * dp1;
        v8 = dchannel_dcolor * dp2;
      }
      const float r = warp_reduce8(v, lane);
#pragma unroll
      for (int o2 = 16; o2 > 0; o2 >>= 1) v8 += __shfl_xor_sync(0xFFFFFFFFu, v8, o2);
      const uint32_t id = sl.id[j];
      if ((lane & 3) == 0) {
        const int k = ((lane & 16) ? 4 : 0) + ((lane & 8) ? 2 : 0) + ((lane & 4) ? 1 : 0);
        float* dst;
        switch (k) {
          case 0: dst = dL_dmean2D + 3 * (size_t)id; break;
          case 1: dst = dL_dmean2D + 3 * (size_t)id + 1; break;
          case 2: dst = dL_dconic + 4 * (size_t)id; break;
          case 3: dst = dL_dconic + 4 * (size_t)id + 1; break;
          case 4: dst = dL_dconic + 4 * (size_t)id + 3; break;
          case 5: dst = dL_dopacity + id; break;
          case 6: dst = dL_dcolors + 3 * (size_t)id; break;
          default: dst = dL_dcolors + 3 * (size_t)id + 1; break;
        }
        atomicAdd(dst, r);
      } else if (lane == 1) {
        atomicAdd(dL_dcolors + 3 * (size_t)id + 2, v8);
      }
    }
    __syncwarp();
  }
}

}  // namespace

int gs_launch_blend_fwd(const GsForwardArgs& a, const GsGeom& g, const GsImage& im, const uint32_t* list,
                        cudaStream_t st) {
  dim3 grid((a.width + GS_TILE - 1) / GS_TILE, (a.height + GS_TILE - 1) / GS_TILE);
  blend_fwd_kernel<<<grid, kWarps * 32, 0, st>>>(im.ranges, list, a.width, a.height, g.rec, a.background, im.final_T,
                                                 im.n_contrib, a.out_color);
  return gs_check(cudaGetLastError());
}

int gs_launch_blend_bwd(const GsBackwardArgs& a, const GsGeom& g, const GsImage& im, const uint32_t* list,
                        cudaStream_t st) {
  dim3 grid((a.width + GS_TILE - 1) / GS_TILE, (a.height + GS_TILE - 1) / GS_TILE);
  blend_bwd_kernel<<<grid, kWarps * 32, 0, st>>>(im.ranges, list, a.width, a.height, a.background, g.rec, im.final_T,
                                                 im.n_contrib, a.dL_dpix, a.dL_dmean2D, a.dL_dconic, a.dL_dopacity,
                                                 a.dL_dcolor);
  return gs_check(cudaGetLastError());
}
