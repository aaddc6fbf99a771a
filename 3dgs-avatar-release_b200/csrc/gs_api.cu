// gs_api.cu -- extern "C" entry points of libgsplat_b200.so (see include/gsplat_b200.h for the reference
// interfaces they replace).  Orchestration only: carve the caller's workspaces, launch the stage kernels on the
// caller's stream, never allocate, never synchronise (except in debug mode, like CHECK_CUDA auxiliary.h:166-173).
#include <stdio.h>
#include <stdlib.h>

#include "gs_common.cuh"

static thread_local int g_last_cuda_error = 0;

// exp of the blend kernels: 0 = CUDA expf, the reference's instruction sequence (default); 1 = the gs_exp polynomial that the
// CPU oracle reproduces bit for bit (parity tests against the oracle).  GS_EXP_MODE in the environment sets the initial value.
static int g_exp_mode = -1;
int gs_exp_mode() {
  if (g_exp_mode < 0) {
    const char* e = getenv("GS_EXP_MODE");
    g_exp_mode = (e && e[0] == '1') ? 1 : 0;
  }
  return g_exp_mode;
}

int gs_check(cudaError_t e) {
  if (e == cudaSuccess) return GS_OK;
  g_last_cuda_error = (int)e;
  return GS_ERR_CUDA;
}

#define GS_TRY(expr)            \
  do {                          \
    int _s = (expr);            \
    if (_s != GS_OK) return _s; \
  } while (0)

// ---- optional per-stage timing (bench.py's roofline leg): CUDA events on the launching stream ----
enum { ST_PRE = 0, ST_BIN, ST_BLEND_F, ST_BLEND_B, ST_PRE_B, ST_POSE_F, ST_POSE_B, ST_COUNT };
static int g_prof_on = 0;
constexpr int GS_MAX_DEV = 64;
struct ProfDev { cudaEvent_t ev[ST_COUNT][2]; int valid[ST_COUNT]; int init; };
static ProfDev g_prof[GS_MAX_DEV];  // events belong to the device that was current when they were created: one set per device
static ProfDev* prof_dev() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= GS_MAX_DEV) return nullptr;
  return &g_prof[dev];
}
void gs_prof_mark(int stage, int end, cudaStream_t st) {
  if (!g_prof_on) return;
  ProfDev* p = prof_dev();
  if (!p) return;
  if (!p->init) {
    for (int i = 0; i < ST_COUNT; i++) { cudaEventCreate(&p->ev[i][0]); cudaEventCreate(&p->ev[i][1]); p->valid[i] = 0; }
    p->init = 1;
  }
  cudaEventRecord(p->ev[stage][end], st);
  if (end) p->valid[stage] = 1;
}

static int debug_sync(int debug, cudaStream_t st) {
  if (!debug) return GS_OK;
  return gs_check(cudaStreamSynchronize(st));
}

extern "C" {

int gs_last_cuda_error(void) { return g_last_cuda_error; }
const char* gs_version(void) { return "gsplat_b200 0.1 (sm_100a)"; }

size_t gs_geom_bytes(int32_t P) { return gs_carve_geom(nullptr, P, nullptr); }
size_t gs_image_bytes(int32_t width, int32_t height) { return gs_carve_image(nullptr, width, height, nullptr); }
size_t gs_list_bytes(int64_t pair_capacity, int32_t width, int32_t height) {
  return gs_carve_list(nullptr, pair_capacity > 0 ? pair_capacity : 1, width, height, nullptr);
}
/* Largest pair capacity (a multiple of 4096) whose list workspace fits list_bytes: inverse of gs_list_bytes on that
 * lattice, where it is strictly increasing.  Lets a caller that only kept the buffer recover the capacity. */
int64_t gs_list_capacity(size_t list_bytes, int32_t width, int32_t height) {
  int64_t lo = 0, hi = (int64_t)(list_bytes / 4 / 4096) + 1;  // in units of 4096 pairs
  while (lo < hi) {
    int64_t mid = (lo + hi + 1) / 2;
    if (gs_carve_list(nullptr, mid * 4096, width, height, nullptr) <= list_bytes) lo = mid; else hi = mid - 1;
  }
  return lo * 4096;
}
size_t gs_scratch_bytes(int32_t P, int32_t width, int32_t height, int64_t pair_capacity) {
  return gs_carve_scratch(nullptr, P, width, height, pair_capacity > 0 ? pair_capacity : 1, nullptr);
}

int gs_forward(const GsForwardArgs* args, void* geom, size_t geom_bytes, void* image, size_t image_bytes, void* list,
               size_t list_bytes, void* scratch, size_t scratch_bytes, int64_t pair_capacity,
               int64_t* num_rendered_pinned_host, void* stream) {
  if (!args || !geom || !image || !list || !scratch) return GS_ERR_BAD_ARG;
  const GsForwardArgs& a = *args;
  cudaStream_t st = (cudaStream_t)stream;
  if (a.P <= 0 || a.width <= 0 || a.height <= 0 || pair_capacity <= 0) return GS_ERR_BAD_ARG;
  if (!a.means3D || !a.opacities || !a.viewmatrix || !a.projmatrix || !a.background || !a.out_color || !a.radii)
    return GS_ERR_BAD_ARG;
  if (!a.colors_precomp && (!a.shs || !a.campos || a.M <= 0)) return GS_ERR_BAD_ARG;
  if (!a.cov3D_precomp && (!a.scales || !a.rotations)) return GS_ERR_BAD_ARG;
  if ((a.width + GS_TILE - 1) / GS_TILE > 65535 || (a.height + GS_TILE - 1) / GS_TILE > 65535) return GS_ERR_BAD_ARG;
  GsGeom g;
  GsImage im;
  GsScratch s;
  if (gs_carve_geom((char*)geom, a.P, &g) > geom_bytes) return GS_ERR_WORKSPACE_SMALL;
  if (gs_carve_image((char*)image, a.width, a.height, &im) > image_bytes) return GS_ERR_WORKSPACE_SMALL;
  if (gs_carve_scratch((char*)scratch, a.P, a.width, a.height, pair_capacity, &s) > scratch_bytes)
    return GS_ERR_WORKSPACE_SMALL;
  GsList l;
  if (gs_carve_list((char*)list, pair_capacity, a.width, a.height, &l) > list_bytes) return GS_ERR_WORKSPACE_SMALL;

  const size_t tiles = (size_t)((a.width + GS_TILE - 1) / GS_TILE) * ((a.height + GS_TILE - 1) / GS_TILE);
  GS_TRY(gs_check(cudaMemsetAsync(g.hdr, 0, sizeof(GsHeader), st)));
  GS_TRY(gs_check(cudaMemsetAsync(s.zero_begin, 0, (size_t)((char*)s.zero_end - (char*)s.zero_begin), st)));
  GS_TRY(gs_check(cudaMemsetAsync(im.ranges, 0, tiles * sizeof(uint2), st)));  // rasterizer_impl.cu:310

  gs_prof_mark(ST_PRE, 0, st);
  GS_TRY(gs_launch_preprocess(a, g, s, st));
  gs_prof_mark(ST_PRE, 1, st);
  GS_TRY(debug_sync(a.debug, st));
  gs_prof_mark(ST_BIN, 0, st);
  GS_TRY(gs_launch_binning(a, g, im, s, l, pair_capacity, st));
  gs_prof_mark(ST_BIN, 1, st);
  GS_TRY(debug_sync(a.debug, st));
  if (num_rendered_pinned_host)
    GS_TRY(gs_check(cudaMemcpyAsync(num_rendered_pinned_host, &g.hdr->num_rendered, 2 * sizeof(int64_t),
                                    cudaMemcpyDeviceToHost, st)));
  gs_prof_mark(ST_BLEND_F, 0, st);
  GS_TRY(gs_launch_blend_fwd(a, g, im, l, s, st));
  gs_prof_mark(ST_BLEND_F, 1, st);
  GS_TRY(debug_sync(a.debug, st));
  return GS_OK;
}

int gs_read_num_rendered(const void* geom, int64_t* num_rendered_host, void* stream) {
  if (!geom || !num_rendered_host) return GS_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long hdr[2];
  GS_TRY(gs_check(cudaMemcpyAsync(hdr, geom, sizeof(hdr), cudaMemcpyDeviceToHost, st)));
  GS_TRY(gs_check(cudaStreamSynchronize(st)));
  num_rendered_host[0] = (int64_t)hdr[0];
  num_rendered_host[1] = (int64_t)hdr[1];  // low 32 bits: overflow flag, high 32 bits: internal error flag
  return GS_OK;
}

int gs_backward(const GsBackwardArgs* args, const void* geom, size_t geom_bytes, const void* image, size_t image_bytes,
                const void* list, size_t list_bytes, int64_t pair_capacity, void* stream) {
  if (!args || !geom || !image || !list || pair_capacity <= 0) return GS_ERR_BAD_ARG;
  const GsBackwardArgs& a = *args;
  cudaStream_t st = (cudaStream_t)stream;
  if (a.P <= 0) return GS_ERR_BAD_ARG;
  if (!a.dL_dpix || !a.dL_dmean2D || !a.dL_dopacity || !a.dL_dcolor || !a.dL_dmean3D || !a.dL_dcov3D)
    return GS_ERR_BAD_ARG;
  if (a.shs && (!a.dL_dsh || !a.campos)) return GS_ERR_BAD_ARG;
  if (a.scales && (!a.rotations || !a.dL_dscale || !a.dL_drot)) return GS_ERR_BAD_ARG;
  GsGeom g;
  GsImage im;
  if (gs_carve_geom((char*)geom, a.P, &g) > geom_bytes) return GS_ERR_WORKSPACE_SMALL;
  if (gs_carve_image((char*)image, a.width, a.height, &im) > image_bytes) return GS_ERR_WORKSPACE_SMALL;
  GsList l;
  if (gs_carve_list((char*)list, pair_capacity, a.width, a.height, &l) > list_bytes) return GS_ERR_WORKSPACE_SMALL;
  gs_prof_mark(ST_BLEND_B, 0, st);
  GS_TRY(gs_launch_blend_bwd(a, g, im, l, st));
  gs_prof_mark(ST_BLEND_B, 1, st);
  GS_TRY(debug_sync(a.debug, st));
  gs_prof_mark(ST_PRE_B, 0, st);
  GS_TRY(gs_launch_preprocess_bwd(a, g, st));
  gs_prof_mark(ST_PRE_B, 1, st);
  GS_TRY(debug_sync(a.debug, st));
  return GS_OK;
}

int gs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                    uint8_t* present, void* stream) {
  (void)projmatrix;
  if (P <= 0) return GS_OK;
  if (!means3D || !viewmatrix || !present) return GS_ERR_BAD_ARG;
  return gs_launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
}

/* sizeof of the ABI structs (0 GsForwardArgs, 1 GsBackwardArgs, 2 PoseArgs, 3 PoseBackwardArgs, 4 NonRigidArgs, 5 SkinArgs, 6 ColorArgs):
 * lets a binding check its own struct declarations against the library it loaded */
size_t gs_sizeof(int32_t which) {
  switch (which) {
    case 0: return sizeof(GsForwardArgs);
    case 1: return sizeof(GsBackwardArgs);
    case 2: return sizeof(PoseArgs);
    case 3: return sizeof(PoseBackwardArgs);
    case 4: return sizeof(NonRigidArgs);
    case 5: return sizeof(SkinArgs);
    case 6: return sizeof(ColorArgs);
    case 7: return sizeof(AdamArgs);
    case 8: return sizeof(PoseEncoderArgs);
    case 9: return sizeof(SlabAllReduceArgs);
    case 10: return sizeof(NonRigidBackwardArgs);
    case 11: return sizeof(SkinBackwardArgs);
    case 12: return sizeof(ColorBackwardArgs);
    case 13: return sizeof(PoseEncoderBackwardArgs);
    default: return 0;
  }
}

void gs_profile_enable(int on) { g_prof_on = on; }

int gs_set_exp_mode(int mode) {
  const int prev = gs_exp_mode();
  g_exp_mode = mode ? 1 : 0;
  return prev;
}

/* counters[0] = E_bwd, the (pixel, entry) pairs the last forward over `geom` blended; counters[1] = its forward work items.
 * Synchronises the stream. */
int gs_read_counters(const void* geom, uint64_t* counters_host, void* stream) {
  if (!geom || !counters_host) return GS_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  GsHeader h;
  GS_TRY(gs_check(cudaMemcpyAsync(&h, geom, sizeof(h), cudaMemcpyDeviceToHost, st)));
  GS_TRY(gs_check(cudaStreamSynchronize(st)));
  counters_host[0] = h.evals_blended;
  counters_host[1] = h.n_fitems;
  return GS_OK;
}

/* ms[7] = preprocess, binning (sorts+emit+ranges), blend fwd, blend bwd, preprocess bwd, pose fwd, pose bwd of the LAST call
 * (-1 where no stage was recorded).  Synchronises on the recorded events. */
int gs_profile_read(float* ms_host) {
  ProfDev* p = prof_dev();  // the CURRENT device's events
  for (int i = 0; i < ST_COUNT; i++) {
    ms_host[i] = -1.f;
    if (p && p->init && p->valid[i]) {
      if (cudaEventSynchronize(p->ev[i][1]) != cudaSuccess) return GS_ERR_CUDA;
      if (cudaEventElapsedTime(&ms_host[i], p->ev[i][0], p->ev[i][1]) != cudaSuccess) return GS_ERR_CUDA;
    }
  }
  return GS_OK;
}

}  // extern "C"

// FP32 FMA throughput probe (bench.py measures the FP32 roofline denominator with it: it is not in MEASURED_PEAKS.json)
namespace {
__global__ void fma_probe_kernel(int iters, float* out) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
  const float m = 0.999f, c = 1e-3f;
  for (int i = 0; i < iters; i++) {
    a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
    a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
  }
  float r = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
  if (r == 123.456f) out[0] = r;
}
}  // namespace
extern "C" int gs_fma_probe(int blocks, int threads, int iters, float* out_dev, void* stream) {
  fma_probe_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, out_dev);
  return gs_check(cudaGetLastError());
}

// ---- debug accessors ----
namespace {
__global__ void dbg_rec_kernel(int which, int P, const GsRec* __restrict__ rec, const int* __restrict__ radii,
                               float* __restrict__ dst) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const bool vis = radii[i] > 0;  // records of culled Gaussians are never written
  GsRec r = rec[i];
  if (which == 5) { dst[2 * i] = vis ? r.x : 0.f; dst[2 * i + 1] = vis ? r.y : 0.f; }
  else if (which == 6) dst[i] = vis ? r.depth : 0.f;
  else if (which == 7) { dst[4 * i] = vis ? r.ca : 0.f; dst[4 * i + 1] = vis ? r.cb : 0.f; dst[4 * i + 2] = vis ? r.cc : 0.f; dst[4 * i + 3] = vis ? r.o : 0.f; }
  else if (which == 9) { dst[3 * i] = vis ? r.r : 0.f; dst[3 * i + 1] = vis ? r.g : 0.f; dst[3 * i + 2] = vis ? r.b : 0.f; }
}
__global__ void dbg_keys_kernel(int tiles, const uint2* __restrict__ ranges, const uint32_t* __restrict__ list,
                                const GsRec* __restrict__ rec, unsigned long long* __restrict__ dst) {
  int t = blockIdx.x;
  if (t >= tiles) return;
  uint2 r = ranges[t];
  for (uint32_t i = r.x + threadIdx.x; i < r.y; i += blockDim.x)
    dst[i] = ((unsigned long long)t << 32) | (unsigned long long)__float_as_uint(rec[list[i]].depth);
}
}  // namespace

extern "C" int gs_debug_read(int32_t which, const GsForwardArgs* args, const void* geom, const void* image,
                             const void* list, int64_t num_rendered, void* dst, size_t dst_bytes, void* stream) {
  if (!args || !geom || !image || !dst) return GS_ERR_BAD_ARG;
  const GsForwardArgs& a = *args;
  cudaStream_t st = (cudaStream_t)stream;
  GsGeom g;
  GsImage im;
  gs_carve_geom((char*)geom, a.P, &g);
  gs_carve_image((char*)image, a.width, a.height, &im);
  const size_t tiles = (size_t)((a.width + GS_TILE - 1) / GS_TILE) * ((a.height + GS_TILE - 1) / GS_TILE);
  const size_t hw = (size_t)a.width * a.height;
  const void* src = nullptr;
  size_t bytes = 0;
  switch (which) {
    case 0: src = im.ranges; bytes = tiles * 8; break;
    case 1: src = im.n_contrib; bytes = hw * 4; break;
    case 2: src = im.final_T; bytes = hw * 4; break;
    case 3: src = list; bytes = (size_t)num_rendered * 4; break;
    case 8: src = g.tiles_touched; bytes = (size_t)a.P * 4; break;
    case 10: src = g.cov3D; bytes = (size_t)a.P * 24; break;
    case 11: src = g.clamped; bytes = (size_t)a.P * 3; break;
    default: break;
  }
  if (src) {
    if (bytes > dst_bytes) return GS_ERR_WORKSPACE_SMALL;
    if (bytes == 0) return GS_OK;
    return gs_check(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st));
  }
  if (which == 4) {
    if ((size_t)num_rendered * 8 > dst_bytes) return GS_ERR_WORKSPACE_SMALL;
    if (num_rendered == 0) return GS_OK;
    dbg_keys_kernel<<<(unsigned)tiles, 128, 0, st>>>((int)tiles, im.ranges, (const uint32_t*)list, g.rec,
                                                     (unsigned long long*)dst);
    return gs_check(cudaGetLastError());
  }
  if (which == 5 || which == 6 || which == 7 || which == 9) {
    size_t per = which == 5 ? 8 : which == 6 ? 4 : which == 7 ? 16 : 12;
    if ((size_t)a.P * per > dst_bytes) return GS_ERR_WORKSPACE_SMALL;
    dbg_rec_kernel<<<(a.P + 255) / 256, 256, 0, st>>>(which, a.P, g.rec, a.radii, (float*)dst);
    return gs_check(cudaGetLastError());
  }
  return GS_ERR_BAD_ARG;
}
