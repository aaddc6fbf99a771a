// gs_common.cuh -- shared declarations of the sm_100a kernels behind libgsplat_b200.so.
//
// Numerical contract ("canonical op order"): every decision-critical FP32 expression (depth bits of the
// sort key, radius, tile rectangle, blend skip / termination tests) is written with explicit
// __fmaf_rn/__fmul_rn/__fadd_rn intrinsics (which the compiler never re-associates or re-contracts) in
// the order nvcc emits for the reference kernels (cuda_rasterizer/forward.cu, auxiliary.h), so the
// integer artefacts are bit-identical to the reference build and to the CPU oracle.  exp() in the blend is CUDA's expf --
// the instruction sequence nvcc emits for the reference's `exp(power)` (forward.cu:344, backward.cu:483), so that the
// alpha >= 1/255 and T < 1e-4 decisions are the reference's bit for bit -- or, in oracle mode (gs_set_exp_mode(1)),
// gs_exp(): a fixed FMA polynomial that the CPU oracle reproduces bit for bit.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/gsplat_b200.h"

#define GS_TILE 16          // BLOCK_X = BLOCK_Y = 16 (cuda_rasterizer/config.h:16-17); defines keys/ranges/n_contrib
#define GS_ALIGN 256
#ifndef GS_SEG
#define GS_SEG 128          // list entries per backward work item = spacing of the forward's per-pixel checkpoints
#endif
#ifndef GS_FWD_CHUNK
#define GS_FWD_CHUNK 256    // list entries per forward work item (tile, pixel block, chunk): a multiple of GS_SEG.  Measured at
#endif                      // cfg-A: 256 entries / 4 warps / 5 CTAs per SM 78 us, 512 / 8 warps / 3 CTAs 88 us (profiles/r2)
#define GS_FWD_NCH_BINS 2048  // plan_kernel sorts tiles by chunk count with a counting sort of this many bins
#define GS_MASK_WORDS (4 * GS_SEG / 32)  // rowmask words per (segment, pixel block): 4 batches x 4 pixel groups of 4x2
#define GS_HAND_WORDS (32 * 6)  // hand-off slot of one (tile, pixel block, chunk): 32 lanes x {T, C0, C1, C2} + {last|flags, C3}

// ---- workspace layouts (private; reached from outside only through gs_debug_read) ----
struct GsHeader {            // first 256 bytes of the geom workspace
  unsigned long long num_rendered;  // N, written by the emit kernel
  unsigned int overflow;            // N > pair_capacity
  unsigned int error;               // look-back spin limit hit etc.
  unsigned long long evals_blended; // E_bwd: (pixel, entry) pairs the forward blended (counted by the forward's chain warps)
  unsigned int n_fitems;            // forward work items (tile, chunk) of the non-empty tiles, written by plan_kernel
  unsigned int n_empty_tiles;       // ... followed by this many single items of the empty tiles
  unsigned int fwd_ticket;          // next work unit of the persistent forward blend
  unsigned int fwd_order;           // order of the forward work items (plan_body): 0 longest-remaining-chain first, 1 chunk level by level
  unsigned int n_nonempty_tiles;    // tiles with a list (level order: hand-off slot of an item = successor's index - this)
  unsigned int pad[53];
};

// One packed record per Gaussian, what both blend kernels gather per list entry (3 x LDG.128).
struct __align__(16) GsRec {
  float x, y, tau, depth;   // pixel-space mean, ln(255*opacity), view depth
  float ca, cb, cc, o;      // conic (inverse 2D covariance) + opacity
  float r, g, b, pad;       // colour used by the blend (colors_precomp or SH result)
};

struct GsGeom {
  GsHeader* hdr;
  GsRec* rec;               // [P]
  uint32_t* rect;           // [P*2]  (xmin | ymin<<16), (xmax | ymax<<16)
  uint32_t* tiles_touched;  // [P]
  float* cov3D;             // [P*6]  (only written when computed from scales/rotations)
  uint8_t* clamped;         // [P*3]  (SH path)
  float* acc;               // [P*12] accumulators of the backward blend, one 48-byte record per Gaussian: raw moments sum g dx,
                            // g dy, g dx^2, g dx dy, g dy^2, g (g = G * dL/dalpha), colour gradient r, g, b, 3 pad.  Zeroed by
                            // preprocess_kernel (forward) and again by preprocess_bwd_kernel after it has consumed them
};
#define GS_ACC 12

struct GsImage {
  uint2* ranges;            // [tiles]
  uint32_t* n_contrib;      // [H*W]
  float* final_T;           // [H*W]
  uint32_t* seg_off;        // [tiles+1] exclusive prefix of ceil(len/GS_SEG) over tiles; [tiles] = number of work items
};

// The pair-capacity-sized workspace kept for backward (the reference's binningBuffer).
struct GsList {
  uint32_t* point_list;     // [cap] Gaussian index per (tile, depth)-sorted pair
  uint2* items;             // [max_items] (tile, segment) work items of the backward blend
  float4* ckpt;             // [(max_items + tiles) * 256] per-pixel (T, C.rgb) at every segment start + final state
  float* ckpt_op;           // same slots: accumulated opacity channel (written only by the fused opacity pass)
  uint32_t* rowmask;        // [max_items][8 pixel blocks][GS_SEG / 32][4 groups] per 32-entry batch and 4x2 pixel group: entries at
                            // least one pixel of the strip blended in the forward pass (what the backward blend visits)
  uint4* fitems;            // [max_fitems] forward work items (tile, chunk, hand-off slot it sends its chain state to or ~0, slot
                            // it receives from or ~0); a predecessor always has a smaller index
  uint32_t* order;          // [tiles] tiles by descending list length (chunk count): launch order of the per-block forward blend
  size_t max_items, max_fitems;
};

struct GsScratch {
  uint32_t* dkey[2];        // [P] depth bits ping-pong
  uint32_t* dval[2];        // [P] Gaussian index ping-pong
  uint32_t* tkey[2];        // [cap] tile id per pair ping-pong
  uint32_t* tval[2];        // [cap] Gaussian index per pair ping-pong (the last pass writes into the list)
  // ---- one contiguous region zeroed by a single memset at the start of every forward ----
  uint32_t* zero_begin;
  uint32_t* hist;           // [8][256] digit histograms (passes 0-3 depth, 4-7 tile id)
  uint32_t* counters;       // [64] tickets: [pass] for the sorts, [8] for the emit kernel
  uint32_t* status_d;       // [4][blocks_d][256] onesweep look-back status, depth passes
  uint32_t* status_t;       // [4][blocks_t][256] onesweep look-back status, tile passes
  unsigned long long* scan_status;  // [blocks_e] emit-kernel look-back
  uint32_t* tile_total;     // [tiles] pairs per tile (direct placement)
  uint32_t* hand;           // [cap / GS_FWD_CHUNK][8 pixel blocks][GS_HAND_WORDS] chain state handed from chunk c to chunk c+1 of
                            // a tile's forward blend; 0 = not yet published (validity is embedded in the vectors)
  uint32_t* zero_end;
  size_t blocks_d, blocks_t, blocks_e;
};

static inline __host__ __device__ size_t gs_align_up(size_t v) { return (v + GS_ALIGN - 1) / GS_ALIGN * GS_ALIGN; }

template <typename T>
static inline __host__ void gs_carve(char*& p, T*& out, size_t count) {
  out = reinterpret_cast<T*>(p);
  p += gs_align_up(count * sizeof(T));
}

static inline __host__ size_t gs_carve_geom(char* base, int P, GsGeom* g) {
  char* p = base;
  GsGeom t;
  gs_carve(p, t.hdr, 1);
  gs_carve(p, t.rec, (size_t)P);
  gs_carve(p, t.rect, (size_t)P * 2);
  gs_carve(p, t.tiles_touched, (size_t)P);
  gs_carve(p, t.cov3D, (size_t)P * 6);
  gs_carve(p, t.clamped, (size_t)P * 3);
  gs_carve(p, t.acc, (size_t)P * GS_ACC);
  if (g) *g = t;
  return (size_t)(p - base);
}

static inline __host__ size_t gs_carve_image(char* base, int W, int H, GsImage* im) {
  char* p = base;
  GsImage t;
  size_t tiles = (size_t)((W + GS_TILE - 1) / GS_TILE) * ((H + GS_TILE - 1) / GS_TILE);
  gs_carve(p, t.ranges, tiles);
  gs_carve(p, t.n_contrib, (size_t)W * H);
  gs_carve(p, t.final_T, (size_t)W * H);
  gs_carve(p, t.seg_off, tiles + 1);
  if (im) *im = t;
  return (size_t)(p - base);
}

static inline __host__ size_t gs_carve_list(char* base, int64_t cap, int W, int H, GsList* l) {
  char* p = base;
  GsList t;
  size_t tiles = (size_t)((W + GS_TILE - 1) / GS_TILE) * ((H + GS_TILE - 1) / GS_TILE);
  t.max_items = (size_t)cap / GS_SEG + tiles;  // sum over tiles of ceil(len/GS_SEG) <= N/GS_SEG + tiles
  t.max_fitems = (size_t)cap / GS_FWD_CHUNK + tiles;  // sum over tiles of max(1, ceil(len/GS_FWD_CHUNK))
  gs_carve(p, t.point_list, (size_t)cap);
  gs_carve(p, t.items, t.max_items);
  gs_carve(p, t.ckpt, (t.max_items + tiles) * 256);
  gs_carve(p, t.ckpt_op, (t.max_items + tiles) * 256);
  gs_carve(p, t.rowmask, t.max_items * 8 * GS_MASK_WORDS);
  gs_carve(p, t.fitems, t.max_fitems);
  gs_carve(p, t.order, tiles);
  if (l) *l = t;
  return (size_t)(p - base);
}

#define GS_SORT_ITEMS 16                                   // keys per thread, large inputs
#ifndef GS_SORT_ITEMS_SMALL
#define GS_SORT_ITEMS_SMALL 4                              // small inputs: more, shorter CTAs (latency-bound regime)
#endif
#define GS_SORT_SMALL_MAX (1 << 19)                        // inputs up to this size use the small tile
#define GS_SORT_THREADS 256
#define GS_SORT_FUSED_ITEMS 2                              // ... with this many keys per thread while that fits (98 CTAs at 50k: 45.2 us of binning against 46.1 with 4, 47.9 with 8)
#define GS_SORT_FUSED_MAX_BLOCKS 148                       // depth chain in one launch up to this many CTAs (co-resident: one per SM)
#define GS_SORT_TILE (GS_SORT_ITEMS * GS_SORT_THREADS)   // 4096 keys per CTA
#define GS_EMIT_THREADS 256
#define GS_L1_MAX_BLOCKS 592                               // fused L1 loss: 4 CTAs per SM
#define GS_BIN_THREADS 1024                                // direct placement: Gaussians per block
#define GS_BIN_FUSED_MAX_BLOCKS 120                        // count+place in one co-resident grid up to this many blocks
#define GS_BIN_TILE_BITS 11                                // log2(GS_BIN_MAX_TILES)
#define GS_BIN_MAX_TILES 2048                              // ... and the largest tile grid it handles (68 B smem per tile)

static inline __host__ __device__ int gs_sort_items(unsigned long long n) { return n <= GS_SORT_SMALL_MAX ? GS_SORT_ITEMS_SMALL : GS_SORT_ITEMS; }
static inline __host__ size_t gs_sort_blocks(unsigned long long n) {
  const size_t tile = (size_t)gs_sort_items(n) * GS_SORT_THREADS;
  return ((size_t)n + tile - 1) / tile;
}

// Direct placement needs the [blocks][tiles] count matrix to fit the pair-sized scratch it borrows.
static inline __host__ bool gs_direct_binning(int P, int tiles, int64_t cap) {
  const int64_t blocks = ((int64_t)P + GS_BIN_THREADS - 1) / GS_BIN_THREADS;
  return tiles <= GS_BIN_MAX_TILES && blocks * tiles <= cap;
}

static inline __host__ size_t gs_carve_scratch(char* base, int P, int W, int H, int64_t cap, GsScratch* s) {
  char* p = base;
  GsScratch t;
  t.blocks_d = gs_sort_blocks((unsigned long long)P) + 1;
  {  // the one-launch depth sort may run 512-key tiles (GS_SORT_FUSED_ITEMS keys per thread): status rows for those too
    const size_t fused = ((size_t)P + GS_SORT_FUSED_ITEMS * GS_SORT_THREADS - 1) / (GS_SORT_FUSED_ITEMS * GS_SORT_THREADS);
    if (fused <= GS_SORT_FUSED_MAX_BLOCKS && fused + 1 > t.blocks_d) t.blocks_d = fused + 1;
  }
  t.blocks_t = gs_sort_blocks((unsigned long long)cap) + 1;
  t.blocks_e = ((size_t)P + GS_EMIT_THREADS - 1) / GS_EMIT_THREADS + 1;
  for (int i = 0; i < 2; i++) gs_carve(p, t.dkey[i], (size_t)P);
  for (int i = 0; i < 2; i++) gs_carve(p, t.dval[i], (size_t)P);
  for (int i = 0; i < 2; i++) gs_carve(p, t.tkey[i], (size_t)cap);
  for (int i = 0; i < 2; i++) gs_carve(p, t.tval[i], (size_t)cap);
  t.zero_begin = reinterpret_cast<uint32_t*>(p);
  gs_carve(p, t.hist, 8 * 256);
  gs_carve(p, t.counters, 64);
  gs_carve(p, t.status_d, 4 * t.blocks_d * 256);
  gs_carve(p, t.status_t, 4 * t.blocks_t * 256);
  gs_carve(p, t.scan_status, t.blocks_e);
  gs_carve(p, t.tile_total, (size_t)((W + GS_TILE - 1) / GS_TILE) * ((H + GS_TILE - 1) / GS_TILE));
  gs_carve(p, t.hand, ((size_t)cap / GS_FWD_CHUNK + 1) * 8 * GS_HAND_WORDS);
  t.zero_end = reinterpret_cast<uint32_t*>(p);
  if (s) *s = t;
  return (size_t)(p - base);
}

// ---- device math helpers ----
#ifdef __CUDACC__
// a*b + c*d + e*f + g as nvcc contracts it for the reference:  g + fma(e,f, fma(a,b, c*d))
__device__ __forceinline__ float gs_dot3p(float a, float b, float c, float d, float e, float f, float g) {
  return __fadd_rn(g, __fmaf_rn(e, f, __fmaf_rn(a, b, __fmul_rn(c, d))));
}
__device__ __forceinline__ float gs_dot3(float a, float b, float c, float d, float e, float f) {
  return __fmaf_rn(e, f, __fmaf_rn(a, b, __fmul_rn(c, d)));
}

// Deterministic exp for x <= 0 (same specification as oracle/gs_oracle.c:gso_exp, written independently).
__device__ __forceinline__ float gs_exp(float x) {
  const float L2E_HI = 0x1.715476p+0f, L2E_LO = 0x1.4ae0cp-26f;
  float xc = fmaxf(x, -87.0f);
  float t = __fmul_rn(xc, L2E_HI);
  float r = __fadd_rn(t, 12582912.0f);
  float n = __fadd_rn(r, -12582912.0f);
  float f = __fmaf_rn(xc, L2E_HI, -n);
  f = __fmaf_rn(xc, L2E_LO, f);
  float p = 0x1.41fbe4p-13f;
  p = __fmaf_rn(p, f, 0x1.5f3e5ap-10f);
  p = __fmaf_rn(p, f, 0x1.3b2d4cp-7f);
  p = __fmaf_rn(p, f, 0x1.c6aee8p-5f);
  p = __fmaf_rn(p, f, 0x1.ebfbdcp-3f);
  p = __fmaf_rn(p, f, 0x1.62e43p-1f);
  p = __fmaf_rn(p, f, 1.0f);
  // r's low mantissa bits hold n + 0x400000 (two's complement for negative n): shift into the exponent
  float res = __uint_as_float(__float_as_uint(p) + (__float_as_uint(r) << 23));
  return (x >= -87.0f) ? res : 0.0f;
}
#endif

// ---- programmatic dependent launch: every hot-path kernel starts with GS_PDL_PROLOGUE() (wait for the previous grid's
// results, then let the next grid be scheduled) and is launched through gs_launch(), so that inside a stream / CUDA graph
// the launch latency of kernel k+1 overlaps the tail of kernel k.  The wait comes first, so no kernel touches memory
// before its producer has completed. ----
#ifdef __CUDACC__
#define GS_PDL_PROLOGUE()                                          \
  do {                                                             \
    asm volatile("griddepcontrol.wait;" ::: "memory");             \
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); \
  } while (0)
// Which kernels run their predecessor-independent prologue BEFORE griddepcontrol.wait (GS_PREWAIT bit mask, default all):
// 1 depth-sort links 1..3, 2 placement, 4 preprocess backward.
static inline int gs_prewait_mask() {
  static const int m = [] { const char* e = getenv("GS_PREWAIT"); return e ? atoi(e) : 7; }();
  return m;
}
template <typename... KArgs, typename... Args>
static inline cudaError_t gs_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif

// ---- launchers (one per translation unit) ----
int gs_launch_preprocess(const GsForwardArgs& a, const GsGeom& g, const GsScratch& s, cudaStream_t st);
int gs_launch_binning(const GsForwardArgs& a, const GsGeom& g, const GsImage& im, const GsScratch& s, const GsList& l,
                      int64_t cap, cudaStream_t st);
int gs_launch_blend_fwd(const GsForwardArgs& a, const GsGeom& g, const GsImage& im, const GsList& l, const GsScratch& s,
                        cudaStream_t st);
int gs_exp_mode();  // gs_api.cu: 0 = CUDA expf (the reference's instruction sequence, default), 1 = gs_exp polynomial (oracle mode)
int gs_launch_blend_bwd(const GsBackwardArgs& a, const GsGeom& g, const GsImage& im, const GsList& l, cudaStream_t st);
int gs_launch_preprocess_bwd(const GsBackwardArgs& a, const GsGeom& g, cudaStream_t st);  // also finishes the blend's raw moments
int gs_launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st);
int gs_check(cudaError_t e);
void gs_prof_mark(int stage, int end, cudaStream_t st);  // stages: see gs_api.cu (5 = pose fwd, 6 = pose bwd)
