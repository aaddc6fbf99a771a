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
#include <stdlib.h>

#include <cstdlib>

#include "gs_common.cuh"

namespace {

constexpr int kWarps = 8;

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

// Gathers the 48-byte record of one list entry (3 x LDG.128).
struct Rec3 { float4 a, b, c; };
__device__ __forceinline__ Rec3 load_rec(const GsRec* __restrict__ rec, uint32_t id) {
  const float4* rp = reinterpret_cast<const float4*>(rec + id);
  Rec3 r;
  r.a = __ldg(rp); r.b = __ldg(rp + 1); r.c = __ldg(rp + 2);
  return r;
}

constexpr int kChunk = 4;  // survivors evaluated together: their exp() chains are independent (ILP), only the T update is serial

// exp of the blend (forward.cu:344, backward.cu:483).  Default: CUDA's expf, i.e. exactly the instructions nvcc emits for the
// reference's `exp(power)` (same compiler, no fast-math on either side), so every alpha >= 1/255 / T < 1e-4 decision is the
// reference's.  POLY (oracle mode, gs_set_exp_mode(1)): the fixed FMA polynomial the CPU oracle reproduces bit for bit.
template <bool POLY>
__device__ __forceinline__ float blend_exp(float x) { return POLY ? gs_exp(x) : expf(x); }

// ---- forward -------------------------------------------------------------------------------------------------------
// The only serial dependency of front-to-back blending is the per-pixel chain T <- T*(1-alpha) (and the termination test on
// it); culling and the alpha evaluation (the exp) do not depend on it.  Avatar frames have few, deep tile lists (146 of 1024
// tiles non-empty at 512x512 / 50k Gaussians, up to 6000 entries), so a CTA per tile -- or even per (tile, 8x4 pixel block)
// with one consumer warp walking the whole list -- leaves the frame waiting for ONE warp's chain.  Here a work item is
// (tile, pixel block, chunk of GS_FWD_CHUNK list entries):
//   1. all 8 warps of the CTA produce the chunk's alpha rows in parallel (32-entry batches round-robin: gather the 48-byte
//      records, cull against the block with the exact minimum of the quadratic, evaluate the reference's tests for every
//      survivor and pixel) into shared memory, with the bit mask of the entries at least one pixel blends;
//   2. warp 0 then runs the chain over the rows -- exactly the reference's operation sequence per pixel, so final_T,
//      n_contrib and the image are bit-identical to a sequential walk -- starting from the state (T, C, last, done) that the
//      previous chunk's CTA published in a global hand-off slot, and publishes its own for the next chunk.
// Chunks of one list thus overlap everything but the chain itself, and the chain runs over ready rows without any
// producer/consumer hand-over.  Work items are ordered by chunk index (plan_kernel), so a CTA only ever waits for a CTA with a
// smaller block index, which the hardware has already dispatched: no deadlock.  A chunk whose predecessor reports "every
// pixel terminated" skips all work and passes the state on.
constexpr int kChunkBatches = GS_FWD_CHUNK / 32;
constexpr int kFwdWarps = kChunkBatches / 2;  // a warp produces two 32-entry batches of the chunk
#ifndef GS_FWD_CTAS
#define GS_FWD_CTAS (GS_FWD_CHUNK <= 256 ? 5 : 3)  // resident CTAs per SM (shared memory: ~38 KB per CTA at 256 entries, ~75 KB at 512)
#endif
constexpr int kPChunk = 4;  // survivors whose alpha is evaluated together (ILP) by a producer warp
static_assert(GS_FWD_CHUNK % GS_SEG == 0, "checkpoints sit at chunk-aligned positions");

struct FwdSmem {
  float alpha[GS_FWD_CHUNK + 1][32];  // [entry of the chunk][pixel lane]: alpha, 0 where the reference's skip tests fail.  Row
                                      // e first stages the entry's (x, y, conic, opacity) until it is evaluated; the last row is
                                      // all zero (padding target of the unrolled chain)
  float4 row[GS_FWD_CHUNK + 1];       // r, g, b, 1-based list position (int bits)
  uint32_t mask[kChunkBatches];       // per 32-entry batch: entries at least one pixel of the block blends
  uint16_t ridx[GS_FWD_CHUNK + 16];   // the chunk's rows in list order (row slots), segments padded to groups of 4
  uint32_t seg_group[5];              // first group of each 128-entry segment in ridx (at most four segments per chunk)
  uint32_t stop;                      // predecessor (or this chunk) reports: every pixel of the block has terminated
  uint32_t ticket;                    // work unit of the CTA's current round
};

__device__ __forceinline__ float4 ld_relaxed_v4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint2 ld_relaxed_v2(const uint32_t* p) {
  uint2 v;
  asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_v4(float* p, float4 v) {
  asm volatile("st.relaxed.gpu.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_relaxed_v2(uint32_t* p, uint2 v) {
  asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// Chain state of one pixel between two chunks.  Slot layout: 32 x float4 {T, C0, C1, C2}, then 32 x uint2 {last | 1<<30 |
// done<<31, C3}.  Both vectors carry their own validity (T is never 0: it is the last transmittance that was still >= 1e-4;
// the flag word is never 0), each is written with ONE aligned vector store, and the slots are zeroed by the memset at the
// start of every forward -- so a reader that sees both non-zero has the complete state without a fence.
#ifdef GS_FWD_TRACE  // opt-in build (profiles/fwd_trace.py): %globaltimer stamps per work unit
__device__ unsigned long long g_fwd_trace[16384 * 4];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TRACE(i) do { if (threadIdx.x == 0 && work < 16384) g_fwd_trace[work * 4 + (i)] = gtime(); } while (0)
#else
#define TRACE(i) do { } while (0)
#endif
#define GS_HAND_DONE 0x80000000u
#define GS_HAND_VALID 0x40000000u
#define GS_FWD_SPIN_LIMIT (1u << 21)

// OP: the fused opacity pass (gaussian_renderer/__init__.py:131-142: a second rasterisation of the same geometry with
// colours = 1, channel 0 kept).  The chain carries one more accumulator, C3 += T * alpha -- what forward.cu:349
// computes for a feature of exactly 1.0 -- so out_opacity is bit-identical to the second call's channel 0.
template <bool OP, bool POLY>
__device__ __forceinline__ void
blend_fwd_work(FwdSmem& sm, const uint32_t work, const uint32_t n_items,
               const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int W, int H,
               const GsRec* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ final_T,
               uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
               const uint32_t* __restrict__ seg_off, float4* __restrict__ ckpt, const uint4* __restrict__ fitems,
               GsHeader* __restrict__ hdr, int grid_x, float* __restrict__ out_opacity,
               float* __restrict__ ckpt_op, uint32_t* __restrict__ rowmask, uint32_t* __restrict__ hand, int poll_ns) {
  if (work >= n_items * 8) {
    // ---- empty tiles: one work unit each, thread = pixel (forward.cu:363-372 with T = 1, C = 0, no contributor) ----
    const uint32_t e = work - n_items * 8;
    const int tile = (int)fitems[n_items + e].x;
    for (int t = threadIdx.x; t < GS_TILE * GS_TILE; t += kFwdWarps * 32) {
      const int px = (tile % grid_x) * GS_TILE + (t & 15), py = (tile / grid_x) * GS_TILE + (t >> 4);
      if (px < W && py < H) {
        const size_t pix = (size_t)py * W + px, hw = (size_t)H * W;
        final_T[pix] = 1.0f;
        n_contrib[pix] = 0u;
        out_color[pix] = __fmaf_rn(1.0f, bg[0], 0.0f);
        out_color[hw + pix] = __fmaf_rn(1.0f, bg[1], 0.0f);
        out_color[2 * hw + pix] = __fmaf_rn(1.0f, bg[2], 0.0f);
        if (OP) out_opacity[pix] = __fmaf_rn(1.0f, bg[0], 0.0f);
      }
    }
    return;
  }
  TRACE(0);
  const uint32_t fi = work >> 3;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int blk = work & 7;  // pixel block of the tile
  const uint4 item = fitems[fi];
  const int tile = (int)item.x;
  const bool has_pred = item.w != 0xFFFFFFFFu, has_succ = item.z != 0xFFFFFFFFu;
  const int tile_x = tile % grid_x, tile_y = tile / grid_x;
  const uint2 range = ranges[tile];
  const int len = (int)(range.y - range.x);
  const int e0 = (int)item.y * GS_FWD_CHUNK;
  const int e1 = has_succ ? e0 + GS_FWD_CHUNK : len;  // the last item of a tile takes what is left (one round, except past the bin clamp)
  const int bx = tile_x * GS_TILE + (blk & 1) * 8, by = tile_y * GS_TILE + (blk >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pfx = (float)px, pfy = (float)py;
  const float X0 = (float)bx, X1 = (float)(bx + 7), Y0 = (float)by, Y1 = (float)(by + 3);
  const uint32_t* const tl = list + range.x;
  uint32_t* const rm = rowmask + ((size_t)seg_off[tile] * 8 + blk) * GS_MASK_WORDS;  // + segment * 8 * GS_MASK_WORDS + batch in segment * 4 + group
  // (item.z / item.w ARE the slot numbers: plan_body numbers the receiving chunks of the frame densely, tile by tile)
  const float* const hin = reinterpret_cast<const float*>(hand + ((size_t)(has_pred ? item.w : 0u) * 8 + blk) * GS_HAND_WORDS);
  float* const hout = reinterpret_cast<float*>(hand + ((size_t)(has_succ ? item.z : 0u) * 8 + blk) * GS_HAND_WORDS);

  // chain state (warp 0).  Tr is the running product of (1 - alpha) over ALL rows; it never increases (0 < 1 - alpha <= 1; a
  // row the pixel skips has alpha = 0 and multiplies by exactly 1), so "the pixel has terminated" (the reference's done flag,
  // set by the first row with T * (1 - alpha) < 1e-4) is simply Tr < 1e-4, and the reference's T is the last Tr that was still
  // >= 1e-4.  A terminated pixel travels between chunks with Tr = 0.
  float T = 1.0f, Tr = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, C3 = 0.f;
  uint32_t last = 0;
  bool have_state = !has_pred;
  auto decode = [&](const float4& v0, const uint2& v1) {
    T = v0.x; C0 = v0.y; C1 = v0.z; C2 = v0.w;
    last = v1.x & ~(GS_HAND_DONE | GS_HAND_VALID);
    C3 = __uint_as_float(v1.y);
    Tr = (v1.x & GS_HAND_DONE) ? 0.0f : T;
  };
  auto publish = [&]() {
    st_relaxed_v4(hout + lane * 4, make_float4(T, C0, C1, C2));
    st_relaxed_v2(reinterpret_cast<uint32_t*>(hout) + 128 + lane * 2,
                  make_uint2(last | GS_HAND_VALID | (Tr < 0.0001f ? GS_HAND_DONE : 0u), __float_as_uint(C3)));
  };
  if (wid == 0) {
    bool stop = false;
    if (has_pred) {  // a look before any work: a predecessor that has already terminated every pixel makes this chunk a no-op
      const float4 v0 = ld_relaxed_v4(hin + lane * 4);
      const uint2 v1 = ld_relaxed_v2(reinterpret_cast<const uint32_t*>(hin) + 128 + lane * 2);
      if (__all_sync(0xFFFFFFFFu, v0.x != 0.0f && v1.x != 0u)) {
        decode(v0, v1);
        have_state = true;
        stop = __all_sync(0xFFFFFFFFu, !inside || Tr < 0.0001f);
      }
    }
    if (lane == 0) sm.stop = stop ? 1u : 0u;
    sm.alpha[GS_FWD_CHUNK][lane] = 0.0f;
    if (lane == 0) sm.row[GS_FWD_CHUNK] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncthreads();
  if (sm.stop) {
    if (wid == 0 && has_succ) publish();
    return;
  }

  const size_t ck_base = ((size_t)seg_off[tile] + tile) * 256 + blk * 32 + lane;
  float4* const ck = ckpt + ck_base;
  float* const ck3 = OP ? ckpt_op + ck_base : nullptr;
  bool all_done = false;
  int r0 = e0;
  do {
    const int r1 = min(e1, r0 + GS_FWD_CHUNK);
    const int nb = (r1 - r0 + 31) >> 5;
    // ------------------------------------------------ rows: all warps ------------------------------------------------
    // a warp owns batches wid and wid + 8 of the chunk: both gathers are in flight before either is used
    static_assert(kChunkBatches == 2 * kFwdWarps, "two batches per warp");
    Rec3 rc[2];
    bool have[2];
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const int e = r0 + (wid + q * kFwdWarps) * 32 + lane;  // position in the tile list
      have[q] = e < r1;
      rc[q].a = rc[q].b = rc[q].c = make_float4(0.f, 0.f, 0.f, 0.f);
      if (have[q]) rc[q] = load_rec(rec, tl[e]);
    }
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const int b = wid + q * kFwdWarps;
      if (b >= nb) break;  // warp-uniform
      const int rowbase = b * 32;
      const Rec3 cur = rc[q];
      const bool keep = have[q] && may_contribute(cur.a.x, cur.a.y, cur.b.x, cur.b.y, cur.b.z, cur.a.z, X0, X1, Y0, Y1);
      if (keep) {
        float* const stage = sm.alpha[rowbase + lane];
        *reinterpret_cast<float4*>(stage) = make_float4(cur.a.x, cur.a.y, cur.b.x, cur.b.y);
        *reinterpret_cast<float2*>(stage + 4) = make_float2(cur.b.z, cur.b.w);
        sm.row[rowbase + lane] = make_float4(cur.c.x, cur.c.y, cur.c.z, __int_as_float(r0 + rowbase + lane + 1));
      }
      __syncwarp();
      uint32_t m = __ballot_sync(0xFFFFFFFFu, keep), kept0 = 0, kept1 = 0, kept2 = 0, kept3 = 0;
      while (m) {
        float al[kPChunk];
        bool ok[kPChunk];
        int jj[kPChunk];
#pragma unroll
        for (int u = 0; u < kPChunk; u++) {
          const bool valid = m != 0;
          const int j = valid ? (__ffs(m) - 1) : 0;
          m &= m - 1;
          jj[u] = valid ? j : -1;
          const float* const stage = sm.alpha[rowbase + j];
          const float4 q0 = *reinterpret_cast<const float4*>(stage);
          const float2 q1 = *reinterpret_cast<const float2*>(stage + 4);
          const float dx = __fsub_rn(q0.x, pfx), dy = __fsub_rn(q0.y, pfy);
          const float power = __fmaf_rn(__fmaf_rn(dx, __fmul_rn(dx, q0.z), __fmul_rn(dy, __fmul_rn(dy, q1.x))), -0.5f,
                                        -__fmul_rn(dy, __fmul_rn(dx, q0.w)));  // SASS: FFMA q*-0.5 - cross
          al[u] = fminf(0.99f, __fmul_rn(q1.y, blend_exp<POLY>(power)));
          ok[u] = valid && !(power > 0.0f) && !(al[u] < 1.0f / 255.0f);
        }
        __syncwarp();  // every lane has read the staged records of this group before their rows are overwritten
#pragma unroll
        for (int u = 0; u < kPChunk; u++) {
          if (jj[u] >= 0) {  // warp-uniform
            sm.alpha[rowbase + jj[u]][lane] = ok[u] ? al[u] : 0.0f;
            const uint32_t who = __ballot_sync(0xFFFFFFFFu, ok[u]);  // rows no pixel of the block blends never reach the chain
            // per 4x2 pixel group (lane = x + 8 y: x < 4 | x >= 4, y < 2 | y >= 2): what the backward visits
            if (who & 0x00000F0Fu) kept0 |= 1u << jj[u];
            if (who & 0x0000F0F0u) kept1 |= 1u << jj[u];
            if (who & 0x0F0F0000u) kept2 |= 1u << jj[u];
            if (who & 0xF0F00000u) kept3 |= 1u << jj[u];
          }
        }
        __syncwarp();
      }
      if (lane == 0) {
        sm.mask[b] = kept0 | kept1 | kept2 | kept3;
        const int bg_ = (r0 >> 5) + b;  // batch index inside the tile list
        *reinterpret_cast<uint4*>(rm + (size_t)(bg_ >> 2) * (8 * GS_MASK_WORDS) + (bg_ & 3) * 4) = make_uint4(kept0, kept1, kept2, kept3);
      }
    }
    __syncthreads();
    // ---- flat list of the chunk's rows in list order (all warps): ridx[k] = row slot; every 128-entry segment is padded to a
    // multiple of 4 rows with the all-zero row, seg_group[s] = first group of 4 of segment s ----
    {
      const uint32_t mk = lane < nb ? sm.mask[lane] : 0u;  // lanes 0..15: the batches' masks
      const int cnt = __popc(mk);
      int inc = cnt;
      int t = __shfl_up_sync(0xFFFFFFFFu, inc, 1);
      if ((lane & 3) >= 1) inc += t;
      t = __shfl_up_sync(0xFFFFFFFFu, inc, 2);
      if ((lane & 3) >= 2) inc += t;
      const int segtot = __shfl_sync(0xFFFFFFFFu, inc, lane | 3);
      const int segpad = (segtot + 3) & ~3;
      const int p0 = __shfl_sync(0xFFFFFFFFu, segpad, 0), p1 = __shfl_sync(0xFFFFFFFFu, segpad, 4);
      const int p2 = __shfl_sync(0xFFFFFFFFu, segpad, 8), p3 = __shfl_sync(0xFFFFFFFFu, segpad, 12);
      const int sg = (lane >> 2) & 3;
      const int segbase = sg == 0 ? 0 : sg == 1 ? p0 : sg == 2 ? p0 + p1 : p0 + p1 + p2;
      const int off = segbase + inc - cnt;
#pragma unroll
      for (int q = 0; q < 2; q++) {
        const int b = wid + q * kFwdWarps;
        const uint32_t m = __shfl_sync(0xFFFFFFFFu, mk, b & 15);
        const int o = __shfl_sync(0xFFFFFFFFu, off, b & 15);
        if (b < nb && ((m >> lane) & 1u)) sm.ridx[o + __popc(m & ((1u << lane) - 1u))] = (uint16_t)(b * 32 + lane);
      }
      if (wid < 4) {  // padding of segment wid
        const int tot = __shfl_sync(0xFFFFFFFFu, segtot, wid * 4), base = __shfl_sync(0xFFFFFFFFu, segbase, wid * 4);
        if (lane < ((tot + 3) & ~3) - tot) sm.ridx[base + tot + lane] = (uint16_t)GS_FWD_CHUNK;
        if (lane == 0) sm.seg_group[wid] = (uint32_t)(base >> 2);
        if (wid == 3 && lane == 0) sm.seg_group[4] = (uint32_t)((p0 + p1 + p2 + p3) >> 2);
      }
    }
    __syncthreads();
    TRACE(1);
    // ------------------------------------------------ chain: warp 0 ------------------------------------------------
    if (wid == 0) {
      if (!have_state) {  // wait for the previous chunk's state
        float4 v0;
        uint2 v1;
        uint32_t spins = 0;
        while (true) {
          v0 = ld_relaxed_v4(hin + lane * 4);
          v1 = ld_relaxed_v2(reinterpret_cast<const uint32_t*>(hin) + 128 + lane * 2);
          if (__all_sync(0xFFFFFFFFu, v0.x != 0.0f && v1.x != 0u)) break;
          if (++spins > GS_FWD_SPIN_LIMIT) {  // never in a healthy run: report instead of hanging the GPU
            if (lane == 0) atomicOr(&hdr->error, 4u);
            v0 = make_float4(1.f, 0.f, 0.f, 0.f);
            v1 = make_uint2(GS_HAND_VALID, 0u);
            break;
          }
          if (poll_ns > 0) __nanosleep(poll_ns);
        }
        decode(v0, v1);
        have_state = true;
      }
      TRACE(2);
      all_done = __all_sync(0xFFFFFFFFu, !inside || Tr < 0.0001f);
      const int nsegs = (nb + GS_SEG / 32 - 1) / (GS_SEG / 32);
      auto load_group = [&](int g, float (&al)[kChunk], float4 (&rw)[kChunk]) {
        const uint2 ri = *reinterpret_cast<const uint2*>(&sm.ridx[4 * g]);  // four 16-bit row slots
        const uint32_t rr[kChunk] = {ri.x & 0xFFFFu, ri.x >> 16, ri.y & 0xFFFFu, ri.y >> 16};
#pragma unroll
        for (int u = 0; u < kChunk; u++) {
          al[u] = sm.alpha[rr[u]][lane];
          rw[u] = sm.row[rr[u]];
        }
      };
      static_assert(kChunk == 4, "the flat row list is padded to groups of 4");
      for (int sgi = 0; sgi < nsegs && !all_done; sgi++) {
        const int sglob = r0 / GS_SEG + sgi;  // checkpoint at every segment start: the backward's work items restart here
        ck[(size_t)sglob * 256] = make_float4(T, C0, C1, C2);
        if (OP) ck3[(size_t)sglob * 256] = C3;
        int g = (int)sm.seg_group[sgi];
        const int gend = (int)sm.seg_group[sgi + 1];
        if (g < gend) {
          float al[kChunk];
          float4 rw[kChunk];
          load_group(g, al, rw);
          for (; g < gend; g++) {
            float aln[kChunk];
            float4 rwn[kChunk];
            load_group(min(g + 1, gend - 1), aln, rwn);  // next group's rows are in flight while this one is applied
#pragma unroll
            for (int u = 0; u < kChunk; u++) {  // forward.cu:335-361 without carried control state, same values in the same order
              const float test_T = __fmul_rn(Tr, __fsub_rn(1.0f, al[u]));
              const bool okc = !(test_T < 0.0001f);
              const float Tw = okc ? Tr : 0.0f;  // branch-free: a terminated pixel adds 0 * (alpha * c) = 0, C keeps its bits
              C0 = __fmaf_rn(Tw, __fmul_rn(al[u], rw[u].x), C0);
              C1 = __fmaf_rn(Tw, __fmul_rn(al[u], rw[u].y), C1);
              C2 = __fmaf_rn(Tw, __fmul_rn(al[u], rw[u].z), C2);
              if (OP) C3 = __fmaf_rn(Tw, al[u], C3);  // colour 1.0: (1 * alpha) * T
              T = okc ? test_T : T;
              last = (okc && al[u] != 0.0f) ? (uint32_t)__float_as_int(rw[u].w) : last;  // contributor, forward.cu:328 (1-based)
              Tr = test_T;
            }
#pragma unroll
            for (int u = 0; u < kChunk; u++) { al[u] = aln[u]; rw[u] = rwn[u]; }
          }
        }
        all_done = __all_sync(0xFFFFFFFFu, !inside || Tr < 0.0001f);
      }
      if (lane == 0) sm.stop = all_done ? 1u : 0u;
      TRACE(3);
    }
    __syncthreads();  // the rows are overwritten by the next round (only past the bin clamp of plan_kernel)
    all_done = sm.stop != 0;
    r0 = r1;
  } while (r0 < e1 && !all_done);

  if (wid != 0) return;
  if (has_succ) publish();  // also when every pixel has terminated: the successors pass it on and leave
  if (has_succ && !all_done) return;
  // final state of the block: the last chunk, or the chunk in which the last pixel terminated
  ck[(size_t)((len + GS_SEG - 1) / GS_SEG) * 256] = make_float4(T, C0, C1, C2);
  if (OP) ck3[(size_t)((len + GS_SEG - 1) / GS_SEG) * 256] = C3;
  if (inside) {
    const size_t pix = (size_t)py * W + px, hw = (size_t)H * W;
    final_T[pix] = T;
    n_contrib[pix] = last;
    out_color[pix] = __fmaf_rn(T, bg[0], C0);
    out_color[hw + pix] = __fmaf_rn(T, bg[1], C1);
    out_color[2 * hw + pix] = __fmaf_rn(T, bg[2], C2);
    if (OP) out_opacity[pix] = __fmaf_rn(T, bg[0], C3);  // channel 0 of the second call: same background
  }
}

// Persistent grid: 3 CTAs per SM take work units -- (item, pixel block) pairs in plan order, then the empty tiles -- from a
// ticket counter in the frame header (zeroed by the memset that starts every forward).  Tickets are handed out in increasing
// order, so whoever waits for a predecessor's hand-off waits for a unit that a running CTA has already taken.
template <bool OP, bool POLY>
__global__ void __launch_bounds__(kFwdWarps * 32, GS_FWD_CTAS)
blend_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int W, int H,
                 const GsRec* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ final_T,
                 uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                 const uint32_t* __restrict__ seg_off, float4* __restrict__ ckpt, const uint4* __restrict__ fitems,
                 GsHeader* __restrict__ hdr, int grid_x, float* __restrict__ out_opacity,
                 float* __restrict__ ckpt_op, uint32_t* __restrict__ rowmask, uint32_t* __restrict__ hand, int poll_ns) {
  GS_PDL_PROLOGUE();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FwdSmem& sm = *reinterpret_cast<FwdSmem*>(smem_raw);
  const uint32_t n_items = hdr->n_fitems;
  const uint32_t n_work = n_items * 8 + hdr->n_empty_tiles;
  while (true) {
    __syncthreads();  // the previous unit is finished by every warp (its rows are no longer read)
    if (threadIdx.x == 0) sm.ticket = atomicAdd(&hdr->fwd_ticket, 1u);
    __syncthreads();
    const uint32_t work = sm.ticket;
    if (work >= n_work) break;
    blend_fwd_work<OP, POLY>(sm, work, n_items, ranges, list, W, H, rec, bg, final_T, n_contrib, out_color, seg_off, ckpt, fitems,
                             hdr, grid_x, out_opacity, ckpt_op, rowmask, hand, poll_ns);
  }
}

// ---- forward, per-block variant: one CTA per (tile, 8x4 pixel block) walks the WHOLE list ------------------------------
// Seven producer warps take the list's 32-entry batches round-robin (records software-pipelined one batch ahead, list indices
// two ahead), cull, evaluate and leave rows alpha[row][pixel] in their own shared-memory buffer; ONE consumer warp applies the
// rows in list order (the same chain as above); full / empty mbarriers per producer buffer.  Producers and consumer overlap
// inside the CTA, so all eight warps stay busy -- the right shape when a frame has thousands of SHALLOW tiles (1M Gaussians at
// 1080p: 8160 tiles of one or two chunks): nothing is gained from cutting chains there, and the chunked kernel's rows-then-chain
// units leave three of four warps idle during the chain (757 us against 531 us for this kernel).  With few deep tiles (avatar
// frames at 512^2) the consumer of the deepest block is the critical path and the chunked kernel wins.  The launcher picks by
// the size of the tile grid.  Same operations in the same order per pixel: both kernels produce identical bits.
constexpr int kBkProducers = 7;
struct BkBuf {             // one per producer warp
  float4 q0[32], q1[32];   // staged records: x, y, conic a, b | conic c, opacity, r, g
  float q2[32];            // b
  float alpha[32][32];     // [row][pixel lane]; 0 = the reference's skip tests failed
  float4 row[32];          // [row] r, g, b, 1-based list position (int bits)
  uint32_t cnt;            // rows in this batch
  uint32_t pad[3];
};
struct BkCtl {
  unsigned long long full[kBkProducers];   // producer -> consumer: rows of the buffer are ready
  unsigned long long empty[kBkProducers];  // consumer -> producer: buffer may be overwritten
  uint32_t done;                           // every pixel of the block has terminated
  uint32_t pad[3];
};
__device__ __forceinline__ uint32_t bk_ld_acquire(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"((uint32_t)__cvta_generic_to_shared(p)) : "memory");
  return v;
}
__device__ __forceinline__ void bk_st_release(uint32_t* p, uint32_t v) {
  asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void bk_mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void bk_mbar_arrive(unsigned long long* bar) {  // release: prior shared-memory writes are visible
  asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void bk_mbar_arrive_relaxed(unsigned long long* bar) {  // WAR hand-back: the reads were consumed
  asm volatile("mbarrier.arrive.relaxed.cta.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void bk_mbar_wait(unsigned long long* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "BKW_%=:\n"
      "mbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%0], %1;\n"
      "@p bra BKD_%=;\n"
      "bra BKW_%=;\n"
      "BKD_%=:\n"
      "}\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ bool bk_mbar_try_wait(unsigned long long* bar, uint32_t parity) {  // bounded suspend, then returns
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%1], %2, %3;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n" : "=r"(ok) : "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(parity), "r"(1000u) : "memory");
  return ok != 0;
}

template <bool OP, bool POLY>
__global__ void __launch_bounds__((kBkProducers + 1) * 32)
blend_fwd_block_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int W, int H, const GsRec* __restrict__ rec,
                       const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                       float* __restrict__ out_color, const uint32_t* __restrict__ seg_off, float4* __restrict__ ckpt,
                       const uint32_t* __restrict__ order, int grid_x, float* __restrict__ out_opacity, float* __restrict__ ckpt_op,
                       uint32_t* __restrict__ rowmask) {
  GS_PDL_PROLOGUE();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BkCtl& ctl = *reinterpret_cast<BkCtl*>(smem_raw);
  BkBuf* const bufs = reinterpret_cast<BkBuf*>(smem_raw + sizeof(BkCtl));
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;  // warp 0 = consumer, 1..kBkProducers = producers
  const int blk = blockIdx.x & 7;                              // pixel block of the tile
  const int tile = (int)order[blockIdx.x >> 3];                // longest lists first
  const int tile_x = tile % grid_x, tile_y = tile / grid_x;
  const uint2 range = ranges[tile];
  const int len = (int)(range.y - range.x);
  const int nb = (len + 31) >> 5;
  const int bx = tile_x * GS_TILE + (blk & 1) * 8, by = tile_y * GS_TILE + (blk >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < W && py < H;
  if (threadIdx.x == 0) {
    ctl.done = 0;
    for (int i = 0; i < kBkProducers; i++) { bk_mbar_init(&ctl.full[i], 1); bk_mbar_init(&ctl.empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (wid > 0) {
    // ------------------------------------------------ producer ------------------------------------------------
    const int p = wid - 1;
    BkBuf& ws = bufs[p];
    const float pfx = (float)px, pfy = (float)py;
    const float X0 = (float)bx, X1 = (float)(bx + 7), Y0 = (float)by, Y1 = (float)(by + 3);
    const uint32_t* const tl = list + range.x;
    uint32_t* const rm = rowmask + ((size_t)seg_off[tile] * 8 + blk) * GS_MASK_WORDS;  // + segment * 8 * GS_MASK_WORDS + batch * 4 + group
    Rec3 cur;
    cur.a = cur.b = cur.c = make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t id_next = 0;
    {
      const int e = p * 32 + lane;
      if (e < len) cur = load_rec(rec, tl[e]);
      if (e + kBkProducers * 32 < len) id_next = tl[e + kBkProducers * 32];
    }
    uint32_t it = 0;
    for (int b = p; b < nb; b += kBkProducers, it++) {
      // software pipeline: records of this warp's next batch, list indices of the one after
      const int e = b * 32 + lane;
      Rec3 nxt = cur;
      if (e + kBkProducers * 32 < len) nxt = load_rec(rec, id_next);
      uint32_t id_nn = 0;
      if (e + 2 * kBkProducers * 32 < len) id_nn = tl[e + 2 * kBkProducers * 32];
      const bool keep = (e < len) && may_contribute(cur.a.x, cur.a.y, cur.b.x, cur.b.y, cur.b.z, cur.a.z, X0, X1, Y0, Y1);
      // the consumer has finished with the previous rows -- or every pixel has terminated: the flag is polled inside the wait
      // because the consumer's wake-up arrivals at termination can put the barrier two phases ahead (parity alias)
      if (it > 0) {
        while (!bk_mbar_try_wait(&ctl.empty[p], (it - 1) & 1u)) {
          if (bk_ld_acquire(&ctl.done)) break;
        }
      }
      __syncwarp();
      {
        uint32_t t = 0;
        if (lane == 0) t = bk_ld_acquire(&ctl.done);
        if (__shfl_sync(0xFFFFFFFFu, t, 0)) break;
      }
      if (e < len) {
        ws.q0[lane] = make_float4(cur.a.x, cur.a.y, cur.b.x, cur.b.y);
        ws.q1[lane] = make_float4(cur.b.z, cur.b.w, cur.c.x, cur.c.y);
        ws.q2[lane] = cur.c.z;
      }
      __syncwarp();
      uint32_t m = __ballot_sync(0xFFFFFFFFu, keep), kept0 = 0, kept1 = 0, kept2 = 0, kept3 = 0;
      int k = 0;
      while (m) {
#pragma unroll
        for (int u = 0; u < kPChunk; u++) {
          const bool valid = m != 0;
          const int j = valid ? (__ffs(m) - 1) : 0;
          m &= m - 1;
          const float4 q0 = ws.q0[j], q1 = ws.q1[j];
          const float dx = __fsub_rn(q0.x, pfx), dy = __fsub_rn(q0.y, pfy);
          const float power = __fmaf_rn(__fmaf_rn(dx, __fmul_rn(dx, q0.z), __fmul_rn(dy, __fmul_rn(dy, q1.x))), -0.5f,
                                        -__fmul_rn(dy, __fmul_rn(dx, q0.w)));  // SASS: FFMA q*-0.5 - cross
          const float alpha = fminf(0.99f, __fmul_rn(q1.y, blend_exp<POLY>(power)));
          const bool ok = valid && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
          const uint32_t who = __ballot_sync(0xFFFFFFFFu, ok);
          if (who) {  // rows no pixel of the block blends are dropped here, off the serial chain
            ws.alpha[k][lane] = ok ? alpha : 0.0f;
            if (lane == 0) ws.row[k] = make_float4(q1.z, q1.w, ws.q2[j], __int_as_float(b * 32 + j + 1));
            // per 4x2 pixel group (lane = x + 8 y: x < 4 | x >= 4, y < 2 | y >= 2): what the backward visits
            if (who & 0x00000F0Fu) kept0 |= 1u << j;
            if (who & 0x0000F0F0u) kept1 |= 1u << j;
            if (who & 0x0F0F0000u) kept2 |= 1u << j;
            if (who & 0xF0F00000u) kept3 |= 1u << j;
            k++;
          }
        }
      }
      for (int kk = k; kk & (kChunk - 1); kk++) {  // the consumer reads whole groups of rows, branch-free: pad rows must be 0 * finite
        ws.alpha[kk][lane] = 0.0f;
        if (lane == 0) ws.row[kk] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      if (lane == 0) {
        ws.cnt = (uint32_t)k;
        *reinterpret_cast<uint4*>(rm + (size_t)(b >> 2) * (8 * GS_MASK_WORDS) + (b & 3) * 4) = make_uint4(kept0, kept1, kept2, kept3);
      }
      __syncwarp();
      if (lane == 0) bk_mbar_arrive(&ctl.full[p]);
      cur = nxt;
      id_next = id_nn;
    }
    return;
  }

  // -------------------------------------------------- consumer --------------------------------------------------
  // checkpoint slots of this tile: one per segment start + one final (the backward's work items restart from them)
  const size_t ck_base = ((size_t)seg_off[tile] + tile) * 256 + blk * 32 + lane;
  float4* const ck = ckpt + ck_base;
  float* const ck3 = OP ? ckpt_op + ck_base : nullptr;
  float T = 1.0f, Tr = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, C3 = 0.f;  // see blend_fwd_work for the chain's state
  uint32_t last = 0;
  for (int b = 0; b < nb; b++) {
    const int p = b % kBkProducers;
    BkBuf& ws = bufs[p];
    bk_mbar_wait(&ctl.full[p], (uint32_t)(b / kBkProducers) & 1u);
    if ((b & (GS_SEG / 32 - 1)) == 0) {
      ck[(size_t)(b / (GS_SEG / 32)) * 256] = make_float4(T, C0, C1, C2);
      if (OP) ck3[(size_t)(b / (GS_SEG / 32)) * 256] = C3;
    }
    const int cnt = (int)ws.cnt;  // rows up to the next multiple of kChunk are zero-padded by the producer
    for (int k0 = 0; k0 < cnt; k0 += kChunk) {
      float al[kChunk];
      float4 rw[kChunk];
#pragma unroll
      for (int u = 0; u < kChunk; u++) {
        al[u] = ws.alpha[k0 + u][lane];
        rw[u] = ws.row[k0 + u];
      }
#pragma unroll
      for (int u = 0; u < kChunk; u++) {  // forward.cu:335-361 without carried control state, same values in the same order
        const float test_T = __fmul_rn(Tr, __fsub_rn(1.0f, al[u]));
        const bool ok = !(test_T < 0.0001f);
        const float Tw = ok ? Tr : 0.0f;
        C0 = __fmaf_rn(Tw, __fmul_rn(al[u], rw[u].x), C0);
        C1 = __fmaf_rn(Tw, __fmul_rn(al[u], rw[u].y), C1);
        C2 = __fmaf_rn(Tw, __fmul_rn(al[u], rw[u].z), C2);
        if (OP) C3 = __fmaf_rn(Tw, al[u], C3);
        T = ok ? test_T : T;
        last = (ok && al[u] != 0.0f) ? (uint32_t)__float_as_int(rw[u].w) : last;
        Tr = test_T;
      }
    }
    const bool all_done = __all_sync(0xFFFFFFFFu, !inside || Tr < 0.0001f);
    if (lane == 0) {
      if (all_done) {  // stop the producers: whoever waits on an empty barrier wakes and sees the flag
        bk_st_release(&ctl.done, 1u);
        for (int i = 0; i < kBkProducers; i++) bk_mbar_arrive(&ctl.empty[i]);
      } else {
        bk_mbar_arrive_relaxed(&ctl.empty[p]);
      }
    }
    if (all_done) break;
  }
  ck[(size_t)((len + GS_SEG - 1) / GS_SEG) * 256] = make_float4(T, C0, C1, C2);
  if (OP) ck3[(size_t)((len + GS_SEG - 1) / GS_SEG) * 256] = C3;
  if (inside) {
    const size_t pix = (size_t)py * W + px, hw = (size_t)H * W;
    final_T[pix] = T;
    n_contrib[pix] = last;
    out_color[pix] = __fmaf_rn(T, bg[0], C0);
    out_color[hw + pix] = __fmaf_rn(T, bg[1], C1);
    out_color[2 * hw + pix] = __fmaf_rn(T, bg[2], C2);
    if (OP) out_opacity[pix] = __fmaf_rn(T, bg[0], C3);  // channel 0 of the second call: same background
  }
}

// Transposing butterfly step: N values per lane -> N/2, exchanging with lane ^ XOR.  After the steps for XOR = 16, 8, 4, 2, 1
// on 32 values, lane L holds the warp total of value L: 31 shuffles for 32 sums instead of 160.
template <int N, int XOR>
__device__ __forceinline__ void xpose_step(float* v, int lane) {
  const bool up = lane & XOR;
#pragma unroll
  for (int k = 0; k < N / 2; k++) {
    const float send = up ? v[k] : v[k + N / 2];
    const float keepv = up ? v[k + N / 2] : v[k];
    v[k] = keepv + __shfl_xor_sync(0xFFFFFFFFu, send, XOR);
  }
}

// One CTA per work item = (tile, segment of GS_SEG list entries), 8 warps = the tile's eight 8x4 pixel blocks.  A pixel
// whose last contributor lies behind the segment restarts from the forward's checkpoint at the segment end:
// T = T(seg end), accumulated colour behind it = (C_final - C(seg end)) / T(seg end); inside the segment the recurrences
// are the reference's (backward.cu:497-534).  Deep tile lists thus spread over many SMs instead of one.
//
// Inside a warp the four 4x2 pixel GROUPS (eight consecutive lanes each) walk their OWN entry sequences -- the forward left one
// mask per group and 32-entry batch: the entries at least one pixel of the group blended -- so one warp instruction evaluates
// four different entries, and the reductions run over 8 lanes instead of 32 (avatar splats cover a handful of pixels of an 8x4
// block: with whole-block visits most lanes were dead, with 8x2 strips still four of five).  Per (pixel, entry) only RAW MOMENTS
// of g = G * dL/dalpha are formed -- sum g, g dx, g dy, g dx^2, g dx dy, g dy^2 and the three colour terms -- reduced with a
// transposing butterfly (2 entries x 8 terms over 8 lanes -> 14 shuffles, a lane ends with two adjacent terms of one entry)
// BEFORE one 8-byte vector reduction per lane; preprocess_bwd_kernel turns the moments into dL/dmean2D, dL/dconic (it reads
// the conic and the opacity of the Gaussian anyway).
#ifndef GS_BWD_MINBLOCKS
#define GS_BWD_MINBLOCKS 3
#endif
struct BwdSlab {  // one warp's segment: GS_SEG staged entries (only those some strip visits)
  float4 q0[GS_SEG];  // x, y, conic a, conic b
  float4 q1[GS_SEG];  // conic c, opacity, r, g
  float q2[GS_SEG];   // b
  uint32_t id[GS_SEG];
};
static_assert(GS_SEG == 128, "the backward blend unrolls four 32-entry batches per segment");

// OP: upstream gradient of the fused opacity image too.  dL/dalpha is linear in the per-channel terms, so the opacity
// channel (colour 1, no colour gradient) is one more term of the same sum and every geometry gradient comes out as the
// sum of the reference's two backward calls.
template <bool OP, bool POLY>
__device__ __forceinline__ void
blend_bwd_item(const uint32_t item_index, BwdSlab& sl, const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int W, int H,
               int grid_x, const uint32_t* __restrict__ seg_off, const uint2* __restrict__ items, const float4* __restrict__ ckpt,
               const float* __restrict__ bg, const GsRec* __restrict__ rec, const float* __restrict__ final_Ts,
               const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels, float* __restrict__ acc /*[P,GS_ACC]*/,
               const float* __restrict__ ckpt_op, const float* __restrict__ dL_dop_pix, const uint32_t* __restrict__ rowmask,
               GsHeader* __restrict__ hdr) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint2 item = items[item_index];
  const int tile = (int)item.x;
  const uint2 range = ranges[tile];
  const int len = (int)(range.y - range.x);
  const int seg_begin = (int)item.y * GS_SEG, seg_end = min(len, seg_begin + GS_SEG);
  const int tx = tile % grid_x, ty = tile / grid_x;
  const int bx = tx * GS_TILE + (w & 1) * 8, by = ty * GS_TILE + (w >> 1) * 4;
  // lane -> pixel of the warp's 8x4 block: the four 4x2 pixel GROUPS are eight consecutive lanes each (the reductions below run
  // inside a group); `flane` is the same pixel's lane in the forward's x + 8 y numbering (checkpoints are stored by it)
  const int grp = lane >> 3, gi = lane & 7;
  const int xl = (grp & 1) * 4 + (gi & 3), yl = (grp >> 1) * 2 + (gi >> 2), flane = xl + 8 * yl;
  const int px = bx + xl, py = by + yl;
  const bool inside = px < W && py < H;
  const float pfx = (float)px, pfy = (float)py;
  const size_t pix = (size_t)py * W + px, hw = (size_t)H * W;
  const int last_contributor = inside ? (int)n_contrib[pix] : 0;
  const int hi = min(last_contributor, seg_end);  // this pixel owns list positions [seg_begin, hi)
  // entries at positions >= max(hi) over the warp are never touched
  int wmax = hi;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xFFFFFFFFu, wmax, o));
  if (wmax <= seg_begin) return;
  // per 4x2 pixel group and batch: the entries at least one pixel of the group blended in the forward pass (written by the
  // forward's producers for every batch of every chunk it ran, i.e. every batch below wmax): the only ones worth visiting
  const uint4* const mp = reinterpret_cast<const uint4*>(rowmask + (((size_t)seg_off[tile] + item.y) * 8 + w) * GS_MASK_WORDS);
  wmax -= seg_begin;  // positions below are relative to the segment start (a multiple of 32)
  const int hi_rel = hi - seg_begin;
  uint32_t mall[4], mine[4];
#pragma unroll
  for (int b = 0; b < 4; b++) {  // batch b: {group 0, 1, 2, 3}
    const uint4 mb = mp[b];
    mall[b] = mb.x | mb.y | mb.z | mb.w;
    mine[b] = grp == 0 ? mb.x : grp == 1 ? mb.y : grp == 2 ? mb.z : mb.w;
  }
#pragma unroll
  for (int b = 0; b < 4; b++) {  // nothing at or behind wmax
    const int n = wmax - b * 32;
    const uint32_t live = n >= 32 ? 0xFFFFFFFFu : n <= 0 ? 0u : ((1u << n) - 1u);
    mall[b] &= live;
    mine[b] &= live;
  }
  if ((mall[0] | mall[1] | mall[2] | mall[3]) == 0u) return;
  // ---- stage the visited entries of the whole segment: every gather of the warp is in flight at once ----
  const uint32_t* const seg_list = list + range.x + seg_begin;
#pragma unroll
  for (int b = 0; b < 4; b++) {
    if ((mall[b] >> lane) & 1u) {
      const uint32_t id = seg_list[b * 32 + lane];
      const Rec3 r = load_rec(rec, id);
      sl.q0[b * 32 + lane] = make_float4(r.a.x, r.a.y, r.b.x, r.b.y);
      sl.q1[b * 32 + lane] = make_float4(r.b.z, r.b.w, r.c.x, r.c.y);
      sl.q2[b * 32 + lane] = r.c.z;
      sl.id[b * 32 + lane] = id;
    }
  }
  const float T_final = inside ? final_Ts[pix] : 0.f;
  float T = T_final;
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, last_alpha = 0.f;
  float acc3 = 0.f, lc3 = 0.f;  // opacity channel: its colour is 1 once a contributor has been seen
  if (last_contributor > seg_end) {  // restart from the checkpoint at the segment end
    const size_t ck_base = ((size_t)seg_off[tile] + tile) * 256 + w * 32 + flane;
    const float4* ck = ckpt + ck_base;
    const float4 c = ck[(size_t)(item.y + 1) * 256];
    const float4 cf = ck[(size_t)((len + GS_SEG - 1) / GS_SEG) * 256];
    T = c.x;
    const float rT = 1.f / T;
    acc0 = (cf.y - c.y) * rT; acc1 = (cf.z - c.z) * rT; acc2 = (cf.w - c.w) * rT;
    if (OP) {
      const float* ck3 = ckpt_op + ck_base;
      acc3 = (ck3[(size_t)((len + GS_SEG - 1) / GS_SEG) * 256] - ck3[(size_t)(item.y + 1) * 256]) * rT;
    }
  }
  float dp0 = 0.f, dp1 = 0.f, dp2 = 0.f, dp3 = 0.f;
  if (inside) { dp0 = dL_dpixels[pix]; dp1 = dL_dpixels[hw + pix]; dp2 = dL_dpixels[2 * hw + pix]; }
  if (OP && inside) dp3 = dL_dop_pix[pix];
  // the second call's image has the same background; only its channel 0 carries an upstream gradient (`[:1]`)
  const float bg_dot_dpixel = bg[0] * dp0 + bg[1] * dp1 + bg[2] * dp2 + (OP ? bg[0] * dp3 : 0.f);
  const float nTfbg = -T_final * bg_dot_dpixel;
  // terms of the Gaussian's 48-byte accumulator record: 0,1 -> sum g dx, g dy; 2,3,4 -> sum g dx^2, g dx dy, g dy^2; 5 -> sum g
  // (= dL_dopacity); 6,7 (+ the separate ninth term at 8) -> colour gradient
  uint32_t nblend = 0;  // blended (pixel, entry) pairs = E_bwd of SURVEY.md 8(d), summed into the frame header
  __syncwarp();

#pragma unroll
  for (int b = 3; b >= 0; b--) {  // back to front
    uint32_t m = mine[b];
    // a strip that has run out of entries (or holds an odd count) evaluates a STAGED entry of the batch with ok = false: an
    // unstaged slot is whatever the previous kernel left in shared memory, and 0 * (NaN or inf) would reach the atomics
    const int j_idle = mall[b] ? (__ffs(mall[b]) - 1) : 0;
    while (__any_sync(0xFFFFFFFFu, m != 0u)) {
      float v[16], v8[2];
      uint32_t ids[2];
      bool any_ok = false;
#pragma unroll
      for (int k = 0; k < 2; k++) {
        const bool valid = m != 0u;
        const int j = valid ? (31 - __clz(m)) : j_idle;
        m &= ~(valid ? (1u << j) : 0u);
        const int e = b * 32 + j;
        ids[k] = sl.id[e];
        const float4 q0 = sl.q0[e], q1 = sl.q1[e];
        const float c2 = sl.q2[e];
        const float dx = __fsub_rn(q0.x, pfx), dy = __fsub_rn(q0.y, pfy);
        const float power = __fmaf_rn(__fmaf_rn(dx, __fmul_rn(dx, q0.z), __fmul_rn(dy, __fmul_rn(dy, q1.x))), -0.5f,
                                      -__fmul_rn(dy, __fmul_rn(dx, q0.w)));
        const float G = blend_exp<POLY>(power);
        const float alpha = fminf(0.99f, __fmul_rn(q1.y, G));
        const bool ok = valid && (e < hi_rel) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        any_ok |= ok;
        nblend += ok ? 1u : 0u;
        // transmittance / colour recurrence of backward.cu:497-534
        const float om = 1.f - alpha;
        float rinv = __fdividef(1.f, om);               // MUFU.RCP + one Newton step (om in [0.01, 1])
        rinv = __fmaf_rn(__fmaf_rn(-om, rinv, 1.f), rinv, rinv);
        const float Tn = T * rinv;
        const float a0 = __fmaf_rn(last_alpha, lc0 - acc0, acc0);
        const float a1 = __fmaf_rn(last_alpha, lc1 - acc1, acc1);
        const float a2 = __fmaf_rn(last_alpha, lc2 - acc2, acc2);
        float dsum = (q1.z - a0) * dp0 + (q1.w - a1) * dp1 + (c2 - a2) * dp2;
        float a3 = 0.f;
        if (OP) {
          a3 = __fmaf_rn(last_alpha, lc3 - acc3, acc3);
          dsum += (1.f - a3) * dp3;
        }
        const float dL_dalpha = __fmaf_rn(dsum, Tn, nTfbg * rinv);
        const float g = ok ? G * dL_dalpha : 0.f;        // a pair that was not blended contributes exact zeros
        const float wgt = ok ? alpha * Tn : 0.f;         // dchannel_dcolor
        const float gdx = g * dx, gdy = g * dy;
        float* vk = v + 8 * k;
        vk[0] = gdx; vk[1] = gdy; vk[2] = gdx * dx; vk[3] = gdx * dy; vk[4] = gdy * dy; vk[5] = g;
        vk[6] = wgt * dp0; vk[7] = wgt * dp1; v8[k] = wgt * dp2;
        if (ok) {
          T = Tn; acc0 = a0; acc1 = a1; acc2 = a2; lc0 = q1.z; lc1 = q1.w; lc2 = c2; last_alpha = alpha;
          if (OP) { acc3 = a3; lc3 = 1.f; }
        }
      }
      if (!__any_sync(0xFFFFFFFFu, any_ok)) continue;
      // group reduction BEFORE the atomics: 2 entries x 8 terms over the 8 lanes of the group -> lane i ends with the totals of
      // terms 2 (i & 3), 2 (i & 3) + 1 of entry (i >> 2): one 8-byte vector reduction per lane
      xpose_step<16, 4>(v, lane);
      xpose_step<8, 2>(v, lane);
      xpose_step<4, 1>(v, lane);
      xpose_step<2, 4>(v8, lane);
      float r8 = v8[0];
      r8 += __shfl_xor_sync(0xFFFFFFFFu, r8, 2);
      r8 += __shfl_xor_sync(0xFFFFFFFFu, r8, 1);
      const uint32_t my_id = (lane & 4) ? ids[1] : ids[0];
      float* const dst = acc + (size_t)GS_ACC * my_id + 2 * (lane & 3);
      if (v[0] != 0.f || v[1] != 0.f) asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst), "f"(v[0]), "f"(v[1]) : "memory");
      if ((lane & 3) == 0 && r8 != 0.f) atomicAdd(acc + (size_t)GS_ACC * my_id + 8, r8);
    }
  }
  nblend = __reduce_add_sync(0xFFFFFFFFu, nblend);
  if (lane == 0 && nblend) atomicAdd(&hdr->evals_blended, (unsigned long long)nblend);
}

// The grid is sized from the pair CAPACITY (the host does not know N), capped at GS_BWD_MAX_GRID CTAs that walk the frame's
// items with a grid stride: a capacity left large by an earlier, bigger frame then costs nothing (it used to launch one CTA
// per possible item -- 69 000 mostly empty CTAs, 130 us instead of 62, after a 1M-Gaussian frame had raised the capacity).
#define GS_BWD_MAX_GRID (148 * GS_BWD_MINBLOCKS * 8)
template <bool OP, bool POLY>
__global__ void __launch_bounds__(kWarps * 32, GS_BWD_MINBLOCKS)
blend_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int W, int H, int grid_x,
                 const uint32_t* __restrict__ seg_off, int tiles, const uint2* __restrict__ items,
                 const float4* __restrict__ ckpt,
                 const float* __restrict__ bg, const GsRec* __restrict__ rec, const float* __restrict__ final_Ts,
                 const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                 float* __restrict__ acc /*[P,GS_ACC]*/,
                 const float* __restrict__ ckpt_op, const float* __restrict__ dL_dop_pix,
                 const uint32_t* __restrict__ rowmask, GsHeader* __restrict__ hdr) {
  GS_PDL_PROLOGUE();
  extern __shared__ __align__(16) unsigned char bwd_smem[];
  BwdSlab& sl = reinterpret_cast<BwdSlab*>(bwd_smem)[threadIdx.x >> 5];
  const uint32_t n_items = seg_off[tiles];
  for (uint32_t it = blockIdx.x; it < n_items; it += gridDim.x) {
    blend_bwd_item<OP, POLY>(it, sl, ranges, list, W, H, grid_x, seg_off, items, ckpt, bg, rec, final_Ts, n_contrib, dL_dpixels, acc,
                             ckpt_op, dL_dop_pix, rowmask, hdr);
    __syncwarp();  // the warp's slab is reused by its next item
  }
}

}  // namespace

int gs_launch_blend_fwd(const GsForwardArgs& a, const GsGeom& g, const GsImage& im, const GsList& l, const GsScratch& s,
                        cudaStream_t st) {
  const int gx = (a.width + GS_TILE - 1) / GS_TILE, gy = (a.height + GS_TILE - 1) / GS_TILE;
  const bool op = a.out_opacity != nullptr, poly = gs_exp_mode() != 0;
  // Which forward blend: the chunked kernel for small tile grids (avatar frames: few deep tiles, the chain is the critical path),
  // the per-block kernel for large ones (thousands of shallow tiles: throughput).  GS_FWD_KERNEL=chunk|block overrides.
  static const int forced = [] { const char* e = getenv("GS_FWD_KERNEL"); return !e ? 0 : (e[0] == 'b' ? 2 : 1); }();
  const bool block_kernel = forced ? forced == 2 : gx * gy > GS_BIN_MAX_TILES;
  if (block_kernel) {
    const size_t smem = sizeof(BkCtl) + kBkProducers * sizeof(BkBuf);
    auto kern = op ? (poly ? blend_fwd_block_kernel<true, true> : blend_fwd_block_kernel<true, false>)
                   : (poly ? blend_fwd_block_kernel<false, true> : blend_fwd_block_kernel<false, false>);
    if (gs_check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != GS_OK) return GS_ERR_CUDA;
    gs_launch(kern, dim3((unsigned)(gx * gy) * 8u), dim3((kBkProducers + 1) * 32), smem, st, im.ranges, l.point_list, a.width, a.height,
              g.rec, a.background, im.final_T, im.n_contrib, a.out_color, im.seg_off, l.ckpt, (const uint32_t*)l.order, gx, a.out_opacity,
              l.ckpt_op, l.rowmask);
    return gs_check(cudaGetLastError());
  }
  const size_t smem = sizeof(FwdSmem);
  auto kern = op ? (poly ? blend_fwd_kernel<true, true> : blend_fwd_kernel<true, false>)
                 : (poly ? blend_fwd_kernel<false, true> : blend_fwd_kernel<false, false>);
  // per device and per instantiation; cheap enough to set on every launch (a process may drive several GPUs)
  if (gs_check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != GS_OK) return GS_ERR_CUDA;
  (void)gy;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  static int sm_count[64];
  if (dev >= 0 && dev < 64) {
    if (!sm_count[dev]) cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    sms = sm_count[dev] ? sm_count[dev] : sms;
  }
  static const int poll_ns = [] { const char* e = getenv("GS_FWD_POLL_NS"); return e ? atoi(e) : 32; }();  // pause between hand-off polls
  const size_t work_max = l.max_fitems * 8;  // upper bound of the frame's work units
  const unsigned grid = (unsigned)(work_max < (size_t)sms * GS_FWD_CTAS ? work_max : (size_t)sms * GS_FWD_CTAS);
  gs_launch(kern, dim3(grid), dim3(kFwdWarps * 32), smem, st, im.ranges, l.point_list, a.width,
            a.height, g.rec, a.background, im.final_T, im.n_contrib, a.out_color, im.seg_off, l.ckpt, (const uint4*)l.fitems,
            g.hdr, gx, a.out_opacity, l.ckpt_op, l.rowmask, s.hand, poll_ns);
  return gs_check(cudaGetLastError());
}

int gs_launch_blend_bwd(const GsBackwardArgs& a, const GsGeom& g, const GsImage& im, const GsList& l, cudaStream_t st) {
  const int gx = (a.width + GS_TILE - 1) / GS_TILE, gy = (a.height + GS_TILE - 1) / GS_TILE;
  const bool op = a.dL_dopacity_pix != nullptr, poly = gs_exp_mode() != 0;
  auto kern = op ? (poly ? blend_bwd_kernel<true, true> : blend_bwd_kernel<true, false>)
                 : (poly ? blend_bwd_kernel<false, true> : blend_bwd_kernel<false, false>);
  const size_t smem = kWarps * sizeof(BwdSlab);
  if (gs_check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != GS_OK) return GS_ERR_CUDA;
  const unsigned grid = (unsigned)(l.max_items < (size_t)GS_BWD_MAX_GRID ? l.max_items : (size_t)GS_BWD_MAX_GRID);
  gs_launch(kern, dim3(grid), dim3(kWarps * 32), smem, st,
      im.ranges, l.point_list, a.width, a.height, gx, im.seg_off, gx * gy, l.items, l.ckpt, a.background, g.rec,
      im.final_T, im.n_contrib, a.dL_dpix, g.acc,
      (const float*)l.ckpt_op, a.dL_dopacity_pix, (const uint32_t*)l.rowmask, g.hdr);
  return gs_check(cudaGetLastError());
}

#ifdef GS_FWD_TRACE
extern "C" int gs_debug_fwd_trace(unsigned long long* host, int n) {
  return gs_check(cudaMemcpyFromSymbol(host, g_fwd_trace, sizeof(unsigned long long) * n));
}
#endif
