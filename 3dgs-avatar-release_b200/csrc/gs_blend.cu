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

#include "gs_common.cuh"

namespace {

constexpr int kWarps = 8;

struct Slab {  // per-warp staging of 32 list entries: 2 x LDS.128 + 1 x LDS.32 broadcast reads per evaluated entry
  float4 q0[32];  // x, y, conic a, conic b
  float4 q1[32];  // conic c, opacity, r, g
  float q2[32];   // b
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

// Gathers the 48-byte record of one list entry (3 x LDG.128).
struct Rec3 { float4 a, b, c; };
__device__ __forceinline__ Rec3 load_rec(const GsRec* __restrict__ rec, uint32_t id) {
  const float4* rp = reinterpret_cast<const float4*>(rec + id);
  Rec3 r;
  r.a = __ldg(rp); r.b = __ldg(rp + 1); r.c = __ldg(rp + 2);
  return r;
}

constexpr int kChunk = 4;  // survivors evaluated together: their exp() chains are independent (ILP), only the T update is serial

// ---- forward -------------------------------------------------------------------------------------------------------
// The only serial dependency of front-to-back blending is the transmittance chain T <- T*(1-alpha); culling and the
// alpha evaluation (the exp) of a batch do not depend on it.  One CTA = one 8x4 pixel block of one tile (the eight
// blocks of a deep tile land on different SMs: one SM's issue slots were the bound), warp-specialised:
//   * kProducers producer warps take the block's 32-entry batches round-robin; each culls its batch against the block,
//     evaluates alpha for every survivor and pixel (the reference's tests, forward.cu:325-345) and leaves the rows in
//     its shared-memory buffer, in list order;
//   * ONE consumer warp keeps (T, C, last, done) in registers and applies the rows batch by batch in list order --
//     exactly the reference's operation sequence, so final_T / n_contrib / image stay bit-identical.
// full/empty mbarriers per producer buffer: waiting warps are suspended by the hardware instead of spinning on issue
// slots.  A tile list of 6000 entries (few tiles, deep lists: the avatar case) costs ~500 cycles per batch on the
// consumer's chain instead of ~2200 in a one-warp-does-everything loop.
#ifndef GS_FWD_PRODUCERS
#define GS_FWD_PRODUCERS 7
#endif
#ifndef GS_FWD_PCHUNK
#define GS_FWD_PCHUNK 4
#endif
constexpr int kProducers = GS_FWD_PRODUCERS;
constexpr int kFwdWarps = kProducers + 1;
constexpr int kPChunk = GS_FWD_PCHUNK;  // survivors whose alpha is evaluated together (ILP) by a producer

struct FwdBuf {            // one per producer warp
  float4 q0[32], q1[32];   // staged records (see Slab)
  float q2[32];
  float alpha[32][32];     // [survivor row][pixel lane]; 0 = the reference's skip tests failed
  float4 row[32];          // [survivor row] r, g, b, 1-based list position (int bits)
  uint32_t cnt;            // rows in this batch
  uint32_t pad[3];
};
struct FwdCtl {
  unsigned long long full[kProducers];   // producer -> consumer: rows of the buffer are ready
  unsigned long long empty[kProducers];  // consumer -> producer: buffer may be overwritten
  uint32_t done;                         // every pixel of the block has terminated
  uint32_t pad[3];
};

__device__ __forceinline__ uint32_t ld_acquire_cta(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"((uint32_t)__cvta_generic_to_shared(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_cta(uint32_t* p, uint32_t v) {
  asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {  // release: prior shared-memory writes are visible
  asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_relaxed(unsigned long long* bar) {  // WAR hand-back: the reads were consumed
  asm volatile("mbarrier.arrive.relaxed.cta.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, uint32_t parity) {  // bounded suspend, then returns
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%1], %2, %3;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n" : "=r"(ok) : "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(parity), "r"(1000u) : "memory");
  return ok != 0;
}
// warp-uniform read of a shared flag: lane 0 loads (acquire), the value is broadcast
__device__ __forceinline__ uint32_t flag_now(const uint32_t* f, int lane) {
  uint32_t t = 0;
  if (lane == 0) t = ld_acquire_cta(f);
  t = __shfl_sync(0xFFFFFFFFu, t, 0);
  return t;
}

#ifdef GS_FWD_TRACE
__device__ long long g_fwd_trace[512 * 8];
#define TRACE(b, i) do { if (trace_on && lane == 0 && (b) < 512) s_trace[(b) * 8 + (i)] = clock64(); } while (0)
#else
#define TRACE(b, i) do { } while (0)
#endif

// OP: the fused opacity pass (gaussian_renderer/__init__.py:131-142: a second rasterisation of the same geometry with
// colours = 1, channel 0 kept).  The consumer carries one more accumulator, C3 += T * alpha -- what forward.cu:349
// computes for a feature of exactly 1.0 -- so out_opacity is bit-identical to the second call's channel 0.
template <bool OP>
__global__ void __launch_bounds__(kFwdWarps * 32)
blend_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int W, int H,
                 const GsRec* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ final_T,
                 uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
                 const uint32_t* __restrict__ seg_off, float4* __restrict__ ckpt, const uint32_t* __restrict__ order,
                 int grid_x, float* __restrict__ out_opacity, float* __restrict__ ckpt_op, uint32_t* __restrict__ rowmask) {
  GS_PDL_PROLOGUE();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FwdCtl& ctl = *reinterpret_cast<FwdCtl*>(smem_raw);
  FwdBuf* const bufs = reinterpret_cast<FwdBuf*>(smem_raw + sizeof(FwdCtl));
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;  // warp 0 = consumer, 1..kProducers = producers
  const int blk = blockIdx.x & 7;                              // pixel block of the tile
  const int tile = (int)order[blockIdx.x >> 3];                // longest lists first
  const int tile_x = tile % grid_x, tile_y = tile / grid_x;
  const uint2 range = ranges[tile];
  const int len = (int)(range.y - range.x);
  const int nb = (len + 31) >> 5;
  const int bx = tile_x * GS_TILE + (blk & 1) * 8, by = tile_y * GS_TILE + (blk >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < W && py < H;
#ifdef GS_FWD_TRACE
  __shared__ long long s_trace[512 * 8];
  const bool trace_on = len >= 6000 && blk == 3;
#endif
  if (threadIdx.x == 0) {
    ctl.done = 0;
    for (int i = 0; i < kProducers; i++) { mbar_init(&ctl.full[i], 1); mbar_init(&ctl.empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (wid > 0) {
    // ------------------------------------------------ producer ------------------------------------------------
    const int p = wid - 1;
    FwdBuf& ws = bufs[p];
    const float pfx = (float)px, pfy = (float)py;
    const float X0 = (float)bx, X1 = (float)(bx + 7), Y0 = (float)by, Y1 = (float)(by + 3);
    const uint32_t* const tl = list + range.x;
    uint32_t* const rm = rowmask + ((size_t)seg_off[tile] * 8 + blk) * (GS_SEG / 32);  // + segment * 32 + batch in segment
    Rec3 cur;
    cur.a = cur.b = cur.c = make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t id_next = 0;
    {
      const int e = p * 32 + lane;
      if (e < len) cur = load_rec(rec, tl[e]);
      if (e + kProducers * 32 < len) id_next = tl[e + kProducers * 32];
    }
    uint32_t it = 0;
    for (int b = p; b < nb; b += kProducers, it++) {
      // software pipeline: records of this warp's next batch, list indices of the one after
      const int e = b * 32 + lane;
      Rec3 nxt = cur;
      if (e + kProducers * 32 < len) nxt = load_rec(rec, id_next);
      uint32_t id_nn = 0;
      if (e + 2 * kProducers * 32 < len) id_nn = tl[e + 2 * kProducers * 32];
      const bool keep = (e < len) && may_contribute(cur.a.x, cur.a.y, cur.b.x, cur.b.y, cur.b.z, cur.a.z, X0, X1, Y0, Y1);
      // the consumer has finished with the previous rows -- or every pixel has terminated: the flag is polled inside the
      // wait because the consumer's wake-up arrivals at termination can put the barrier two phases ahead (parity alias)
      if (it > 0) {
        while (!mbar_try_wait(&ctl.empty[p], (it - 1) & 1u)) {
          if (ld_acquire_cta(&ctl.done)) break;
        }
      }
      __syncwarp();
      if (flag_now(&ctl.done, lane)) break;
      TRACE(b, 0);
      if (e < len) {
        ws.q0[lane] = make_float4(cur.a.x, cur.a.y, cur.b.x, cur.b.y);
        ws.q1[lane] = make_float4(cur.b.z, cur.b.w, cur.c.x, cur.c.y);
        ws.q2[lane] = cur.c.z;
      }
      __syncwarp();
      const uint32_t mask = __ballot_sync(0xFFFFFFFFu, keep);
      uint32_t m = mask, kept = 0;
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
          const float alpha = fminf(0.99f, __fmul_rn(q1.y, gs_exp(power)));
          const bool ok = valid && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
          if (__any_sync(0xFFFFFFFFu, ok)) {  // rows no pixel of the block blends are dropped here, off the serial chain
            ws.alpha[k][lane] = ok ? alpha : 0.0f;
            if (lane == 0) ws.row[k] = make_float4(q1.z, q1.w, ws.q2[j], __int_as_float(b * 32 + j + 1));
            kept |= 1u << j;
            k++;
          }
        }
      }
      for (int kk = k; kk & (kChunk - 1); kk++) {  // the consumer reads whole chunks, branch-free: pad rows must be 0 * finite
        ws.alpha[kk][lane] = 0.0f;
        if (lane == 0) ws.row[kk] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      if (lane == 0) {
        ws.cnt = (uint32_t)k;
        rm[(size_t)(b >> 2) * 32 + (b & 3)] = kept;  // the backward blend visits exactly these entries
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl.full[p]);
      TRACE(b, 1);
      cur = nxt;
      id_next = id_nn;
    }
    return;
  }

  // -------------------------------------------------- consumer --------------------------------------------------
  // checkpoint slots of this tile: one per segment start + one final (backward work items restart from them)
  const size_t ck_base = ((size_t)seg_off[tile] + tile) * 256 + blk * 32 + lane;
  float4* const ck = ckpt + ck_base;
  float* const ck3 = OP ? ckpt_op + ck_base : nullptr;
  // Front-to-back blending of one pixel, forward.cu:335-361, without carried control state.  Tr is the running product
  // of (1 - alpha) over ALL rows; it never increases (0 < 1 - alpha <= 1; a row the pixel skips has alpha = 0 and
  // multiplies by exactly 1), so "the pixel has terminated" (the reference's done flag, set by the first row with
  // T * (1 - alpha) < 1e-4) is simply Tr < 1e-4, and the reference's T is the last Tr that was still >= 1e-4.  A live row
  // with alpha = 0 adds Tr * (0 * c) = 0 to C, leaving its bits unchanged (finite colours), so only n_contrib needs the
  // alpha != 0 test.
  // Same values in the same order as the reference; 12 instructions per row on the consumer's chain instead of 24.
  float T = 1.0f, Tr = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, C3 = 0.f;
  uint32_t last = 0;
  for (int b = 0; b < nb; b++) {
    const int p = b % kProducers;
    FwdBuf& ws = bufs[p];
    TRACE(b, 2);
    mbar_wait(&ctl.full[p], (uint32_t)(b / kProducers) & 1u);
    TRACE(b, 3);
    if ((b & (GS_SEG / 32 - 1)) == 0) {
      ck[(size_t)(b / (GS_SEG / 32)) * 256] = make_float4(T, C0, C1, C2);
      if (OP) ck3[(size_t)(b / (GS_SEG / 32)) * 256] = C3;
    }
    const int cnt = (int)ws.cnt;  // rows up to the next multiple of kChunk are zero-padded by the producer
#ifdef GS_FWD_TRACE
    if (trace_on && lane == 0 && b < 512) s_trace[b * 8 + 5] = cnt;
#endif
    for (int k0 = 0; k0 < cnt; k0 += kChunk) {
      float al[kChunk];
      float4 rw[kChunk];
#pragma unroll
      for (int u = 0; u < kChunk; u++) {
        al[u] = ws.alpha[k0 + u][lane];
        rw[u] = ws.row[k0 + u];
      }
#pragma unroll
      for (int u = 0; u < kChunk; u++) {
        const float test_T = __fmul_rn(Tr, __fsub_rn(1.0f, al[u]));
        const bool ok = !(test_T < 0.0001f);
        const float Tw = ok ? Tr : 0.0f;  // branch-free: a terminated pixel adds 0 * (alpha * c) = 0, C keeps its bits
        C0 = __fmaf_rn(Tw, __fmul_rn(al[u], rw[u].x), C0);
        C1 = __fmaf_rn(Tw, __fmul_rn(al[u], rw[u].y), C1);
        C2 = __fmaf_rn(Tw, __fmul_rn(al[u], rw[u].z), C2);
        if (OP) C3 = __fmaf_rn(Tw, al[u], C3);  // colour 1.0: (1 * alpha) * T
        T = ok ? test_T : T;
        last = (ok && al[u] != 0.0f) ? (uint32_t)__float_as_int(rw[u].w) : last;  // contributor, forward.cu:328 (1-based)
        Tr = test_T;
      }
    }
    const bool all_done = __all_sync(0xFFFFFFFFu, !inside || Tr < 0.0001f);
    if (lane == 0) {
      if (all_done) {  // stop the producers: whoever waits on an empty barrier wakes and sees the flag
        st_release_cta(&ctl.done, 1u);
        for (int i = 0; i < kProducers; i++) mbar_arrive(&ctl.empty[i]);
      } else {
        mbar_arrive_relaxed(&ctl.empty[p]);
      }
    }
    TRACE(b, 4);
    if (all_done) break;
  }
#ifdef GS_FWD_TRACE
  __syncwarp();
  if (trace_on) for (int i = lane; i < 512 * 8; i += 32) g_fwd_trace[i] = s_trace[i];
#endif
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

// ---- forward, two-kernel variant -------------------------------------------------------------------------------------
// The trace of the one-kernel blend (profiles/tools/fwd_trace.py) shows the deepest tile's consumer spending ~70 % of its
// time on producer/consumer hand-overs, not on the transmittance chain.  Here the two halves are separate grids:
//   blend_rows_kernel   fully parallel over (tile, 128-entry segment) x 8 pixel blocks: cull + the reference's alpha tests,
//                       surviving rows written to fixed (tile, block, batch) slots (L2-resident between the two grids);
//   blend_chain_kernel  ONE warp per (tile, pixel block), longest lists first: streams its rows through a cp.async ring
//                       (no cross-warp synchronisation at all) and runs only the serial part -- the same operations in the
//                       same order as the consumer above, so final_T / n_contrib / image stay bit-identical.
__device__ __forceinline__ size_t rows_batch_index(uint32_t seg_base, uint32_t nseg, int blk, int b) {
  return ((size_t)seg_base * 8 + (size_t)blk * nseg) * (GS_SEG / 32) + (size_t)b;  // [tile][pixel block][batch]
}

__global__ void __launch_bounds__(kWarps * 32)
blend_rows_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int grid_x,
                  const uint32_t* __restrict__ seg_off, int tiles, const uint2* __restrict__ items,
                  const GsRec* __restrict__ rec, float* __restrict__ rows_alpha, float4* __restrict__ rows_data,
                  uint8_t* __restrict__ rows_cnt, uint32_t* __restrict__ rowmask) {
  GS_PDL_PROLOGUE();
  __shared__ Slab s_slab[kWarps];
  if (blockIdx.x >= seg_off[tiles]) return;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Slab& sl = s_slab[w];
  const uint2 item = items[blockIdx.x];
  const int tile = (int)item.x, seg = (int)item.y;
  const uint2 range = ranges[tile];
  const int len = (int)(range.y - range.x);
  const int seg_begin = seg * GS_SEG, seg_end = min(len, seg_begin + GS_SEG);
  const uint32_t seg_base = seg_off[tile], nseg = (uint32_t)((len + GS_SEG - 1) / GS_SEG);
  const int tx = tile % grid_x, ty = tile / grid_x;
  const int bx = tx * GS_TILE + (w & 1) * 8, by = ty * GS_TILE + (w >> 1) * 4;
  const float pfx = (float)(bx + (lane & 7)), pfy = (float)(by + (lane >> 3));
  const float X0 = (float)bx, X1 = (float)(bx + 7), Y0 = (float)by, Y1 = (float)(by + 3);
  const uint32_t* const sl_list = list + range.x + seg_begin;
  constexpr int NB = GS_SEG / 32;
  Rec3 r[NB];  // all gathers of the segment in flight at once
#pragma unroll
  for (int bb = 0; bb < NB; bb++) {
    r[bb].a = r[bb].b = r[bb].c = make_float4(0.f, 0.f, 0.f, 0.f);
    if (seg_begin + bb * 32 + lane < seg_end) r[bb] = load_rec(rec, sl_list[bb * 32 + lane]);
  }
#pragma unroll
  for (int bb = 0; bb < NB; bb++) {
    const bool valid_e = seg_begin + bb * 32 + lane < seg_end;
    const int b = seg * NB + bb;  // batch index inside the tile list
    const size_t q = rows_batch_index(seg_base, nseg, w, b);
    const Rec3 cur = r[bb];
    const bool keep = valid_e && may_contribute(cur.a.x, cur.a.y, cur.b.x, cur.b.y, cur.b.z, cur.a.z, X0, X1, Y0, Y1);
    if (valid_e) {
      sl.q0[lane] = make_float4(cur.a.x, cur.a.y, cur.b.x, cur.b.y);
      sl.q1[lane] = make_float4(cur.b.z, cur.b.w, cur.c.x, cur.c.y);
      sl.q2[lane] = cur.c.z;
    }
    __syncwarp();
    uint32_t m = __ballot_sync(0xFFFFFFFFu, keep), kept = 0;
    float* const arow = rows_alpha + q * 1024 + lane;
    float4* const drow = rows_data + q * 32;
    int k = 0;
    while (m) {
#pragma unroll
      for (int u = 0; u < kPChunk; u++) {
        const bool valid = m != 0;
        const int j = valid ? (__ffs(m) - 1) : 0;
        m &= m - 1;
        const float4 q0 = sl.q0[j], q1 = sl.q1[j];
        const float dx = __fsub_rn(q0.x, pfx), dy = __fsub_rn(q0.y, pfy);
        const float power = __fmaf_rn(__fmaf_rn(dx, __fmul_rn(dx, q0.z), __fmul_rn(dy, __fmul_rn(dy, q1.x))), -0.5f,
                                      -__fmul_rn(dy, __fmul_rn(dx, q0.w)));
        const float alpha = fminf(0.99f, __fmul_rn(q1.y, gs_exp(power)));
        const bool ok = valid && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        if (__any_sync(0xFFFFFFFFu, ok)) {  // rows no pixel of the block blends never reach the chain
          arow[k * 32] = ok ? alpha : 0.0f;
          if (lane == 0) drow[k] = make_float4(q1.z, q1.w, sl.q2[j], __int_as_float(b * 32 + j + 1));
          kept |= 1u << j;
          k++;
        }
      }
    }
    for (int kk = k; kk & (kChunk - 1); kk++) {  // the chain reads whole chunks, branch-free: pad rows are 0 * finite
      arow[kk * 32] = 0.0f;
      if (lane == 0) drow[kk] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (lane == 0) {
      rows_cnt[q] = (uint8_t)k;
      rowmask[(((size_t)seg_base + seg) * 8 + w) * NB + bb] = kept;
    }
    __syncwarp();  // the slab is reused by the next batch
  }
}

constexpr int kRowDepth = 4;  // batches in flight per chain warp
struct RowBuf {
  float alpha[32][32];
  float4 row[32];
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <bool OP>
__global__ void __launch_bounds__(32)
blend_chain_kernel(const uint2* __restrict__ ranges, int W, int H, const float* __restrict__ bg, float* __restrict__ final_T,
                   uint32_t* __restrict__ n_contrib, float* __restrict__ out_color, const uint32_t* __restrict__ seg_off,
                   float4* __restrict__ ckpt, const uint32_t* __restrict__ order, int grid_x, float* __restrict__ out_opacity,
                   float* __restrict__ ckpt_op, const float* __restrict__ rows_alpha, const float4* __restrict__ rows_data,
                   const uint8_t* __restrict__ rows_cnt, uint32_t max_items) {
  GS_PDL_PROLOGUE();
  __shared__ __align__(16) RowBuf ring[kRowDepth];
  const int lane = threadIdx.x;
  const int blk = blockIdx.x & 7;
  const int tile = (int)order[blockIdx.x >> 3];  // longest lists first
  const int tile_x = tile % grid_x, tile_y = tile / grid_x;
  const uint2 range = ranges[tile];
  const int len = (int)(range.y - range.x);
  const uint32_t seg_base = seg_off[tile], nseg = (uint32_t)((len + GS_SEG - 1) / GS_SEG);
  int nb = (len + 31) >> 5;
  if (seg_base + nseg > max_items) nb = 0;  // overflowed frame (invalid anyway, rerun by the caller): stay in bounds
  const int bx = tile_x * GS_TILE + (blk & 1) * 8, by = tile_y * GS_TILE + (blk >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < W && py < H;
  const size_t q0 = rows_batch_index(seg_base, nseg, blk, 0);
  // row counts of 64 batches in two registers per lane (lane j: batch chunk_base + j and chunk_base + 32 + j)
  int chunk_base = 0;
  uint32_t cA = lane < nb ? rows_cnt[q0 + lane] : 0u, cB = 32 + lane < nb ? rows_cnt[q0 + 32 + lane] : 0u;
  auto count_of = [&](int b) -> int {
    const int rel = b - chunk_base;
    return (int)__shfl_sync(0xFFFFFFFFu, rel < 32 ? cA : cB, rel & 31);
  };
  auto issue = [&](int b) {  // copies of batch b into its ring slot (one commit group per batch, possibly empty)
    if (b < nb) {
      const int cp = (count_of(b) + kChunk - 1) & ~(kChunk - 1);
      RowBuf& buf = ring[b % kRowDepth];
      const float* const srcA = rows_alpha + (q0 + b) * 1024;
      for (int r0 = 0; r0 < cp; r0 += 4)  // 4 rows x 128 B per warp instruction
        cp_async16(&buf.alpha[r0 + (lane >> 3)][(lane & 7) * 4], srcA + (r0 + (lane >> 3)) * 32 + (lane & 7) * 4);
      if (lane < cp) cp_async16(&buf.row[lane], rows_data + (q0 + b) * 32 + lane);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int d = 0; d < kRowDepth; d++) issue(d);

  const size_t ck_base = ((size_t)seg_base + tile) * 256 + blk * 32 + lane;
  float4* const ck = ckpt + ck_base;
  float* const ck3 = OP ? ckpt_op + ck_base : nullptr;
  float T = 1.0f, Tr = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, C3 = 0.f;  // see the consumer of blend_fwd_kernel
  uint32_t last = 0;
  for (int b = 0; b < nb; b++) {
    if (b - chunk_base == 32) {
      chunk_base += 32;
      cA = cB;
      cB = chunk_base + 32 + lane < nb ? rows_cnt[q0 + chunk_base + 32 + lane] : 0u;
    }
    cp_async_wait<kRowDepth - 1>();
    __syncwarp();
    if ((b & (GS_SEG / 32 - 1)) == 0) {
      ck[(size_t)(b / (GS_SEG / 32)) * 256] = make_float4(T, C0, C1, C2);
      if (OP) ck3[(size_t)(b / (GS_SEG / 32)) * 256] = C3;
    }
    const int cnt = count_of(b);
    const RowBuf& ws = ring[b % kRowDepth];
    for (int k0 = 0; k0 < cnt; k0 += kChunk) {
      float al[kChunk];
      float4 rw[kChunk];
#pragma unroll
      for (int u = 0; u < kChunk; u++) {
        al[u] = ws.alpha[k0 + u][lane];
        rw[u] = ws.row[k0 + u];
      }
#pragma unroll
      for (int u = 0; u < kChunk; u++) {
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
    if (__all_sync(0xFFFFFFFFu, !inside || Tr < 0.0001f)) break;  // every pixel of the block has terminated
    __syncwarp();  // every lane has finished reading the slot that is refilled now
    issue(b + kRowDepth);
  }
  cp_async_wait<0>();
  ck[(size_t)((len + GS_SEG - 1) / GS_SEG) * 256] = make_float4(T, C0, C1, C2);
  if (OP) ck3[(size_t)((len + GS_SEG - 1) / GS_SEG) * 256] = C3;
  if (inside) {
    const size_t pix = (size_t)py * W + px, hw = (size_t)H * W;
    final_T[pix] = T;
    n_contrib[pix] = last;
    out_color[pix] = __fmaf_rn(T, bg[0], C0);
    out_color[hw + pix] = __fmaf_rn(T, bg[1], C1);
    out_color[2 * hw + pix] = __fmaf_rn(T, bg[2], C2);
    if (OP) out_opacity[pix] = __fmaf_rn(T, bg[0], C3);
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
#ifndef GS_BWD_MINBLOCKS
#define GS_BWD_MINBLOCKS 2
#endif
// OP: upstream gradient of the fused opacity image too.  dL/dalpha is linear in the per-channel terms, so the opacity
// channel (colour 1, no colour gradient) is one more term of the same sum and every geometry gradient comes out as the
// sum of the reference's two backward calls.
template <bool OP>
__global__ void __launch_bounds__(kWarps * 32, GS_BWD_MINBLOCKS)
blend_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ list, int W, int H, int grid_x,
                 const uint32_t* __restrict__ seg_off, int tiles, const uint2* __restrict__ items,
                 const float4* __restrict__ ckpt,
                 const float* __restrict__ bg, const GsRec* __restrict__ rec, const float* __restrict__ final_Ts,
                 const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                 float* __restrict__ dL_dmean2D /*[P,3]*/, float* __restrict__ dL_dconic /*[P,4]*/,
                 float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolors,
                 const float* __restrict__ ckpt_op, const float* __restrict__ dL_dop_pix,
                 const uint32_t* __restrict__ rowmask) {
  GS_PDL_PROLOGUE();
  __shared__ Slab s_slab[kWarps];
  if (blockIdx.x >= seg_off[tiles]) return;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Slab& sl = s_slab[w];
  const uint2 item = items[blockIdx.x];
  const int tile = (int)item.x;
  const uint2 range = ranges[tile];
  const int len = (int)(range.y - range.x);
  const int seg_begin = (int)item.y * GS_SEG, seg_end = min(len, seg_begin + GS_SEG);
  const int tx = tile % grid_x, ty = tile / grid_x;
  const int bx = tx * GS_TILE + (w & 1) * 8, by = ty * GS_TILE + (w >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
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
  // entries of this segment that at least one pixel of this block blended in the forward pass (written by the forward's
  // producers for every batch its consumer reached, i.e. every batch below wmax): the only ones worth visiting
  const uint4 mk = *reinterpret_cast<const uint4*>(rowmask + (((size_t)seg_off[tile] + item.y) * 8 + w) * (GS_SEG / 32));
  auto mask_of = [&](int bq) -> uint32_t { return bq == 0 ? mk.x : bq == 1 ? mk.y : bq == 2 ? mk.z : mk.w; };
  const float T_final = inside ? final_Ts[pix] : 0.f;
  float T = T_final;
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, last_alpha = 0.f;
  float acc3 = 0.f, lc3 = 0.f;  // opacity channel: its colour is 1 once a contributor has been seen
  if (last_contributor > seg_end) {  // restart from the checkpoint at the segment end
    const size_t ck_base = ((size_t)seg_off[tile] + tile) * 256 + threadIdx.x;
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
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;  // backward.cu:460-461
  // which accumulator this lane feeds after the transposing reduction: entry (lane >> 3) of the chunk, component lane & 7
  const int comp = lane & 7;
  float* const comp_base = comp < 2 ? dL_dmean2D + comp : comp < 5 ? dL_dconic + (comp == 4 ? 3 : comp - 2)
                                    : comp == 5 ? dL_dopacity : dL_dcolors + (comp - 6);
  const int comp_stride = comp < 2 ? 3 : comp < 5 ? 4 : comp == 5 ? 1 : 3;

  // batches aligned to the start of the range so that positions map to lanes: pos = b*32 + lane; software pipeline as
  // in the forward kernel (records of batch b and indices of batch b-1 in registers while batch b is processed)
  const uint32_t* const seg_list = list + range.x + seg_begin;
  wmax -= seg_begin;  // positions below are relative to the segment start (a multiple of 32)
  const int hi_rel = hi - seg_begin;
  int b = (wmax - 1) >> 5;
  Rec3 cur;
  cur.a = cur.b = cur.c = make_float4(0.f, 0.f, 0.f, 0.f);
  uint32_t id_cur = 0, id_next = 0;
  {
    const int pos = b * 32 + lane;
    if (pos < wmax) { id_cur = seg_list[pos]; if ((mask_of(b) >> lane) & 1u) cur = load_rec(rec, id_cur); }
    if (b > 0) id_next = seg_list[pos - 32];
  }
  for (; b >= 0; b--) {
    const int pos = b * 32 + lane;
    bool keep = false;
    if (pos < wmax && ((mask_of(b) >> lane) & 1u)) {
      keep = true;
      sl.q0[lane] = make_float4(cur.a.x, cur.a.y, cur.b.x, cur.b.y);
      sl.q1[lane] = make_float4(cur.b.z, cur.b.w, cur.c.x, cur.c.y);
      sl.q2[lane] = cur.c.z;
      sl.id[lane] = id_cur;
    }
    Rec3 nxt = cur;
    if (b > 0 && ((mask_of(b - 1) >> lane) & 1u)) nxt = load_rec(rec, id_next);
    uint32_t id_nn = 0;
    if (b > 1) id_nn = seg_list[pos - 64];
    __syncwarp();
    uint32_t mask = __ballot_sync(0xFFFFFFFFu, keep);
    while (mask) {
      float v[32], v8[kChunk];
      uint32_t ids[kChunk];
      bool any_ok = false;
      float G_[kChunk], al[kChunk], rinv[kChunk], dx_[kChunk], dy_[kChunk];
      float4 Q0[kChunk], Q1[kChunk];
      float cb2[kChunk];
      bool okk[kChunk];
#pragma unroll
      for (int k = 0; k < kChunk; k++) {  // independent per-entry math (ILP)
        const bool valid = mask != 0;
        const int j = valid ? (31 - __clz(mask)) : 0;  // back to front
        mask &= ~(1u << j);
        if (!valid) mask = 0;
        const int epos = b * 32 + j;
        ids[k] = sl.id[j];
        const float4 q0 = sl.q0[j], q1 = sl.q1[j];
        Q0[k] = q0; Q1[k] = q1; cb2[k] = sl.q2[j];
        const float dx = __fsub_rn(q0.x, pfx), dy = __fsub_rn(q0.y, pfy);
        const float power = __fmaf_rn(__fmaf_rn(dx, __fmul_rn(dx, q0.z), __fmul_rn(dy, __fmul_rn(dy, q1.x))), -0.5f,
                                      -__fmul_rn(dy, __fmul_rn(dx, q0.w)));
        const float G = gs_exp(power);
        const float alpha = fminf(0.99f, __fmul_rn(q1.y, G));
        okk[k] = valid && (epos < hi_rel) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        any_ok |= okk[k];
        G_[k] = G; al[k] = alpha; dx_[k] = dx; dy_[k] = dy;
        rinv[k] = __frcp_rn(1.f - alpha);
      }
      if (!__any_sync(0xFFFFFFFFu, any_ok)) continue;
#pragma unroll
      for (int k = 0; k < kChunk; k++) {  // serial part: transmittance / colour recurrence of backward.cu:497-534
        float* vk = v + 8 * k;
        const bool ok = okk[k];
        const float cx = Q0[k].z, cy = Q0[k].w, cz = Q1[k].x, o = Q1[k].y;
        const float c0 = Q1[k].z, c1 = Q1[k].w, c2 = cb2[k];
        const float Tn = T * rinv[k];
        const float a0 = last_alpha * lc0 + (1.f - last_alpha) * acc0;
        const float a1 = last_alpha * lc1 + (1.f - last_alpha) * acc1;
        const float a2 = last_alpha * lc2 + (1.f - last_alpha) * acc2;
        float dsum = (c0 - a0) * dp0 + (c1 - a1) * dp1 + (c2 - a2) * dp2;
        float a3 = 0.f;
        if (OP) {
          a3 = last_alpha * lc3 + (1.f - last_alpha) * acc3;
          dsum += (1.f - a3) * dp3;
        }
        float dL_dalpha = dsum * Tn + nTfbg * rinv[k];
        const float dchannel_dcolor = al[k] * Tn;
        const float dL_dG = o * dL_dalpha;
        const float gdx = G_[k] * dx_[k], gdy = G_[k] * dy_[k];
        const float dG_ddelx = -gdx * cx - gdy * cy;
        const float dG_ddely = -gdy * cz - gdx * cy;
        vk[0] = ok ? dL_dG * dG_ddelx * ddelx_dx : 0.f;
        vk[1] = ok ? dL_dG * dG_ddely * ddely_dy : 0.f;
        vk[2] = ok ? -0.5f * gdx * dx_[k] * dL_dG : 0.f;
        vk[3] = ok ? -0.5f * gdx * dy_[k] * dL_dG : 0.f;
        vk[4] = ok ? -0.5f * gdy * dy_[k] * dL_dG : 0.f;
        vk[5] = ok ? G_[k] * dL_dalpha : 0.f;
        vk[6] = ok ? dchannel_dcolor * dp0 : 0.f;
        vk[7] = ok ? dchannel_dcolor * dp1 : 0.f;
        v8[k] = ok ? dchannel_dcolor * dp2 : 0.f;
        if (ok) {
          T = Tn; acc0 = a0; acc1 = a1; acc2 = a2; lc0 = c0; lc1 = c1; lc2 = c2; last_alpha = al[k];
          if (OP) { acc3 = a3; lc3 = 1.f; }
        }
      }
      // warp reduction BEFORE the atomics: lane L ends with the total of v[L] (entry L>>3, component L&7)
      xpose_step<32, 16>(v, lane);
      xpose_step<16, 8>(v, lane);
      xpose_step<8, 4>(v, lane);
      xpose_step<4, 2>(v, lane);
      xpose_step<2, 1>(v, lane);
      xpose_step<4, 16>(v8, lane);
      xpose_step<2, 8>(v8, lane);
      float r8 = v8[0];
      r8 += __shfl_xor_sync(0xFFFFFFFFu, r8, 4);
      r8 += __shfl_xor_sync(0xFFFFFFFFu, r8, 2);
      r8 += __shfl_xor_sync(0xFFFFFFFFu, r8, 1);
      static_assert(kChunk == 4, "the reduction layout assumes 4 entries x 8 components");
      const int ke = lane >> 3;
      const uint32_t my_id = ke == 0 ? ids[0] : ke == 1 ? ids[1] : ke == 2 ? ids[2] : ids[3];
      const float r = v[0];
      if (r != 0.f) atomicAdd(comp_base + (size_t)comp_stride * my_id, r);
      if (comp == 0 && r8 != 0.f) atomicAdd(dL_dcolors + 3 * (size_t)my_id + 2, r8);
    }
    __syncwarp();
    cur = nxt;
    id_cur = id_next;
    id_next = id_nn;
  }
}

}  // namespace

int gs_launch_blend_fwd(const GsForwardArgs& a, const GsGeom& g, const GsImage& im, const GsList& l, const GsScratch& s,
                        cudaStream_t st) {
  dim3 grid((a.width + GS_TILE - 1) / GS_TILE, (a.height + GS_TILE - 1) / GS_TILE);
  if (s.rows_alpha) {  // GS_FWD_TWO_KERNEL=1: rows + chain variant (the scratch carries its row storage only then)
    const int gx2 = (int)grid.x, tiles = (int)(grid.x * grid.y);
    gs_launch(blend_rows_kernel, dim3((unsigned)l.max_items), dim3(kWarps * 32), 0, st, im.ranges, l.point_list, gx2, im.seg_off,
              tiles, l.items, g.rec, s.rows_alpha, s.rows_data, s.rows_cnt, l.rowmask);
    gs_launch(a.out_opacity ? blend_chain_kernel<true> : blend_chain_kernel<false>, dim3((unsigned)tiles * kWarps), dim3(32), 0, st,
              im.ranges, a.width, a.height, a.background, im.final_T, im.n_contrib, a.out_color, im.seg_off, l.ckpt, im.order, gx2,
              a.out_opacity, l.ckpt_op, (const float*)s.rows_alpha, (const float4*)s.rows_data, (const uint8_t*)s.rows_cnt,
              (uint32_t)l.max_items);
    return gs_check(cudaGetLastError());
  }
#ifndef GS_FWD_SMEM_PAD
#define GS_FWD_SMEM_PAD 0
#endif
  const size_t smem = sizeof(FwdCtl) + kProducers * sizeof(FwdBuf) + GS_FWD_SMEM_PAD;
  const int gx = (int)grid.x;
  grid = dim3(grid.x * grid.y * kWarps);  // one CTA per 8x4 pixel block of a tile, tiles in im.order
  static bool configured = false;
  if (!configured) {
    if (gs_check(cudaFuncSetAttribute(blend_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != GS_OK ||
        gs_check(cudaFuncSetAttribute(blend_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != GS_OK)
      return GS_ERR_CUDA;
    configured = true;
  }
  gs_launch(a.out_opacity ? blend_fwd_kernel<true> : blend_fwd_kernel<false>, grid, dim3(kFwdWarps * 32), smem, st, im.ranges,
            l.point_list, a.width, a.height, g.rec, a.background, im.final_T, im.n_contrib, a.out_color, im.seg_off, l.ckpt,
            im.order, gx, a.out_opacity, l.ckpt_op, l.rowmask);
  return gs_check(cudaGetLastError());
}

int gs_launch_blend_bwd(const GsBackwardArgs& a, const GsGeom& g, const GsImage& im, const GsList& l, cudaStream_t st) {
  const int gx = (a.width + GS_TILE - 1) / GS_TILE, gy = (a.height + GS_TILE - 1) / GS_TILE;
  gs_launch(a.dL_dopacity_pix ? blend_bwd_kernel<true> : blend_bwd_kernel<false>, dim3((unsigned)l.max_items), dim3(kWarps * 32), 0, st,
      im.ranges, l.point_list, a.width, a.height, gx, im.seg_off, gx * gy, l.items, l.ckpt, a.background, g.rec,
      im.final_T, im.n_contrib, a.dL_dpix, a.dL_dmean2D, a.dL_dconic, a.dL_dopacity, a.dL_dcolor,
      (const float*)l.ckpt_op, a.dL_dopacity_pix, (const uint32_t*)l.rowmask);
  return gs_check(cudaGetLastError());
}

#ifdef GS_FWD_TRACE
extern "C" int gs_debug_fwd_trace(long long* host, int n) {
  return gs_check(cudaMemcpyFromSymbol(host, g_fwd_trace, sizeof(long long) * n));
}
#endif
