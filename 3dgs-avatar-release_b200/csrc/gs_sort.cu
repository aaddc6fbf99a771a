// gs_sort.cu -- binning: replaces InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs +
// identifyTileRanges (cuda_rasterizer/rasterizer_impl.cu:70-138, 277-318).
//
// The reference sorts N (Gaussian, tile) pairs by the 64-bit key tile<<32 | depth_bits with a stable LSD radix sort
// (6 x 8-bit passes over 12-byte pairs).  The same final order -- ascending (tile, depth bits, Gaussian index) --
// is produced here with far less traffic and without the host round trip for N:
//   1. the P Gaussians are sorted by their 32-bit depth key (4 onesweep passes over P, not N, items);
//   2. one kernel walks the depth-sorted Gaussians, does the tiles_touched prefix sum as a decoupled-look-back
//      chained scan and emits (tile id, Gaussian index) pairs in that order -- N is produced on the device;
//   3. the pairs are stably sorted by tile id only (ceil(bits/8) = 2 onesweep passes for up to 65536 tiles);
//   4. tile ranges are read off the sorted tile ids.
// Stability of step 3 preserves the depth order of step 1 inside each tile; ties in depth keep ascending Gaussian
// index because step 1 is stable too, which is exactly the tie order of the reference's stable sort.
//
// When the tile grid is small enough for per-warp tile counters in shared memory (<= GS_BIN_MAX_TILES tiles: every avatar
// configuration), steps 2-4 are replaced by DIRECT placement, with no pair sort at all: blocks of 1024 depth-sorted
// Gaussians count their pairs per tile (count kernel), and the place kernel writes every Gaussian index straight to
//   tile_start[tile] + (pairs of earlier blocks) + (pairs of earlier warps) + (rank inside the warp),
// walking the Gaussians of a warp one after the other so that the order inside a tile stays the depth order.
//
// onesweep pass: one kernel per 8-bit digit; warp-level multi-split with __match_any_sync, per-warp digit counters
// in shared memory, CTA-wide digit offsets, decoupled look-back across CTAs on a (flag|count) status word per
// (CTA, digit); digit histograms come for free from the producing kernels (preprocess / emit).
#include <cstdlib>

#include "gs_common.cuh"
namespace {

#define FLAG_LOCAL 0x40000000u
#define FLAG_INCL 0x80000000u
#define VALUE_MASK 0x3FFFFFFFu
#define SPIN_LIMIT (1u << 26)

// Lanes holding the same BITS-bit value (valid lanes only), from BITS ballots.  MATCH.ANY has a long, serialised
// latency; with 32 warps per SM ranking at once (direct placement) the ballots are no slower and do not contend.
template <int BITS>
__device__ __forceinline__ uint32_t match_ballot(uint32_t v, bool valid) {
  uint32_t m = __ballot_sync(0xFFFFFFFFu, valid);
#pragma unroll
  for (int b = 0; b < BITS; b++) {
    const bool bit = (v >> b) & 1u;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, bit);
    m &= bit ? bal : ~bal;
  }
  return m;
}

__device__ __forceinline__ uint32_t ld_volatile(const uint32_t* p) { return *reinterpret_cast<const volatile uint32_t*>(p); }
__device__ __forceinline__ void st_volatile(uint32_t* p, uint32_t v) { *reinterpret_cast<volatile uint32_t*>(p) = v; }

// One stable 8-bit radix pass.  n is read from n_ptr when non-null (clipped to n_cap), else n = n_cap.
// A pass whose digit is the same for every key (e.g. the top byte of avatar depths) degenerates to a straight copy.
// chain_pass >= 0 (depth sort): the pass is link `chain_pass` of a chain over the ping-pong buffers (keys_in, keys_out) /
// (vals_in, vals_out) = (buffer 0, buffer 1); constant-digit passes are skipped WITHOUT a copy, so a pass reads the buffer
// selected by the number of effective passes before it (all histograms, `hist_chain`, exist before the first pass) and the
// consumers locate the result with depth_result_buffer().  chain_pass < 0: fixed buffers, constant digit = plain copy.
template <int ITEMS, bool EARLY>
__global__ void __launch_bounds__(GS_SORT_THREADS)
onesweep_pass_kernel(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                     const unsigned long long* __restrict__ n_ptr, unsigned long long n_cap, int shift,
                     const uint32_t* __restrict__ hist, uint32_t* __restrict__ status, uint32_t* __restrict__ ticket,
                     unsigned int* __restrict__ error_flag, int chain_pass, const uint32_t* __restrict__ hist_chain) {
  // Links 1..3 of the depth chain read only what the kernels BEFORE the chain produced up to their key loads (n, the digit
  // histograms of preprocess, counters zeroed at the start of the call) and their direct predecessor is always link p-1, which
  // has itself waited for its predecessor before letting this grid start: the prologue below (histogram words, ticket,
  // shared-memory clear) runs while link p-1 is still scattering, and the wait moves down to the first key load (EARLY).
  // A template parameter, not a run-time branch: loads through const __restrict__ pointers are invariant loads, which the
  // compiler hoists above a CONDITIONAL wait -- link 0 then read the histograms while preprocess was still adding to them.
  if (!EARLY) GS_PDL_PROLOGUE();
  constexpr int LB = ITEMS <= 8 ? 32 : 8;  // look-back width: predecessors fetched per round trip
  constexpr int WARPS = GS_SORT_THREADS / 32;
  __shared__ uint32_t s_warp_hist[WARPS][256];
  __shared__ uint32_t s_digit_base[256];
  __shared__ uint32_t s_scan[WARPS];
  __shared__ uint32_t s_ticket;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  unsigned long long n = n_cap;
  if (n_ptr) { unsigned long long nn = *n_ptr; n = nn < n_cap ? nn : n_cap; }
  // histogram words first: their latency overlaps the ticket round trip
  const uint32_t h_own = hist[tid];
  uint32_t h_prev[3] = {0u, 0u, 0u};
#pragma unroll
  for (int q = 0; q < 3; q++) if (q < chain_pass) h_prev[q] = hist_chain[q * 256 + tid];
  if (tid == 0) s_ticket = atomicAdd(ticket, 1u);
#pragma unroll
  for (int i = 0; i < WARPS; i++) s_warp_hist[i][tid] = 0;
  __syncthreads();
  if (EARLY) GS_PDL_PROLOGUE();  // before any exit: a grid that returns without waiting would release ITS successor early
  const uint32_t tile = s_ticket;
  const unsigned long long base = (unsigned long long)tile * (ITEMS * GS_SORT_THREADS);
  if (base >= n) return;
  const bool constant = __syncthreads_or((unsigned long long)h_own == n);
  if (chain_pass >= 0) {
    if (constant) return;
    int effective = 0;
#pragma unroll
    for (int q = 0; q < 3; q++)
      if (q < chain_pass) effective += __syncthreads_or((unsigned long long)h_prev[q] == n) ? 0 : 1;
    if (effective & 1) {
      const uint32_t* t1 = keys_in; keys_in = keys_out; keys_out = const_cast<uint32_t*>(t1);
      const uint32_t* t2 = vals_in; vals_in = vals_out; vals_out = const_cast<uint32_t*>(t2);
    }
  }

  // ---- load keys (warp-striped inside a warp-contiguous chunk) ----
  uint32_t key[ITEMS], val[ITEMS], rank[ITEMS];
  const unsigned long long wbase = base + (unsigned long long)w * (32 * ITEMS);
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    unsigned long long idx = wbase + i * 32 + lane;
    key[i] = idx < n ? keys_in[idx] : 0xFFFFFFFFu;
    val[i] = idx < n ? vals_in[idx] : 0u;  // payload loaded up front: its latency hides behind the ranking
  }
  // constant digit: the pass is the identity permutation
  if (constant) {
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      unsigned long long idx = wbase + i * 32 + lane;
      if (idx < n) {
        if (keys_out) keys_out[idx] = key[i];
        vals_out[idx] = val[i];
      }
    }
    return;
  }
  const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    unsigned long long idx = wbase + i * 32 + lane;
    bool valid = idx < n;
    uint32_t digit = (key[i] >> shift) & 255u;
    uint32_t m = __match_any_sync(0xFFFFFFFFu, valid ? digit : 256u);
    int leader = __ffs(m) - 1;
    uint32_t prev = 0;
    if (valid && lane == leader) {
      prev = s_warp_hist[w][digit];
      s_warp_hist[w][digit] = prev + __popc(m);
    }
    prev = __shfl_sync(0xFFFFFFFFu, prev, leader);
    rank[i] = prev + __popc(m & lt_mask);
    __syncwarp();
  }
  __syncthreads();

  // ---- per digit (thread d = digit d): warp-exclusive offsets, CTA count, look-back ----
  const int d = tid;
  uint32_t run = 0;
#pragma unroll
  for (int i = 0; i < WARPS; i++) {
    uint32_t c = s_warp_hist[i][d];
    s_warp_hist[i][d] = run;
    run += c;
  }
  uint32_t* my_status = status + (size_t)tile * 256 + d;
  st_volatile(my_status, (tile == 0 ? FLAG_INCL : FLAG_LOCAL) | run);
  uint32_t excl = 0;
  if (tile > 0) {
    // Decoupled look-back, LB predecessors per round trip: when all CTAs start together nobody has an inclusive prefix
    // yet and a serial walk costs one L2 latency per predecessor (the whole pass at small sizes).
    long long t = (long long)tile - 1;
    uint32_t spins = 0;
    bool finished = false;
    while (!finished && t >= 0) {
      uint32_t sv[LB];
#pragma unroll
      for (int k = 0; k < LB; k++) sv[k] = (t - k >= 0) ? ld_volatile(status + (size_t)(t - k) * 256 + d) : FLAG_INCL;
      int used = LB;
#pragma unroll
      for (int k = 0; k < LB; k++) {
        if (k < used) {
          if (sv[k] == 0) { used = k; }           // not published yet: retry from this predecessor
          else {
            excl += sv[k] & VALUE_MASK;
            if (sv[k] & FLAG_INCL) { finished = true; used = k; }
          }
        }
      }
      if (!finished) {
        t -= used;
        if (used == 0 && ++spins > SPIN_LIMIT) { atomicExch(error_flag, 1u); break; }
      }
    }
    st_volatile(my_status, FLAG_INCL | (excl + run));
  }
  // exclusive scan of the global digit histogram over the 256 digits
  uint32_t h = h_own;
  uint32_t incl = h;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t2 = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += t2;
  }
  if (lane == 31) s_scan[w] = incl;
  __syncthreads();
  uint32_t wpre = 0;
#pragma unroll
  for (int i = 0; i < WARPS; i++) wpre += (i < w) ? s_scan[i] : 0u;
  s_digit_base[d] = wpre + incl - h + excl;
  __syncthreads();

  // ---- scatter ----
#pragma unroll
  for (int i = 0; i < ITEMS; i++) {
    unsigned long long idx = wbase + i * 32 + lane;
    if (idx < n) {
      uint32_t digit = (key[i] >> shift) & 255u;
      uint32_t pos = s_digit_base[digit] + s_warp_hist[w][digit] + rank[i];
      if (keys_out) keys_out[pos] = key[i];
      vals_out[pos] = val[i];
    }
  }
}

// The whole depth chain in ONE launch (small inputs: every CTA of the grid is resident at once).  The passes over non-constant
// digits run back to back inside the kernel, separated by a grid-wide barrier (one atomic arrive + spin per CTA, ~1.5 us)
// instead of a kernel boundary (~3 us of drain + dependent wake-up each, and a whole launch for a skipped digit).  Tile = block
// index: all predecessors of the look-back are running.  Keys and payloads written by another SM one pass ago are read with
// ld.global.cg -- the ping-pong returns to a buffer this SM may still hold in L1 from two passes ago.  Same ranking, same
// look-back, same scatter as onesweep_pass_kernel: identical permutation; the result sits where depth_result_buffer() says.
template <int ITEMS>
__global__ void __launch_bounds__(GS_SORT_THREADS)
depth_sort_fused_kernel(uint32_t* key0, uint32_t* key1, uint32_t* val0, uint32_t* val1, unsigned long long n,
                        const uint32_t* __restrict__ hist_chain, uint32_t* __restrict__ status, size_t status_stride,
                        uint32_t* grid_bar, unsigned int* __restrict__ error_flag) {
  GS_PDL_PROLOGUE();
  constexpr int LB = 32;  // predecessors per look-back round trip (measured: 64 -> binning 52 us; 16 or 8 -> the stage alone 42 us but the whole step +11 us; 32 -> 45 us)
  constexpr int WARPS = GS_SORT_THREADS / 32;
  __shared__ uint32_t s_warp_hist[WARPS][256];
  __shared__ uint32_t s_digit_base[256];
  __shared__ uint32_t s_scan[WARPS];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t tile = blockIdx.x;
  const unsigned long long base = (unsigned long long)tile * (ITEMS * GS_SORT_THREADS);
  const unsigned long long wbase = base + (unsigned long long)w * (32 * ITEMS);
  uint32_t hd[4];
#pragma unroll
  for (int p = 0; p < 4; p++) hd[p] = hist_chain[p * 256 + tid];
  const uint32_t lt_mask = (1u << lane) - 1u;
  int cur = 0;
  uint32_t phase = 0;
#pragma unroll 1
  for (int p = 0; p < 4; p++) {
    const uint32_t h_own = p == 0 ? hd[0] : p == 1 ? hd[1] : p == 2 ? hd[2] : hd[3];
    if (__syncthreads_or((unsigned long long)h_own == n)) continue;  // constant digit: identity, skipped without a copy
    if (phase > 0) {  // grid-wide barrier: the previous effective pass has been scattered by every CTA
      __syncthreads();
      if (tid == 0) {
        __threadfence();
        atomicAdd(grid_bar, 1u);
        uint32_t spins = 0;
        while (ld_volatile(grid_bar) < gridDim.x * phase)
          if (++spins > SPIN_LIMIT) { atomicExch(error_flag, 1u); break; }
        __threadfence();
      }
      __syncthreads();
    }
    phase++;
    const uint32_t* keys_in = cur ? key1 : key0;
    const uint32_t* vals_in = cur ? val1 : val0;
    uint32_t* keys_out = cur ? key0 : key1;
    uint32_t* vals_out = cur ? val0 : val1;
    cur ^= 1;
    if (base >= n) continue;  // (block-uniform) a spare CTA only keeps the barriers company
    const int shift = 8 * p;
#pragma unroll
    for (int i = 0; i < WARPS; i++) s_warp_hist[i][tid] = 0;
    uint32_t key[ITEMS], val[ITEMS], rank[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const unsigned long long idx = wbase + i * 32 + lane;
      key[i] = idx < n ? __ldcg(keys_in + idx) : 0xFFFFFFFFu;
      val[i] = idx < n ? __ldcg(vals_in + idx) : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const unsigned long long idx = wbase + i * 32 + lane;
      const bool valid = idx < n;
      const uint32_t digit = (key[i] >> shift) & 255u;
      const uint32_t m = __match_any_sync(0xFFFFFFFFu, valid ? digit : 256u);
      const int leader = __ffs(m) - 1;
      uint32_t prev = 0;
      if (valid && lane == leader) {
        prev = s_warp_hist[w][digit];
        s_warp_hist[w][digit] = prev + __popc(m);
      }
      prev = __shfl_sync(0xFFFFFFFFu, prev, leader);
      rank[i] = prev + __popc(m & lt_mask);
      __syncwarp();
    }
    __syncthreads();
    const int d = tid;
    uint32_t run = 0;
#pragma unroll
    for (int i = 0; i < WARPS; i++) {
      const uint32_t c = s_warp_hist[i][d];
      s_warp_hist[i][d] = run;
      run += c;
    }
    uint32_t* const st = status + (size_t)p * status_stride;
    uint32_t* const my_status = st + (size_t)tile * 256 + d;
    st_volatile(my_status, (tile == 0 ? FLAG_INCL : FLAG_LOCAL) | run);
    uint32_t excl = 0;
    if (tile > 0) {
      long long t = (long long)tile - 1;
      uint32_t spins = 0;
      bool finished = false;
      while (!finished && t >= 0) {
        uint32_t sv[LB];
#pragma unroll
        for (int k = 0; k < LB; k++) sv[k] = (t - k >= 0) ? ld_volatile(st + (size_t)(t - k) * 256 + d) : FLAG_INCL;
        int used = LB;
#pragma unroll
        for (int k = 0; k < LB; k++) {
          if (k < used) {
            if (sv[k] == 0) { used = k; }
            else {
              excl += sv[k] & VALUE_MASK;
              if (sv[k] & FLAG_INCL) { finished = true; used = k; }
            }
          }
        }
        if (!finished) {
          t -= used;
          if (used == 0 && ++spins > SPIN_LIMIT) { atomicExch(error_flag, 1u); break; }
        }
      }
      st_volatile(my_status, FLAG_INCL | (excl + run));
    }
    uint32_t incl = h_own;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t2 = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += t2;
    }
    if (lane == 31) s_scan[w] = incl;
    __syncthreads();
    uint32_t wpre = 0;
#pragma unroll
    for (int i = 0; i < WARPS; i++) wpre += (i < w) ? s_scan[i] : 0u;
    s_digit_base[d] = wpre + incl - h_own + excl;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const unsigned long long idx = wbase + i * 32 + lane;
      if (idx < n) {
        const uint32_t digit = (key[i] >> shift) & 255u;
        const uint32_t pos = s_digit_base[digit] + s_warp_hist[w][digit] + rank[i];
        keys_out[pos] = key[i];
        vals_out[pos] = val[i];
      }
    }
    __syncthreads();  // s_warp_hist / s_digit_base are rewritten by the next pass
  }
}

// Index (0/1) of the depth ping-pong buffer that holds the sorted result: the number of effective (non-constant) passes,
// mod 2.  Block-wide call (>= 256 threads, all threads).
__device__ __forceinline__ int depth_result_buffer(const uint32_t* __restrict__ hist_chain, unsigned long long n, int tid) {
  int effective = 0;
  for (int q = 0; q < 4; q++)
    effective += __syncthreads_or(tid < 256 && (unsigned long long)hist_chain[q * 256 + tid] == n) ? 0 : 1;
  return effective & 1;
}

// ---- direct placement (small tile grids) ----
// A warp owns 32 depth-consecutive Gaussians.  Their (Gaussian, tile) pairs, in (Gaussian, row-major tile) order, are
// handled 32 at a time, one pair per lane: the lane finds its Gaussian from the start positions of the rectangles inside
// the 32-pair window (one warp-wide OR + popc), equal tiles inside the window are ranked with __match_any_sync, and the
// first lane of every tile bumps this warp's per-tile counter (plain shared-memory read-modify-write).
// COUNT: the counters end as the warp's pairs per tile.  PLACE: the counters start as the warp's exclusive offset inside
// the block and every pair gets its slot.
struct BinWarp {
  uint32_t total;   // pairs of this warp
  uint32_t excl;    // this lane: first pair index of its Gaussian
  bool nz;          // this lane's Gaussian touches tiles
};

__device__ __forceinline__ BinWarp bin_load(int P, int i, const uint32_t* __restrict__ sorted_gid,
                                            const uint32_t* __restrict__ tiles_touched, const uint32_t* __restrict__ rect,
                                            uint4* __restrict__ stage_w, int lane) {
  uint32_t gid = 0, cnt = 0;
  uint2 r = make_uint2(0u, 0u);
  if (i < P) {
    gid = __ldcg(sorted_gid + i);  // NOT an invariant load: it must stay behind the caller's griddepcontrol.wait
    cnt = tiles_touched[gid];
    r = *reinterpret_cast<const uint2*>(rect + 2 * (size_t)gid);  // unconditional (in bounds; unused when cnt == 0): both loads in one round trip
  }
  uint32_t incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += v;
  }
  BinWarp bw;
  bw.total = __shfl_sync(0xFFFFFFFFu, incl, 31);
  bw.excl = incl - cnt;
  bw.nz = cnt != 0;
  const uint32_t todo = __ballot_sync(0xFFFFFFFFu, bw.nz);
  if (bw.nz) stage_w[__popc(todo & ((1u << lane) - 1u))] = make_uint4(bw.excl, gid, r.x, r.y);  // compacted
  __syncwarp();
  return bw;
}

template <bool PLACE>
__device__ __forceinline__ void bin_warp_walk(const BinWarp bw, const uint4* __restrict__ stage_w, int grid_x,
                                              uint16_t* __restrict__ s_off_w, const uint32_t* __restrict__ s_base,
                                              uint32_t* __restrict__ list, unsigned long long cap, int lane) {
  const uint32_t lt_mask = (1u << lane) - 1u;
  for (uint32_t B = 0; B < bw.total; B += 32) {
    const uint32_t before = __popc(__ballot_sync(0xFFFFFFFFu, bw.nz && bw.excl < B));
    const uint32_t bits = __reduce_or_sync(0xFFFFFFFFu, (bw.nz && bw.excl >= B && bw.excl < B + 32) ? (1u << (bw.excl - B)) : 0u);
    const uint32_t k = B + lane;
    const bool valid = k < bw.total;
    uint32_t t = 0xFFFFFFFFu, gg = 0;
    if (valid) {
      const uint4 e = stage_w[before + __popc(bits & ((2u << lane) - 1u)) - 1];
      const int x0 = e.z & 0xFFFF, y0 = e.z >> 16, wdt = (int)(e.w & 0xFFFF) - x0, j = (int)(k - e.x);
      const int yy = __float2int_rz(__fdividef((float)j + 0.5f, (float)wdt));  // exact for these small integers
      t = (uint32_t)((y0 + yy) * grid_x + x0 + (j - yy * wdt));
      gg = e.y;
    }
    const uint32_t m = match_ballot<GS_BIN_TILE_BITS>(t, valid);
    const int leader = __ffs(m) - 1;
    uint32_t o = 0;
    if (valid && lane == leader) {
      o = s_off_w[t];
      s_off_w[t] = (uint16_t)(o + __popc(m));
    }
    if (PLACE) {
      o = __shfl_sync(0xFFFFFFFFu, o, leader) + __popc(m & lt_mask);
      if (valid) {
        const unsigned long long pos = (unsigned long long)s_base[t] + o;
        if (pos < cap) list[pos] = gg;
      }
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(GS_BIN_THREADS)
bin_count_kernel(int P, const uint32_t* __restrict__ dval0, const uint32_t* __restrict__ dval1,
                 const uint32_t* __restrict__ hist_chain, const uint32_t* __restrict__ tiles_touched,
                 const uint32_t* __restrict__ rect, int grid_x, int tiles, uint32_t* __restrict__ blockcount,
                 uint32_t* __restrict__ tile_total) {
  GS_PDL_PROLOGUE();
  extern __shared__ __align__(16) unsigned char s_raw[];
  uint16_t* s_off = reinterpret_cast<uint16_t*>(s_raw);  // [32][tiles]
  __shared__ uint4 s_stage[32][32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  for (int k = tid; k < 16 * tiles; k += GS_BIN_THREADS) reinterpret_cast<uint32_t*>(s_off)[k] = 0;
  const uint32_t* sorted_gid = depth_result_buffer(hist_chain, (unsigned long long)P, tid) ? dval1 : dval0;  // + barrier
  const BinWarp bw = bin_load(P, blockIdx.x * GS_BIN_THREADS + tid, sorted_gid, tiles_touched, rect, s_stage[w], lane);
  bin_warp_walk<false>(bw, s_stage[w], grid_x, s_off + (size_t)w * tiles, nullptr, nullptr, 0ull, lane);
  __syncthreads();
  for (int t = tid; t < tiles; t += GS_BIN_THREADS) {
    uint32_t sum = 0;
#pragma unroll 8
    for (int k = 0; k < 32; k++) sum += s_off[(size_t)k * tiles + t];
    blockcount[(size_t)blockIdx.x * tiles + t] = sum;
    if (sum) atomicAdd(&tile_total[t], sum);
  }
}

// Work plans of the two blend kernels.  One CTA.
//  * backward: every tile list is cut into segments of GS_SEG entries; seg_off = exclusive prefix of the segment counts over
//    tiles (seg_off[tiles] = number of items), items[] = (tile, segment);
//  * forward: every non-empty tile list is cut into chunks of GS_FWD_CHUNK entries; fitems[] = (tile, chunk, index of the
//    successor item, index of the predecessor item), ordered by the number of chunks that REMAIN in the tile's chain
//    (this one included), longest first -- "longest remaining path first": the deep tiles, whose chains are the critical path
//    of avatar frames, start at once, every last chunk comes at the end, and an item's predecessor (one more remaining)
//    always has a smaller index: CTAs are dispatched in index order, which makes the hand-off wait of blend_fwd_kernel
//    deadlock-free.  With nch(t) = chunks of tile t, pos(t) = rank of t in descending nch order (counting sort),
//    ge[k] = #tiles with nch > k and base = exclusive prefix of ge, item (t, c) with r = nch - c remaining sits at
//    n_fitems - base[r] + pos(t).  The empty tiles follow as one item each (their pixels still get the background).
__device__ __forceinline__ uint32_t block_scan_incl_1024(uint32_t v, uint32_t* s_w, uint32_t* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += t;
  }
  __syncthreads();  // s_w may still be read from the previous scan
  if (lane == 31) s_w[w] = incl;
  __syncthreads();
  uint32_t wv = s_w[lane], wincl = wv;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xFFFFFFFFu, wincl, o);
    if (lane >= o) wincl += t;
  }
  *total = __shfl_sync(0xFFFFFFFFu, wincl, 31);
  return incl + __shfl_sync(0xFFFFFFFFu, wincl - wv, w);
}

// `ranges` may live in global memory (plan_kernel) or in shared memory (the plan block of bin_place_kernel).  1024 threads.
__device__ __forceinline__ void plan_body(int tiles, const uint2* ranges, uint32_t* __restrict__ seg_off, uint2* __restrict__ items,
                                          uint32_t max_items, uint4* __restrict__ fitems, uint32_t max_fitems,
                                          GsHeader* __restrict__ hdr, int fwd_order, uint32_t* __restrict__ order) {
  __shared__ uint32_t s_w[32];
  __shared__ uint32_t s_hist[GS_FWD_NCH_BINS];  // tiles per chunk count -> cursor of the counting sort
  __shared__ uint32_t s_base[GS_FWD_NCH_BINS];  // base[r] = sum over k < r of ge[k]
  __shared__ uint32_t s_total, s_nonempty, s_maxch;
  constexpr int MIX_D = 32;
  __shared__ uint32_t s_mix_first[2 * MIX_D], s_mix_off[2 * MIX_D][MIX_D + 1], s_mix_ge[MIX_D + 1];
  const int tid = threadIdx.x;
  for (int k = tid; k < GS_FWD_NCH_BINS; k += 1024) s_hist[k] = 0;
  if (tid == 0) s_maxch = 0;
  __syncthreads();
  uint32_t carry = 0, rcarry = 0;
  uint32_t roff_l[64];  // this thread's tiles (t = t0 + tid in both loops): first hand-off slot of the tile; 65535 tiles at most
  for (int t0 = 0; t0 < tiles; t0 += 1024) {
    const int t = t0 + tid;
    uint32_t nseg = 0, nrecv = 0;
    if (t < tiles) {
      const uint2 r = ranges[t];
      const uint32_t len = r.y - r.x;
      nseg = (len + GS_SEG - 1) / GS_SEG;
      const uint32_t nch = min((len + GS_FWD_CHUNK - 1) / GS_FWD_CHUNK, (uint32_t)GS_FWD_NCH_BINS - 1);
      atomicAdd(&s_hist[nch], 1u);
      atomicMax(&s_maxch, nch);
      nrecv = nch ? nch - 1 : 0;  // chunks that receive a chain state
    }
    uint32_t total, rtotal;
    const uint32_t off = carry + block_scan_incl_1024(nseg, s_w, &total) - nseg;
    roff_l[(t0 >> 10) & 63] = rcarry + block_scan_incl_1024(nrecv, s_w, &rtotal) - nrecv;
    if (t < tiles) {
      seg_off[t] = off;
      for (uint32_t k = 0; k < nseg; k++)
        if (off + k < max_items) items[off + k] = make_uint2((uint32_t)t, k);
    }
    carry += total;
    rcarry += rtotal;
  }
  if (tid == 0) seg_off[tiles] = carry < max_items ? carry : max_items;
  __syncthreads();
  // ge[k] = #tiles with chunk count > k = tiles - inclusive prefix of hist; base = exclusive prefix of ge.
  // Two elements per thread (GS_FWD_NCH_BINS = 2 * 1024).
  static_assert(GS_FWD_NCH_BINS == 2048, "two bins per thread");
  {
    const uint32_t h0 = s_hist[2 * tid], h1 = s_hist[2 * tid + 1];
    uint32_t total;
    const uint32_t incl = block_scan_incl_1024(h0 + h1, s_w, &total);
    const uint32_t ge0 = (uint32_t)tiles - (incl - h1), ge1 = (uint32_t)tiles - incl;
    uint32_t total2;
    const uint32_t incl2 = block_scan_incl_1024(ge0 + ge1, s_w, &total2);
    __syncthreads();
    s_hist[2 * tid] = ge0; s_hist[2 * tid + 1] = ge1;                          // cursor: tiles with a larger chunk count come first
    s_base[2 * tid] = incl2 - ge0 - ge1; s_base[2 * tid + 1] = incl2 - ge1;
    if (tid == 0) {
      s_total = total2;                                                        // items of the non-empty tiles
      hdr->n_fitems = total2 < max_fitems ? total2 : max_fitems;
      s_nonempty = ge0;                                                        // ge[0]
      hdr->n_empty_tiles = h0;                                                 // hist[0]
      hdr->n_nonempty_tiles = ge0;
    }
  }
  __syncthreads();
  const int D = (int)s_maxch;
  if (fwd_order >= 2 && D > MIX_D) fwd_order = 0;  // very deep lists (> 8192 entries in a tile): the closed-form order
  if (tid == 0) hdr->fwd_order = (unsigned)fwd_order;
  // mixed order (fwd_order 2): bucket K = 2 c + (D - nch) holds chunk c of the tiles with nch chunks -- half-way between level
  // order (K = c) and longest-remaining-first (K = c + D - nch): the shallow tiles' chunks fill the machine while the deep chains
  // run, instead of all arriving behind them.  s_mix_off[K][n] = items of bucket K that belong to tiles deeper than n (deeper
  // tiles come first inside a bucket; [K][0] = the whole bucket), s_mix_first[K] = first item of the bucket.
  if (fwd_order == 2) {
    if (tid < 2 * D) {  // one thread per bucket: suffix sums over the chunk counts
      const int K = tid;
      uint32_t o = 0;
      for (int n = D; n >= 0; n--) {
        s_mix_off[K][n] = o;
        const int e = K - D + n;  // chunk c = e / 2 of the tiles with n chunks lies in this bucket if it is an integer in [0, n)
        if (n >= 1 && e >= 0 && !(e & 1) && (e >> 1) < n) o += s_hist[n - 1] - s_hist[n];  // hist[n] = ge[n-1] - ge[n]
      }
    }
    if (tid <= MIX_D) s_mix_ge[tid] = s_hist[tid];  // ge[n] before the cursors below move
    __syncthreads();
    if (tid == 0) {
      uint32_t run = 0;
      for (int K = 0; K < 2 * D; K++) { s_mix_first[K] = run; run += s_mix_off[K][0]; }
    }
    __syncthreads();
  }
  const uint32_t n_items = s_total, n_nonempty = s_nonempty;
  for (int t = tid; t < tiles; t += 1024) {
    const uint2 r = ranges[t];
    const uint32_t len = r.y - r.x;
    const uint32_t nch = min((len + GS_FWD_CHUNK - 1) / GS_FWD_CHUNK, (uint32_t)GS_FWD_NCH_BINS - 1);
    const uint32_t pos = atomicAdd(&s_hist[nch], 1u);  // rank of the tile by descending chunk count (empty tiles last)
    order[pos] = (uint32_t)t;
    const uint32_t roff = roff_l[(t >> 10) & 63];
    if (nch == 0) {  // empty tile: one item behind everything else
      const uint32_t fi = n_items + (pos - n_nonempty);
      if (fi < max_fitems) fitems[fi] = make_uint4((uint32_t)t, 0u, 0xFFFFFFFFu, 0xFFFFFFFFu);
      continue;
    }
    // item = (tile, chunk, hand-off slot it SENDS its chain state to or ~0, slot it RECEIVES from or ~0); a predecessor always
    // has a smaller index in every order, which is what makes the forward's wait deadlock-free
    for (uint32_t c = 0; c < nch; c++) {
      const uint32_t rem = nch - c;  // chunks remaining in this tile's chain, this one included
      uint32_t fi;
      if (fwd_order == 0) fi = n_items - s_base[rem] + pos;  // longest remaining chain first
      else if (fwd_order == 1) fi = s_base[c] + pos;         // level by level, deepest tiles first inside a level
      else {  // mixed
        const int K = 2 * (int)c + D - (int)nch;
        fi = s_mix_first[K] + s_mix_off[K][nch] + (pos - s_mix_ge[nch]);  // deeper tiles first, then the rank among equals
      }
      if (fi < max_fitems) fitems[fi] = make_uint4((uint32_t)t, c, rem > 1 ? roff + c : 0xFFFFFFFFu, c > 0 ? roff + c - 1 : 0xFFFFFFFFu);
    }
  }
}

__global__ void __launch_bounds__(1024)
plan_kernel(int tiles, const uint2* __restrict__ ranges, uint32_t* __restrict__ seg_off, uint2* __restrict__ items,
            uint32_t max_items, uint4* __restrict__ fitems, uint32_t max_fitems, GsHeader* __restrict__ hdr, int fwd_order,
            uint32_t* __restrict__ order) {
  GS_PDL_PROLOGUE();
  plan_body(tiles, ranges, seg_off, items, max_items, fitems, max_fitems, hdr, fwd_order, order);
}

// tiles <= 2 * GS_BIN_THREADS.  Block 0 also publishes the tile ranges (identifyTileRanges, rasterizer_impl.cu:116-138:
// empty tiles stay (0,0)) and the pair count.  FUSED: the grid is small enough to be co-resident (the launcher checks), the
// kernel does the counting itself and crosses a grid-wide barrier instead of a kernel boundary; otherwise
// bin_count_kernel has run before.
template <bool FUSED, bool PRE>
__global__ void __launch_bounds__(GS_BIN_THREADS)
bin_place_kernel(int P, const uint32_t* __restrict__ dval0, const uint32_t* __restrict__ dval1,
                 const uint32_t* __restrict__ hist_chain, const uint32_t* __restrict__ tiles_touched,
                 const uint32_t* __restrict__ rect, int grid_x, int tiles, uint32_t* blockcount, uint32_t* tile_total,
                 unsigned long long cap, uint32_t* __restrict__ list, uint2* __restrict__ ranges, GsHeader* __restrict__ hdr,
                 uint32_t* grid_bar, uint32_t* __restrict__ seg_off, uint2* __restrict__ items, uint32_t max_items,
                 uint4* __restrict__ fitems, uint32_t max_fitems, int fwd_order, uint32_t* __restrict__ order) {
  if (!PRE) GS_PDL_PROLOGUE();
  extern __shared__ __align__(16) unsigned char s_raw[];
  uint16_t* s_off = reinterpret_cast<uint16_t*>(s_raw);                                  // [32][tiles]
  uint32_t* s_base = reinterpret_cast<uint32_t*>(s_raw + (size_t)64 * tiles);            // [tiles]
  __shared__ uint32_t s_wsum[32];
  __shared__ uint4 s_stage[32][32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  // FUSED: the LAST block of the grid places nothing.  It waits for the counts like everybody else, derives the tile ranges in
  // shared memory and builds the work plans of the two blend kernels (plan_body) WHILE the other blocks place the pairs: the
  // separate plan_kernel launch (8 us behind the placement) disappears from the step.
  const bool plan_block = FUSED && blockIdx.x == gridDim.x - 1;
  const unsigned nblocks = FUSED ? gridDim.x - 1 : gridDim.x;
  BinWarp bw;
  bw.total = 0; bw.excl = 0; bw.nz = false;
  // Before the wait: the counter clear and the choice of the sorted buffer (digit histograms of preprocess; the direct
  // predecessor of this grid is always the last link of the depth chain, which waited for everything before it).
  const uint32_t* sorted_gid = nullptr;
  if (!plan_block) {  // block-uniform
    for (int k = tid; k < 16 * tiles; k += GS_BIN_THREADS) reinterpret_cast<uint32_t*>(s_off)[k] = 0;
    sorted_gid = depth_result_buffer(hist_chain, (unsigned long long)P, tid) ? dval1 : dval0;  // + barrier
  }
  if (PRE) GS_PDL_PROLOGUE();
  if (!plan_block) {
    bw = bin_load(P, blockIdx.x * GS_BIN_THREADS + tid, sorted_gid, tiles_touched, rect, s_stage[w], lane);
    bin_warp_walk<false>(bw, s_stage[w], grid_x, s_off + (size_t)w * tiles, nullptr, nullptr, 0ull, lane);
  }
  if (FUSED) {
    __syncthreads();
    if (!plan_block)
      for (int t = tid; t < tiles; t += GS_BIN_THREADS) {
        uint32_t sum = 0;
#pragma unroll 8
        for (int k = 0; k < 32; k++) sum += s_off[(size_t)k * tiles + t];
        blockcount[(size_t)blockIdx.x * tiles + t] = sum;
        if (sum) atomicAdd(&tile_total[t], sum);
      }
    __syncthreads();
    if (tid == 0) {  // grid-wide barrier over the counting blocks (the plan block only waits)
      if (!plan_block) {
        __threadfence();
        atomicAdd(grid_bar, 1u);
      }
      uint32_t spins = 0;
      while (ld_volatile(grid_bar) < nblocks)
        if (++spins > SPIN_LIMIT) { atomicExch(&hdr->error, 3u); break; }
      __threadfence();
    }
    __syncthreads();
  }

  // tile starts: exclusive scan of the totals over tiles, two consecutive tiles per thread
  const int ta = 2 * tid, tb = 2 * tid + 1;
  const uint32_t ca = ta < tiles ? __ldcg(tile_total + ta) : 0u, cb = tb < tiles ? __ldcg(tile_total + tb) : 0u;
  uint32_t pa = 0, pb = 0;  // pairs of the earlier blocks
  const unsigned earlier = plan_block ? 0u : blockIdx.x;
  for (unsigned b0 = 0; b0 < earlier; b0 += 16) {  // 32 independent loads in flight per round trip
    uint32_t va[16], vb[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
      const bool in = b0 + k < earlier;
      va[k] = (in && ta < tiles) ? __ldcg(blockcount + (size_t)(b0 + k) * tiles + ta) : 0u;
      vb[k] = (in && tb < tiles) ? __ldcg(blockcount + (size_t)(b0 + k) * tiles + tb) : 0u;
    }
#pragma unroll
    for (int k = 0; k < 16; k++) { pa += va[k]; pb += vb[k]; }
  }
  uint32_t incl = ca + cb;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_wsum[w] = incl;
  __syncthreads();  // also: every warp finished counting
  uint32_t wpre = 0, total = 0;
#pragma unroll
  for (int k = 0; k < 32; k++) { const uint32_t v = s_wsum[k]; if (k < w) wpre += v; total += v; }
  const uint32_t sa = wpre + incl - (ca + cb), sb = sa + ca;
  if (plan_block) {  // the ranges block 0 publishes, kept in shared memory (the count area is free here), then the plans
    const uint32_t lim = (unsigned long long)total > cap ? (uint32_t)cap : total;
    uint2* const s_range = reinterpret_cast<uint2*>(s_raw);  // [tiles] <= 64 * tiles bytes
    if (ta < tiles) s_range[ta] = ca ? make_uint2(min(sa, lim), min(sa + ca, lim)) : make_uint2(0u, 0u);
    if (tb < tiles) s_range[tb] = cb ? make_uint2(min(sb, lim), min(sb + cb, lim)) : make_uint2(0u, 0u);
    __syncthreads();
    plan_body(tiles, s_range, seg_off, items, max_items, fitems, max_fitems, hdr, fwd_order, order);
    return;
  }
  if (ta < tiles) s_base[ta] = sa + pa;
  if (tb < tiles) s_base[tb] = sb + pb;
  if (blockIdx.x == 0) {
    const uint32_t lim = (unsigned long long)total > cap ? (uint32_t)cap : total;
    if (ta < tiles) ranges[ta] = ca ? make_uint2(min(sa, lim), min(sa + ca, lim)) : make_uint2(0u, 0u);
    if (tb < tiles) ranges[tb] = cb ? make_uint2(min(sb, lim), min(sb + cb, lim)) : make_uint2(0u, 0u);
    if (tid == 0) {
      hdr->num_rendered = total;
      hdr->overflow = (unsigned long long)total > cap ? 1u : 0u;
    }
  }
  // per-warp counts -> exclusive offsets over the warps of this block
  for (int t = tid; t < tiles; t += GS_BIN_THREADS) {
    uint32_t run = 0;
#pragma unroll 8
    for (int k = 0; k < 32; k++) {
      const uint32_t c = s_off[(size_t)k * tiles + t];
      s_off[(size_t)k * tiles + t] = (uint16_t)run;
      run += c;
    }
  }
  __syncthreads();
  bin_warp_walk<true>(bw, s_stage[w], grid_x, s_off + (size_t)w * tiles, s_base, list, cap, lane);
}

// Chained scan of tiles_touched over the depth-sorted Gaussians + pair emission (duplicateWithKeys,
// rasterizer_impl.cu:70-111, without the 64-bit key: the depth order is already in the traversal order).
__global__ void __launch_bounds__(GS_EMIT_THREADS)
emit_kernel(int P, const uint32_t* __restrict__ dval0, const uint32_t* __restrict__ dval1,
            const uint32_t* __restrict__ hist_chain, const uint32_t* __restrict__ tiles_touched,
            const uint32_t* __restrict__ rect, int grid_x, unsigned long long cap, uint32_t* __restrict__ tkey,
            uint32_t* __restrict__ tval, uint32_t* __restrict__ hist /* [4][256] tile-id digits */, int tile_passes,
            unsigned long long* __restrict__ scan_status, uint32_t* __restrict__ ticket, GsHeader* __restrict__ hdr) {
  GS_PDL_PROLOGUE();
  constexpr int WARPS = GS_EMIT_THREADS / 32;
  __shared__ uint32_t s_hist[4 * 256];
  __shared__ unsigned long long s_wsum[WARPS];
  __shared__ unsigned long long s_base;
  __shared__ uint32_t s_ticket;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (tid == 0) s_ticket = atomicAdd(ticket, 1u);
  for (int i = tid; i < tile_passes * 256; i += GS_EMIT_THREADS) s_hist[i] = 0;
  __syncthreads();
  const uint32_t blk = s_ticket;
  const uint32_t* sorted_gid = depth_result_buffer(hist_chain, (unsigned long long)P, tid) ? dval1 : dval0;
  const int nblocks = (P + GS_EMIT_THREADS - 1) / GS_EMIT_THREADS;
  const int i = blk * GS_EMIT_THREADS + tid;
  uint32_t gid = 0, cnt = 0;
  if (i < P) { gid = sorted_gid[i]; cnt = tiles_touched[gid]; }
  // block-exclusive scan of cnt
  uint32_t incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t2 = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += t2;
  }
  if (lane == 31) s_wsum[w] = incl;
  __syncthreads();
  unsigned long long wpre = 0, total = 0;
#pragma unroll
  for (int k = 0; k < WARPS; k++) { unsigned long long v = s_wsum[k]; if (k < w) wpre += v; total += v; }
  if (w == 0) {  // warp-parallel decoupled look-back: 32 predecessors per step
    const unsigned long long F_LOCAL = 1ull << 62, F_INCL = 2ull << 62, VMASK = (1ull << 62) - 1;
    volatile unsigned long long* st = scan_status;
    if (lane == 0) st[blk] = (blk == 0 ? F_INCL : F_LOCAL) | total;
    unsigned long long excl = 0;
    if (blk > 0) {
      long long top = (long long)blk - 1;
      uint32_t spins = 0;
      while (true) {
        const long long idx = top - lane;
        const unsigned long long sv = idx >= 0 ? st[idx] : F_INCL;  // virtual inclusive 0 in front of block 0
        const uint32_t ready = __ballot_sync(0xFFFFFFFFu, sv != 0);
        const uint32_t incl_m = __ballot_sync(0xFFFFFFFFu, (sv & F_INCL) != 0);
        const int fi = incl_m ? (__ffs(incl_m) - 1) : 32;  // nearest predecessor with an inclusive prefix
        const uint32_t need = fi < 32 ? ((2u << fi) - 1u) : 0xFFFFFFFFu;
        if ((ready & need) != need) {
          if (++spins > SPIN_LIMIT) { if (lane == 0) atomicExch(&hdr->error, 2u); break; }
          continue;
        }
        unsigned long long v = (lane <= fi) ? (sv & VMASK) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
        excl += v;
        if (fi < 32) break;
        top -= 32;
      }
      if (lane == 0) st[blk] = F_INCL | (excl + total);
    }
    if (lane == 0) {
      s_base = excl;
      if ((int)blk == nblocks - 1) {
        hdr->num_rendered = excl + total;
        hdr->overflow = (excl + total) > cap ? 1u : 0u;
      }
    }
  }
  __syncthreads();
  if (cnt) {
    unsigned long long off = s_base + wpre + (incl - cnt);
    uint32_t r0 = rect[2 * (size_t)gid], r1 = rect[2 * (size_t)gid + 1];
    int x0 = r0 & 0xFFFF, y0 = r0 >> 16, x1 = r1 & 0xFFFF, y1 = r1 >> 16;
    for (int y = y0; y < y1; y++)
      for (int x = x0; x < x1; x++) {
        uint32_t t = (uint32_t)(y * grid_x + x);
        if (off < cap) {
          tkey[off] = t;
          tval[off] = gid;
          for (int p = 0; p < tile_passes; p++) atomicAdd(&s_hist[p * 256 + ((t >> (8 * p)) & 255u)], 1u);
        }
        off++;
      }
  }
  __syncthreads();
  for (int k = tid; k < tile_passes * 256; k += GS_EMIT_THREADS) {
    uint32_t c = s_hist[k];
    if (c) atomicAdd(&hist[k], c);
  }
}

// identifyTileRanges, rasterizer_impl.cu:116-138 (ranges zeroed beforehand, :310)
__global__ void tile_ranges_kernel(const uint32_t* __restrict__ tkey_sorted, const unsigned long long* __restrict__ n_ptr,
                                   unsigned long long cap, uint2* __restrict__ ranges) {
  GS_PDL_PROLOGUE();
  unsigned long long n = *n_ptr;
  if (n > cap) n = cap;
  for (unsigned long long idx = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; idx < n;
       idx += (unsigned long long)gridDim.x * blockDim.x) {
    uint32_t cur = tkey_sorted[idx];
    if (idx == 0) ranges[cur].x = 0;
    else {
      uint32_t prev = tkey_sorted[idx - 1];
      if (cur != prev) { ranges[prev].y = (uint32_t)idx; ranges[cur].x = (uint32_t)idx; }
    }
    if (idx == n - 1) ranges[cur].y = (uint32_t)n;
  }
}

uint32_t higher_msb(uint32_t n) {  // rasterizer_impl.cu:35-50
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

}  // namespace

int gs_launch_binning(const GsForwardArgs& a, const GsGeom& g, const GsImage& im, const GsScratch& s, const GsList& l,
                      int64_t cap, cudaStream_t st) {
  uint32_t* list = l.point_list;
  // plan_body's order of the forward work items: 2 (default) mixed, 0 longest remaining chain first, 1 level by level
  static const int fwd_order = [] { const char* e = getenv("GS_FWD_ORDER"); return e ? atoi(e) : 2; }();
  const int P = a.P;
  const int gx = (a.width + GS_TILE - 1) / GS_TILE, gy = (a.height + GS_TILE - 1) / GS_TILE;
  const int tiles = gx * gy;
  // 1. depth sort of the P Gaussians: up to 4 onesweep passes over the dkey/dval ping-pong; constant-digit passes are
  //    skipped and the consumers find the result with depth_result_buffer().  (A single-cluster sort -- 8 CTAs, digit
  //    counts over distributed shared memory, cluster barriers instead of kernel boundaries -- was measured at 40-58 us
  //    for 50k keys against 26 us for these passes: 8 SMs at ~1 IPC lose to a latency chain spread over 49 SMs.)
  const int blocks_d = (int)gs_sort_blocks((unsigned long long)P);
  const bool small_d = gs_sort_items((unsigned long long)P) == GS_SORT_ITEMS_SMALL;
  const int prewait = gs_prewait_mask();
  // one launch when every CTA of the chain can be resident at once (GS_SORT_FUSED=0: the four-launch chain)
  static const bool sort_fused = [] { const char* e = getenv("GS_SORT_FUSED"); return e ? atoi(e) != 0 : true; }();
  const int blocks_f = (P + GS_SORT_FUSED_ITEMS * GS_SORT_THREADS - 1) / (GS_SORT_FUSED_ITEMS * GS_SORT_THREADS);
  if (sort_fused && small_d && blocks_f <= GS_SORT_FUSED_MAX_BLOCKS) {  // shorter tiles while the grid stays co-resident
    gs_launch(depth_sort_fused_kernel<GS_SORT_FUSED_ITEMS>, dim3(blocks_f), dim3(GS_SORT_THREADS), 0, st, s.dkey[0], s.dkey[1], s.dval[0],
              s.dval[1], (unsigned long long)P, (const uint32_t*)s.hist, s.status_d, (size_t)s.blocks_d * 256, s.counters + 12,
              &g.hdr->error);
  } else if (sort_fused && small_d && blocks_d <= GS_SORT_FUSED_MAX_BLOCKS) {
    gs_launch(depth_sort_fused_kernel<GS_SORT_ITEMS_SMALL>, dim3(blocks_d), dim3(GS_SORT_THREADS), 0, st, s.dkey[0], s.dkey[1], s.dval[0],
              s.dval[1], (unsigned long long)P, (const uint32_t*)s.hist, s.status_d, (size_t)s.blocks_d * 256, s.counters + 12,
              &g.hdr->error);
  } else
  for (int p = 0; p < 4; p++) {
    const bool early = (prewait & 1) && p >= 1;  // link 0 follows preprocess, the producer of everything it reads
    auto kern = small_d ? (early ? onesweep_pass_kernel<GS_SORT_ITEMS_SMALL, true> : onesweep_pass_kernel<GS_SORT_ITEMS_SMALL, false>)
                        : (early ? onesweep_pass_kernel<GS_SORT_ITEMS, true> : onesweep_pass_kernel<GS_SORT_ITEMS, false>);
    gs_launch(kern, dim3(blocks_d), dim3(GS_SORT_THREADS), 0, st,
        (const uint32_t*)s.dkey[0], s.dkey[1], (const uint32_t*)s.dval[0], s.dval[1], (const unsigned long long*)nullptr,
        (unsigned long long)P, 8 * p, (const uint32_t*)(s.hist + p * 256), s.status_d + (size_t)p * s.blocks_d * 256,
        s.counters + p, &g.hdr->error, p, (const uint32_t*)s.hist);
  }
  const int blocks_c = (P + GS_BIN_THREADS - 1) / GS_BIN_THREADS;
  if (gs_direct_binning(P, tiles, cap)) {
    // 2-4 (direct): count pairs per (block of 1024 Gaussians, tile), then place every pair; no pair sort
    const size_t smem_c = (size_t)64 * tiles, smem_p = (size_t)68 * tiles;
    {  // per call: function attributes are per DEVICE, and one process may drive several
      const int max_smem = 68 * GS_BIN_MAX_TILES;
      if (cudaFuncSetAttribute(bin_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem) != cudaSuccess ||
          cudaFuncSetAttribute(bin_place_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem) != cudaSuccess ||
          cudaFuncSetAttribute(bin_place_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem) != cudaSuccess ||
          cudaFuncSetAttribute(bin_place_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem) != cudaSuccess)
        return gs_check(cudaGetLastError());
    }
    uint32_t* blockcount = s.tkey[0];  // [blocks_c][tiles] <= cap words
    const bool fused = blocks_c <= GS_BIN_FUSED_MAX_BLOCKS;  // co-resident: one 1024-thread block per SM
    if (!fused)
      gs_launch(bin_count_kernel, dim3(blocks_c), dim3(GS_BIN_THREADS), smem_c, st, P, (const uint32_t*)s.dval[0],
                (const uint32_t*)s.dval[1], (const uint32_t*)s.hist, (const uint32_t*)g.tiles_touched, (const uint32_t*)g.rect,
                gx, tiles, blockcount, s.tile_total);
    const bool pre = fused && (prewait & 2);
    gs_launch(fused ? (pre ? bin_place_kernel<true, true> : bin_place_kernel<true, false>) : bin_place_kernel<false, false>,
              dim3(blocks_c + (fused ? 1 : 0)), dim3(GS_BIN_THREADS), smem_p,
              st, P, (const uint32_t*)s.dval[0], (const uint32_t*)s.dval[1], (const uint32_t*)s.hist, (const uint32_t*)g.tiles_touched,
              (const uint32_t*)g.rect, gx, tiles, blockcount, s.tile_total, (unsigned long long)cap, list, im.ranges, g.hdr,
              s.counters + 16, im.seg_off, l.items, (uint32_t)l.max_items, (uint4*)l.fitems, (uint32_t)l.max_fitems, fwd_order, l.order);
    if (fused) return gs_check(cudaGetLastError());  // the plans came from the placement grid's last block
  } else {
    // 2. chained scan + pair emission in depth order
    int bits = (int)higher_msb((uint32_t)tiles);
    int tile_passes = (bits + 7) / 8;
    if (tile_passes < 1) tile_passes = 1;
    if (tile_passes > 4) tile_passes = 4;
    uint32_t* vbuf[2] = {s.tval[0], s.tval[1]};
    const int blocks_e = (P + GS_EMIT_THREADS - 1) / GS_EMIT_THREADS;
    gs_launch(emit_kernel, dim3(blocks_e), dim3(GS_EMIT_THREADS), 0, st, P, (const uint32_t*)s.dval[0],
              (const uint32_t*)s.dval[1], (const uint32_t*)s.hist, (const uint32_t*)g.tiles_touched, (const uint32_t*)g.rect, gx,
              (unsigned long long)cap, s.tkey[0], vbuf[0], s.hist + 4 * 256, tile_passes, s.scan_status, s.counters + 8, g.hdr);
    // 3. stable sort of the pairs by tile id; the last pass writes into `list`
    const int blocks_t = (int)gs_sort_blocks((unsigned long long)cap);
    const bool small_t = gs_sort_items((unsigned long long)cap) == GS_SORT_ITEMS_SMALL;
    for (int p = 0; p < tile_passes; p++) {
      uint32_t* vout = (p == tile_passes - 1) ? list : vbuf[(p + 1) & 1];
      auto kern = small_t ? onesweep_pass_kernel<GS_SORT_ITEMS_SMALL, false> : onesweep_pass_kernel<GS_SORT_ITEMS, false>;
      gs_launch(kern, dim3(blocks_t), dim3(GS_SORT_THREADS), 0, st,
          (const uint32_t*)s.tkey[p & 1], s.tkey[(p + 1) & 1], (const uint32_t*)vbuf[p & 1], vout,
          (const unsigned long long*)&g.hdr->num_rendered, (unsigned long long)cap, 8 * p,
          (const uint32_t*)(s.hist + (4 + p) * 256), s.status_t + (size_t)p * s.blocks_t * 256, s.counters + 4 + p,
          &g.hdr->error, -1, (const uint32_t*)nullptr);
    }
    // 4. tile ranges
    int rb = (int)((cap + 255) / 256);
    if (rb > 148 * 16) rb = 148 * 16;
    if (rb < 1) rb = 1;
    gs_launch(tile_ranges_kernel, dim3(rb), dim3(256), 0, st, (const uint32_t*)s.tkey[tile_passes & 1],
              (const unsigned long long*)&g.hdr->num_rendered, (unsigned long long)cap, im.ranges);
  }
  // 5. backward work plan (segments of GS_SEG entries per tile)
  gs_launch(plan_kernel, dim3(1), dim3(1024), 0, st, tiles, im.ranges, im.seg_off, l.items, (uint32_t)l.max_items, l.fitems,
            (uint32_t)l.max_fitems, g.hdr, fwd_order, l.order);
  return gs_check(cudaGetLastError());
}
