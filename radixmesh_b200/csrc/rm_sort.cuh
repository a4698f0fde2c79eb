// Stable LSD radix sort of (u64 key, u32 value) pairs on the device, 8-bit digits.
//
// Used to (1) group a match batch's split requests by (node, split point), (2) group an insert
// batch by first token while keeping arrival order inside a group, (3) order eviction candidates
// by (last-access stamp, depth) — the "device radix sort over last-access stamps" of north_star.
// Three kernels per pass: per-tile digit histogram, exclusive scan of the digit-major histogram
// table, stable scatter (warp-level multisplit with __match_any_sync; per-warp digit bases in
// shared memory).  Only the low `bits` bits of the key take part, so a 29-bit key costs 4 passes.
#pragma once
#include "rm_common.cuh"

#define RM_SORT_TPB 256
#define RM_SORT_WARPS (RM_SORT_TPB / 32)
#define RM_SORT_MIN_TILE 1024      // 8 warps x 4 rounds x 32 lanes: keeps the histogram table (256 x tiles) small
#define RM_SORT_MAX_TILE 4096      // 8 warps x 16 rounds x 32 lanes
#define RM_SORT_ROUNDS (RM_SORT_MAX_TILE / RM_SORT_TPB)   // rounds of 32 items per warp in the largest tile
#define RM_SORT_SOLO 2048          // up to here one block alone is faster than the grid with its barriers (measured:
                                   // 2000 items 23 us alone, 4097 items 29 us on the grid; tools/sort_probe.cu)
#ifndef RM_SORT_FUSED_TILES
#define RM_SORT_FUSED_TILES 148    // largest grid of the one-launch form (blocks must be resident together)
#endif
#define RM_SORT_TABLE_OFF 8        // hist[0..1]: grid barrier of the one-launch form (zero between launches); tables after

// Items per block: small inputs get small tiles so that the grid still covers the 148 SMs (a 64K-item
// sort with 4096-item tiles ran on 16 blocks and was latency bound); large inputs get 4096-item tiles.
static inline int rm_sort_tile(int n) {
  int t = (n + 147) / 148;                       // aim at >= 1 block per SM
  t = (t + RM_SORT_MIN_TILE - 1) / RM_SORT_MIN_TILE * RM_SORT_MIN_TILE;
  if (t < RM_SORT_MIN_TILE) t = RM_SORT_MIN_TILE;
  if (t > RM_SORT_MAX_TILE) t = RM_SORT_MAX_TILE;
  return t;
}
static inline int rm_sort_nblocks(int n) { int t = rm_sort_tile(n); return (n + t - 1) / t; }
// u32 words of histogram scratch rm_sort_pairs needs for up to n items (worst case: smallest tiles)
static inline size_t rm_sort_scratch_words(int n) {
  size_t nb_small = ((size_t)n + RM_SORT_MIN_TILE - 1) / RM_SORT_MIN_TILE;
  size_t nb = nb_small < 148 ? nb_small : 148;
  size_t nb_big = ((size_t)n + RM_SORT_MAX_TILE - 1) / RM_SORT_MAX_TILE;
  if (nb_big > nb) nb = nb_big;
  return (size_t)256 * (nb + 2) + 8 + RM_SORT_TABLE_OFF;
}

#ifndef RM_EMU

static __global__ void __launch_bounds__(RM_SORT_TPB) rm_sort_hist_kernel(const u64* __restrict__ keys, int n_max,
                                                                  const int* __restrict__ n_dev, int shift,
                                                                  unsigned* __restrict__ ghist, int nblocks, int tile) {
  __shared__ unsigned h[256];
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  if ((int)blockIdx.x * tile >= n) return;   // tiles past the (device-side) count: nothing to do, nothing read later
  h[threadIdx.x] = 0;
  __syncthreads();
  int base = blockIdx.x * tile;
  int end = min(base + tile, n);
  for (int i = base + threadIdx.x; i < end; i += RM_SORT_TPB) atomicAdd(&h[(unsigned)(keys[i] >> shift) & 255u], 1u);
  __syncthreads();
  ghist[threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// Exclusive scan, in place, of every ROW of the digit-major histogram table data[256][nb] (row d = counts of
// digit d per tile), one block per row, and the row's total into data[256 * nb_stride + d].  The scatter
// kernel adds the digit bases itself (an exclusive scan of the 256 totals, per block).  (One block scanning
// the whole table took 24 us per pass at 254 tiles: five passes of a 1 M-key sort spent 120 of 245 us there.)
static __global__ void __launch_bounds__(RM_SORT_TPB) rm_scan_rows_kernel(unsigned* __restrict__ data, int nb_stride, int n_max,
                                                                  const int* __restrict__ n_dev, int tile) {
  __shared__ unsigned wtot[RM_SORT_WARPS];
  __shared__ unsigned carry_s;
  const int n_items = n_dev ? min(*n_dev, n_max) : n_max;
  const int nb = (n_items + tile - 1) / tile;   // active tiles; rows keep their stride of nb_stride
  unsigned* __restrict__ row = data + (size_t)blockIdx.x * nb_stride;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int j0 = 0; j0 < nb; j0 += RM_SORT_TPB) {
    const int j = j0 + (int)threadIdx.x;
    const unsigned v = j < nb ? row[j] : 0u;
    unsigned inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += x;
    }
    if (lane == 31) wtot[wid] = inc;
    __syncthreads();
    unsigned before = carry_s;
    for (int w = 0; w < wid; ++w) before += wtot[w];
    if (j < nb) row[j] = before + inc - v;
    __syncthreads();
    if (threadIdx.x == RM_SORT_TPB - 1) carry_s = before + inc;
    __syncthreads();
  }
  if (threadIdx.x == 0) data[(size_t)256 * nb_stride + blockIdx.x] = carry_s;
}

static __global__ void __launch_bounds__(RM_SORT_TPB) rm_sort_scatter_kernel(const u64* __restrict__ kin, const unsigned* __restrict__ vin,
                                                                     u64* __restrict__ kout, unsigned* __restrict__ vout, int n_max,
                                                                     const int* __restrict__ n_dev, int shift,
                                                                     const unsigned* __restrict__ gscan, int nblocks, int tile) {
  __shared__ unsigned wcnt[RM_SORT_WARPS][256];
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  if ((int)blockIdx.x * tile >= n) return;
  const int per_warp = tile / RM_SORT_WARPS;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < RM_SORT_WARPS * 256; i += RM_SORT_TPB) (&wcnt[0][0])[i] = 0;
  __syncthreads();
  const int wbase = blockIdx.x * tile + wid * per_warp;
  const unsigned lt = (1u << lane) - 1u;
  // phase 1: digit counts of this warp's contiguous run
  for (int r = 0; r < per_warp; r += 32) {
    int i = wbase + r + lane;
    unsigned d = (i < n) ? ((unsigned)(kin[i] >> shift) & 255u) : 0xFFFFu;
    unsigned m = __match_any_sync(0xffffffffu, d);
    if (d != 0xFFFFu && (m & lt) == 0) wcnt[wid][d] += __popc(m);
    __syncwarp();
  }
  __syncthreads();
  // phase 2: thread d turns the per-warp counts of digit d into global start positions: digit base (exclusive
  // scan of the row totals) + the digit's count in the tiles before this one
  {
    __shared__ unsigned dtot[RM_SORT_WARPS];
    unsigned d = threadIdx.x;
    const unsigned tot = gscan[(size_t)256 * nblocks + d];
    unsigned inc = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += x;
    }
    if (lane == 31) dtot[wid] = inc;
    __syncthreads();
    unsigned run = inc - tot + gscan[d * nblocks + blockIdx.x];
    for (int w = 0; w < wid; ++w) run += dtot[w];
#pragma unroll
    for (int w = 0; w < RM_SORT_WARPS; ++w) {
      unsigned c = wcnt[w][d];
      wcnt[w][d] = run;
      run += c;
    }
  }
  __syncthreads();
  // phase 3: same walk, now scattering; equal digits keep their order (lane order inside a
  // round, rounds in order, warps in order, blocks in order) => the sort is stable
  for (int r = 0; r < per_warp; r += 32) {
    int i = wbase + r + lane;
    bool ok = i < n;
    u64 k = ok ? kin[i] : 0;
    unsigned d = ok ? ((unsigned)(k >> shift) & 255u) : 0xFFFFu;
    unsigned m = __match_any_sync(0xffffffffu, d);
    unsigned dst = 0;
    if (ok) dst = wcnt[wid][d] + __popc(m & lt);
    __syncwarp();
    if (ok && (m & lt) == 0) wcnt[wid][d] += __popc(m);
    __syncwarp();
    if (ok) {
      kout[dst] = k;
      vout[dst] = vin[i];
    }
  }
}

// All passes in ONE launch, for inputs of at most 148 tiles (every sort of a match or insert batch).
// A pass of the three-kernel form costs three dependent launches of ~5 us each even when the input is a few
// hundred items (the split requests on shared edges of a 64K-query batch: 12 launches, 65 us for ~500 items).
// Here a pass is: per-warp digit counts in shared memory -> tile totals to ghist[tile][256] -> grid barrier ->
// every block sums the totals it needs itself (digit base + tiles before it; <= 148 coalesced loads per
// thread) -> stable scatter -> grid barrier.  Inputs of at most RM_SORT_SOLO items (a grid-uniform
// test on the device-side count) are sorted by block 0 alone with no grid barrier at all.
// The grid barrier is a counter in hist[0] (arrivals; the blocks of one launch, at most one per SM, are all
// resident or become resident as soon as an SM is free, and wait for nobody but each other); the last block
// to leave the kernel puts hist[0] and hist[1] back to zero.  (A cooperative launch costs ~20 us more than a
// plain one here, which is the whole budget of this sort.)
static __device__ __forceinline__ void rm_sort_grid_barrier(unsigned* bar, unsigned target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1u);
    while (*(volatile unsigned*)bar < target) {
    }
    __threadfence();
  }
  __syncthreads();
}
// The sort proper, run by the blocks `bid` = 0..nblk-1 of RM_SORT_TPB threads of ONE launch (all resident).  With
// n <= RM_SORT_SOLO only bid 0 works and no barrier is used.  `arrivals` is the caller's running count of barrier
// arrivals on hist[0] (a kernel that has other phases around the sort shares the barrier with it).
static __device__ __forceinline__ void rm_sort_body(u64* k0, unsigned* v0, u64* k1, unsigned* v1, int n, int passes,
                                                    unsigned* hist, int tile, const int bid, const int nblk,
                                                    unsigned& arrivals) {
  __shared__ unsigned wcnt[RM_SORT_WARPS][256];
  __shared__ unsigned wtot[RM_SORT_WARPS];
  const bool solo = n <= RM_SORT_SOLO;
  if (solo && bid != 0) return;
  unsigned* ghist = hist + RM_SORT_TABLE_OFF;
  const int my_tile = solo ? (n + 255) / 256 * 256 : tile;
  const int nb = solo ? 1 : (n + tile - 1) / tile;
  const bool active = bid < nb;
  const int per_warp = my_tile / RM_SORT_WARPS;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int wbase = bid * my_tile + wid * per_warp;
  const unsigned lt = (1u << lane) - 1u;
  for (int p = 0; p < passes; ++p) {
    const int shift = 8 * p;
    const u64* kin = (p & 1) ? k1 : k0;
    const unsigned* vin = (p & 1) ? v1 : v0;
    u64* kout = (p & 1) ? k0 : k1;
    unsigned* vout = (p & 1) ? v0 : v1;
    // this warp's run of the tile, loaded up front: the loads of all rounds are in flight together (round
    // after round, each waiting for its own L2 round trip, a 4096-item single-block sort took 87 us)
    u64 kreg[RM_SORT_ROUNDS];
    unsigned vreg[RM_SORT_ROUNDS];
#pragma unroll
    for (int r = 0; r < RM_SORT_ROUNDS; ++r) {
      const int i = wbase + r * 32 + lane;
      const bool ok = active && r * 32 < per_warp && i < n;
      kreg[r] = ok ? __ldcg(kin + i) : 0ull;
      vreg[r] = ok ? __ldcg(vin + i) : 0u;
    }
    for (int i = threadIdx.x; i < RM_SORT_WARPS * 256; i += RM_SORT_TPB) (&wcnt[0][0])[i] = 0;
    __syncthreads();
    if (active) {
#pragma unroll
      for (int r = 0; r < RM_SORT_ROUNDS; ++r) {
        if (r * 32 < per_warp) {
          const int i = wbase + r * 32 + lane;
          unsigned d = (i < n) ? ((unsigned)(kreg[r] >> shift) & 255u) : 0xFFFFu;
          unsigned m = __match_any_sync(0xffffffffu, d);
          if (d != 0xFFFFu && (m & lt) == 0) wcnt[wid][d] += __popc(m);
          __syncwarp();
        }
      }
    }
    __syncthreads();
    const unsigned d = threadIdx.x;
    unsigned total = 0, before = 0;
#pragma unroll
    for (int w = 0; w < RM_SORT_WARPS; ++w) total += wcnt[w][d];
    if (!solo) {
      if (active) ghist[bid * 256 + d] = total;
      arrivals += (unsigned)nblk;
      rm_sort_grid_barrier(hist, arrivals);
      total = 0;
#pragma unroll 16
      for (int j = 0; j < nb; ++j) {
        unsigned x = __ldcg(ghist + j * 256 + d);
        total += x;
        if (j < bid) before += x;
      }
    }
    // exclusive scan of the digit totals over the 256 threads
    unsigned inc = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += x;
    }
    if (lane == 31) wtot[wid] = inc;
    __syncthreads();
    unsigned run = inc - total + before;
    for (int w = 0; w < wid; ++w) run += wtot[w];
#pragma unroll
    for (int w = 0; w < RM_SORT_WARPS; ++w) {
      unsigned c = wcnt[w][d];
      wcnt[w][d] = run;
      run += c;
    }
    __syncthreads();
    if (active) {
#pragma unroll
      for (int r = 0; r < RM_SORT_ROUNDS; ++r) {
        if (r * 32 < per_warp) {
          const int i = wbase + r * 32 + lane;
          const bool ok = i < n;
          unsigned dd = ok ? ((unsigned)(kreg[r] >> shift) & 255u) : 0xFFFFu;
          unsigned m = __match_any_sync(0xffffffffu, dd);
          unsigned dst = 0;
          if (ok) dst = wcnt[wid][dd] + __popc(m & lt);
          __syncwarp();
          if (ok && (m & lt) == 0) wcnt[wid][dd] += __popc(m);
          __syncwarp();
          if (ok) {
            kout[dst] = kreg[r];
            vout[dst] = vreg[r];
          }
        }
      }
    }
    if (p + 1 < passes) {
      if (solo) {
        __syncthreads();
      } else {
        arrivals += (unsigned)nblk;
        rm_sort_grid_barrier(hist, arrivals);
      }
    }
  }
}

// The same stable LSD sort run by ONE block of RM_SORT_TPB threads on any number of items, with shared memory as
// its only scratch: per pass a digit histogram of all items, an exclusive scan, then tile after tile (2048 items) the
// warp-level stable multisplit with running digit bases.  For the rare big groups of the split stage (rm_match.cu),
// where the block that sorts is the last one alive of its launch.  Returns 0 if the result is in (k0, v0), 1 if in
// (k1, v1).  Ends with a __syncthreads().
static __device__ __forceinline__ int rm_sort_block(u64* k0, unsigned* v0, u64* k1, unsigned* v1, int n, int passes) {
  __shared__ unsigned wcnt[RM_SORT_WARPS][256];
  __shared__ unsigned base[256];
  __shared__ unsigned wtot[RM_SORT_WARPS];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  const int tile = RM_SORT_TPB * 8;
  for (int p = 0; p < passes; ++p) {
    const int shift = 8 * p;
    const u64* kin = (p & 1) ? k1 : k0;
    const unsigned* vin = (p & 1) ? v1 : v0;
    u64* kout = (p & 1) ? k0 : k1;
    unsigned* vout = (p & 1) ? v0 : v1;
    base[threadIdx.x] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += RM_SORT_TPB) atomicAdd(&base[(unsigned)(__ldcg(kin + i) >> shift) & 255u], 1u);
    __syncthreads();
    {
      const unsigned total = base[threadIdx.x];
      unsigned inc = total;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        unsigned x = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += x;
      }
      if (lane == 31) wtot[wid] = inc;
      __syncthreads();
      unsigned run = inc - total;
      for (int w = 0; w < wid; ++w) run += wtot[w];
      base[threadIdx.x] = run;
    }
    __syncthreads();
    for (int t0 = 0; t0 < n; t0 += tile) {
      const int wbase = t0 + wid * (tile / RM_SORT_WARPS);
      for (int i = threadIdx.x; i < RM_SORT_WARPS * 256; i += RM_SORT_TPB) (&wcnt[0][0])[i] = 0;
      __syncthreads();
      u64 kreg[8];
      unsigned vreg[8];
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int i = wbase + r * 32 + lane;
        const bool ok = i < n;
        kreg[r] = ok ? __ldcg(kin + i) : 0ull;
        vreg[r] = ok ? __ldcg(vin + i) : 0u;
      }
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int i = wbase + r * 32 + lane;
        unsigned d = (i < n) ? ((unsigned)(kreg[r] >> shift) & 255u) : 0xFFFFu;
        unsigned m = __match_any_sync(0xffffffffu, d);
        if (d != 0xFFFFu && (m & lt) == 0) wcnt[wid][d] += __popc(m);
        __syncwarp();
      }
      __syncthreads();
      {
        const unsigned d = threadIdx.x;
        unsigned run = base[d];
#pragma unroll
        for (int w = 0; w < RM_SORT_WARPS; ++w) {
          unsigned c = wcnt[w][d];
          wcnt[w][d] = run;
          run += c;
        }
        base[d] = run;
      }
      __syncthreads();
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int i = wbase + r * 32 + lane;
        const bool ok = i < n;
        unsigned dd = ok ? ((unsigned)(kreg[r] >> shift) & 255u) : 0xFFFFu;
        unsigned m = __match_any_sync(0xffffffffu, dd);
        unsigned dst = 0;
        if (ok) dst = wcnt[wid][dd] + __popc(m & lt);
        __syncwarp();
        if (ok && (m & lt) == 0) wcnt[wid][dd] += __popc(m);
        __syncwarp();
        if (ok) {
          kout[dst] = kreg[r];
          vout[dst] = vreg[r];
        }
      }
      __syncthreads();
    }
    __threadfence_block();
    __syncthreads();
  }
  return passes & 1;
}

// the last block out of a launch that used the barrier puts hist[0] and hist[1] back to zero
static __device__ __forceinline__ void rm_sort_barrier_rearm(unsigned* hist, int nblk) {
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(hist + 1, 1u) == (unsigned)nblk - 1) {
      hist[0] = 0;
      hist[1] = 0;
      __threadfence();
    }
  }
}

static __global__ void __launch_bounds__(RM_SORT_TPB) rm_sort_fused_kernel(u64* k0, unsigned* v0, u64* k1, unsigned* v1, int n_max,
                                                                   const int* n_dev, int passes, unsigned* hist, int tile) {
  rm_grid_dependency_trigger();
  rm_grid_dependency_wait();     // RM_LAUNCH_PDL: the producer of the items may still be running
  const int n = n_dev ? min(*n_dev, n_max) : n_max;
  if (n <= 0) return;
  if (n_dev) {
    // the grid was sized from the bound n_max; the tile follows the real count (rm_sort_tile), but never so small
    // that the grid could not hold the tiles
    int tl = ((n + 147) / 148 + RM_SORT_MIN_TILE - 1) / RM_SORT_MIN_TILE * RM_SORT_MIN_TILE;
    const int need = ((n + (int)gridDim.x - 1) / (int)gridDim.x + RM_SORT_MIN_TILE - 1) / RM_SORT_MIN_TILE * RM_SORT_MIN_TILE;
    tl = max(max(tl, need), RM_SORT_MIN_TILE);
    tile = min(tl, RM_SORT_MAX_TILE);
  }
  unsigned arrivals = 0;
  rm_sort_body(k0, v0, k1, v1, n, passes, hist, tile, (int)blockIdx.x, (int)gridDim.x, arrivals);
  if (n > RM_SORT_SOLO) rm_sort_barrier_rearm(hist, (int)gridDim.x);
}

// Sorts n pairs by the low `bits` bits of the key.  Returns 0 if the result is in (k0, v0), 1 if in
// (k1, v1).  hist must hold rm_sort_scratch_words(n) u32.  If n_dev is not NULL the item count is read on
// the device (min(*n_dev, n)) and n only bounds the grid, so no host synchronisation is needed.
static inline int rm_sort_pairs(u64* k0, unsigned* v0, u64* k1, unsigned* v1, int n, int bits, unsigned* hist,
                                cudaStream_t s, const int* n_dev = nullptr) {
  if (n <= 1 || bits <= 0) return 0;
  int passes = (bits + 7) / 8;
  int tile = rm_sort_tile(n);
  int nb = (n + tile - 1) / tile;
  if (nb <= RM_SORT_FUSED_TILES) {   // one launch, all blocks resident
    RM_LAUNCH_PDL(rm_sort_fused_kernel, nb, RM_SORT_TPB, s, k0, v0, k1, v1, n, n_dev, passes, hist, tile);
    return passes & 1;
  }
  hist += RM_SORT_TABLE_OFF;
  int cur = 0;
  for (int p = 0; p < passes; ++p) {
    u64* ki = cur ? k1 : k0;
    unsigned* vi = cur ? v1 : v0;
    u64* ko = cur ? k0 : k1;
    unsigned* vo = cur ? v0 : v1;
    RM_LAUNCH(rm_sort_hist_kernel, nb, RM_SORT_TPB, s, ki, n, n_dev, p * 8, hist, nb, tile);
    RM_LAUNCH(rm_scan_rows_kernel, 256, RM_SORT_TPB, s, hist, nb, n, n_dev, tile);
    RM_LAUNCH(rm_sort_scatter_kernel, nb, RM_SORT_TPB, s, ki, vi, ko, vo, n, n_dev, p * 8, hist, nb, tile);
    cur ^= 1;
  }
  return cur;
}

#else  // RM_EMU: the emulation build sorts on the host (the sort kernels need real block-level parallelism)

#include <vector>
static inline int rm_sort_pairs(u64* k0, unsigned* v0, u64* k1, unsigned* v1, int n, int bits, unsigned*, cudaStream_t,
                                const int* n_dev = nullptr) {
  if (n_dev && *n_dev < n) n = *n_dev;
  if (bits <= 0) return 0;
  if (n <= 1) {
    int passes1 = (bits + 7) / 8;
    if (passes1 % 2 == 1 && n == 1) {
      k1[0] = k0[0];
      v1[0] = v0[0];
    }
    return passes1 % 2;
  }
  int eff = ((bits + 7) / 8) * 8;
  u64 mask = eff >= 64 ? ~0ull : ((1ull << eff) - 1ull);
  std::vector<int> ord(n);
  for (int i = 0; i < n; ++i) ord[i] = i;
  std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) { return (k0[a] & mask) < (k0[b] & mask); });
  for (int i = 0; i < n; ++i) {
    k1[i] = k0[ord[i]];
    v1[i] = v0[ord[i]];
  }
  int passes = (bits + 7) / 8;
  if (passes % 2 == 0) {  // mirror the ping-pong parity of the device version
    memcpy(k0, k1, sizeof(u64) * n);
    memcpy(v0, v1, sizeof(unsigned) * n);
    return 0;
  }
  return 1;
}
#endif
