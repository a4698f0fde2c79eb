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
#define RM_SORT_TILE 4096          // items per block: 8 warps x 16 rounds x 32 lanes
#define RM_SORT_WARPS (RM_SORT_TPB / 32)
#define RM_SORT_PER_WARP (RM_SORT_TILE / RM_SORT_WARPS)

static inline int rm_sort_nblocks(int n) { return (n + RM_SORT_TILE - 1) / RM_SORT_TILE; }
// u32 words of histogram scratch rm_sort_pairs needs for up to n items
static inline size_t rm_sort_scratch_words(int n) { return (size_t)256 * (size_t)(rm_sort_nblocks(n) + 1) + 8; }

#ifndef RM_EMU

static __global__ void __launch_bounds__(RM_SORT_TPB) rm_sort_hist_kernel(const u64* __restrict__ keys, int n, int shift,
                                                                  unsigned* __restrict__ ghist, int nblocks) {
  __shared__ unsigned h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  int base = blockIdx.x * RM_SORT_TILE;
  int end = min(base + RM_SORT_TILE, n);
  for (int i = base + threadIdx.x; i < end; i += RM_SORT_TPB) atomicAdd(&h[(unsigned)(keys[i] >> shift) & 255u], 1u);
  __syncthreads();
  ghist[threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// Exclusive scan of `len` u32 values in place, one block of 1024 threads, 8 values per thread per chunk.
static __global__ void __launch_bounds__(1024) rm_scan_u32_kernel(unsigned* __restrict__ data, int len) {
  __shared__ unsigned warp_tot[32];
  __shared__ unsigned carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int chunk = 0; chunk < len; chunk += 1024 * 8) {
    int i0 = chunk + threadIdx.x * 8;
    unsigned v[8], sum = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      v[k] = (i0 + k < len) ? data[i0 + k] : 0u;
      sum += v[k];
    }
    unsigned inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      unsigned o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    if (wid == 0) {
      unsigned w = warp_tot[lane], winc = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        unsigned o = __shfl_up_sync(0xffffffffu, winc, d);
        if (lane >= d) winc += o;
      }
      warp_tot[lane] = winc - w;  // exclusive offset of each warp
    }
    __syncthreads();
    unsigned carry = carry_s;
    unsigned excl = carry + warp_tot[wid] + (inc - sum);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (i0 + k < len) data[i0 + k] = excl;
      excl += v[k];
    }
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl;  // last thread ends at the chunk total
    __syncthreads();
  }
}

static __global__ void __launch_bounds__(RM_SORT_TPB) rm_sort_scatter_kernel(const u64* __restrict__ kin, const unsigned* __restrict__ vin,
                                                                     u64* __restrict__ kout, unsigned* __restrict__ vout, int n,
                                                                     int shift, const unsigned* __restrict__ gscan, int nblocks) {
  __shared__ unsigned wcnt[RM_SORT_WARPS][256];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < RM_SORT_WARPS * 256; i += RM_SORT_TPB) (&wcnt[0][0])[i] = 0;
  __syncthreads();
  const int wbase = blockIdx.x * RM_SORT_TILE + wid * RM_SORT_PER_WARP;
  const unsigned lt = (1u << lane) - 1u;
  // phase 1: digit counts of this warp's contiguous run
  for (int r = 0; r < RM_SORT_PER_WARP; r += 32) {
    int i = wbase + r + lane;
    unsigned d = (i < n) ? ((unsigned)(kin[i] >> shift) & 255u) : 0xFFFFu;
    unsigned m = __match_any_sync(0xffffffffu, d);
    if (d != 0xFFFFu && (m & lt) == 0) wcnt[wid][d] += __popc(m);
    __syncwarp();
  }
  __syncthreads();
  // phase 2: thread d turns the per-warp counts of digit d into global start positions
  {
    unsigned d = threadIdx.x;
    unsigned run = gscan[d * nblocks + blockIdx.x];
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
  for (int r = 0; r < RM_SORT_PER_WARP; r += 32) {
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

// Sorts n pairs by the low `bits` bits of the key.  Returns 0 if the result is in (k0, v0), 1 if in
// (k1, v1).  hist must hold rm_sort_scratch_words(n) u32.
static inline int rm_sort_pairs(u64* k0, unsigned* v0, u64* k1, unsigned* v1, int n, int bits, unsigned* hist,
                                cudaStream_t s) {
  if (n <= 1 || bits <= 0) return 0;
  int passes = (bits + 7) / 8;
  int nb = rm_sort_nblocks(n);
  int cur = 0;
  for (int p = 0; p < passes; ++p) {
    u64* ki = cur ? k1 : k0;
    unsigned* vi = cur ? v1 : v0;
    u64* ko = cur ? k0 : k1;
    unsigned* vo = cur ? v0 : v1;
    RM_LAUNCH(rm_sort_hist_kernel, nb, RM_SORT_TPB, s, ki, n, p * 8, hist, nb);
    RM_LAUNCH(rm_scan_u32_kernel, 1, 1024, s, hist, 256 * nb);
    RM_LAUNCH(rm_sort_scatter_kernel, nb, RM_SORT_TPB, s, ki, vi, ko, vo, n, p * 8, hist, nb);
    cur ^= 1;
  }
  return cur;
}

#else  // RM_EMU: the emulation build sorts on the host (the sort kernels need real block-level parallelism)

#include <vector>
static inline int rm_sort_pairs(u64* k0, unsigned* v0, u64* k1, unsigned* v1, int n, int bits, unsigned*, cudaStream_t) {
  if (n <= 1 || bits <= 0) return 0;
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
