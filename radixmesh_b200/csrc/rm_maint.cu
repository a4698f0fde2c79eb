// Reset, K5 lock refs, K6 LRU eviction, K7 GC-duplicate candidate scan (sm_100a).
//
//   reset   RadixCache.reset                     radix_cache.py:118-125
//   K5      inc_lock_ref / dec_lock_ref          radix_cache.py:204-230, radix_mesh.py:168-179
//   K6      evict + _collect_leaves + _delete_leaf   radix_cache.py:179-202, 366-377, 347-352
//   K7      RadixMesh._gc candidate scan         radix_mesh.py:151-156
#include <vector>

#include "rm_internal.h"
#include "rm_sort.cuh"

#ifdef RM_EMU
#define RM_MT_TPB 64
#else
#define RM_MT_TPB 256
#endif
#define RM_INT_MAX 0x7fffffff

// ------------------------------------------------------------------ reset
__global__ void rm_reset_kernel(RmTree t, int* lock_first, int* lock_cnt) {
  const i64 n = t.max_nodes;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 x = (i64)blockIdx.x * blockDim.x + threadIdx.x; x < n; x += stride) {
    t.free_ids[x] = (int)(x + 1);       // ring position x hands out id x+1 (id 0 is the root)
    t.parent[x] = -1;
    t.edge_len[x] = 0;
    t.edge_off[x] = 0;
    t.lock_ref[x] = (x == 0) ? 1 : 0;   // root.lock_ref = 1 (radix_cache.py:122)
    t.stamp[x] = 0;
    t.rank[x] = (t.mode == RM_MODE_BASE) ? -1 : 0;   // mesh roots belong to the master rank (radix_mesh.py:243-245)
    t.depth_end[x] = 0;
    t.nchild[x] = 0;
    t.gen[x] = 0;
    t.parked[x] = -1;
    t.born[x] = 0;
    t.ord[x] = 0;
    lock_first[x] = RM_INT_MAX;
    lock_cnt[x] = 0;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    RmCounters c;
    c.free_head = 0;
    c.free_tail = (u64)(t.max_nodes - 1);
    c.arena_top = 0;
    c.evictable = 0;
    c.protected_ = 0;
    c.n_nodes = 0;
    c.node_hwm = 1;
    c.err = 0;
    c.n_dups = 0;
    c.n_req = 0;
    c.evict_n = 0;
    c.split_done = 0;
    c.n_tomb = 0;
    c.n_kids = 0;
    c.n_big = 0;
    c.dbg_max_group = 0;
    c.dbg_groups = 0;
    c.ev_total = 0;
    c.ev_k0 = 0;
    c.ins_ordered = 0;
    c.ins_resumed = 0;
    *t.ctr = c;
  }
}

struct RmLockScratch {
  int* first;   // per node: lowest op index of the batch that touches it
  int* cnt;     // per node: ops of the batch that touch it
  int* flag;    // batch needs the exact sequential path
};
static inline RmLockScratch rm_lock_scratch(rm_index* idx) {
  RmLockScratch ls = {idx->sc.lock_first, idx->sc.lock_cnt, idx->sc.lock_flag};
  return ls;
}

int rm_launch_reset(rm_index* idx, cudaStream_t s) {
  RmTree& t = idx->t;
  RM_CUDA(cudaMemsetAsync(t.table, 0xFF, sizeof(RmSlot) * ((size_t)t.table_mask + 1), s));
  RM_CUDA(cudaMemsetAsync(t.dup_key, 0xFF, sizeof(u64) * ((size_t)t.dup_mask + 1), s));
  RM_CUDA(cudaMemsetAsync(t.req_head, 0xFF, sizeof(int) * (size_t)t.max_nodes, s));
  RM_CUDA(cudaMemsetAsync(idx->sc.ins_mark, 0, sizeof(int) * (size_t)t.max_nodes, s));
  RM_CUDA(cudaMemsetAsync(t.ts, 0xFF, sizeof(int2) * ((size_t)t.max_tokens + RM_ARENA_PAD), s));   // token -1, owner -1 (free)
  RM_CUDA(cudaMemsetAsync(t.req_cnt, 0, sizeof(int) * (size_t)t.max_nodes, s));
  RmLockScratch ls = rm_lock_scratch(idx);
  RM_CUDA(cudaMemsetAsync(ls.flag, 0, sizeof(int), s));
  int grid = (int)min((i64)1184, ((i64)t.max_nodes + RM_MT_TPB - 1) / RM_MT_TPB);
  RM_LAUNCH(rm_reset_kernel, grid, RM_MT_TPB, s, t, ls.first, ls.cnt);
  RM_CUDA(cudaGetLastError());
  idx->op_seq = 1;   // reset itself is operation 1 of the logical clock
  idx->ins_records = 0;
  idx->hwm_exact = 1;
  idx->hwm_seq = idx->op_seq;
  return RM_OK;
}

// ------------------------------------------------------------------ K5 lock refs
// A handle (node id, generation) taken before a locked swap stands for the parked object.
__device__ __forceinline__ int rm_resolve(const RmTree& t, int node, const int* gens, int i) {
  if (gens == nullptr || node <= 0) return node;
  int want = gens[i];
  if (want < 0 || t.gen[node] == want) return node;
  int d = t.parked[node];
  while (d >= 0) {
    if (t.gen[d] == -(want + 1)) return d;
    d = t.parked[d];
  }
  return node;
}

__global__ void rm_lock_mark_kernel(RmTree t, const int* nodes, const int* gens, int B, RmLockScratch ls) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  int x = rm_resolve(t, nodes[i], gens, i);
  while (x > 0) {
    atomicAdd(&ls.cnt[x], 1);
    atomicMin(&ls.first[x], i);
    x = t.parent[x];
  }
}

// len(node.value): edge length, except router values whose len() is 1 (radix_mesh.py:62-63)
__device__ __forceinline__ i64 rm_value_len(const RmTree& t, int x) {
  return t.mode == RM_MODE_ROUTER ? 1 : (i64)t.edge_len[x];
}

__global__ void rm_lock_delta_kernel(RmTree t, const int* nodes, const int* gens, int B, int dir, RmLockScratch ls,
                                     i64* out_delta) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  int x = rm_resolve(t, nodes[i], gens, i);
  i64 delta = 0;
  while (x > 0) {
    const int r0 = t.lock_ref[x];
    if (dir > 0) {
      if (r0 == 0 && ls.first[x] == i) delta -= rm_value_len(t, x);
    } else {
      if (r0 == 1 && ls.first[x] == i) delta += rm_value_len(t, x);
      if (r0 >= 2 && ls.cnt[x] >= r0) atomicExch(ls.flag, 1);   // a later op of the batch hits 1 -> 0
    }
    x = t.parent[x];
  }
  out_delta[i] = delta;
}

__global__ void rm_lock_apply_kernel(RmTree t, const int* nodes, const int* gens, int B, int dir, RmLockScratch ls,
                                     const i64* out_delta) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  if (*ls.flag) return;
  int x = rm_resolve(t, nodes[i], gens, i);
  while (x > 0) {
    const int par = t.parent[x];
    if (ls.first[x] == i) {
      const int r0 = t.lock_ref[x], c = ls.cnt[x];
      int r1 = dir > 0 ? r0 + c : r0 - c;
      if (dir < 0 && t.mode != RM_MODE_BASE && r1 < 0) r1 = 0;   // radix_mesh.py:175-177
      t.lock_ref[x] = r1;
    }
    x = par;
  }
  const i64 d = out_delta[i];
  if (d != 0) {
    rm_add_i64(&t.ctr->evictable, d);
    rm_add_i64(&t.ctr->protected_, -d);
  }
}

__global__ void rm_lock_clear_kernel(RmTree t, const int* nodes, const int* gens, int B, RmLockScratch ls) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  int x = rm_resolve(t, nodes[i], gens, i);
  while (x > 0) {
    ls.cnt[x] = 0;
    ls.first[x] = RM_INT_MAX;
    x = t.parent[x];
  }
}

// Exact fallback: the batch in order, one thread (only when several decrements of one batch
// take the same node from >= 2 down to 0).
__global__ void rm_lock_seq_kernel(RmTree t, const int* nodes, const int* gens, int B, int dir, RmLockScratch ls,
                                   i64* out_delta) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  if (*ls.flag == 0) return;
  *ls.flag = 0;
  i64 ev = 0;
  for (int i = 0; i < B; ++i) {
    int x = rm_resolve(t, nodes[i], gens, i);
    i64 delta = 0;
    while (x > 0) {
      int r = t.lock_ref[x];
      if (dir > 0) {
        if (r == 0) delta -= rm_value_len(t, x);
        t.lock_ref[x] = r + 1;
      } else {
        if (r == 1) delta += rm_value_len(t, x);
        if (t.mode == RM_MODE_BASE || r > 0) t.lock_ref[x] = r - 1;
      }
      x = t.parent[x];
    }
    out_delta[i] = delta;
    ev += delta;
  }
  rm_add_i64(&t.ctr->evictable, ev);
  rm_add_i64(&t.ctr->protected_, -ev);
}

int rm_launch_lock_g(rm_index* idx, const int* nodes, const int* gens, int B, int dir, i64* out_delta, cudaStream_t s) {
  if (B <= 0) return RM_OK;
  RmLockScratch ls = rm_lock_scratch(idx);
  const int grid = (B + RM_MT_TPB - 1) / RM_MT_TPB;
  RM_LAUNCH(rm_lock_mark_kernel, grid, RM_MT_TPB, s, idx->t, nodes, gens, B, ls);
  RM_LAUNCH(rm_lock_delta_kernel, grid, RM_MT_TPB, s, idx->t, nodes, gens, B, dir, ls, out_delta);
  RM_LAUNCH(rm_lock_apply_kernel, grid, RM_MT_TPB, s, idx->t, nodes, gens, B, dir, ls, out_delta);
  RM_LAUNCH(rm_lock_clear_kernel, grid, RM_MT_TPB, s, idx->t, nodes, gens, B, ls);
  RM_LAUNCH(rm_lock_seq_kernel, 1, 1, s, idx->t, nodes, gens, B, dir, ls, out_delta);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_launch_lock(rm_index* idx, const int* nodes, int B, int dir, i64* out_delta, cudaStream_t s) {
  return rm_launch_lock_g(idx, nodes, nullptr, B, dir, out_delta, s);
}

// ------------------------------------------------------------------ K6 eviction
// Closed form of the heap loop: with one stamp per operation a parent's stamp is
// never older than a child's, so leaves leave the heap in ascending (stamp, deeper first) order and
// the evicted set is the shortest prefix of that order whose value lengths sum to >= num_tokens.
// A node can only be evicted after all its descendants, so a locked node pins every ancestor, whatever
// the ancestors' own lock_ref (inc/dec pairs keep lock_ref monotone along a path, but nothing forces
// callers to pair them: the heap loop of radix_cache.py:187-202 simply never reaches such a parent).
// (The pinning is the last step of rm_evict_count_kernel.)

// Lazy stamps.  Walks and inserts stamp only the deepest node(s) they reach (one atomic per operation instead
// of one per hop); every other node the reference would have stamped is an ancestor of such an endpoint.
// The stamp the reference holds for a node is therefore the maximum of the raw stamps in its subtree,
//     eff(x) = max(raw(x), max over children c of eff(c)),
// (deleting a node would lower its ancestors' eff, so eviction pushes eff(x) into raw(parent) first).
//
// Computing eff for the whole tree bottom-up is a chain of dependent atomics as long as the tree is deep
// (0.5 ms on a 1 M-node tree).  Only the nodes that CAN be evicted by this call need it:
//   1. take the evictable LEAVES alone, keyed (raw stamp, deeper first) - a leaf's eff is its raw stamp - and
//      find the frontier key K0 where their lengths, in key order, reach num_tokens (a radix select, below);
//   2. the true cut lies at or before K0 (inner nodes only add tokens earlier), and a node's key is never
//      below a descendant's, so every node the call can evict has its whole subtree at keys <= K0;
//   3. so: the leaves up to K0 climb; a parent joins the candidates when ALL its children have reported
//      (they are candidates themselves) and its own key is still <= K0, and carries on upward;
//   4. candidates are sorted again and cut exactly.
// The climbs are as short as the evicted region is shallow.
__global__ void rm_evict_count_kernel(RmTree t, int n, u64* __restrict__ eff, int* __restrict__ rem, int* __restrict__ pinned) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n || x >= t.ctr->node_hwm) return;          // n may be a host-side bound above the mark
  eff[x] = t.stamp[x];
  if (x == 0) return;
  const int par = t.parent[x];
  if (par < 0 || t.gen[x] < 0) return;
  if (par > 0) atomicAdd(&rem[par], 1);
  // a locked node pins every ancestor (note above; same pass over the node records)
  if (t.lock_ref[x] > 0) {
    int y = par;
    while (y > 0 && atomicExch(&pinned[y], 1) == 0) y = t.parent[y];
  }
}

__device__ __forceinline__ u64 rm_evict_key(u64 stamp, int depth_end, int depth_bits) {
  const u64 dmask = (1ull << depth_bits) - 1ull;
  return (stamp << depth_bits) | (dmask - (u64)(unsigned)depth_end);
}

// keys of the evictable leaves (everything else sorts last)
__global__ void rm_evict_keys_kernel(RmTree t, int n, int depth_bits, const int* __restrict__ pinned,
                                     const int* __restrict__ rem, u64* key, unsigned* val) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n) return;
  u64 k = RM_NO_REQ;
  if (x > 0 && x < t.ctr->node_hwm && t.parent[x] >= 0 && t.gen[x] >= 0 && t.lock_ref[x] == 0 && pinned[x] == 0 && rem[x] == 0)
    k = rm_evict_key(t.stamp[x], t.depth_end[x], depth_bits);
  key[x] = k;
  val[x] = (unsigned)x;
}

// Step 1 without sorting a million keys to use a few thousand of them: a weighted radix SELECT.  The frontier key K0 is
// the smallest key such that the value lengths of the evictable leaves at or below it reach num_tokens.  Twelve bits at
// a time from the top: a histogram of token weight per digit over the keys that share the prefix chosen so far, then
// one block picks the digit where the running weight crosses what is still needed.  ceil(bits / 12) passes over the
// key array (each a few microseconds at 1 M nodes) instead of 5 sort passes of three kernels each; the leaves at or
// below K0 are then compacted, unordered, for the climb (the candidates are sorted afterwards anyway).
#define RM_SEL_BITS 12
#define RM_SEL_BINS (1 << RM_SEL_BITS)
#define RM_SEL_PICK_TPB 64        // (emulation build only: the device picks inside the level kernel)
struct RmEvictSel {
  u64 prefix;      // digits chosen so far (high bits of K0)
  i64 need;        // tokens still to be covered inside the prefix
  int all;         // the evictable leaves do not hold num_tokens: take every one of them
  unsigned done;   // blocks of the current level's histogram that have added their part (the last one picks the digit)
};

__global__ void rm_evict_sel_init_kernel(RmTree t, int n_bound, RmEvictSel* st, i64 num_tokens, u64* __restrict__ hw,
                                         unsigned* __restrict__ hc) {
  for (int b = threadIdx.x; b < RM_SEL_BINS; b += blockDim.x) {
    hw[b] = 0;
    hc[b] = 0;
  }
  if (threadIdx.x == 0) {
    st->prefix = 0;
    st->need = num_tokens;
    st->all = 0;
    st->done = 0;
    t.ctr->ev_total = 0;
    t.ctr->ev_over = t.ctr->node_hwm > n_bound ? 1 : 0;      // nodes beyond the host's bound: this pass evicts nothing
  }
}

// up > 0: the last level when bits is not a multiple of 12 (the remaining low bits are shifted UP into the digit).
#ifdef RM_EMU
// Emulation build (threads run one after another): the level in two plain steps, histogram and pick.
__global__ void rm_evict_sel_hist_kernel(RmTree t, const u64* __restrict__ key, int n, int shift, int up, const RmEvictSel* st,
                                         u64* __restrict__ hw) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n || st->all) return;
  const u64 k = key[x];
  const u64 kk = up ? (k << up) : (k >> shift);
  if (k != RM_NO_REQ && (kk >> RM_SEL_BITS) == st->prefix) atomicAdd(&hw[(unsigned)kk & (RM_SEL_BINS - 1)], (u64)t.edge_len[x]);
}

// the digit where the running weight reaches `need`; clears the histogram for the next level
__global__ void rm_evict_sel_pick_kernel(RmEvictSel* st, u64* __restrict__ hw, unsigned* __restrict__ hc) {
  if (threadIdx.x == 0 && !st->all) {
    i64 run = 0;
    int b = 0;
    for (; b < RM_SEL_BINS; ++b) {
      if (run + (i64)hw[b] >= st->need) break;
      run += (i64)hw[b];
    }
    if (b == RM_SEL_BINS) {
      st->all = 1;
    } else {
      st->need -= run;
      st->prefix = (st->prefix << RM_SEL_BITS) | (u64)b;
    }
  }
  for (int b = threadIdx.x; b < RM_SEL_BINS; b += blockDim.x) {
    hw[b] = 0;
    hc[b] = 0;
  }
}
#endif

#ifndef RM_EMU
// One level of the select in ONE launch: a few blocks per SM stride over the keys (a block per 256 keys spent its
// time clearing and flushing 4096 shared bins: 15-26 us per level at 1 M nodes), and the last block to add its part
// picks the digit (a launch of its own: 7 us per level).
#define RM_SEL_BPT (RM_SEL_BINS / RM_MT_TPB)     // bins per thread in the pick
__global__ void __launch_bounds__(RM_MT_TPB)
rm_evict_sel_level_kernel(RmTree t, const u64* __restrict__ key, int n, int shift, int up, RmEvictSel* st, u64* __restrict__ hw) {
  // 32-bit bins (the launcher keeps a block's keys x the longest edge below 2^31; a 64-bit shared atomic is a
  // compare-and-swap loop), and ONE atomic for a warp whose leaves all fall into the same digit: on the first level
  // the leaves of a served tree share their top digits, and 32 lanes adding to one address serialise (26 us)
  __shared__ unsigned sh[RM_SEL_BINS];
  __shared__ i64 wtot[RM_MT_TPB / 32];
  __shared__ int last_s, found_s;
  const bool all = st->all != 0;
  const u64 prefix = st->prefix;
  for (int b = threadIdx.x; b < RM_SEL_BINS; b += RM_MT_TPB) sh[b] = 0;
  __syncthreads();
  if (!all) {
    n = min(n, t.ctr->node_hwm);
    for (int base = blockIdx.x * RM_MT_TPB; base < n; base += gridDim.x * RM_MT_TPB) {
      const int x = base + (int)threadIdx.x;
      bool in = x < n;
      unsigned bin = 0, w = 0;
      if (in) {
        const u64 k = key[x];
        const u64 kk = up ? (k << up) : (k >> shift);
        in = k != RM_NO_REQ && (kk >> RM_SEL_BITS) == prefix;
        bin = (unsigned)kk & (RM_SEL_BINS - 1);
        if (in) w = (unsigned)t.edge_len[x];
      }
      const unsigned act = __ballot_sync(0xffffffffu, in);
      if (act == 0) continue;
      const int lead = __ffs(act) - 1;
      const unsigned b0 = __shfl_sync(0xffffffffu, bin, lead);
      if (__all_sync(0xffffffffu, !in || bin == b0)) {
        const unsigned tot = __reduce_add_sync(0xffffffffu, w);
        if ((int)(threadIdx.x & 31) == lead) atomicAdd(&sh[b0], tot);
      } else if (in) {
        atomicAdd(&sh[bin], w);
      }
    }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < RM_SEL_BINS; b += RM_MT_TPB) {
    const unsigned v = sh[b];
    if (v) atomicAdd(&hw[b], (u64)v);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last_s = atomicAdd(&st->done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last_s) return;
  __threadfence();
  // the pick: exclusive scan of the 4096 weights over the block, RM_SEL_BPT consecutive bins per thread
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) found_s = 0;
  const i64 need = st->need;
  i64 wv[RM_SEL_BPT];
  i64 mine = 0;
#pragma unroll
  for (int j = 0; j < RM_SEL_BPT; ++j) {
    wv[j] = (i64)__ldcg(&hw[RM_SEL_BPT * threadIdx.x + j]);
    mine += wv[j];
  }
  i64 inc = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const i64 o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) wtot[wid] = inc;
  __syncthreads();
  i64 before = inc - mine;
  for (int w = 0; w < wid; ++w) before += wtot[w];
  if (!all && before < need && before + mine >= need) {     // the crossing is inside my bins
    i64 run = before;
    int j = 0;
#pragma unroll
    for (int j2 = 0; j2 < RM_SEL_BPT; ++j2) {
      if (j == j2 && run + wv[j2] < need) {
        run += wv[j2];
        ++j;
      }
    }
    st->need = need - run;
    st->prefix = (prefix << RM_SEL_BITS) | (u64)(RM_SEL_BPT * threadIdx.x + j);
    found_s = 1;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (!all && !found_s) st->all = 1;
    st->done = 0;
  }
#pragma unroll
  for (int j = 0; j < RM_SEL_BPT; ++j) hw[RM_SEL_BPT * threadIdx.x + j] = 0;
}
#endif

// the leaves at or below K0 (K0 = the full prefix after the last level; shift_out drops padding bits of the last digit)
__global__ void rm_evict_sel_compact_kernel(RmTree t, const u64* __restrict__ key, int n, const RmEvictSel* st, int pad_bits,
                                            u64* __restrict__ skey, unsigned* __restrict__ sval) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x == 0 && blockIdx.x == 0) t.ctr->ev_k0 = st->all ? RM_NO_REQ - 1 : (st->prefix >> pad_bits);
  if (x >= n) return;
  const u64 k = key[x];
  if (k == RM_NO_REQ) return;
  if (!st->all && k > (st->prefix >> pad_bits)) return;
  const int slot = atomicAdd(&t.ctr->ev_total, 1);
  skey[slot] = k;
  sval[slot] = (unsigned)x;
}

__global__ void rm_evict_sel_seal_kernel(RmTree t) {
  if (threadIdx.x == 0 && blockIdx.x == 0) t.ctr->evict_n = t.ctr->ev_total;   // evict_n doubles as "leaves among the candidates" until the cut
}

// step 3: parents exposed by the candidate leaves are appended behind them
__global__ void rm_evict_climb_kernel(RmTree t, int n, int depth_bits, const int* __restrict__ pinned, u64* __restrict__ eff,
                                      int* __restrict__ rem, u64* skey, unsigned* sval, int m_leaves_bound) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int m_leaves = min(t.ctr->evict_n, m_leaves_bound);       // sealed before this launch (rm_evict_sel_seal_kernel)
  if (i >= m_leaves) return;
  const u64 k0 = t.ctr->ev_k0;
  int y = (int)sval[i];
  u64 v = eff[y];
  while (true) {
    const int p = t.parent[y];
    if (p <= 0) break;                      // the root is never evicted
    atomicMax(&eff[p], v);
    __threadfence();
    if (atomicSub(&rem[p], 1) != 1) break;  // a sibling subtree is not (or not yet) among the candidates
    if (t.lock_ref[p] != 0 || pinned[p] != 0) break;
    v = *reinterpret_cast<volatile u64*>(&eff[p]);
    const u64 kp = rm_evict_key(v, t.depth_end[p], depth_bits);
    if (kp > k0) break;
    const int slot = atomicAdd(&t.ctr->ev_total, 1);
    if (slot < n) {
      skey[slot] = kp;
      sval[slot] = (unsigned)p;
    }
    y = p;
  }
}

// One block: walk the sorted candidates in chunks, prefix-sum their value lengths, stop at the first
// position where the running total reaches num_tokens.  Writes the count to ctr->evict_n.
__global__ void rm_evict_cut_kernel(RmTree t, const u64* __restrict__ skey, const unsigned* __restrict__ sval, int m_bound,
                                    i64 num_tokens, int cap) {
  // the candidate count stays on the device; more candidates than were sorted (or nodes the pass did not look at):
  // nothing is evicted and the host repeats the call with exact sizes
  const int n = t.ctr->ev_total;
  if (t.ctr->ev_over != 0 || n > m_bound) {
    if (threadIdx.x == 0) {
      t.ctr->evict_n = 0;
      t.ctr->ev_over = 1;
    }
    return;
  }
#ifdef RM_EMU
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  i64 run = 0;
  int k = 0;
  while (k < n && k < cap && run < num_tokens && skey[k] != RM_NO_REQ) {
    run += t.edge_len[sval[k]];
    ++k;
  }
  t.ctr->evict_n = k;
#else
  // RM_CUT_PER consecutive candidates per thread: a usual call (a few thousand candidates) is one round of the block
  // (one candidate per thread: four block barriers per 1024 candidates, 11 us)
#define RM_CUT_PER 8
  __shared__ i64 warp_tot[32];
  __shared__ i64 carry_s;
  __shared__ int cut_s;
  if (threadIdx.x == 0) {
    carry_s = 0;
    cut_s = RM_INT_MAX;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int lim = min(n, cap);
  for (int base = 0; base < lim; base += 1024 * RM_CUT_PER) {
    const i64 carry = carry_s;
    const int j0 = base + (int)threadIdx.x * RM_CUT_PER;
    i64 v[RM_CUT_PER];
    bool ok[RM_CUT_PER];
    i64 sum = 0;
#pragma unroll
    for (int c = 0; c < RM_CUT_PER; ++c) {
      const int j = j0 + c;
      ok[c] = j < lim && skey[j] != RM_NO_REQ;
      v[c] = ok[c] ? (i64)t.edge_len[sval[j]] : 0;
      sum += v[c];
    }
    i64 inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      i64 o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    if (wid == 0) {
      i64 w = warp_tot[lane], winc = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        i64 o = __shfl_up_sync(0xffffffffu, winc, d);
        if (lane >= d) winc += o;
      }
      warp_tot[lane] = winc - w;
    }
    __syncthreads();
    // the list ends at the first non-candidate; the cut falls after the first candidate whose running total
    // reaches num_tokens
    i64 run = carry + warp_tot[wid] + inc - sum;    // total before this thread's candidates
    int mycut = RM_INT_MAX;
#pragma unroll
    for (int c = 0; c < RM_CUT_PER; ++c) {
      if (mycut == RM_INT_MAX) {
        run += v[c];
        if (!ok[c]) mycut = j0 + c;
        else if (run >= num_tokens) mycut = j0 + c + 1;
      }
    }
    if (mycut != RM_INT_MAX) atomicMin(&cut_s, mycut);
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = carry + warp_tot[wid] + inc;
    const int cut = cut_s;
    __syncthreads();
    if (cut != RM_INT_MAX) break;
  }
  if (threadIdx.x == 0) t.ctr->evict_n = min(cut_s, lim);
#endif
}

// G lanes per evicted node: lane 0 does the bookkeeping, all lanes clear the owner words of the run (a thread clearing
// its ~230 words alone made this kernel 42 us for 4 K nodes: 41 stalls per issue on the store queue)
template <int G>
__device__ __forceinline__ void rm_evict_apply_one(RmTree& t, RmGroup<G>& g, int k, const unsigned* __restrict__ sval,
                                                   const u64* __restrict__ eff, int* out_nodes, i64* out_offs, int* out_lens) {
  const int x = (int)sval[k];
  const int par = t.parent[x];
  const i64 off = t.edge_off[x];
  const int len = t.edge_len[x];
  g.sync();                                                     // every lane has read the record before lane 0 clears it
  for (int k2 = g.lane; k2 < len; k2 += G) rm_set_seg(t, off + k2, -1);   // the run is free again
  if (g.lane != 0) return;
  out_nodes[k] = x;
  out_offs[k] = off;
  out_lens[k] = len;
  RmSlot s;
  int sidx = rm_find<true>(t, rm_key(par, rm_tok(t, off)), s);
  if (sidx >= 0) {
    *reinterpret_cast<volatile u64*>(&t.table[sidx].key) = RM_TOMB_KEY;
    atomicAdd(&t.ctr->n_tomb, 1);
  }
  if (par > 0) atomicMax(&t.stamp[par], eff[x]);            // lazy stamps: the parent keeps what it saw through x
  atomicSub(&t.nchild[par], 1);
  t.parent[x] = -1;
  t.edge_len[x] = 0;
  // _delete_leaf: evictable_size_ -= len(node.key) (radix_cache.py:352)
  rm_add_i64(&t.ctr->evictable, -(i64)len);
  rm_add_i64(&t.ctr->n_nodes, (i64)-1);
  u64 pos = atomicAdd(&t.ctr->free_tail, 1ull);
  t.free_ids[pos % (u64)t.max_nodes] = x;
}

template <int G>
__global__ void rm_evict_apply_kernel(RmTree t, const unsigned* __restrict__ sval, const u64* __restrict__ eff,
                                      int* out_nodes, i64* out_offs, int* out_lens) {
  RmGroup<G> g;
  const int n = t.ctr->evict_n;
  const int stride = (int)((gridDim.x * (unsigned)blockDim.x) / G);       // the grid is sized from a bound, not from n
  for (int k = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G); k < n; k += stride)
    rm_evict_apply_one<G>(t, g, k, sval, eff, out_nodes, out_offs, out_lens);
}

// rm_read_runs: value runs of the arena, packed one after another (a group of lanes per run)
template <int G>
__global__ void rm_copy_runs_kernel(RmTree t, const i64* __restrict__ off, const int* __restrict__ len, const i64* __restrict__ at,
                                    int n, i64* __restrict__ out) {
  RmGroup<G> g;
  const int r = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G);
  if (r >= n) return;
  const i64 o = off[r], a = at[r];
  const int m = len[r];
  for (int x = g.lane; x < m; x += G) out[a + x] = t.val[o + x];
}

int rm_launch_copy_runs(rm_index* idx, const i64* d_off, const int* d_len, const i64* d_at, int n, i64* d_out, cudaStream_t s) {
  if (n <= 0) return RM_OK;
#ifdef RM_EMU
  RM_LAUNCH(rm_copy_runs_kernel<1>, (n + RM_MT_TPB - 1) / RM_MT_TPB, RM_MT_TPB, s, idx->t, d_off, d_len, d_at, n, d_out);
#else
  RM_LAUNCH(rm_copy_runs_kernel<32>, (int)(((long long)n * 32 + RM_MT_TPB - 1) / RM_MT_TPB), RM_MT_TPB, s, idx->t, d_off, d_len, d_at,
            n, d_out);
#endif
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

static int rm_bits_for(u64 v) {
  int b = 1;
  while ((v >> b) != 0 && b < 63) ++b;
  return b;
}

// One pass of rm_evict over the nodes [0, n) with room for m_bound candidates in the exact sort: every count the
// kernels need stays on the device, and the host synchronises once, for the results.  Bounds that turn out too small
// (c.ev_over) leave the tree untouched.
static int rm_evict_pass(rm_index* idx, int n, int m_bound, i64 num_tokens, int cap, int* out_nodes_h, i64* out_offs_h,
                         int* out_lens_h, RmCounters& c, cudaStream_t s) {
  RmTree& t = idx->t;
  RmScratch& sc = idx->sc;
  const int depth_bits = idx->len_bits;
  const int bits = rm_bits_for(idx->op_seq * 2 + 1) + depth_bits + 1;
  int* pinned = sc.ev_lens;   // free until rm_evict_apply_kernel writes the evicted lengths
  const int grid_n = (n + RM_MT_TPB - 1) / RM_MT_TPB;
#ifdef RM_EMU
  const int cut_tpb = 1;
#else
  const int cut_tpb = 1024;
#endif
  RM_CUDA(cudaMemsetAsync(pinned, 0, sizeof(int) * (size_t)n, s));
  RM_CUDA(cudaMemsetAsync(sc.ev_rem, 0, sizeof(int) * (size_t)n, s));
  RM_LAUNCH(rm_evict_count_kernel, grid_n, RM_MT_TPB, s, t, n, sc.ev_eff, sc.ev_rem, pinned);
  RM_LAUNCH(rm_evict_keys_kernel, grid_n, RM_MT_TPB, s, t, n, depth_bits, pinned, sc.ev_rem, sc.ev_key[0], sc.ev_val[0]);
  // 1. frontier key by weighted radix select over the leaves' keys, 2. the leaves at or below it, compacted
  RmEvictSel* st = reinterpret_cast<RmEvictSel*>(sc.ev_selw + RM_SEL_BINS);
  const int levels = (bits + RM_SEL_BITS - 1) / RM_SEL_BITS;
  const int pad_bits = levels * RM_SEL_BITS - bits;                 // the key is padded at the LOW end of the last digit
  RM_LAUNCH(rm_evict_sel_init_kernel, 1, RM_MT_TPB, s, t, n, st, num_tokens, sc.ev_selw, sc.ev_selc);
#ifndef RM_EMU
  // blocks of the select: four per SM, more when a block's share of the keys times the longest edge would not fit
  // its 32-bit shared bins
  int sel_grid = grid_n < 148 * 4 ? grid_n : 148 * 4;
  {
    long long per_block = (1ll << 31) / (long long)(t.max_key_len > 0 ? t.max_key_len : 1);
    if (per_block < RM_MT_TPB) per_block = RM_MT_TPB;
    const long long need_blocks = ((long long)n + per_block - 1) / per_block;
    if (need_blocks > sel_grid) sel_grid = (int)(need_blocks < grid_n ? need_blocks : grid_n);
  }
#endif
  for (int l = 0; l < levels; ++l) {
    const int shift = (levels - 1 - l) * RM_SEL_BITS - pad_bits;    // may be negative on the last level: keys are shifted up instead
#ifdef RM_EMU
    RM_LAUNCH(rm_evict_sel_hist_kernel, grid_n, RM_MT_TPB, s, t, sc.ev_key[0], n, shift >= 0 ? shift : 0, shift >= 0 ? 0 : -shift, st,
              sc.ev_selw);
    RM_LAUNCH(rm_evict_sel_pick_kernel, 1, RM_SEL_PICK_TPB, s, st, sc.ev_selw, sc.ev_selc);
#else
    RM_LAUNCH(rm_evict_sel_level_kernel, sel_grid, RM_MT_TPB, s, t, sc.ev_key[0], n, shift >= 0 ? shift : 0,
              shift >= 0 ? 0 : -shift, st, sc.ev_selw);
#endif
  }
  int w = 1;
  RM_LAUNCH(rm_evict_sel_compact_kernel, grid_n, RM_MT_TPB, s, t, sc.ev_key[0], n, st, pad_bits, sc.ev_key[w], sc.ev_val[w]);
  RM_LAUNCH(rm_evict_sel_seal_kernel, 1, 1, s, t);
  // 3. climb, 4. exact order and cut over the candidates: ctr->ev_total of them, at most m_bound sorted
  RM_LAUNCH(rm_evict_climb_kernel, grid_n, RM_MT_TPB, s, t, n, depth_bits, pinned, sc.ev_eff, sc.ev_rem, sc.ev_key[w], sc.ev_val[w], n);
  {
    const int w2 = rm_sort_pairs(sc.ev_key[w], sc.ev_val[w], sc.ev_key[w ^ 1], sc.ev_val[w ^ 1], m_bound, bits, sc.ev_hist, s,
                                 &t.ctr->ev_total);
    w ^= w2;
  }
  RM_LAUNCH(rm_evict_cut_kernel, 1, cut_tpb, s, t, sc.ev_key[w], sc.ev_val[w], m_bound, num_tokens, cap);
  // outputs are staged in the other ping-pong buffer (large enough: n entries of 8/4 bytes)
  int* d_nodes = reinterpret_cast<int*>(sc.ev_val[w ^ 1]);
  i64* d_offs = reinterpret_cast<i64*>(sc.ev_key[w ^ 1]);
  int* d_lens = sc.ev_lens;
  const int lim = min(m_bound, cap);
#ifdef RM_EMU
  RM_LAUNCH(rm_evict_apply_kernel<1>, (lim + RM_MT_TPB - 1) / RM_MT_TPB, RM_MT_TPB, s, t, sc.ev_val[w], sc.ev_eff, d_nodes, d_offs,
            d_lens);
#else
  {
    // a warp per evicted node, over a grid that covers a usual call (a few thousand nodes) in one round
    const long long want = ((long long)lim * 32 + RM_MT_TPB - 1) / RM_MT_TPB;
    const int grid = (int)(want < 148 * 16 ? want : 148 * 16);
    RM_LAUNCH(rm_evict_apply_kernel<32>, grid, RM_MT_TPB, s, t, sc.ev_val[w], sc.ev_eff, d_nodes, d_offs, d_lens);
  }
#endif
  RM_CUDA(cudaGetLastError());
  // the count and the first results travel together through pinned memory: one synchronisation for the usual call
  const int first = lim < RM_EVICT_FIRST ? lim : RM_EVICT_FIRST;
  char* mail = static_cast<char*>(idx->mail);
  int* m_nodes = reinterpret_cast<int*>(mail + 4096);
  int* m_lens = m_nodes + RM_EVICT_FIRST;
  i64* m_offs = reinterpret_cast<i64*>(m_lens + RM_EVICT_FIRST);
  RM_CUDA(cudaMemcpyAsync(mail, t.ctr, sizeof(c), cudaMemcpyDeviceToHost, s));
  if (first > 0) {
    RM_CUDA(cudaMemcpyAsync(m_nodes, d_nodes, sizeof(int) * first, cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaMemcpyAsync(m_lens, d_lens, sizeof(int) * first, cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaMemcpyAsync(m_offs, d_offs, sizeof(i64) * first, cudaMemcpyDeviceToHost, s));
  }
  RM_CUDA(cudaStreamSynchronize(s));
  memcpy(&c, mail, sizeof(c));
  if (!c.ev_over) {
    const int got = c.evict_n < first ? c.evict_n : first;
    memcpy(out_nodes_h, m_nodes, sizeof(int) * (size_t)got);
    memcpy(out_lens_h, m_lens, sizeof(int) * (size_t)got);
    memcpy(out_offs_h, m_offs, sizeof(i64) * (size_t)got);
  }
  if (c.ev_over) return RM_OK;
  const int k = c.evict_n;
  if (k > first) {
    RM_CUDA(cudaMemcpyAsync(out_nodes_h + first, d_nodes + first, sizeof(int) * (k - first), cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaMemcpyAsync(out_offs_h + first, d_offs + first, sizeof(i64) * (k - first), cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaMemcpyAsync(out_lens_h + first, d_lens + first, sizeof(int) * (k - first), cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaStreamSynchronize(s));
  }
  return RM_OK;
}

// Candidates the first pass sorts: an eviction that frees a few batches' worth of tokens takes a few thousand nodes
#define RM_EVICT_M_OPT 65536

int rm_launch_evict(rm_index* idx, i64 num_tokens, int cap, int* out_nodes_h, i64* out_offs_h, int* out_lens_h,
                    int* out_n_h, cudaStream_t s) {
  RmTree& t = idx->t;
  idx->op_seq += 1;   // evict is one operation of the logical clock (it stamps nothing)
  *out_n_h = 0;
  if (num_tokens <= 0 || cap <= 0) return RM_OK;
  // No question to the device before the work: the node high-water mark is bounded from the last exact reading (an
  // operation of the logical clock creates at most three nodes: split, leaf, parked copy), the candidate count by
  // RM_EVICT_M_OPT (environment override: tests force the second pass).  The pass checks both on the device and
  // evicts nothing if either was too small; the call then runs again with exact sizes.
  long long nb = (long long)idx->hwm_exact + 3ll * (long long)(idx->op_seq - idx->hwm_seq);
  if (nb > t.max_nodes) nb = t.max_nodes;
  int n = (int)nb;
  if (n <= 1) return RM_OK;
  RmCounters c;
  int m_opt = RM_EVICT_M_OPT;
  if (const char* e = getenv("RM_EVICT_M_OPT")) m_opt = atoi(e) > 1 ? atoi(e) : 2;
  int rc = rm_evict_pass(idx, n, n < m_opt ? n : m_opt, num_tokens, cap, out_nodes_h, out_offs_h, out_lens_h, c, s);
  if (rc != RM_OK) return rc;
  if (c.ev_over) {
    n = c.node_hwm;
    rc = rm_evict_pass(idx, n, n, num_tokens, cap, out_nodes_h, out_offs_h, out_lens_h, c, s);
    if (rc != RM_OK) return rc;
    if (c.ev_over) return RM_E_NOSPACE;   // not reachable: the sizes were exact
  }
  idx->hwm_exact = c.node_hwm;
  idx->hwm_seq = idx->op_seq;
  *out_n_h = c.evict_n;
  if ((u64)c.n_tomb * 4 > (u64)t.table_mask + 1) return rm_table_rebuild(idx, c.node_hwm, s);
  return RM_OK;
}

// ------------------------------------------------------------------ K7 duplicate-node table scan
__global__ void rm_dup_scan_kernel(RmTree t, int only_candidates, int cap, int* out, int* out_n) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > t.dup_mask) return;
  u64 k = t.dup_key[i];
  if (k == RM_EMPTY_KEY || k == RM_TOMB_KEY) return;   // tombstone: an entry rm_gc_exec_h removed
  int parked = t.dup_parked[i];
  int lk = parked >= 0 ? t.lock_ref[parked] : 0;
  if (only_candidates && !(parked >= 0 && lk == 0)) return;   // radix_mesh.py:152-156
  int slot = atomicAdd(out_n, 1);
  if (slot >= cap) return;
  out[slot * 4 + 0] = (int)(k >> 32);
  out[slot * 4 + 1] = (int)(unsigned)(k & 0xffffffffu) - 1;
  out[slot * 4 + 2] = parked;
  out[slot * 4 + 3] = lk;
}

int rm_launch_dups(rm_index* idx, int only_candidates, int cap, int* out_h, int* out_n_h, cudaStream_t s) {
  RmTree& t = idx->t;
  RmScratch& sc = idx->sc;
  *out_n_h = 0;
  if (cap <= 0) return RM_OK;
  int dcap = (int)min((i64)cap, (i64)t.dup_mask + 1);
  RM_CUDA(cudaMemsetAsync(sc.dup_out_n, 0, sizeof(int), s));
  RM_LAUNCH(rm_dup_scan_kernel, (int)((t.dup_mask + RM_MT_TPB) / RM_MT_TPB), RM_MT_TPB, s, t, only_candidates, dcap, sc.dup_out, sc.dup_out_n);
  RM_CUDA(cudaGetLastError());
  int n = 0;
  RM_CUDA(cudaMemcpyAsync(&n, sc.dup_out_n, sizeof(int), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  if (n > dcap) n = dcap;
  if (n > 0) {
    RM_CUDA(cudaMemcpyAsync(out_h, sc.dup_out, sizeof(int) * 4 * n, cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaStreamSynchronize(s));
  }
  *out_n_h = n;
  return RM_OK;
}

// ------------------------------------------------------------------ duplicate-node GC: agree / execute
// RadixMesh._gc_handle (radix_mesh.py:362-389).  A key travels between replicas as (full key tokens, rank); on a
// replica it is located read-only (the walk kernel, no stamps, no split requests) and stands for (tree node id, rank)
// in the duplicate table.
__device__ int rm_dup_find(const RmTree& t, int node, int rank) {
  const u64 key = ((u64)(unsigned)node << 32) | (unsigned)(rank + 1);
  unsigned i = (rm_hash2((unsigned)(key >> 32), (unsigned)key, 0) >> 7) & t.dup_mask;
  for (unsigned probes = 0; probes <= t.dup_mask; ++probes) {
    const u64 cur = *reinterpret_cast<volatile u64*>(&t.dup_key[i]);
    if (cur == key) return (int)i;
    if (cur == RM_EMPTY_KEY) return -1;
    i = (i + 1) & t.dup_mask;
  }
  return -1;
}

// the node whose full key IS the query (matched to its end, at a node boundary), else -1
__global__ void rm_gc_locate_kernel(RmTree t, const i64* __restrict__ off, int n, const int* __restrict__ wlen,
                                    const int* __restrict__ wnode, int* __restrict__ out_node) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int L = (int)(off[i + 1] - off[i]);
  const int node = wnode[i];
  out_node[i] = (L > 0 && wlen[i] == L && node > 0 && t.depth_end[node] == L) ? node : -1;
}

// radix_mesh.py:386-387: key in dup_nodes and (entry is None or entry.lock_ref == 0)
__global__ void rm_gc_agree_kernel(RmTree t, const int* __restrict__ nodes, const int* __restrict__ ranks, int n,
                                   unsigned char* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned char a = 0;
  if (nodes[i] > 0) {
    const int sl = rm_dup_find(t, nodes[i], ranks[i]);
    if (sl >= 0) {
      const int parked = t.dup_parked[sl];
      a = (parked < 0 || t.lock_ref[parked] == 0) ? 1 : 0;
    }
  }
  out[i] = a;
}

// GC_EXEC: dup_nodes.pop(k, None) (radix_mesh.py:363-366).  On the replica that asked (free_parked), the parked
// node's run is handed back (the caller frees its KV indices, :375-377) and the node is released.  One thread:
// a GC round is a handful of keys, and keys of one tree node share its chain of parked nodes.
__global__ void rm_gc_exec_kernel(RmTree t, const int* __restrict__ nodes, const int* __restrict__ ranks, int n,
                                  int free_parked, i64* __restrict__ out_off, int* __restrict__ out_len) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  for (int i = 0; i < n; ++i) {
    out_off[i] = -1;
    out_len[i] = 0;
    if (nodes[i] <= 0) continue;
    const int sl = rm_dup_find(t, nodes[i], ranks[i]);
    if (sl < 0) continue;
    const int d = t.dup_parked[sl];
    t.dup_key[sl] = RM_TOMB_KEY;
    t.dup_parked[sl] = -1;
    t.ctr->n_dups -= 1;
    if (d < 0 || !free_parked) continue;
    const i64 off = t.edge_off[d];
    const int len = t.edge_len[d];
    out_off[i] = off;
    out_len[i] = len;
    // unlink d from the chain of parked predecessors of its tree node
    for (int x = nodes[i], guard = 0; x >= 0 && guard < t.max_nodes; ++guard) {
      const int nx = t.parked[x];
      if (nx == d) {
        t.parked[x] = t.parked[d];
        break;
      }
      x = nx;
    }
    for (int k = 0; k < len; ++k)
      if ((rm_seg(t, off + k) & 0x3FFFFFFF) == d) rm_set_seg(t, off + k, -1);
    t.parked[d] = -1;
    t.parent[d] = -1;
    t.edge_len[d] = 0;
    t.gen[d] = 0;
    const u64 pos = atomicAdd(&t.ctr->free_tail, 1ull);
    t.free_ids[pos % (u64)t.max_nodes] = d;
  }
}

// full key of tree nodes, one row of `stride` tokens per node
__global__ void rm_node_key_kernel(RmTree t, const int* __restrict__ nodes, int n, int stride, int* __restrict__ out_tok,
                                   int* __restrict__ out_len) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int v = nodes[i];
  const int L = (v > 0 && t.parent[v] >= 0) ? t.depth_end[v] : 0;
  out_len[i] = L;
  if (L > stride) return;
  int* __restrict__ row = out_tok + (size_t)i * stride;
  while (v > 0) {
    const int e = t.edge_len[v];
    const i64 off = t.edge_off[v];
    const int at = t.depth_end[v] - e;
    for (int k = 0; k < e; ++k) row[at + k] = rm_tok(t, off + k);
    v = t.parent[v];
  }
}

int rm_launch_walk_readonly(rm_index* idx, const int* tokens, const i64* offsets, int N, int* out_len, int* out_node,
                            cudaStream_t s);

struct RmTmp {          // device scratch of a control-plane call (GC rounds are rare: plain cudaMalloc / cudaFree)
  void* p[8];
  int n = 0;
  template <class T>
  cudaError_t get(T** out, size_t count) {
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(out), sizeof(T) * (count ? count : 1));
    if (e == cudaSuccess) p[n++] = *out;
    return e;
  }
  ~RmTmp() {
    for (int i = 0; i < n; ++i) cudaFree(p[i]);
  }
};

// tokens/offsets/ranks on the host -> device copies + located node ids
static int rm_gc_stage(rm_index* idx, RmTmp& tmp, const int* tokens_h, const i64* offsets_h, const int* ranks_h, int n,
                       int** d_nodes, int** d_ranks, cudaStream_t s) {
  const i64 total = offsets_h[n] - offsets_h[0];
  int *d_tok, *d_len;
  i64* d_off;
  RM_CUDA(tmp.get(&d_tok, (size_t)total));
  RM_CUDA(tmp.get(&d_off, (size_t)n + 1));
  RM_CUDA(tmp.get(&d_len, (size_t)n));
  RM_CUDA(tmp.get(d_nodes, (size_t)n));
  RM_CUDA(tmp.get(d_ranks, (size_t)n));
  int* d_wnode;
  RM_CUDA(tmp.get(&d_wnode, (size_t)n));
  RM_CUDA(cudaMemcpyAsync(d_tok, tokens_h + offsets_h[0], sizeof(int) * (size_t)total, cudaMemcpyHostToDevice, s));
  std::vector<i64> rel((size_t)n + 1);
  for (int i = 0; i <= n; ++i) rel[i] = offsets_h[i] - offsets_h[0];
  RM_CUDA(cudaMemcpyAsync(d_off, rel.data(), sizeof(i64) * ((size_t)n + 1), cudaMemcpyHostToDevice, s));
  RM_CUDA(cudaMemcpyAsync(*d_ranks, ranks_h, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, s));
  RM_CUDA(cudaStreamSynchronize(s));     // rel[] goes out of scope
  int rc = rm_launch_walk_readonly(idx, d_tok, d_off, n, d_len, d_wnode, s);
  if (rc) return rc;
  RM_LAUNCH(rm_gc_locate_kernel, (n + RM_MT_TPB - 1) / RM_MT_TPB, RM_MT_TPB, s, idx->t, d_off, n, d_len, d_wnode, *d_nodes);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_launch_gc_agree(rm_index* idx, const int* tokens_h, const i64* offsets_h, const int* ranks_h, int n,
                       unsigned char* out_agree_h, cudaStream_t s) {
  if (n <= 0) return RM_OK;
  RmTmp tmp;
  int *d_nodes, *d_ranks;
  int rc = rm_gc_stage(idx, tmp, tokens_h, offsets_h, ranks_h, n, &d_nodes, &d_ranks, s);
  if (rc) return rc;
  unsigned char* d_out;
  RM_CUDA(tmp.get(&d_out, (size_t)n));
  RM_LAUNCH(rm_gc_agree_kernel, (n + RM_MT_TPB - 1) / RM_MT_TPB, RM_MT_TPB, s, idx->t, d_nodes, d_ranks, n, d_out);
  RM_CUDA(cudaGetLastError());
  RM_CUDA(cudaMemcpyAsync(out_agree_h, d_out, (size_t)n, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  return RM_OK;
}

int rm_launch_gc_exec(rm_index* idx, const int* tokens_h, const i64* offsets_h, const int* ranks_h, int n, int free_parked,
                      i64* out_off_h, int* out_len_h, cudaStream_t s) {
  if (n <= 0) return RM_OK;
  RmTmp tmp;
  int *d_nodes, *d_ranks;
  int rc = rm_gc_stage(idx, tmp, tokens_h, offsets_h, ranks_h, n, &d_nodes, &d_ranks, s);
  if (rc) return rc;
  i64* d_off;
  int* d_len;
  RM_CUDA(tmp.get(&d_off, (size_t)n));
  RM_CUDA(tmp.get(&d_len, (size_t)n));
  RM_LAUNCH(rm_gc_exec_kernel, 1, 1, s, idx->t, d_nodes, d_ranks, n, free_parked, d_off, d_len);
  RM_CUDA(cudaGetLastError());
  RM_CUDA(cudaMemcpyAsync(out_off_h, d_off, sizeof(i64) * (size_t)n, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaMemcpyAsync(out_len_h, d_len, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  return RM_OK;
}

int rm_launch_gc_keys(rm_index* idx, const int* nodes_h, int n, int stride, int* out_tokens_h, int* out_lens_h,
                      cudaStream_t s) {
  if (n <= 0) return RM_OK;
  RmTmp tmp;
  int *d_nodes, *d_tok, *d_len;
  RM_CUDA(tmp.get(&d_nodes, (size_t)n));
  RM_CUDA(tmp.get(&d_tok, (size_t)n * (size_t)stride));
  RM_CUDA(tmp.get(&d_len, (size_t)n));
  RM_CUDA(cudaMemcpyAsync(d_nodes, nodes_h, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, s));
  RM_LAUNCH(rm_node_key_kernel, (n + RM_MT_TPB - 1) / RM_MT_TPB, RM_MT_TPB, s, idx->t, d_nodes, n, stride, d_tok, d_len);
  RM_CUDA(cudaGetLastError());
  RM_CUDA(cudaMemcpyAsync(out_tokens_h, d_tok, sizeof(int) * (size_t)n * (size_t)stride, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaMemcpyAsync(out_lens_h, d_len, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  return RM_OK;
}

// ------------------------------------------------------------------ arena compaction + table rebuild
// Eviction frees node ids and tombstones table slots, but the edge arenas are bump allocated: under
// streaming inserts at steady occupancy (BASELINE configs[2]) the space of evicted runs has to be
// reclaimed.  rm_compact moves every live run (tree nodes and parked duplicates) to the front of fresh
// arenas, heavy paths contiguous (below), rewrites edge_off and the owner map, and rebuilds the child
// table from the node pool (which also drops the tombstones).  A maintenance call: it allocates the new arenas and synchronises.
__device__ __forceinline__ bool rm_node_has_run(const RmTree& t, int x) {
  return x > 0 && (t.parent[x] >= 0 || t.gen[x] < 0) && t.edge_len[x] > 0;
}

// Layout of the new arenas: HEAVY PATHS are contiguous.  A walk follows the arena for as long as the next node of
// its path starts where the previous one ended (RM_SEG_CONT) and goes back to the child table otherwise; bump
// allocation only gives that adjacency to the halves of a split, so a query that runs down a hot document pays a
// table probe at every point where a later insert extended it.  Compaction therefore lays every node out directly
// behind its parent when it is the parent's heaviest child (most nodes below it: where most lookups go), top-down
// along each such chain, and flags it contiguous.  Three passes over the node pool: subtree weights (every node adds
// itself to its ancestors), heaviest child per node (atomic max of (weight, id)), chain head per node (climb while I
// am the heaviest child).  The sort key (chain head, depth) then puts chains together, top first.
__device__ __forceinline__ bool rm_is_tree_node(const RmTree& t, int x) { return x > 0 && t.parent[x] >= 0 && t.gen[x] >= 0; }

__global__ void rm_compact_weight_kernel(RmTree t, int n, int* __restrict__ weight) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n || !rm_is_tree_node(t, x)) return;
  atomicAdd(&weight[x], 1);
  for (int p = t.parent[x], guard = 0; p > 0 && guard < t.max_nodes; p = t.parent[p], ++guard) atomicAdd(&weight[p], 1);
}

__global__ void rm_compact_heavy_kernel(RmTree t, int n, const int* __restrict__ weight, u64* __restrict__ best) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n || !rm_is_tree_node(t, x)) return;
  const int p = t.parent[x];
  if (p > 0) atomicMax(&best[p], ((u64)(unsigned)weight[x] << 32) | (unsigned)x);
}

__device__ __forceinline__ bool rm_is_heavy_child(const RmTree& t, const u64* __restrict__ best, int x) {
  const int p = t.parent[x];
  return p > 0 && (int)(unsigned)best[p] == x;
}

__global__ void rm_compact_keys_kernel(RmTree t, int n, const u64* __restrict__ best, u64* __restrict__ key,
                                       unsigned* __restrict__ val) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n) return;
  val[x] = (unsigned)x;
  if (!rm_node_has_run(t, x)) {
    key[x] = (1ull << 62) - 1ull;
    return;
  }
  int head = x;
  if (rm_is_tree_node(t, x))
    for (int guard = 0; rm_is_heavy_child(t, best, head) && guard < t.max_nodes; ++guard) head = t.parent[head];
  key[x] = ((u64)(unsigned)head << 32) | (unsigned)t.depth_end[x];   // parked duplicates: a chain of their own
}

__global__ void rm_compact_len_kernel(RmTree t, int n, const unsigned* __restrict__ sval, i64* __restrict__ len_out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int x = (int)sval[i];
  len_out[i] = rm_node_has_run(t, x) ? (i64)t.edge_len[x] : 0;
}

// exclusive scan of n i64 values in place, one block (maintenance path: simplicity over speed)
__global__ void rm_scan_i64_kernel(i64* __restrict__ data, int n, i64* __restrict__ total) {
#ifdef RM_EMU
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  i64 run = 0;
  for (int i = 0; i < n; ++i) {
    i64 v = data[i];
    data[i] = run;
    run += v;
  }
  *total = run;
#else
  __shared__ i64 warp_tot[32];
  __shared__ i64 carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int base = 0; base < n; base += 1024) {
    const i64 carry = carry_s;
    const int j = base + threadIdx.x;
    const i64 v = j < n ? data[j] : 0;
    i64 inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      i64 o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    if (wid == 0) {
      i64 w = warp_tot[lane], winc = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        i64 o = __shfl_up_sync(0xffffffffu, winc, d);
        if (lane >= d) winc += o;
      }
      warp_tot[lane] = winc - w;
    }
    __syncthreads();
    const i64 incl = carry + warp_tot[wid] + inc;
    if (j < n) data[j] = incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry_s;
#endif
}

// one group of lanes per run, in arena order: copy tokens, values and owner map, then point the node at
// its new place (only this group reads edge_off[x])
template <int G>
__global__ void rm_compact_copy_kernel(RmTree t, int n, const unsigned* __restrict__ sval, const i64* __restrict__ new_off,
                                       const u64* __restrict__ best, int2* __restrict__ nts, i64* __restrict__ nval,
                                       int with_values) {
  RmGroup<G> g;
  const int i = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G);
  if (i >= n) return;
  const int x = (int)sval[i];
  if (!rm_node_has_run(t, x)) return;
  const i64 src = t.edge_off[x], dst = new_off[i];
  const int len = t.edge_len[x];
  // owner words are rewritten: the node's id, flagged contiguous on the first position iff the node sits directly
  // behind its parent in the new layout (heaviest child of a parent that has a run)
  const int first = (rm_is_tree_node(t, x) && rm_is_heavy_child(t, best, x)) ? (x | RM_SEG_CONT) : x;
  for (int k = g.lane; k < len; k += G) {
    nts[dst + k] = make_int2(t.ts[src + k].x, k == 0 ? first : x);
    if (with_values) nval[dst + k] = t.val[src + k];
  }
  g.sync();
  if (g.lane == 0) t.edge_off[x] = dst;
}

// after the arenas were swapped: one table slot per tree node
__global__ void rm_compact_relink_kernel(RmTree t, int n) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n || !rm_node_has_run(t, x)) return;
  if (t.gen[x] >= 0) {   // parked duplicates are not reachable from the table
    const i64 off = t.edge_off[x];
    rm_table_insert(t, rm_key(t.parent[x], rm_tok(t, off)), x, t.edge_len[x], off, t.rank[x]);
  }
}

// Rebuild the child table from the node pool: one slot per live tree node, no tombstones.  The reference deletes a
// child from a dict (radix_cache.py:347-352) and the dict shrinks; an open-addressed table only gets its empty slots
// back this way.  rm_evict calls it once tombstones hold a quarter of the table; rm_compact always does.
int rm_table_rebuild(rm_index* idx, int n_nodes_hwm, cudaStream_t s) {
  RmTree& t = idx->t;
  RM_CUDA(cudaMemsetAsync(t.table, 0xFF, sizeof(RmSlot) * ((size_t)t.table_mask + 1), s));
  RM_CUDA(cudaMemsetAsync(&t.ctr->n_tomb, 0, sizeof(int), s));
  if (n_nodes_hwm > 0) RM_LAUNCH(rm_compact_relink_kernel, (n_nodes_hwm + RM_MT_TPB - 1) / RM_MT_TPB, RM_MT_TPB, s, t, n_nodes_hwm);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

static int rm_bits_u64(u64 v) {
  int b = 1;
  while ((v >> b) != 0 && b < 63) ++b;
  return b;
}

extern "C" int rm_compact(rm_index* idx, void* stream) {
  if (!idx) return RM_E_INVALID;
  RM_CUDA(cudaSetDevice(idx->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  RmTree& t = idx->t;
  RmScratch& sc = idx->sc;
  RmCounters c;
  RM_CUDA(cudaMemcpyAsync(&c, t.ctr, sizeof(c), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  const int n = c.node_hwm;
  const bool with_values = idx->cfg.mode != RM_MODE_ROUTER;
  // runs in heavy-path order
  int* weight = sc.ev_rem;
  u64* best = sc.ev_eff;
  const int grid_n = (n + RM_MT_TPB - 1) / RM_MT_TPB;
  RM_CUDA(cudaMemsetAsync(weight, 0, sizeof(int) * (size_t)n, s));
  RM_CUDA(cudaMemsetAsync(best, 0, sizeof(u64) * (size_t)n, s));
  RM_LAUNCH(rm_compact_weight_kernel, grid_n, RM_MT_TPB, s, t, n, weight);
  RM_LAUNCH(rm_compact_heavy_kernel, grid_n, RM_MT_TPB, s, t, n, weight, best);
  RM_LAUNCH(rm_compact_keys_kernel, grid_n, RM_MT_TPB, s, t, n, best, sc.ev_key[0], sc.ev_val[0]);
  const int w = rm_sort_pairs(sc.ev_key[0], sc.ev_val[0], sc.ev_key[1], sc.ev_val[1], n, 62, sc.ev_hist, s);
  i64* new_off = reinterpret_cast<i64*>(sc.ev_key[w ^ 1]);   // the spare ping-pong buffer, n x 8 bytes
  i64* total_d = reinterpret_cast<i64*>(sc.sufmax);
  RM_LAUNCH(rm_compact_len_kernel, (n + RM_MT_TPB - 1) / RM_MT_TPB, RM_MT_TPB, s, t, n, sc.ev_val[w], new_off);
#ifdef RM_EMU
  RM_LAUNCH(rm_scan_i64_kernel, 1, 1, s, new_off, n, total_d);
#else
  RM_LAUNCH(rm_scan_i64_kernel, 1, 1024, s, new_off, n, total_d);
#endif
  int2* nts = nullptr;
  i64* nval = nullptr;
  const size_t cap = (size_t)t.max_tokens + RM_ARENA_PAD;
  RM_CUDA(cudaMalloc(reinterpret_cast<void**>(&nts), sizeof(int2) * cap));
  if (with_values) RM_CUDA(cudaMalloc(reinterpret_cast<void**>(&nval), sizeof(i64) * cap));
  RM_CUDA(cudaMemsetAsync(nts, 0xFF, sizeof(int2) * cap, s));
#ifdef RM_EMU
  auto kc = rm_compact_copy_kernel<1>;
  const long long cthreads = n;
#else
  auto kc = rm_compact_copy_kernel<32>;
  const long long cthreads = (long long)n * 32;
#endif
  RM_LAUNCH(kc, (int)((cthreads + RM_MT_TPB - 1) / RM_MT_TPB), RM_MT_TPB, s, t, n, sc.ev_val[w], new_off, best, nts, nval,
            with_values ? 1 : 0);
  RM_CUDA(cudaGetLastError());
  RM_CUDA(cudaStreamSynchronize(s));
  // swap the arenas in, drop the old ones
  cudaFree(t.ts);
  t.ts = nts;
  if (with_values) {
    cudaFree(t.val);
    t.val = nval;
  }
  RM_CUDA(cudaMemsetAsync(t.table, 0xFF, sizeof(RmSlot) * ((size_t)t.table_mask + 1), s));
  RM_CUDA(cudaMemsetAsync(&t.ctr->n_tomb, 0, sizeof(int), s));
  RM_LAUNCH(rm_compact_relink_kernel, (n + RM_MT_TPB - 1) / RM_MT_TPB, RM_MT_TPB, s, t, n);
  i64 total = 0;
  RM_CUDA(cudaMemcpyAsync(&total, total_d, sizeof(i64), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  const u64 top = (u64)total;
  RM_CUDA(cudaMemcpyAsync(&t.ctr->arena_top, &top, sizeof(u64), cudaMemcpyHostToDevice, s));
  RM_CUDA(cudaStreamSynchronize(s));
  RM_CUDA(cudaGetLastError());
  (void)rm_bits_u64;
  return RM_OK;
}

// ------------------------------------------------------------------ TreeNode proxy support (O(1) in the tree size)
// The drop-in's single-key surface (RadixCache.match_prefix -> MatchResult.last_device_node, TreeNode.parent /
// .lock_ref / .children / .key / .value; radix_cache.py:35-64) reads the handful of nodes it touches, not the pool.
__global__ void rm_node_info_kernel(RmTree t, const int* __restrict__ ids, int n, rm_node_info_t* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int x = ids[i];
  rm_node_info_t r;
  if (x < 0 || x >= t.max_nodes) {
    r.parent = -2;
    r.edge_len = r.lock_ref = r.rank = r.depth_end = r.nchild = r.gen = 0;
    r.parked = -1;
    r.edge_off = 0;
    r.stamp = 0;
  } else {
    r.parent = t.parent[x];
    r.edge_len = t.edge_len[x];
    r.lock_ref = t.lock_ref[x];
    r.rank = t.rank[x];
    r.depth_end = t.depth_end[x];
    r.nchild = t.nchild[x];
    r.gen = t.gen[x];
    r.parked = t.parked[x];
    r.edge_off = t.edge_off[x];
    r.stamp = t.stamp[x];
  }
  out[i] = r;
}

// children of one node: (child id, first token of its edge, generation, dict-order key) per child, unordered
__global__ void rm_node_children_kernel(RmTree t, int n, int node, int cap, int* __restrict__ out_id, int* __restrict__ out_tok,
                                        int* __restrict__ out_gen, u64* __restrict__ out_ord) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= n || x == 0 || t.parent[x] != node || t.gen[x] < 0 || t.edge_len[x] <= 0) return;
  const int k = atomicAdd(&t.ctr->n_kids, 1);
  if (k >= cap) return;
  out_id[k] = x;
  out_tok[k] = rm_tok(t, t.edge_off[x]);
  out_gen[k] = t.gen[x];
  out_ord[k] = t.ord[x];
}

// RadixCache.total_size (radix_cache.py:354-364): sum of len(value) over the tree
__global__ void rm_total_size_kernel(RmTree t, int n, u64* __restrict__ out) {
  i64 local = 0;
  for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < n; x += gridDim.x * blockDim.x)
    if (x > 0 && t.parent[x] >= 0 && t.gen[x] >= 0) local += t.mode == RM_MODE_ROUTER ? 1 : t.edge_len[x];
#ifndef RM_EMU
  for (int d = 16; d > 0; d >>= 1) local += __shfl_xor_sync(0xffffffffu, local, d);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(out, (u64)local);
#else
  if (local) atomicAdd(out, (u64)local);
#endif
}

#define RM_INFO_CHUNK 256
int rm_launch_node_info(rm_index* idx, const int* ids_h, int n, rm_node_info_t* out_h, cudaStream_t s) {
  RmScratch& sc = idx->sc;
  for (int a = 0; a < n; a += RM_INFO_CHUNK) {
    const int m = n - a < RM_INFO_CHUNK ? n - a : RM_INFO_CHUNK;
    RM_CUDA(cudaMemcpyAsync(sc.info_ids, ids_h + a, sizeof(int) * (size_t)m, cudaMemcpyHostToDevice, s));
    RM_LAUNCH(rm_node_info_kernel, 1, RM_INFO_CHUNK, s, idx->t, sc.info_ids, m, sc.info_rows);
    RM_CUDA(cudaMemcpyAsync(out_h + a, sc.info_rows, sizeof(rm_node_info_t) * (size_t)m, cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaStreamSynchronize(s));
  }
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_launch_node_info_dev(rm_index* idx, const int* ids_d, int n, cudaStream_t s) {
  RM_LAUNCH(rm_node_info_kernel, 1, RM_INFO_CHUNK, s, idx->t, ids_d, n, idx->sc.info_rows);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_launch_node_children(rm_index* idx, int node, int cap, int* ids_h, int* toks_h, int* gens_h, u64* ords_h, int* out_n_h,
                            cudaStream_t s) {
  RmTree& t = idx->t;
  RmScratch& sc = idx->sc;
  RmCounters c;
  RM_CUDA(cudaMemcpyAsync(&c, t.ctr, sizeof(c), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  const int n = c.node_hwm;
  RM_CUDA(cudaMemsetAsync(&t.ctr->n_kids, 0, sizeof(int), s));
  // the eviction scratch is free between calls (one host thread at a time per index): ids / tokens / gens / order keys
  int* d_id = reinterpret_cast<int*>(sc.ev_val[0]);
  int* d_tok = reinterpret_cast<int*>(sc.ev_val[1]);
  int* d_gen = sc.ev_lens;
  u64* d_ord = sc.ev_key[0];
  RM_LAUNCH(rm_node_children_kernel, (n + RM_MT_TPB - 1) / RM_MT_TPB, RM_MT_TPB, s, t, n, node, t.max_nodes, d_id, d_tok, d_gen, d_ord);
  RM_CUDA(cudaGetLastError());
  RM_CUDA(cudaMemcpyAsync(&c, t.ctr, sizeof(c), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  int k = c.n_kids;
  *out_n_h = k;
  if (k > cap) k = cap;
  if (k > 0) {
    RM_CUDA(cudaMemcpyAsync(ids_h, d_id, sizeof(int) * (size_t)k, cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaMemcpyAsync(toks_h, d_tok, sizeof(int) * (size_t)k, cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaMemcpyAsync(gens_h, d_gen, sizeof(int) * (size_t)k, cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaMemcpyAsync(ords_h, d_ord, sizeof(u64) * (size_t)k, cudaMemcpyDeviceToHost, s));
    RM_CUDA(cudaStreamSynchronize(s));
  }
  return RM_OK;
}

int rm_launch_total_size(rm_index* idx, i64* out_h, cudaStream_t s) {
  RmTree& t = idx->t;
  RmCounters c;
  RM_CUDA(cudaMemcpyAsync(&c, t.ctr, sizeof(c), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  u64* acc = reinterpret_cast<u64*>(idx->sc.info_rows);
  RM_CUDA(cudaMemsetAsync(acc, 0, sizeof(u64), s));
  const int n = c.node_hwm;
  int grid = (n + RM_MT_TPB - 1) / RM_MT_TPB;
  if (grid > 1184) grid = 1184;
  RM_LAUNCH(rm_total_size_kernel, grid, RM_MT_TPB, s, t, n, acc);
  RM_CUDA(cudaGetLastError());
  u64 v = 0;
  RM_CUDA(cudaMemcpyAsync(&v, acc, sizeof(u64), cudaMemcpyDeviceToHost, s));
  RM_CUDA(cudaStreamSynchronize(s));
  *out_h = (i64)v;
  return RM_OK;
}
