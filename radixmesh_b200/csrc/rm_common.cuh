// Shared device-side types and helpers of the radix-tree index (sm_100a).
//
// HBM layout (DESIGN.md "Data layout"):
//   child table  open-addressed, 32-byte slots (= one DRAM sector) keyed (parent id, first token);
//                a slot carries everything one hop of a walk needs: child id, edge length, edge
//                offset, owner rank.  Replaces TreeNode.children (radix_cache.py:39).
//   node pool    structure of arrays indexed by node id (TreeNode fields, radix_cache.py:35-53).
//   edge arenas  ts[{tok i32, seg i32}] and val[i64] in parallel: a node's key and value are the same
//                [edge_off, edge_off+edge_len) run of both, so _split_node (radix_cache.py:277-294)
//                is zero-copy.  A position's token and its owner word (seg, below) sit in one 8-byte
//                record: a step of the walk reads both with ONE 64-bit load per lane.
#pragma once
#include <stdint.h>
#ifdef RM_EMU
#include "rm_emu.h"
#define RM_G_DEFAULT 1
#else
#include <cuda_runtime.h>
#include <stdlib.h>
extern unsigned long long g_rm_launches;   // kernels launched by this library (rm_launch_count)
#define RM_LAUNCH(kernel, grid, block, stream, ...) \
  do {                                               \
    ++g_rm_launches;                                 \
    kernel<<<(grid), (block), 0, (stream)>>>(__VA_ARGS__); \
  } while (0)
// Programmatic dependent launch: the kernel may be scheduled while its predecessor in the stream is still running
// and MUST call rm_grid_dependency_wait() before it touches anything the predecessor writes.  Hides the launch
// latency between the short dependent kernels of a match batch (~10 us per launch otherwise).  RM_PDL=0 turns it off.
static inline bool rm_pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("RM_PDL");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on != 0;
}
#define RM_LAUNCH_PDL(kernel, grid, block, strm_, ...)                                    \
  do {                                                                                    \
    ++g_rm_launches;                                                                      \
    cudaLaunchConfig_t _cfg = {};                                                         \
    _cfg.gridDim = dim3((unsigned)(grid));                                                \
    _cfg.blockDim = dim3((unsigned)(block));                                              \
    _cfg.dynamicSmemBytes = 0;                                                            \
    _cfg.stream = (strm_);                                                                \
    cudaLaunchAttribute _attr[1];                                                         \
    _attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                     \
    _attr[0].val.programmaticStreamSerializationAllowed = rm_pdl_enabled() ? 1 : 0;       \
    _cfg.attrs = _attr;                                                                   \
    _cfg.numAttrs = 1;                                                                    \
    cudaLaunchKernelEx(&_cfg, kernel, __VA_ARGS__);                                       \
  } while (0)
__device__ __forceinline__ void rm_grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void rm_grid_dependency_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif
#include "../../include/radixmesh_b200.h"

typedef unsigned long long u64;
typedef long long i64;

#define RM_EMPTY_KEY 0xFFFFFFFFFFFFFFFFull
#define RM_TOMB_KEY 0xFFFFFFFFFFFFFFFEull
#define RM_NO_REQ 0xFFFFFFFFFFFFFFFFull

struct __align__(32) RmSlot {
  u64 key;            // (parent id << 32) | first token of the child's edge
  int child;          // child node id
  unsigned len_rank;  // edge length (low 24 bits) | (owner rank + 1) << 24
  i64 edge_off;       // start of the edge in the arenas
  int tok1, tok2;     // 2nd and 3rd token of the edge, inline: edges of <= 3 tokens (96 % of the hops of a
                      // Zipf workload once first-sight splits have happened) are verified from the slot alone
};
#define RM_MAX_EDGE_LEN 0xFFFFFF
// seg[a] = owner node id of arena position a, with RM_SEG_CONT set on the FIRST position of a node that is the
// contiguous child of the node owning position a-1 (the lower half of a split).  A walk that has matched a
// node in full can then carry on along the same arena run, comparing tokens and reading seg[] side by side,
// without a child-table probe per node: first-sight splits turn hot prefixes into long chains of one-token
// nodes, but those chains stay contiguous in the arena.
#define RM_SEG_CONT 0x40000000
#define RM_ARENA_PAD 128   // walks read up to RM_RUN_DEPTH warp widths past a run position (rm_match.cu)
__device__ __forceinline__ int rm_seg_owner(int sg) { return sg < 0 ? -1 : (sg & 0x3FFFFFFF); }
__device__ __forceinline__ int rm_slot_len(unsigned w) { return (int)(w & 0xFFFFFFu); }
__device__ __forceinline__ int rm_slot_rank(unsigned w) { return (int)(w >> 24) - 1; }
__device__ __forceinline__ unsigned rm_slot_pack(int len, int rank) { return ((unsigned)len & 0xFFFFFFu) | ((unsigned)(rank + 1) << 24); }

struct RmCounters {
  u64 free_head;      // node ids handed out  (ring of free ids)
  u64 free_tail;      // node ids returned
  u64 arena_top;      // tokens handed out of the arenas
  i64 evictable;      // RadixCache.evictable_size_
  i64 protected_;     // RadixCache.protected_size_
  i64 n_nodes;        // live non-root nodes
  int node_hwm;       // high-water node id + 1
  int err;            // sticky rm_status raised on the device
  int n_dups;
  int n_req;          // scratch: split requests of the current match batch
  int n_big;          // scratch: of those, the requests on edges that more than one query asked to split
                      // (kept next to n_req: one memset clears both)
  int evict_n;        // scratch: nodes evicted by the current rm_evict
  unsigned split_done; // scratch: blocks of the split kernel that have finished (the last one runs the rare big groups)
  int n_kids;         // scratch: children found by rm_node_children
  int n_tomb;         // tombstoned child-table slots (evicted nodes); empty slots = capacity - n_nodes - n_tomb
  int dbg_max_group;  // largest split-request group seen (diagnostics)
  int dbg_groups;     // split-request groups seen (diagnostics)
  int ev_total;       // scratch: eviction candidates (leaves up to the frontier key + the parents they expose)
  int ev_over;        // scratch: the current rm_evict pass ran with bounds that were too small (nothing was evicted)
  u64 ev_k0;          // scratch: frontier key of the current rm_evict (rm_maint.cu)
  u64 ins_ordered;    // diagnostics: mesh insert records applied on the ordered (first-token) path
  u64 ins_resumed;    // diagnostics: mesh insert records applied on the resumed (divergence-group) path
};

struct RmTree {
  int max_key_len;   // bound on key / edge length (rm_config.max_key_len)
  RmSlot* table;
  unsigned table_mask;
  unsigned table_shift;   // 32 - log2(capacity)
  int* parent;
  i64* edge_off;
  int* edge_len;
  int* lock_ref;
  u64* stamp;
  int* rank;
  int* depth_end;
  int* nchild;
  int* gen;        // generation of the object a tree id stands for (bumped by a locked swap); parked: -(gen)-1
  int* parked;     // tree node: most recently parked predecessor (or -1); parked node: the one parked before it
  u64* born;       // op seq of the insert record that created the node (0 = older than any batch in flight)
  u64* ord;        // place among the parent's children in the reference's dict order (radix_cache.py:39, 239-248):
                   // op seq of the insert that added the child to its parent; a split's upper node inherits it
  int2* ts;        // per arena position: .x = token, .y = seg = id of the node whose edge covers it (-1 = free), see RM_SEG_CONT
  i64* val;
  int* req_head;   // per node: last split request of the current match batch on it (an index into the batch's request list; -1 between batches)
  int* req_cnt;    // per node: how many split requests of the current match batch it received (0 between batches)
  int* req_link;   // per request of the batch: the request on the same node before it (-1 = none)
  int* free_ids;
  RmCounters* ctr;
  int max_nodes;
  i64 max_tokens;
  int mode, n_prefill, n_decode, self_rank;
  u64 qpol;        // L2 evict-first cache-policy descriptor for query loads (createpolicy, evaluated once at rm_create)
  // duplicate-node table (RadixMesh.dup_nodes): open addressed, key (tree node << 32 | rank+1)
  u64* dup_key;
  int* dup_parked;
  unsigned dup_mask;
};

// arena accessors (coherent loads: the kernels that call them may have written the arena themselves)
__device__ __forceinline__ int rm_tok(const RmTree& t, i64 a) { return __ldcg(&t.ts[a].x); }
__device__ __forceinline__ int rm_seg(const RmTree& t, i64 a) { return __ldcg(&t.ts[a].y); }
__device__ __forceinline__ void rm_set_seg(const RmTree& t, i64 a, int v) { t.ts[a].y = v; }

// Table hash of (parent id, first token): Fibonacci-style multiplicative hash, top bits.  It sits on
// the dependent chain of every hop of the walk, so it is two independent IMADs, an add and a shift
// (`shift` = 32 - log2(table capacity)).
__device__ __forceinline__ unsigned rm_hash2(unsigned parent, unsigned token, unsigned shift) {
  return (parent * 0x9E3779B1u + token * 0x85EBCA77u) >> shift;
}
__device__ __forceinline__ unsigned rm_hash(u64 k, unsigned shift) { return rm_hash2((unsigned)(k >> 32), (unsigned)k, shift); }

__device__ __forceinline__ u64 rm_key(int parent, int token) { return ((u64)(unsigned)parent << 32) | (unsigned)token; }

// Lane bookkeeping for a cooperating group of G lanes (G = 32: one warp per unit of work;
// smaller G: several independent walks per warp to keep more dependent-load chains in flight).
template <int G>
struct RmGroup {
  int lane;
  unsigned mask;
  __device__ __forceinline__ RmGroup() {
#ifdef RM_EMU
    lane = 0;
    mask = 1u;
#else
    lane = threadIdx.x & (G - 1);
    unsigned base = (threadIdx.x & 31) & ~(G - 1);
    mask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << base);
#endif
  }
  __device__ __forceinline__ unsigned ballot(int pred) const {
#ifdef RM_EMU
    return pred ? 1u : 0u;
#else
    if (G == 1) return pred ? 1u : 0u;   // one lane per unit of work: no cross-lane traffic at all
    unsigned b = __ballot_sync(mask, pred);
    return (G == 32) ? b : ((b & mask) >> ((threadIdx.x & 31) & ~(G - 1)));
#endif
  }
  template <class T>
  __device__ __forceinline__ T bcast(T v, int src) const {
#ifdef RM_EMU
    return v;
#else
    if (G == 1) return v;
    return __shfl_sync(mask, v, src, G);
#endif
  }
  template <class T>
  __device__ __forceinline__ T shfl_up(T v, int d) const {
#ifdef RM_EMU
    return v;
#else
    if (G == 1) return v;
    return __shfl_up_sync(mask, v, d, G);
#endif
  }
  __device__ __forceinline__ int reduce_max(int v) const {
#ifndef RM_EMU
#pragma unroll
    for (int d = G >> 1; d > 0; d >>= 1) {
      const int o = __shfl_xor_sync(mask, v, d, G);
      v = o > v ? o : v;
    }
#endif
    return v;
  }
  __device__ __forceinline__ void sync() const {
    if (G > 1) __syncwarp(mask);
  }
  static __device__ __forceinline__ unsigned full() { return (G == 32) ? 0xffffffffu : ((1u << G) - 1u); }
};

// ---- cache-policy loads (sm_100a).  The tree must stay L2 resident while every batch streams tens
// of MB of query tokens through the same L2, so table slots are read with L2::evict_last (one
// 256-bit LDG.E.ELL2 per slot) and query tokens with an evict-first policy.
struct RmSlotRegs {
  int4 a, b;   // a = {token, parent, child, len_rank}; b = {edge_off lo, edge_off hi, tok1, tok2}
};
__device__ __forceinline__ RmSlotRegs rm_ld_slot(const RmSlot* p) {
  RmSlotRegs r;
#ifdef RM_EMU
  const int4* q = reinterpret_cast<const int4*>(p);
  r.a = q[0];
  r.b = q[1];
#else
  asm volatile("ld.global.nc.L2::evict_last.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.a.x), "=r"(r.a.y), "=r"(r.a.z), "=r"(r.a.w), "=r"(r.b.x), "=r"(r.b.y), "=r"(r.b.z), "=r"(r.b.w)
               : "l"(p));
#endif
  return r;
}
__device__ __forceinline__ u64 rm_policy_evict_first() {
#ifdef RM_EMU
  return 0;
#else
  u64 pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
#endif
}
// query token: cached in L1 (a lane re-reads its own 32-byte sector for several hops), first out of L2
__device__ __forceinline__ int rm_ld_query(const int* p, u64 pol) {
#ifdef RM_EMU
  return *p;
#else
  int v;
  asm volatile("ld.global.nc.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
  return v;
#endif
}

// Probe the child table for `key`.  Returns the slot index or -1.  Every lane of the calling
// group passes the same key, so the loads are broadcasts (one 32-byte sector per probe step).
// COHERENT = true reads through L2 (ld.cg) for kernels that mutate the table themselves.
// A probe sequence ends at an EMPTY slot; tombstones do not end it.  Every loop over the table is bounded by the
// table size and raises the sticky RM_E_TABLE instead of spinning if it has seen every slot (rm_evict rebuilds the
// table when tombstones have taken a quarter of it, so this is a safety net, not a code path).
template <bool COHERENT>
__device__ __forceinline__ int rm_find(const RmTree& t, u64 key, RmSlot& out) {
  unsigned i = rm_hash(key, t.table_shift);
  for (unsigned steps = 0; steps <= t.table_mask; ++steps) {
    const int4* p = reinterpret_cast<const int4*>(t.table + i);
    int4 a = COHERENT ? __ldcg(p) : __ldg(p);
    u64 k = ((u64)(unsigned)a.y << 32) | (unsigned)a.x;
    if (k == key) {
      int4 b = COHERENT ? __ldcg(p + 1) : __ldg(p + 1);
      out.key = k;
      out.child = a.z;
      out.len_rank = (unsigned)a.w;
      out.edge_off = (i64)(((u64)(unsigned)b.y << 32) | (unsigned)b.x);
      out.tok1 = b.z;
      out.tok2 = b.w;
      return (int)i;
    }
    if (k == RM_EMPTY_KEY) return -1;
    i = (i + 1) & t.table_mask;
  }
  atomicMax(&t.ctr->err, (int)RM_E_TABLE);
  return -1;
}

__device__ __forceinline__ void rm_slot_store_toks(RmSlot* s, u64 key, int child, int edge_len, i64 edge_off, int rank, int tok1, int tok2) {
  // payload first, key last.  No fence: inside one kernel only the warp that inserts a key looks it
  // up again (program order + __syncwarp cover that); every other reader runs in a later kernel.
  s->child = child;
  s->len_rank = rm_slot_pack(edge_len, rank);
  s->edge_off = edge_off;
  s->tok1 = tok1;
  s->tok2 = tok2;
  *reinterpret_cast<volatile u64*>(&s->key) = key;
}
__device__ __forceinline__ void rm_slot_store(const RmTree& t, RmSlot* s, u64 key, int child, int edge_len, i64 edge_off, int rank) {
  // the edge's tokens are in the arena already
  rm_slot_store_toks(s, key, child, edge_len, edge_off, rank, edge_len > 1 ? rm_tok(t, edge_off + 1) : 0,
                     edge_len > 2 ? rm_tok(t, edge_off + 2) : 0);
}
// in-place edits of a live slot (its key and edge start do not change, so the inline tokens stay valid)
__device__ __forceinline__ void rm_slot_set_len(RmSlot* s, int len) { s->len_rank = (s->len_rank & 0xFF000000u) | ((unsigned)len & 0xFFFFFFu); }
__device__ __forceinline__ void rm_slot_set_rank(RmSlot* s, int rank) { s->len_rank = (s->len_rank & 0xFFFFFFu) | ((unsigned)(rank + 1) << 24); }

// Claim a free (empty or tombstoned) slot for `key` and return it reserved: the caller fills it (rm_slot_store*).
// Caller guarantees the key is not present.  One lane calls this.  The claim is the FIRST access to a slot: a
// compare-and-swap that expects EMPTY (most probes end there at load factor <= 0.5) and tells what the slot holds
// otherwise, so the common case is one memory round trip, not a load followed by the swap.
__device__ __forceinline__ RmSlot* rm_table_claim(const RmTree& t, u64 key, int child) {
  unsigned i = rm_hash(key, t.table_shift);
  // reserve with a key nobody looks up (parent id 0xFFFFFFFD never exists), then publish
  const u64 reserve = 0xFFFFFFFD00000000ull | (unsigned)child;
  for (unsigned steps = 0; steps <= t.table_mask;) {
    u64* kp = &t.table[i].key;
    u64 cur = atomicCAS(kp, RM_EMPTY_KEY, reserve);
    if (cur == RM_EMPTY_KEY) return &t.table[i];
    if (cur == RM_TOMB_KEY) {
      if (atomicCAS(kp, RM_TOMB_KEY, reserve) == RM_TOMB_KEY) {
        atomicSub(&t.ctr->n_tomb, 1);
        return &t.table[i];
      }
      continue;  // lost the race for this slot: look at it again
    }
    i = (i + 1) & t.table_mask;
    ++steps;
  }
  atomicMax(&t.ctr->err, (int)RM_E_TABLE);   // every slot is taken: the node stays unreachable, the error is sticky
  return nullptr;
}
__device__ __forceinline__ void rm_table_insert(const RmTree& t, u64 key, int child, int edge_len, i64 edge_off, int rank) {
  RmSlot* s = rm_table_claim(t, key, child);
  if (s) rm_slot_store(t, s, key, child, edge_len, edge_off, rank);
}
__device__ __forceinline__ void rm_table_insert_toks(const RmTree& t, u64 key, int child, int edge_len, i64 edge_off, int rank,
                                                     int tok1, int tok2) {
  RmSlot* s = rm_table_claim(t, key, child);
  if (s) rm_slot_store_toks(s, key, child, edge_len, edge_off, rank, tok1, tok2);
}

// Pop `n` node ids from the free ring.  One lane calls this; returns the ring position of the
// first id (ids are free_ids[(pos + k) % max_nodes]) or sets the sticky error and returns ~0.
__device__ __forceinline__ u64 rm_alloc_nodes(const RmTree& t, int n) {
  u64 pos = atomicAdd(&t.ctr->free_head, (u64)n);
  u64 tail = *reinterpret_cast<volatile u64*>(&t.ctr->free_tail);
  if (pos + (u64)n > tail) {
    atomicMax(&t.ctr->err, (int)RM_E_NOSPACE);
    return ~0ull;
  }
  return pos;
}
__device__ __forceinline__ int rm_free_id_at(const RmTree& t, u64 pos) { return t.free_ids[pos % (u64)t.max_nodes]; }

__device__ __forceinline__ i64 rm_alloc_arena(const RmTree& t, int n) {
  u64 off = atomicAdd(&t.ctr->arena_top, (u64)n);
  if (off + (u64)n > (u64)t.max_tokens) {
    atomicMax(&t.ctr->err, (int)RM_E_NOSPACE);
    return -1;
  }
  return (i64)off;
}

// 64-bit signed add (CUDA only has the unsigned overload; two's complement makes them the same)
__device__ __forceinline__ void rm_add_i64(i64* p, i64 v) { atomicAdd(reinterpret_cast<u64*>(p), (u64)v); }

// Host-side launchers implemented in the .cu files (rm_api.cu calls them).
struct RmScratch;
