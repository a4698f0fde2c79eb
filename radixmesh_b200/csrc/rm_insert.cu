// K4 ordered batched insert apply (sm_100a).
//
// Replaces RadixCache._insert_helper (radix_cache.py:296-327) and, in the mesh modes,
// RadixMesh._insert_helper + NodeRankConflictResolver.keep + _swap_node (radix_mesh.py:273-323,
// 466-495; policy/conflict_resolve.py:1-6): what oplog_received does with a run of INSERT oplogs.
//
// Order equivalence.  A read-only pre-pass (the walk kernel, rm_match.cu) takes every record of the batch to its
// deepest fully matched node in the tree as it stood.  Two records can then only interact if they start changing
// the tree at the same node through the same token, so the batch is stably sorted by a hash of (start node, next
// token) and one warp per group applies its records one after another in arrival order, each from its own start
// node, while different groups proceed in parallel.  The start node is the divergence node in RadixCache mode; in the
// mesh modes it may lie higher (see "Mesh modes" below).  Inside a warp the lanes cooperate on the edge compares and
// on copying the new leaf's key/value run into the arenas; node ids and arena space come from device-wide
// allocators (one atomic per record).
#include "rm_internal.h"
#include "rm_sort.cuh"

#ifndef RM_EMU
#define RM_INS_TPB 128
#define RM_INS_G 32
#else
#define RM_INS_TPB 64
#define RM_INS_G 1
#endif

// RadixCache semantics, fast path: group key of a record after the read-only pre-pass.  Two records can only
// interact if they leave the old tree at the same node through the same token, so the key is a hash of
// (deepest fully matched node, next token); unrelated records that collide merely share a warp.
// Records that matched in full need nothing more: insert returns len(key) (radix_cache.py:327).
__global__ void rm_insert_keys2_kernel(RmTree t, const int* __restrict__ tok, const i64* __restrict__ off, int N,
                                       const int* __restrict__ xnode, u64* __restrict__ key, unsigned* __restrict__ val,
                                       int* __restrict__ out_prefix) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  const i64 o = off[r];
  const int n = (int)(off[r + 1] - o);
  const int x = xnode[r];
  const int pos = x > 0 ? t.depth_end[x] : 0;
  u64 k = 0;                                   // 0 = nothing left to do
  if (pos < n) k = 1ull + (u64)(rm_hash2((unsigned)x, (unsigned)tok[o + pos], 0) >> 12);   // 20-bit group hash
  else out_prefix[r] = n;
  key[r] = k;
  val[r] = (unsigned)r;
}

// Record a duplicate-node entry: dup_nodes[(full_key(node), rank)] = parked (radix_mesh.py:310, 476, 493).
// The tree node id stands for its full key (stable under splits: the lower half keeps the id).
__device__ __forceinline__ void rm_dup_put(const RmTree& t, int node, int rank, int parked) {
  const u64 key = ((u64)(unsigned)node << 32) | (unsigned)(rank + 1);
  const unsigned home = (rm_hash2((unsigned)(key >> 32), (unsigned)key, 0) >> 7) & t.dup_mask;
  for (int attempt = 0; attempt < 64; ++attempt) {
    unsigned i = home;
    int tomb = -1;                 // first slot an rm_gc_exec_h removal left behind: reused before a fresh one
    bool retry = false;
    for (unsigned probes = 0; probes <= t.dup_mask; ++probes) {
      const u64 cur = *reinterpret_cast<volatile u64*>(&t.dup_key[i]);
      if (cur == key) {
        t.dup_parked[i] = parked;  // dict assignment overwrites
        return;
      }
      if (cur == RM_TOMB_KEY && tomb < 0) tomb = (int)i;
      if (cur == RM_EMPTY_KEY) {
        const unsigned at = tomb >= 0 ? (unsigned)tomb : i;
        const u64 expect = tomb >= 0 ? RM_TOMB_KEY : RM_EMPTY_KEY;
        const u64 old = atomicCAS(&t.dup_key[at], expect, key);
        if (old == expect || old == key) {
          t.dup_parked[at] = parked;
          if (old == expect) atomicAdd(&t.ctr->n_dups, 1);
          return;
        }
        retry = true;              // somebody else took the slot: look again from the start (it may hold this key)
        break;
      }
      i = (i + 1) & t.dup_mask;
    }
    if (!retry && tomb >= 0) {     // no empty slot anywhere, but a removed one
      const u64 old = atomicCAS(&t.dup_key[tomb], RM_TOMB_KEY, key);
      if (old == RM_TOMB_KEY || old == key) {
        t.dup_parked[tomb] = parked;
        if (old == RM_TOMB_KEY) atomicAdd(&t.ctr->n_dups, 1);
        return;
      }
      retry = true;
    }
    if (!retry) break;             // the table is full
  }
  atomicMax(&t.ctr->err, (int)RM_E_NOSPACE);
}

// MESH: 0 = RadixCache semantics, 1 = prefill/decode replica (values compared), 2 = router replica
// (owner ranks compared, no value arena traffic).
template <int G, int MESH>
__global__ void __launch_bounds__(RM_INS_TPB)
rm_insert_kernel(RmTree t, const int* __restrict__ rtok, const i64* __restrict__ rval, const i64* __restrict__ roff,
                 const int* __restrict__ rrank, int N, const u64* __restrict__ skey, const unsigned* __restrict__ sval,
                 u64 seq_base, int* __restrict__ out_prefix, const int* __restrict__ start) {
  // start[r]: the node record r starts at (the pre-pass already stamped the path down to its deepest fully matched
  // node); groups are keyed by a hash of (that node, next token).
  RmGroup<G> g;
  const int i = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G);
  if (i >= N) return;
  const u64 gk = skey[i];
  if (i > 0 && skey[i - 1] == gk) return;  // one group per distinct key
  if (gk == 0) return;                      // records the pre-pass finished (the whole key is in the tree, empty keys)
  const bool resume = (gk >> 40) == 0;      // bit 40: records longer than max_key_len, reported below
  for (int j = i; j < N && skey[j] == gk; ++j) {
    const int r = (int)sval[j];
    const i64 o = roff[r];
    const int n = (int)(roff[r + 1] - o);
    const int* __restrict__ key = rtok + o;
    const i64* __restrict__ val = (MESH == 2) ? nullptr : (rval + o);
    const int owner = (MESH == 0) ? -1 : (rrank ? rrank[r] : t.self_rank);
    // Stamp encoding.  The reference stamps with a monotonic clock, so inside ONE operation a node
    // visited earlier is older.  All nodes an operation stamps lie on one root path (where the
    // tree shape already forces child-before-parent eviction) except the lower half of a node it
    // split: that half is stamped on the way down, before the new upper node and the new leaf
    // exist, and ends up as a SIBLING of the new leaf.  Stamps are therefore 2*op+1, and 2*op for
    // that lower half, which keeps "older than the leaf of the same operation" (radix_cache.py:306-324).
    const u64 seq = ((seq_base + (u64)r) << 1) | 1ull;
    if (n > t.max_key_len) {
      if (g.lane == 0) {
        atomicMax(&t.ctr->err, (int)RM_E_INVALID);
        out_prefix[r] = 0;
      }
      continue;
    }
    int node = 0, pos = 0, total = 0;
    if (resume) {
      node = start[r];
      pos = total = node > 0 ? t.depth_end[node] : 0;
    }
    while (pos < n) {
      RmSlot s;
      const int sidx = rm_find<true>(t, rm_key(node, key[pos]), s);
      if (sidx < 0) break;
      int c = s.child;
      const int elen = rm_slot_len(s.len_rank);
      const int srank = rm_slot_rank(s.len_rank);
      const i64 eoff = s.edge_off;
      const int lim = min(elen, n - pos);
      // longest common prefix of the edge and the rest of the key (radix_cache.py:14-20)
      int m = lim;
      for (int base = 1; base < lim; base += G) {
        int x = base + g.lane;
        bool ne = x < lim && rm_tok(t, eoff + x) != key[pos + x];
        unsigned mb = g.ballot(ne);
        if (mb) {
          m = base + __ffs(mb) - 1;
          break;
        }
      }
      // mesh: does the incoming value agree with the stored one over the matched part?
      bool lost = false, swap = false;
      if (MESH == 1) {
        bool diff = false;
        for (int base = 0; base < m; base += G) {
          int x = base + g.lane;
          bool ne = x < m && __ldcg(t.val + eoff + x) != val[pos + x];
          if (g.ballot(ne)) {
            diff = true;
            break;
          }
        }
        if (diff) {
          if (srank <= owner) lost = true;   // keep(now, new): the lower rank wins
          else swap = true;
        }
      } else if (MESH == 2) {
        if (srank != owner) {
          if (srank <= owner) lost = true;
          else swap = true;
        }
      }
      total += m;
      // stamp: 2*op+1, or 2*op for the lower half of a node this record splits (see above).  With the
      // pre-pass other records of the batch may already have stamped c, so the stamp is a max.
      const u64 c_before = t.stamp[c];
      if (g.lane == 0) atomicMax(&t.stamp[c], (m < elen) ? seq - 1 : seq);
      int target = c;
      if (m < elen) {
        // _split_node (radix_cache.py:277-294): new upper node [0, m), c keeps [m, elen)
        u64 apos = 0;
        if (g.lane == 0) apos = rm_alloc_nodes(t, 1);
        apos = g.bcast(apos, 0);
        if (apos == ~0ull) break;
        const int u = rm_free_id_at(t, apos);
        // owner map: [eoff, eoff+m) now belongs to u (the first position keeps its flag), c starts at
        // eoff+m as the contiguous child of u
        for (int x = g.lane; x < m; x += G) rm_set_seg(t, eoff + x, x == 0 ? ((rm_seg(t, eoff) & RM_SEG_CONT) | u) : u);
        if (g.lane == 0) {
          rm_set_seg(t, eoff + m, c | RM_SEG_CONT);
          const int par = t.parent[c];
          t.parent[u] = par;
          t.edge_off[u] = eoff;
          t.edge_len[u] = m;
          t.depth_end[u] = t.depth_end[c] - elen + m;
          t.lock_ref[u] = t.lock_ref[c];
          t.rank[u] = srank;
          t.nchild[u] = 1;
          t.gen[u] = 0;
          t.parked[u] = -1;
          t.born[u] = seq_base + (u64)r;
          t.ord[u] = t.ord[c];       // u takes c's place in the parent's dict (radix_cache.py:292); c is u's first child
          t.ord[c] = 0;
          t.stamp[u] = c_before > seq ? c_before : seq;   // later records that passed c in full pass u as well
          atomicMax(&t.ctr->node_hwm, u + 1);
          t.table[sidx].child = u;
          rm_slot_set_len(&t.table[sidx], m);
          rm_table_insert(t, rm_key(u, rm_tok(t, eoff + m)), c, elen - m, eoff + m, srank);
          t.parent[c] = u;
          t.edge_off[c] = eoff + m;
          t.edge_len[c] = elen - m;
          rm_add_i64(&t.ctr->n_nodes, (i64)1);
          __threadfence();
        }
        g.sync();
        target = u;
      }
      if (MESH != 0 && swap) {
        // _swap_node (radix_mesh.py:466-495): the incoming (lower-rank) value takes the node
        const int old_rank = srank;
        if (MESH == 2) {
          if (g.lane == 0) {
            t.rank[target] = owner;
            rm_slot_set_rank(&t.table[sidx], owner);
          }
        } else {
          const int lk = __ldcg(t.lock_ref + target);
          if (lk == 0) {
            for (int x = g.lane; x < m; x += G) t.val[eoff + x] = val[pos + x];
            if (g.lane == 0) {
              t.rank[target] = owner;
              rm_slot_set_rank(&t.table[sidx], owner);
              rm_dup_put(t, target, old_rank, -1);
            }
          } else {
            // locked: the tree node keeps its id with the new value in a fresh arena run and
            // lock_ref 0; the old value, rank and lock move to a parked node that dup_nodes owns.
            // gen[target] is bumped so that handles taken before the swap resolve to the parked
            // node (DESIGN.md "locked swap").
            u64 apos = 0;
            i64 noff = 0;
            if (g.lane == 0) {
              apos = rm_alloc_nodes(t, 1);
              noff = rm_alloc_arena(t, m);
            }
            apos = g.bcast(apos, 0);
            noff = g.bcast(noff, 0);
            if (apos == ~0ull || noff < 0) break;
            const int d = rm_free_id_at(t, apos);
            for (int x = g.lane; x < m; x += G) {
              t.ts[noff + x] = make_int2(rm_tok(t, eoff + x), target);   // moved away from its parent's run: no contiguity flag
              t.val[noff + x] = val[pos + x];
              rm_set_seg(t, eoff + x, d);   // the old run stays with the parked node
            }
            if (g.lane == 0) {
              const int nx = rm_seg(t, eoff + m);  // a contiguous child now follows the PARKED run, not its parent
              if (nx >= 0 && (nx & RM_SEG_CONT)) rm_set_seg(t, eoff + m, nx & ~RM_SEG_CONT);
            }
            if (g.lane == 0) {
              t.parent[d] = t.parent[target];
              t.edge_off[d] = eoff;
              t.edge_len[d] = m;
              t.depth_end[d] = t.depth_end[target];
              t.lock_ref[d] = lk;
              t.rank[d] = old_rank;
              t.nchild[d] = 0;
              t.gen[d] = -(t.gen[target] + 1);   // parked marker: -(generation it was known under) - 1
              t.born[d] = seq_base + (u64)r;
              t.stamp[d] = t.stamp[target];
              atomicMax(&t.ctr->node_hwm, d + 1);
              t.edge_off[target] = noff;
              t.lock_ref[target] = 0;
              t.rank[target] = owner;
              t.gen[target] = t.gen[target] + 1;
              t.stamp[target] = seq;
              t.table[sidx].edge_off = noff;
              rm_slot_set_rank(&t.table[sidx], owner);
              t.parked[d] = t.parked[target];
              t.parked[target] = d;
              // like the reference, neither size counter moves (radix_mesh.py:478-495)
              rm_dup_put(t, target, old_rank, d);
            }
          }
        }
        g.sync();
      }
      if (MESH != 0 && lost && g.lane == 0) rm_dup_put(t, target, owner, -1);
      pos += m;
      node = target;
      if (m < elen) break;  // the key diverged (or ended) inside this edge
    }
    if (pos < n && node >= 0) {
      // append the remainder as a new leaf (radix_cache.py:319-326)
      const int rem = n - pos;
      u64 apos = 0;
      i64 noff = 0;
      if (g.lane == 0) {
        apos = rm_alloc_nodes(t, 1);
        noff = rm_alloc_arena(t, rem);
      }
      apos = g.bcast(apos, 0);
      noff = g.bcast(noff, 0);
      if (apos != ~0ull && noff >= 0) {
        const int leaf = rm_free_id_at(t, apos);
        for (int x = g.lane; x < rem; x += G) {
          t.ts[noff + x] = make_int2(key[pos + x], leaf);
          if (MESH != 2) t.val[noff + x] = val[pos + x];
        }
        g.sync();   // the slot's inline tokens are read back from the arena
        if (g.lane == 0) {
          t.parent[leaf] = node;
          t.edge_off[leaf] = noff;
          t.edge_len[leaf] = rem;
          t.depth_end[leaf] = n;
          t.lock_ref[leaf] = 0;
          t.rank[leaf] = owner;
          t.nchild[leaf] = 0;
          t.gen[leaf] = 0;
          t.parked[leaf] = -1;
          t.born[leaf] = seq_base + (u64)r;
          t.ord[leaf] = seq_base + (u64)r;   // appended to the parent's dict by this record (radix_cache.py:323)
          t.stamp[leaf] = seq;
          atomicMax(&t.ctr->node_hwm, leaf + 1);
          atomicAdd(&t.nchild[node], 1);
          rm_table_insert(t, rm_key(node, key[pos]), leaf, rem, noff, owner);
          // evictable_size_ += len(value): 1 for a router value (radix_mesh.py:62-63, 321)
          rm_add_i64(&t.ctr->evictable, (i64)(MESH == 2 ? 1 : rem));
          rm_add_i64(&t.ctr->n_nodes, (i64)1);
          __threadfence();
        }
      }
      g.sync();
    }
    if (g.lane == 0) out_prefix[r] = total;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Mesh modes.  A record whose value DIFFERS from a stored node it passes either LOSES there (stored rank <= its own:
// nothing in the tree changes, dup_nodes gets (node, its rank), radix_mesh.py:300-310) or REPLACES the stored value
// (_swap_node, radix_mesh.py:466-495).  A replacement changes what every later record that reaches the node compares
// against, so where records start is decided by a classification pass between the pre-pass and the apply:
//   * rm_insert_classify_kernel marks, per record, the SHALLOWEST node at which it would replace the stored value, as
//     the tree stands before the batch.  Nodes above every mark of their root path are static for the batch: the first
//     replacement at a node would have been predicted from the old tree, so no mark means no replacement.
//   * rm_insert_keys3_kernel starts every record at the parent of the shallowest marked node T on its path, or at its
//     divergence node when its path holds no mark.  Records that can see each other's replacements share a marked
//     node, hence T, hence the group; a record that leaves the tree inside T's edge has (T's parent, T's first token)
//     as its divergence point, the same key.  (History: one group per first token serialised a Zipf-skewed stream on
//     its hottest prompt, 21 K records/s; keeping whole first-token groups on the path from the root whenever one
//     of their records replaced something still did under alternating owner ranks, 14 ms per 1,024-record batch.)
//   * what lies ABOVE a record's start node is static, so its losses there are booked AFTER the apply
//     (rm_insert_lost_kernel), walking up from the start node: an upper node created by a split of this same batch
//     counts as a node of its own only if the record that made it came EARLIER (born[] holds the creating record's op
//     seq); made later, it was still part of the node below it when this record passed, so its comparison is folded
//     into that node's.
template <int G, int MESH>
__device__ __forceinline__ bool rm_ins_differs(const RmTree& t, const RmGroup<G>& g, int v, const i64* __restrict__ val, int owner) {
  if (MESH == 2) return t.rank[v] != owner;
  const int e = t.edge_len[v];
  const i64 off = t.edge_off[v];
  const int at = t.depth_end[v] - e;
  for (int base = 0; base < e; base += G) {
    const int x = base + g.lane;
    const bool ne = x < e && __ldcg(t.val + off + x) != val[at + x];
    if (g.ballot(ne)) return true;
  }
  return false;
}

// The record of a node as the bottom-up walks below need it.  A walk asks for its parent's record while it still
// compares values against the node itself: one memory round trip per node instead of four (rank, edge, parent, the
// parent's birth stamp, each waiting for the one before: 336 us per 1,024-record batch in the losses kernel).
struct RmNodeF {
  int rank, elen, dend, parent;
  i64 eoff;
};
__device__ __forceinline__ RmNodeF rm_node_f(const RmTree& t, int v) {
  RmNodeF f;
  f.rank = t.rank[v];
  f.elen = t.edge_len[v];
  f.dend = t.depth_end[v];
  f.parent = t.parent[v];
  f.eoff = t.edge_off[v];
  return f;
}
template <int G, int MESH>
__device__ __forceinline__ bool rm_ins_differs_f(const RmTree& t, const RmGroup<G>& g, const RmNodeF& f, const i64* __restrict__ val,
                                                 int owner) {
  if (MESH == 2) return f.rank != owner;
  const int at = f.dend - f.elen;
  for (int base = 0; base < f.elen; base += G) {
    const int x = base + g.lane;
    const bool ne = x < f.elen && __ldcg(t.val + f.eoff + x) != val[at + x];
    if (g.ballot(ne)) return true;
  }
  return false;
}

template <int G, int MESH>
__global__ void __launch_bounds__(RM_INS_TPB)
rm_insert_classify_kernel(RmTree t, const int* __restrict__ rtok, const i64* __restrict__ rval, const i64* __restrict__ roff,
                          const int* __restrict__ rrank, int N, const int* __restrict__ xnode, int* __restrict__ wn,
                          int* __restrict__ low, int* __restrict__ mark) {
  // wn[r]: the SHALLOWEST node at which record r would replace a stored value, as the tree stands before the batch
  // (0: nowhere); low[r]: the deepest node it touches (the child it leaves the tree on, else xnode[r])
  RmGroup<G> g;
  const int r = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G);
  if (r >= N) return;
  const i64 o = roff[r];
  const int n = (int)(roff[r + 1] - o);
  const int* __restrict__ key = rtok + o;
  const i64* __restrict__ val = (MESH == 2) ? nullptr : (rval + o);
  const int owner = rrank ? rrank[r] : t.self_rank;
  const int x = xnode[r];
  int w = 0, lo = x;
  if (n <= t.max_key_len) {
    const int pos = x > 0 ? t.depth_end[x] : 0;
    if (pos < n) {                       // the edge the record leaves the tree on, matched in part
      RmSlot sl;
      const int sidx = rm_find<true>(t, rm_key(x, key[pos]), sl);
      if (sidx >= 0) {
        const int elen = rm_slot_len(sl.len_rank);
        const int srank = rm_slot_rank(sl.len_rank);
        const i64 eoff = sl.edge_off;
        const int lim = min(elen, n - pos);
        int m = lim;
        for (int base = 1; base < lim; base += G) {
          const int xx = base + g.lane;
          const bool ne = xx < lim && rm_tok(t, eoff + xx) != key[pos + xx];
          const unsigned mb = g.ballot(ne);
          if (mb) {
            m = base + __ffs(mb) - 1;
            break;
          }
        }
        bool diff = false;
        if (MESH == 2) {
          diff = srank != owner;
        } else {
          for (int base = 0; base < m; base += G) {
            const int xx = base + g.lane;
            const bool ne = xx < m && __ldcg(t.val + eoff + xx) != val[pos + xx];
            if (g.ballot(ne)) {
              diff = true;
              break;
            }
          }
        }
        lo = sl.child;
        if (diff && srank > owner) w = sl.child;
      }
    }
    // the nodes it matched in full, bottom-up: the last hit is the shallowest
    RmNodeF f = {};
    if (x > 0) f = rm_node_f(t, x);
    for (int v = x; v > 0;) {
      const int p = f.parent;
      RmNodeF fp = {};
      if (p > 0) fp = rm_node_f(t, p);
      if (f.rank > owner && rm_ins_differs_f<G, MESH>(t, g, f, val, owner)) w = v;
      v = p;
      f = fp;
    }
  }
  if (g.lane == 0) {
    wn[r] = w;
    low[r] = lo;
    if (w > 0) mark[w] = 1;
  }
}

// Group keys after the classification.  T = the shallowest MARKED node on the record's path (a node where some record
// of the batch replaces the stored value): from T down, what the record finds depends on the records before it, so it
// starts at T's parent, through T's first token - together with every other record that reaches T, whose path holds
// the same T, and with the records that leave the tree inside T's edge, whose divergence point is that same (parent,
// token).  A record with no marked node on its path starts at its divergence node as before.  Key: hash of (start
// node, next token) + 1, or 0 when there is nothing to do (whole key in the tree, nothing replaced on the way).
__global__ void rm_insert_keys3_kernel(RmTree t, const int* __restrict__ tok, const i64* __restrict__ off, int N,
                                       int* __restrict__ xnode, const int* __restrict__ low, const int* __restrict__ mark,
                                       u64* __restrict__ key, unsigned* __restrict__ val, int* __restrict__ out_prefix) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  const i64 o = off[r];
  const int n = (int)(off[r + 1] - o);
  u64 k = 0;
  if (n > t.max_key_len) {
    k = (1ull << 40) | (1ull << 33);              // error record: the kernel reports it
    atomicAdd(&t.ctr->ins_resumed, 1ull);
  } else {
    int top = 0;
    for (int v = low[r]; v > 0; v = t.parent[v])
      if (mark[v]) top = v;
    const int x = top > 0 ? t.parent[top] : xnode[r];
    const int pos = x > 0 ? t.depth_end[x] : 0;
    if (pos < n) k = 1ull + (u64)(rm_hash2((unsigned)x, (unsigned)tok[o + pos], 0) >> 12);
    else out_prefix[r] = n;
    xnode[r] = x;                                 // from here on: the node the record starts at
    atomicAdd(top > 0 ? &t.ctr->ins_ordered : &t.ctr->ins_resumed, 1ull);
  }
  key[r] = k;
  val[r] = (unsigned)r;
}

__global__ void rm_insert_unmark_kernel(int N, const int* __restrict__ wn, int* __restrict__ mark) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < N && wn[r] > 0) mark[wn[r]] = 0;
}

// losses on the nodes ABOVE a record's start node (none of them is replaced by this batch), booked on the tree as the
// apply left it (see above)
template <int G, int MESH>
__global__ void __launch_bounds__(RM_INS_TPB)
rm_insert_lost_kernel(RmTree t, const i64* __restrict__ rval, const i64* __restrict__ roff, const int* __restrict__ rrank,
                      int N, const int* __restrict__ xnode, u64 seq_base) {
  RmGroup<G> g;
  const int r = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G);
  if (r >= N) return;
  const i64 o = roff[r];
  const int n = (int)(roff[r + 1] - o);
  if (n > t.max_key_len) return;
  const i64* __restrict__ val = (MESH == 2) ? nullptr : (rval + o);
  const int owner = rrank ? rrank[r] : t.self_rank;
  const u64 mine = seq_base + (u64)r;
  int v = xnode[r];
  if (v <= 0) return;
  RmNodeF f = rm_node_f(t, v);
  // The duplicate-table entries of a record's losses are independent of each other ((node, rank) keys, all with the
  // same payload): the walk only notes the nodes, one per lane, and the lanes book them side by side - a probe
  // sequence with its compare-and-swap is 3-4 dependent round trips, and a record that differs from what is stored
  // loses at every node of its path.
  int my_loss = -1, n_loss = 0;
  while (v > 0) {
    const int base = v;
    const int rk = f.rank;
    int p = f.parent;
    RmNodeF fp = {};
    u64 born_p = 0;
    if (p > 0) {                                 // asked for before this node's values are compared
      fp = rm_node_f(t, p);
      born_p = t.born[p];
    }
    bool diff = rm_ins_differs_f<G, MESH>(t, g, f, val, owner);
    while (p > 0 && born_p > mine) {             // made by a later record: still part of `base` when this one passed
      if (!diff) diff = rm_ins_differs_f<G, MESH>(t, g, fp, val, owner);
      p = fp.parent;
      if (p > 0) {
        fp = rm_node_f(t, p);
        born_p = t.born[p];
      }
    }
    if (diff && rk <= owner) {
      if (g.lane == n_loss) my_loss = base;
      if (++n_loss == G) {
        rm_dup_put(t, my_loss, owner, -1);
        my_loss = -1;
        n_loss = 0;
      }
    }
    v = p;
    f = fp;
  }
  if (my_loss >= 0) rm_dup_put(t, my_loss, owner, -1);
}

int rm_launch_walk_records(rm_index* idx, const int* tokens, const i64* offsets, int N, u64 seq_base, int* out_len,
                           int* out_node, cudaStream_t s);

static int rm_launch_insert_one(rm_index* idx, const int* tokens, const i64* values, const i64* offsets, const int* ranks, int N,
                                int* out_prefix_len, cudaStream_t s);

// A batch is applied in arrival order inside every group of records that leave the tree at the same node through the
// same token.  Into a nearly EMPTY tree that is the whole batch per hot document: every record of the document lands in
// one group and re-walks what the records before it built (34.8 ms for the first 4,096-record batch, against 0.18 ms
// in steady state).  While the index has seen fewer records than two such batches, a batch is therefore applied as
// consecutive chunks of 256 records - same order, same result: a chunk's groups are short, and the chunks after it
// find the documents' paths in the tree and spread over many groups.
#define RM_COLD_CHUNK 256
int rm_launch_insert(rm_index* idx, const int* tokens, const i64* values, const i64* offsets, const int* ranks, int N,
                     int* out_prefix_len, cudaStream_t s) {
  if (N <= 0) return RM_OK;
  // (short keys have nothing long to re-walk: a batch known to average under 64 tokens per record is applied whole)
  const bool short_keys = idx->hint_tokens > 0 && idx->hint_tokens < 64ll * N;
  idx->hint_tokens = 0;
  const u64 seen0 = idx->ins_records;
  const bool cold = N > RM_COLD_CHUNK && seen0 < 2ull * (u64)N && !short_keys;
  idx->ins_records += (u64)N;
  if (!cold) return rm_launch_insert_one(idx, tokens, values, offsets, ranks, N, out_prefix_len, s);
  // a chunk is as large as what the tree has already seen (256, 256, 512, 1024, ...): the documents' paths are there
  // after a few hundred records, and every chunk costs a fixed ~0.15 ms of launches
  static const bool grow = !(getenv("RM_COLD_GROW") && atoi(getenv("RM_COLD_GROW")) == 0);
  for (int a = 0; a < N;) {
    u64 want = grow ? seen0 + (u64)a : (u64)RM_COLD_CHUNK;
    if (want < RM_COLD_CHUNK) want = RM_COLD_CHUNK;
    const int n = (u64)(N - a) < want ? N - a : (int)want;
    // offsets are absolute positions in tokens / values: a chunk is a window of the offsets, ranks and outputs
    int rc = rm_launch_insert_one(idx, tokens, values, offsets + a, ranks ? ranks + a : nullptr, n, out_prefix_len + a, s);
    if (rc) return rc;
    a += n;
  }
  return RM_OK;
}

static int rm_launch_insert_one(rm_index* idx, const int* tokens, const i64* values, const i64* offsets, const int* ranks, int N,
                                int* out_prefix_len, cudaStream_t s) {
  if (N <= 0) return RM_OK;
  RmScratch& sc = idx->sc;
  const u64 seq_base = idx->op_seq + 1;
  idx->op_seq += (u64)N;
  const long long threads = (long long)N * RM_INS_G;
  const int grid = (int)((threads + RM_INS_TPB - 1) / RM_INS_TPB);
  if (idx->cfg.mode == RM_MODE_BASE) {
    // RadixCache semantics: parallel read-only walk of all records, then in-order apply per
    // (divergence node, next token) group, resumed at the divergence node
    int* xnode = sc.krank;
    int* xlen = sc.uniq_pos;
    int rc = rm_launch_walk_records(idx, tokens, offsets, N, seq_base, xlen, xnode, s);
    if (rc) return rc;
    RM_LAUNCH(rm_insert_keys2_kernel, (N + 255) / 256, 256, s, idx->t, tokens, offsets, N, xnode, sc.key[0], sc.val[0],
              out_prefix_len);
    const int w = rm_sort_pairs(sc.key[0], sc.val[0], sc.key[1], sc.val[1], N, 21, sc.hist, s);
    auto k = rm_insert_kernel<RM_INS_G, 0>;
    RM_LAUNCH(k, grid, RM_INS_TPB, s, idx->t, tokens, values, offsets, ranks, N, sc.key[w], sc.val[w], seq_base,
              out_prefix_len, xnode);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
  }
  // mesh modes: pre-pass, conflict classification, one launch for both kinds of groups, losses of the resumed ones
  int* xnode = sc.krank;
  int* xlen = sc.uniq_pos;
  int* wn = reinterpret_cast<int*>(sc.uniq_min);
  int* low = reinterpret_cast<int*>(sc.sufmax);
  int rc = rm_launch_walk_records(idx, tokens, offsets, N, seq_base, xlen, xnode, s);
  if (rc) return rc;
  const bool router = idx->cfg.mode == RM_MODE_ROUTER;
  if (router) {
    auto kc = rm_insert_classify_kernel<RM_INS_G, 2>;
    RM_LAUNCH(kc, grid, RM_INS_TPB, s, idx->t, tokens, values, offsets, ranks, N, xnode, wn, low, sc.ins_mark);
  } else {
    auto kc = rm_insert_classify_kernel<RM_INS_G, 1>;
    RM_LAUNCH(kc, grid, RM_INS_TPB, s, idx->t, tokens, values, offsets, ranks, N, xnode, wn, low, sc.ins_mark);
  }
  RM_LAUNCH(rm_insert_keys3_kernel, (N + 255) / 256, 256, s, idx->t, tokens, offsets, N, xnode, low, sc.ins_mark, sc.key[0],
            sc.val[0], out_prefix_len);
  RM_LAUNCH(rm_insert_unmark_kernel, (N + 255) / 256, 256, s, N, wn, sc.ins_mark);
  RM_CUDA(cudaGetLastError());
  const int w = rm_sort_pairs(sc.key[0], sc.val[0], sc.key[1], sc.val[1], N, 41, sc.hist, s);
  RM_CUDA(cudaGetLastError());
  if (router) {
    auto k = rm_insert_kernel<RM_INS_G, 2>;
    RM_LAUNCH(k, grid, RM_INS_TPB, s, idx->t, tokens, values, offsets, ranks, N, sc.key[w], sc.val[w], seq_base,
              out_prefix_len, xnode);
    auto kl = rm_insert_lost_kernel<RM_INS_G, 2>;
    RM_LAUNCH(kl, grid, RM_INS_TPB, s, idx->t, values, offsets, ranks, N, xnode, seq_base);
  } else {
    auto k = rm_insert_kernel<RM_INS_G, 1>;
    RM_LAUNCH(k, grid, RM_INS_TPB, s, idx->t, tokens, values, offsets, ranks, N, sc.key[w], sc.val[w], seq_base,
              out_prefix_len, xnode);
    auto kl = rm_insert_lost_kernel<RM_INS_G, 1>;
    RM_LAUNCH(kl, grid, RM_INS_TPB, s, idx->t, values, offsets, ranks, N, xnode, seq_base);
  }
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}
