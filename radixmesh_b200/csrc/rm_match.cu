// K1 batched longest-prefix match, K3 deferred split apply, K2 value gather (sm_100a).
//
// Replaces RadixCache._match_prefix_helper / _key_match_page_size1 / _split_node
// (radix_cache.py:252-294, 14-20) and RadixMesh._match_prefix_helper (radix_mesh.py:247-271).
//
// K1 is read-only on the tree structure: one group of G lanes walks one query (rm_match_rung_kernel follows
// contiguous arena runs; rm_match_kernel, kept for A/B runs, goes hop by hop through the child table).
// Partial edge matches in the splitting modes are recorded as (node, split point) requests; K3 (rm_split_kernel)
// creates each distinct upper node once and derives every segment's last-access stamp in closed form so that
// the result equals the reference processing the batch query by query.
#include <string.h>

#include "rm_internal.h"
#include "rm_sort.cuh"

#ifndef RM_EMU
#define RM_MATCH_TPB 256
#else
#define RM_MATCH_TPB 64
#endif

struct RmRouteOut {
  int* prefill;
  int* decode;
  unsigned char* err;
};


// One split request of a match batch (K1 -> K3): appended to the batch's compact request list, chained to the other
// requests on the same node and counted per node, so that K3 knows without sorting whether a node has one request
// (the common case), a few (ordered in registers by one warp) or many (the rare sorted path).
__device__ __forceinline__ void rm_record_request(const RmTree& t, u64 req, int q, int len_bits, u64* __restrict__ req_key,
                                                  unsigned* __restrict__ req_val) {
  const int slot = atomicAdd(&t.ctr->n_req, 1);
  req_key[slot] = req;
  req_val[slot] = (unsigned)q;
  const int c = (int)(req >> len_bits);
  t.req_link[slot] = atomicExch(&t.req_head[c], slot);
  atomicAdd(&t.req_cnt[c], 1);
}


template <int G, bool SPLIT, bool ROUTER>
__global__ void __launch_bounds__(RM_MATCH_TPB)
rm_match_kernel(RmTree t, const int* __restrict__ qtok, const i64* __restrict__ qoff, int B, u64 seq_base, int len_bits,
                int* __restrict__ out_len, int* __restrict__ out_node, u64* __restrict__ req_key,
                unsigned* __restrict__ req_val, RmRouteOut ro) {
  RmGroup<G> g;
  const int q = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G);
  if (q >= B) return;
  const i64 o0 = qoff[q];
  const int L = (int)(qoff[q + 1] - o0);
  const int* __restrict__ key = qtok + o0;
  const u64 seq = ((seq_base + (u64)q) << 1) | 1ull;   // stamp encoding: see rm_insert.cu

  int node = 0, pos = 0;
  u64 req = RM_NO_REQ;
  // router pick state (radix_mesh.py:219-238), O(1) per hop
  int cnt = 0, deepest = -1, last_prefill = -1;
  bool bad_below = false;

  // slot of the hop being examined, as two 16-byte halves
  int4 sa, sb;
  u64 want = 0;
  unsigned si = 0;
  bool have = L > 0;
  if (have) {
    want = rm_key(0, __ldg(key));
    si = rm_hash(want, t.table_shift);
    const int4* p = reinterpret_cast<const int4*>(t.table + si);
    sa = __ldg(p);
    sb = __ldg(p + 1);
  }
  while (have) {
    // resolve the probe: linear probing past foreign keys (rare at load factor <= 0.5)
    u64 k = ((u64)(unsigned)sa.y << 32) | (unsigned)sa.x;
    for (unsigned steps = 0; k != want && k != RM_EMPTY_KEY; ++steps) {
      if (steps > t.table_mask) {          // saw every slot: no empty one is left (see rm_find)
        atomicMax(&t.ctr->err, (int)RM_E_TABLE);
        break;
      }
      si = (si + 1) & t.table_mask;
      const int4* p = reinterpret_cast<const int4*>(t.table + si);
      sa = __ldg(p);
      sb = __ldg(p + 1);
      k = ((u64)(unsigned)sa.y << 32) | (unsigned)sa.x;
    }
    if (k != want) break;  // no child starts with this token
    const int c = sa.z, elen = rm_slot_len((unsigned)sa.w), crank = rm_slot_rank((unsigned)sa.w);
    const i64 eoff = (i64)(((u64)(unsigned)sb.y << 32) | (unsigned)sb.x);
    const int n = min(elen, L - pos);
    // issue the next hop's probe now (speculative: only used if this edge matches in full)
    const int npos = pos + elen;
    const bool have_next = npos < L;
    int4 na = sa, nb = sb;
    u64 nwant = 0;
    unsigned ni = 0;
    if (have_next) {
      nwant = rm_key(c, __ldg(key + npos));
      ni = rm_hash(nwant, t.table_shift);
      const int4* p = reinterpret_cast<const int4*>(t.table + ni);
      na = __ldg(p);
      nb = __ldg(p + 1);
    }
    // compare edge tokens 1..n-1 (token 0 is equal by construction of the table key)
    int m = n;
    {
      const int2* __restrict__ e = t.ts + eoff;
      const int* __restrict__ kq = key + pos;
      for (int base = 1; base < n; base += 4 * G) {
        int i0 = base + g.lane, i1 = i0 + G, i2 = i1 + G, i3 = i2 + G;
        int a0 = 0, b0 = 0, a1 = 0, b1 = 0, a2 = 0, b2 = 0, a3 = 0, b3 = 0;
        if (i0 < n) { a0 = __ldg(&e[i0].x); b0 = __ldg(kq + i0); }
        if (i1 < n) { a1 = __ldg(&e[i1].x); b1 = __ldg(kq + i1); }
        if (i2 < n) { a2 = __ldg(&e[i2].x); b2 = __ldg(kq + i2); }
        if (i3 < n) { a3 = __ldg(&e[i3].x); b3 = __ldg(kq + i3); }
        unsigned m0 = g.ballot(a0 != b0), m1 = g.ballot(a1 != b1), m2 = g.ballot(a2 != b2), m3 = g.ballot(a3 != b3);
        if (m0 | m1 | m2 | m3) {
          int first = m0 ? (__ffs(m0) - 1) : m1 ? (G + __ffs(m1) - 1) : m2 ? (2 * G + __ffs(m2) - 1) : (3 * G + __ffs(m3) - 1);
          m = base + first;
          break;
        }
      }
    }
    const bool partial = m < elen;
    if (SPLIT && partial) {
      // the reference would split child c at m and return the new upper node (radix_cache.py:262-266)
      pos += m;
      req = ((u64)(unsigned)c << len_bits) | (u64)(unsigned)m;
      break;
    }
    // child c is reached: stamp it (radix_cache.py:260 / radix_mesh.py:257); the max over the batch is
    // the stamp the last query in order leaves behind
    if (ROUTER) {
      ++cnt;
      deepest = crank;
      if (crank < t.n_prefill) {
        last_prefill = crank;
        bad_below = false;
      } else if (crank >= t.n_prefill + t.n_decode) {
        bad_below = true;
      }
    }
    pos += m;
    node = c;
    if (partial || !have_next) break;
    sa = na;
    sb = nb;
    want = nwant;
    si = ni;
  }
  if (g.lane == 0) {
    if (node > 0) atomicMax(&t.stamp[node], seq);   // endpoint stamp (lazy stamps, see rm_maint.cu rm_evict_eff_*)
    out_len[q] = pos;
    out_node[q] = node;   // K3 overwrites this for queries that asked for a split
    if (SPLIT && req != RM_NO_REQ) {
      rm_record_request(t, req, q, len_bits, req_key, req_val);
    }
    if (ROUTER) {
      int pf = -1, dc = -1, er = 0;
      if (cnt == 1) {
        pf = deepest;
      } else if (cnt > 1) {
        if (deepest < t.n_prefill) {
          pf = deepest;
        } else if (bad_below || last_prefill < 0) {
          er = 1;  // the reference asserts (radix_mesh.py:233-237)
        } else {
          pf = last_prefill;
          dc = deepest;
        }
      }
      ro.prefill[q] = pf;
      ro.decode[q] = dc;
      ro.err[q] = (unsigned char)er;
    }
  }
}

// K1, run-following form: G lanes per query, one arena position per lane and step.  A split is zero-copy, so the
// pieces of a split edge stay adjacent in the arena: after ONE child-table probe the walk streams the arena - lane j
// loads the 8-byte record {token, owner word} of position a+j (one coalesced 64-bit load per lane) - and crosses
// node boundaries from the owner words alone, for as long as the chain stays contiguous (RM_SEG_CONT).  The loads
// of DEPTH such steps are in flight together.  One ballot per step finds the first position where the match stops
// (token mismatch, end of the query, or a boundary that is not a contiguous child); nothing else is tracked per
// step: the node the walk ended in is read off the owner words at the stop, and for a partial match inside a node
// its parent and edge start are loaded once, there.  (The ROUTER variant keeps the rank-pick state of
// radix_mesh.py:219-238 per boundary it crosses.)
//
// (A variant that staged the query in shared memory with cp.async - 256-token chunks, two in flight - so that a step
// waited for its arena records only was built and measured: 52 us against 45 on C2, 72 against 62 on the target
// shape.  With 64 warps per SM the query loads are already hidden, and the ring bookkeeping cost issue slots the
// kernel does not have; it was removed.  profiles/README.md, round 2.)
//
// (So was a form in which a warp takes several queries - q, q + W, q + 2 W, ... - and asks for the first token of its
// next query and the offsets of the one after while it walks the current one, to take the first two of a query's ~6
// dependent round trips off its chain: 40 registers instead of 32, i.e. 48 warps per SM instead of 64, and the walk
// went from 36.3 us to 44.6 (2 queries per warp), 48.6 (4), 54.7 (7) on C2.  Warps in flight beat a shorter chain.)
template <int G, bool SPLIT, bool ROUTER, int DEPTH>
__global__ void __launch_bounds__(RM_MATCH_TPB, 8)
rm_match_rung_kernel(RmTree t, const int* __restrict__ qtok, const i64* __restrict__ qoff, int B, u64 seq_base,
                     int len_bits, int* __restrict__ out_len, int* __restrict__ out_node, u64* __restrict__ req_key,
                     unsigned* __restrict__ req_val, RmRouteOut ro, int flags) {
  static_assert(DEPTH * G <= RM_ARENA_PAD, "a walk reads DEPTH steps past a run position");
  RmGroup<G> g;
  rm_grid_dependency_trigger();        // the split kernel that follows may be scheduled now; it waits for this grid
  const int q = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G);
  if (q >= B) return;
  const i64 o0 = qoff[q];
  const int L = (int)(qoff[q + 1] - o0);
  const int* __restrict__ key = qtok + o0;
  const u64 seq = ((seq_base + (u64)q) << 1) | 1ull;
  const RmSlot* __restrict__ table = t.table;
  const unsigned tmask = t.table_mask, tshift = t.table_shift;
  const int2* __restrict__ ts = t.ts;
  u64* __restrict__ stamp = t.stamp;
  const u64 qpol = t.qpol;               // L2 evict-first descriptor, made once per index: a kernel parameter stays in uniform registers
  const int lane = g.lane;

  int node = 0, pos = 0;               // `node` is matched in full through `pos` (same in every lane)
  u64 req = RM_NO_REQ;
  int cnt = 0, deepest = -1, last_prefill = -1;
  bool bad_below = false;
  bool done = L == 0;

  int tk_next = 0;                     // key[pos] when the step that ended the previous run already read it
  bool have_tk = false;
  while (!done) {
    // ---- child of `node` that starts with key[pos]: child-table probe (same address in every lane)
    int tk = tk_next;
    if (!have_tk) tk = rm_ld_query(key + pos, qpol);
    have_tk = false;
    int cur = -1;
    i64 cs = 0;
    {
      unsigned si = rm_hash2((unsigned)node, (unsigned)tk, tshift);
      unsigned steps = 0;
      while (true) {
        const RmSlotRegs S = rm_ld_slot(table + si);
        if ((unsigned)S.a.x == (unsigned)tk && (unsigned)S.a.y == (unsigned)node) {
          cur = S.a.z;
          cs = (i64)(((u64)(unsigned)S.b.y << 32) | (unsigned)S.b.x);
          break;
        }
        if ((S.a.x & S.a.y) == -1) break;
        si = (si + 1) & tmask;
        if (++steps > tmask) {             // saw every slot: no empty one is left (see rm_find)
          atomicMax(&t.ctr->err, (int)RM_E_TABLE);
          break;
        }
      }
    }
    if (cur < 0) break;                 // the walk ends at `node`
    // ---- stream along the arena run: position a + lane against query token qi + lane.  Per-lane running
    // pointers: a round costs DEPTH loads with immediate offsets, not DEPTH 64-bit address computations.
    const int2* __restrict__ tp = ts + (cs + 1 + lane);
    int qi = pos + 1;
    int lastown = cur;                  // owner of the position before the round
    bool leave = false;
    while (!leave && !done) {
      int2 tv[DEPTH];
#pragma unroll
      for (int u = 0; u < DEPTH; ++u) tv[u] = __ldg(tp + u * G);
      const int lim = L - qi - lane;    // this lane's position u*G is inside the query iff u*G < lim
      int kjv[DEPTH];
      {
        const int* __restrict__ kp = key + qi + lane;
#pragma unroll
        for (int u = 0; u < DEPTH; ++u) kjv[u] = u * G < lim ? rm_ld_query(kp + u * G, qpol) : 0;
      }
      bool all = true;
#pragma unroll
      for (int u = 0; u < DEPTH; ++u) {
        const int tj = tv[u].x, sj = tv[u].y;
        const bool inq = u * G < lim;
        const int kj = kjv[u];
        const int own = sj < 0 ? -1 : (sj & 0x3FFFFFFF);
        int prev = g.shfl_up(own, 1);
        if (lane == 0) prev = lastown;
        const bool bnd = own != prev;                                   // node `prev` ends at this position
        const bool badb = bnd && (sj < 0 || !(sj & RM_SEG_CONT));       // ... and the chain does not go on
        const bool stop = badb || !inq || tj != kj;
        const unsigned sm = g.ballot(stop);
        const int s = sm ? (__ffs(sm) - 1) : G;                         // first position where the match stops
        if (ROUTER) {
          const bool commit = bnd && lane <= s;                         // `prev` is matched in full
          const unsigned cm = g.ballot(commit);
          const int rk = commit ? t.rank[prev] : 0;
          const unsigned pm = g.ballot(commit && rk < t.n_prefill);
          const unsigned im = g.ballot(commit && rk >= t.n_prefill + t.n_decode);
          if (cm) {
            const int jd = 31 - __clz((int)cm);
            cnt += __popc(cm);
            deepest = g.bcast(rk, jd);
            if (pm) {
              const int jp = 31 - __clz((int)pm);
              last_prefill = g.bcast(rk, jp);
              const unsigned above = jp >= 31 ? 0u : (~0u << (jp + 1));
              bad_below = (im & above) != 0;
            } else if (im) {
              bad_below = true;
            }
          }
        }
        if (s == G) {                      // everything matched: next G positions
          lastown = g.bcast(own, G - 1);
          continue;
        }
        // the match stops at position s of step u
        all = false;
        const int own_s = g.bcast(own, s);
        const int prev_s = g.bcast(prev, s);
        pos = qi + u * G + s;
        if (own_s != prev_s) {
          // node prev_s ended exactly there and is matched in full; either the query is exhausted or the next node
          // is not a contiguous match: back to the child table
          node = prev_s;
          if (pos >= L) {
            done = true;
          } else {
            leave = true;
            tk_next = g.bcast(kj, s);      // key[pos]: the token the next probe hashes
            have_tk = true;
          }
        } else {
          // the query ends, or differs, inside node own_s: partial match
          if (SPLIT) {
            // the reference splits own_s here and returns the new upper node (radix_cache.py:262-266); K3 does it.
            // Until then the walk's endpoint is the node above.
            node = t.parent[own_s];
            const i64 eo = t.edge_off[own_s];
            const i64 apos = (i64)((tp - lane) - ts) + u * G + s;      // arena position of the stop
            req = ((u64)(unsigned)own_s << len_bits) | (u64)(unsigned)(apos - eo);
          } else {
            if (ROUTER) {
              const int rk = t.rank[own_s];
              ++cnt;
              deepest = rk;
              if (rk < t.n_prefill) {
                last_prefill = rk;
                bad_below = false;
              } else if (rk >= t.n_prefill + t.n_decode) {
                bad_below = true;
              }
            }
            node = own_s;                  // no-split modes return the partially matched child
          }
          done = true;
        }
        break;
      }
      if (all) {
        tp += DEPTH * G;
        qi += DEPTH * G;
      }
    }
  }
  if (lane == 0) {
    if (node > 0 && !(flags & 1)) atomicMax(stamp + node, seq);   // lazy stamps: the endpoint only
    out_len[q] = pos;
    out_node[q] = node;
    if (SPLIT && req != RM_NO_REQ && !(flags & 4)) rm_record_request(t, req, q, len_bits, req_key, req_val);
    if (ROUTER) {
      int pf = -1, dc = -1, er = 0;
      if (cnt == 1) {
        pf = deepest;
      } else if (cnt > 1) {
        if (deepest < t.n_prefill) {
          pf = deepest;
        } else if (bad_below || last_prefill < 0) {
          er = 1;
        } else {
          pf = last_prefill;
          dc = deepest;
        }
      }
      ro.prefill[q] = pf;
      ro.decode[q] = dc;
      ro.err[q] = (unsigned char)er;
    }
  }
}

#define RM_FILL_COOP 48   // owner-map fills longer than this are done by the whole warp
#define RM_SMALL_MAX 32   // requests on one edge that a warp orders among its lanes (more: the sorted path)
#define RM_THREAD_MAX 8   // requests on one edge that a single thread orders by itself


// K3: one group per edge that received split requests.  `skey` is sorted by (node, split point) and,
// inside equal keys, by query index (the sort is stable), `sval` holds the query index.
// (a device function: the stand-alone kernel and the one-launch form rm_split_fused_kernel share it.  The sorted
// list may have been written earlier in the SAME launch, by other blocks: it is read with ld.cg.)
template <int G>
__device__ __forceinline__ void rm_split_apply_body(const RmTree& t, const u64* skey_, const unsigned* sval_, int Bmax, int len_bits,
                                                    u64 seq_base, int* __restrict__ out_node, u64* __restrict__ sufmax,
                                                    int* __restrict__ krank, int* __restrict__ uniq_pos,
                                                    unsigned* __restrict__ uniq_min, const int group0, const int gstride) {
  RmGroup<G> g;
  struct CgU64 {
    const u64* p;
    __device__ __forceinline__ u64 operator[](int i) const { return __ldcg(p + i); }
  } skey = {skey_};
  struct CgU32 {
    const unsigned* p;
    __device__ __forceinline__ unsigned operator[](int i) const { return __ldcg(p + i); }
  } sval = {sval_};
  const int nbig_now = *(volatile int*)&t.ctr->n_big;
  const int B = min(nbig_now, Bmax);   // requests on shared edges only (the single-request ones are done)
  const u64 pmask = (1ull << len_bits) - 1ull;
  // one group of lanes per edge, found as the first entry of its run in the sorted list; grid-stride, the
  // count lives on the device
  for (int i = group0; i < B; i += gstride) {
  const u64 ki = skey[i];
  const int c = (int)(ki >> len_bits);
  if (i > 0 && (int)(skey[i - 1] >> len_bits) == c) continue;   // not the first request of its edge
  // pass 1 (forward): extent of the group, distinct split points, rank of each request's point
  int e = i, U = 0;
  for (int base = i;; base += G) {
    int j = base + g.lane;
    u64 kj = (j < B) ? skey[j] : RM_NO_REQ;
    bool valid = kj != RM_NO_REQ && (int)(kj >> len_bits) == c;
    unsigned vb = g.ballot(valid);
    int nvalid = (vb == g.full()) ? G : (__ffs(~vb) - 1);
    bool mine = g.lane < nvalid;
    bool head = mine && (j == i || kj != skey[j - 1]);
    unsigned hb = g.ballot(head);
    if (mine) {
      unsigned upto = (g.lane == 31) ? 0xffffffffu : ((2u << g.lane) - 1u);
      int kk = U + __popc(hb & upto) - 1;
      krank[j] = kk;
      if (head) {
        uniq_pos[i + kk] = j;
        uniq_min[i + kk] = sval[j];
      }
    }
    g.sync();
    // the query that created a boundary is the lowest-numbered one that asked for it
    if (mine && !head) atomicMin(&uniq_min[i + krank[j]], sval[j]);
    U += __popc(hb);
    e = base + nvalid;
    if (nvalid < G) break;
  }
  if (g.lane == 0) {
    atomicMax(&t.ctr->dbg_max_group, e - i);
    atomicAdd(&t.ctr->dbg_groups, 1);
    t.req_head[c] = -1;                  // ready for the next batch (the split kernel left that to this path)
    t.req_cnt[c] = 0;
  }
  g.sync();
  // pass 2 (backward): sufmax[j] = max op seq over requests j' >= j of this edge
  {
    u64 carry = 0;
    for (int hi = e; hi > i; hi -= G) {
      int j = hi - 1 - g.lane;
      u64 v = (j >= i) ? (((seq_base + (u64)sval[j]) << 1) | 1ull) : 0ull;
#pragma unroll
      for (int d = 1; d < G; d <<= 1) {
        u64 o = g.shfl_up(v, d);
        if (g.lane >= d) v = (o > v) ? o : v;
      }
      v = (carry > v) ? carry : v;
      if (j >= i) sufmax[j] = v;
      carry = g.bcast(v, G - 1);
    }
  }
  g.sync();
  u64 apos = 0;
  if (g.lane == 0) apos = rm_alloc_nodes(t, U);
  apos = g.bcast(apos, 0);
  if (apos == ~0ull) continue;
  const int par = t.parent[c];
  const i64 off = t.edge_off[c];
  const int len = t.edge_len[c];
  const int dend = t.depth_end[c];
  const int lk = t.lock_ref[c];
  const int rk = t.rank[c];
  const u64 st_c = t.stamp[c];
  const u64 ord_c = t.ord[c];
  g.sync();
  // pass 3: one upper node per distinct split point, chained parent -> u_0 -> ... -> u_{U-1} -> c
  for (int kk0 = 0; kk0 < U; kk0 += G) {
    const int kk = kk0 + g.lane;
    i64 fill_off = 0;
    int fill_n = 0, fill_id = 0;
    if (kk < U) {
    const int j = uniq_pos[i + kk];
    const int p = (int)(skey[j] & pmask);
    const int jprev = kk > 0 ? uniq_pos[i + kk - 1] : -1;
    const int a = kk > 0 ? (int)(skey[jprev] & pmask) : 0;
    const int u = rm_free_id_at(t, apos + (u64)kk);
    const int uprev = kk > 0 ? rm_free_id_at(t, apos + (u64)kk - 1) : par;
    // stamp of segment [a, p): queries that went on past p, the query that created boundary a
    // (it stamped the whole remainder of the edge before splitting it), and full-edge passers
    u64 st = st_c;
    if (sufmax[j] > st) st = sufmax[j];
    if (kk > 0) {
      u64 smin_prev = ((seq_base + (u64)uniq_min[i + kk - 1]) << 1) | 1ull;
      if (smin_prev > st) st = smin_prev;
    }
    t.parent[u] = uprev;
    t.edge_off[u] = off + a;
    t.edge_len[u] = p - a;
    t.depth_end[u] = dend - len + p;
    t.lock_ref[u] = lk;       // radix_cache.py:283
    t.rank[u] = rk;
    t.nchild[u] = 1;
    t.gen[u] = 0;
    t.parked[u] = -1;
    t.stamp[u] = st;
    t.ord[u] = kk == 0 ? ord_c : 0;    // the first upper node takes c's place in the parent's dict; the rest are only children
    if (kk == U - 1) t.ord[c] = 0;
    atomicMax(&t.ctr->node_hwm, u + 1);
    // owner map of segment [a, p): the first position keeps the edge's own flag (kk == 0) or marks u as the
    // contiguous child of the previous segment
    {
      const int first = kk == 0 ? ((rm_seg(t, off) & RM_SEG_CONT) | u) : (u | RM_SEG_CONT);
      rm_set_seg(t, off + a, first);
      fill_off = off + a;          // positions a+1 .. p-1: filled by the whole group below
      fill_n = p - a;
      fill_id = u;
      if (kk == U - 1) rm_set_seg(t, off + p, c | RM_SEG_CONT);
    }
    if (kk == 0) {
      RmSlot s;
      int sidx = rm_find<true>(t, rm_key(par, rm_tok(t, off)), s);
      if (sidx < 0) {
        atomicMax(&t.ctr->err, (int)RM_E_INVALID);
      } else {
        t.table[sidx].child = u;
        rm_slot_set_len(&t.table[sidx], p);
      }
    } else {
      rm_table_insert(t, rm_key(uprev, rm_tok(t, off + a)), u, p - a, off + a, rk);
    }
    if (kk == U - 1) {
      rm_table_insert(t, rm_key(u, rm_tok(t, off + p)), c, len - p, off + p, rk);
      t.parent[c] = u;
      t.edge_off[c] = off + p;
      t.edge_len[c] = len - p;
      u64 smin = ((seq_base + (u64)uniq_min[i + kk]) << 1) | 1ull;
      t.stamp[c] = smin > st_c ? smin : st_c;
    }
    }
    if (fill_n <= RM_FILL_COOP) {
      for (int x = 1; x < fill_n; ++x) rm_set_seg(t, fill_off + x, fill_id);
      fill_n = 0;
    }
    unsigned todo = g.ballot(fill_n > 1);
    while (todo) {
      const int l = __ffs(todo) - 1;
      todo &= todo - 1;
      const i64 o = g.bcast(fill_off, l);
      const int n = g.bcast(fill_n, l);
      const int id = g.bcast(fill_id, l);
      for (int x = 1 + g.lane; x < n; x += G) rm_set_seg(t, o + x, id);
    }
  }
  if (g.lane == 0) rm_add_i64(&t.ctr->n_nodes, (i64)U);
  // pass 4: every request's last node is the upper node that ends at its split point
  for (int j = i + g.lane; j < e; j += G) out_node[sval[j]] = rm_free_id_at(t, apos + (u64)krank[j]);
  }
}

#ifdef RM_EMU
// the emulation build sorts the big groups on the host and applies them with this kernel; on the GPU the last block
// of rm_split_kernel does both
template <int G>
__global__ void __launch_bounds__(RM_MATCH_TPB)
rm_split_apply_kernel(RmTree t, const u64* skey, const unsigned* sval, int Bmax, int len_bits, u64 seq_base,
                      int* __restrict__ out_node, u64* __restrict__ sufmax, int* __restrict__ krank,
                      int* __restrict__ uniq_pos, unsigned* __restrict__ uniq_min) {
  rm_split_apply_body<G>(t, skey, sval, Bmax, len_bits, seq_base, out_node, sufmax, krank, uniq_pos, uniq_min,
                         (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G), (int)((gridDim.x * (unsigned)blockDim.x) / G));
}
#endif

// K3, a few requests on one edge (2..RM_SMALL_MAX; the chains of a served tree are 2-5 long): the warp of the request
// that heads the node's chain walks the chain, and the lanes order the split points among themselves - for every
// request the rank of its point among the edge's distinct points, the lowest query that asked for the same point
// (it created the boundary), the highest query at or beyond it (suffix maximum of the closed-form stamps)
// and the point before it - with n^2 <= 1024 comparisons on values kept in shared memory.  Then one lane per
// distinct point creates its upper node exactly as the sorted path does (rm_split_apply_body pass 3).  No sort, no
// second launch.  G = 32 on the device (lane j owns request j), 1 in the host emulation (one lane loops).
template <int G>
__device__ __forceinline__ void rm_split_small_group(const RmTree& t, const u64* rkey, const unsigned* rval, const int c,
                                                     const int head_slot, const int n, const int len_bits, const u64 seq_base,
                                                     int* __restrict__ out_node, int* sp, unsigned* sq, unsigned char* srep) {
  RmGroup<G> g;
  const u64 pmask = (1ull << len_bits) - 1ull;
  {
    int x = head_slot;
    for (int j = 0; j < n; ++j) {
      if ((j & (G - 1)) == g.lane) {
        sp[j] = (int)(__ldcg(rkey + x) & pmask);
        sq[j] = __ldcg(rval + x);
      }
      x = __ldcg(t.req_link + x);
    }
  }
  g.sync();
  for (int j = g.lane; j < n; j += G) {         // first occurrence of its split point?
    const int pj = sp[j];
    bool rep = true;
    for (int k = 0; k < j; ++k) rep = rep && sp[k] != pj;
    srep[j] = rep ? 1 : 0;
  }
  g.sync();
  int U = 0;
  for (int k = 0; k < n; ++k) U += srep[k];
  u64 apos = 0;
  if (g.lane == 0) apos = rm_alloc_nodes(t, U);
  apos = g.bcast(apos, 0);
  if (apos == ~0ull) return;
  const int par = t.parent[c];
  const i64 off = t.edge_off[c];
  const int len = t.edge_len[c];
  const int dend = t.depth_end[c];
  const int lk = t.lock_ref[c];
  const int rk = t.rank[c];
  const u64 st_c = t.stamp[c];
  const u64 ord_c = t.ord[c];
  const int seg0 = rm_seg(t, off);
  g.sync();
  for (int j0 = 0; j0 < n; j0 += G) {
    const int j = j0 + g.lane;
    i64 fill_off = 0;
    int fill_n = 0, fill_id = 0;
    if (j < n) {
      const int pj = sp[j];
      int kk = 0, a = 0;
      unsigned sufq = 0, minq = 0xFFFFFFFFu;
      for (int k = 0; k < n; ++k) {
        const int pk = sp[k];
        const unsigned qk = sq[k];
        if (pk >= pj && qk > sufq) sufq = qk;
        if (pk == pj && qk < minq) minq = qk;
        if (pk < pj) {
          kk += srep[k];
          if (pk > a) a = pk;
        }
      }
      const int u = rm_free_id_at(t, apos + (u64)kk);
      out_node[sq[j]] = u;                         // the upper node that ends at this request's split point
      if (srep[j]) {
        unsigned minq_prev = 0xFFFFFFFFu;
        for (int k = 0; k < n; ++k)
          if (sp[k] == a && sq[k] < minq_prev) minq_prev = sq[k];
        const int p = pj;
        const int uprev = kk > 0 ? rm_free_id_at(t, apos + (u64)kk - 1) : par;
        // stamp of segment [a, p): queries that went on past p, the query that created boundary a (it stamped the
        // whole remainder of the edge before splitting it), and full-edge passers
        u64 st = st_c;
        const u64 ssuf = ((seq_base + (u64)sufq) << 1) | 1ull;
        if (ssuf > st) st = ssuf;
        if (kk > 0) {
          const u64 smin_prev = ((seq_base + (u64)minq_prev) << 1) | 1ull;
          if (smin_prev > st) st = smin_prev;
        }
        t.parent[u] = uprev;
        t.edge_off[u] = off + a;
        t.edge_len[u] = p - a;
        t.depth_end[u] = dend - len + p;
        t.lock_ref[u] = lk;       // radix_cache.py:283
        t.rank[u] = rk;
        t.nchild[u] = 1;
        t.gen[u] = 0;
        t.parked[u] = -1;
        t.stamp[u] = st;
        t.ord[u] = kk == 0 ? ord_c : 0;
        if (kk == U - 1) t.ord[c] = 0;
        atomicMax(&t.ctr->node_hwm, u + 1);
        rm_set_seg(t, off + a, kk == 0 ? ((seg0 & RM_SEG_CONT) | u) : (u | RM_SEG_CONT));
        fill_off = off + a;
        fill_n = p - a;
        fill_id = u;
        if (kk == 0) {
          RmSlot s;
          const int sidx = rm_find<true>(t, rm_key(par, rm_tok(t, off)), s);
          if (sidx < 0) {
            atomicMax(&t.ctr->err, (int)RM_E_INVALID);
          } else {
            t.table[sidx].child = u;
            rm_slot_set_len(&t.table[sidx], p);
          }
        } else {
          rm_table_insert(t, rm_key(uprev, rm_tok(t, off + a)), u, p - a, off + a, rk);
        }
        if (kk == U - 1) {
          rm_set_seg(t, off + p, c | RM_SEG_CONT);
          rm_table_insert(t, rm_key(u, rm_tok(t, off + p)), c, len - p, off + p, rk);
          t.parent[c] = u;
          t.edge_off[c] = off + p;
          t.edge_len[c] = len - p;
          const u64 smin = ((seq_base + (u64)minq) << 1) | 1ull;
          t.stamp[c] = smin > st_c ? smin : st_c;
        }
      }
    }
    if (fill_n <= RM_FILL_COOP) {
      for (int x = 1; x < fill_n; ++x) rm_set_seg(t, fill_off + x, fill_id);
      fill_n = 0;
    }
    unsigned todo = g.ballot(fill_n > 1);
    while (todo) {
      const int l = __ffs(todo) - 1;
      todo &= todo - 1;
      const i64 o = g.bcast(fill_off, l);
      const int fn = g.bcast(fill_n, l);
      const int id = g.bcast(fill_id, l);
      for (int x = 1 + g.lane; x < fn; x += G) rm_set_seg(t, o + x, id);
    }
  }
  if (g.lane == 0) rm_add_i64(&t.ctr->n_nodes, (i64)U);
  g.sync();   // sp / sq / srep are reused by the next group of this warp
}

// K3, the whole split stage in ONE launch.  One thread per request of K1's compact, unsorted list:
//   * alone on its node (~95 % in steady state): the thread splits the edge at once, 32 independent splits sharing
//     every memory round trip of a warp;
//   * 2..small_max requests on the node: the head of the node's chain hands the node to its warp
//     (rm_split_small_group), the other requests of the node do nothing;
//   * more: every request copies itself to a second list, which the LAST block to finish sorts by
//     (node, split point) and applies a warp per edge (rm_split_apply_body) - thousands of requests on one edge only
//     happen when a batch meets a long unsplit edge for the first time.
// That last block also zeroes the batch counters, so a match batch needs no memset.
template <int G>
__global__ void __launch_bounds__(RM_MATCH_TPB)
rm_split_kernel(RmTree t, u64* rkey, unsigned* rval, int Bmax, int len_bits, int passes, int small_max, int thread_max, u64 seq_base,
                int* __restrict__ out_node, u64* mkey, unsigned* mval, u64* __restrict__ sufmax, int* __restrict__ krank,
                int* __restrict__ uniq_pos, unsigned* __restrict__ uniq_min) {
  __shared__ int sp[RM_MATCH_TPB / 32 > 0 ? RM_MATCH_TPB / 32 : 1][RM_SMALL_MAX];
  __shared__ unsigned sq[RM_MATCH_TPB / 32 > 0 ? RM_MATCH_TPB / 32 : 1][RM_SMALL_MAX];
  __shared__ unsigned char srep[RM_MATCH_TPB / 32 > 0 ? RM_MATCH_TPB / 32 : 1][RM_SMALL_MAX];
  rm_grid_dependency_wait();               // launched while the walk kernel may still be running (RM_LAUNCH_PDL)
  RmGroup<G> g;
  const int wid = G == 1 ? 0 : (int)(threadIdx.x >> 5);
  const int n_req_now = *(volatile int*)&t.ctr->n_req;
  const int B = min(n_req_now, Bmax);
  // Two roles when the grid allows it: even blocks apply the single requests, odd blocks the small groups - each set
  // walks the whole request list, so a group's chain of dependent memory round trips (about twice a single's) runs
  // beside the singles instead of after them in the same threads.
  const bool roles = thread_max > 1 && gridDim.x >= 2 && (gridDim.x & 1) == 0 && G > 1;
  const int role = roles ? (int)(blockIdx.x & 1) : -1;            // 0: singles, 1: groups, -1: both
  const int gthreads = (int)((roles ? gridDim.x >> 1 : gridDim.x) * blockDim.x);
  for (int base = (int)((roles ? blockIdx.x >> 1 : blockIdx.x) * blockDim.x); base < B; base += gthreads) {
    const int i = base + (int)threadIdx.x;
    // owner-map fill of a single split's upper node, done by the whole warp after the per-thread part (a thread
    // writing its own p-1 words one after another made this stage 57 us on 4096-token keys)
    i64 fill_off = 0;
    int fill_n = 0, fill_id = 0;
    int grp_c = -1, grp_n = 0;             // set by the head of a small group
    // ---- per request: classify
    bool single = false;
    u64 ki = 0;
    int c = 0, par = 0, len = 0, dend = 0, lk = 0, rk = 0;
    unsigned q = 0;
    i64 off = 0;
    u64 st_c = 0, ord_c = 0;
    if (i < B) {
      ki = rkey[i];
      c = (int)(ki >> len_bits);
      q = rval[i];
      // everything that depends only on (c, q) is loaded at once, before any of it is tested: one memory round trip
      const int head = t.req_head[c];
      const int n = t.req_cnt[c];
      par = t.parent[c];
      off = t.edge_off[c];
      len = t.edge_len[c];
      dend = t.depth_end[c];
      lk = t.lock_ref[c];
      rk = t.rank[c];
      st_c = t.stamp[c];
      ord_c = t.ord[c];
      const bool is_head = head == i;
      if (n > small_max && n > 1) {                                         // many: sorted, then a warp per edge
        if (role != 1) {                                                    // (that path re-arms head and count itself)
          const int slot = atomicAdd(&t.ctr->n_big, 1);
          mkey[slot] = ki;
          mval[slot] = q;
        }
      } else if (n > 1) {
        if (is_head && role != 0) {
          grp_c = c;
          grp_n = n;
          t.req_head[c] = -1;                                               // ready for the next batch
          t.req_cnt[c] = 0;
        }
      } else if (is_head && n == 1 && role != 1) {
        single = true;
        t.req_head[c] = -1;
        t.req_cnt[c] = 0;
      }
    }
    // ---- single requests: node ids for the whole warp with ONE atomic on the free ring (the counters are single
    // words in L2; an atomic per request on each of them serialised the stage)
    const unsigned smask = g.ballot(single);
    int made_hwm = 0;
    if (smask) {
      const int leader = __ffs(smask) - 1;
      u64 wbase = 0;
      if (g.lane == leader) wbase = rm_alloc_nodes(t, __popc(smask));
      wbase = g.bcast(wbase, leader);
      if (single && wbase != ~0ull) {
        const u64 apos = wbase + (u64)__popc(smask & ((1u << g.lane) - 1u));
        const int p = (int)(ki & ((1ull << len_bits) - 1ull));
        const u64 seq = ((seq_base + (u64)q) << 1) | 1ull;
        // second round trip: the new id, the edge's tokens and its owner word, together
        const int2 r0 = __ldcg(t.ts + off);
        const int tokp = rm_tok(t, off + p);
        const int tokp1 = len - p > 1 ? rm_tok(t, off + p + 1) : 0;     // inline tokens of the lower node's new slot
        const int tokp2 = len - p > 2 ? rm_tok(t, off + p + 2) : 0;
        const int u = rm_free_id_at(t, apos);
        // third: the parent's slot
        RmSlot s;
        const int sidx = rm_find<true>(t, rm_key(par, r0.x), s);
        const u64 st = seq > st_c ? seq : st_c;       // the query stamped the whole edge, then split it
        t.parent[u] = par;
        t.edge_off[u] = off;
        t.edge_len[u] = p;
        t.depth_end[u] = dend - len + p;
        t.lock_ref[u] = lk;       // radix_cache.py:283
        t.rank[u] = rk;
        t.nchild[u] = 1;
        t.gen[u] = 0;
        t.parked[u] = -1;
        t.stamp[u] = st;
        t.ord[u] = ord_c;           // u takes c's place in the parent's dict; c is u's only child
        t.ord[c] = 0;
        made_hwm = u + 1;
        rm_set_seg(t, off, (r0.y & RM_SEG_CONT) | u);          // owner map: [off, off+p) is u's, c starts at off+p
        rm_set_seg(t, off + p, c | RM_SEG_CONT);
        fill_off = off;
        fill_n = p;
        fill_id = u;
        if (sidx < 0) {
          atomicMax(&t.ctr->err, (int)RM_E_INVALID);
        } else {
          t.table[sidx].child = u;
          rm_slot_set_len(&t.table[sidx], p);
        }
        rm_table_insert_toks(t, rm_key(u, tokp), c, len - p, off + p, rk, tokp1, tokp2);
        t.parent[c] = u;
        t.edge_off[c] = off + p;
        t.edge_len[c] = len - p;
        t.stamp[c] = st;
        out_node[q] = u;
      }
      const int hw = g.reduce_max(made_hwm);
      if (g.lane == leader && wbase != ~0ull) {
        rm_add_i64(&t.ctr->n_nodes, (i64)__popc(smask));
        atomicMax(&t.ctr->node_hwm, hw);
      }
    }
    // short fills stay with their thread (32 of them proceed side by side); only long ones are worth
    // serialising the warp for
    if (fill_n <= RM_FILL_COOP) {
      for (int x = 1; x < fill_n; ++x) rm_set_seg(t, fill_off + x, fill_id);
      fill_n = 0;
    }
    unsigned todo = g.ballot(fill_n > 1);
    while (todo) {
      const int l = __ffs(todo) - 1;
      todo &= todo - 1;
      const i64 o = g.bcast(fill_off, l);
      const int fn = g.bcast(fill_n, l);
      const int id = g.bcast(fill_id, l);
      for (int x = 1 + g.lane; x < fn; x += G) rm_set_seg(t, o + x, id);
    }
    // nodes with a handful of requests (2..thread_max, nearly all shared edges of a served tree): the head's own
    // thread orders them, one lane working like a group of one, so the heads of a warp - they cluster at the end of
    // the request list, being the last of their node to finish the walk - proceed side by side instead of queueing
    // for the warp
    if (G > 1 && grp_n > 0 && grp_n <= thread_max) {
      int lsp[RM_THREAD_MAX];
      unsigned lsq[RM_THREAD_MAX];
      unsigned char lrep[RM_THREAD_MAX];
      rm_split_small_group<1>(t, rkey, rval, grp_c, i, grp_n, len_bits, seq_base, out_node, lsp, lsq, lrep);
      grp_n = 0;
    }
    // up to small_max requests: the warp takes the nodes one after another
    todo = g.ballot(grp_n > 0);
    while (todo) {
      const int l = __ffs(todo) - 1;
      todo &= todo - 1;
      const int c = g.bcast(grp_c, l);
      const int n = g.bcast(grp_n, l);
      const int hs = g.bcast(i, l);
      rm_split_small_group<G>(t, rkey, rval, c, hs, n, len_bits, seq_base, out_node, sp[wid], sq[wid], srep[wid]);
    }
  }
#ifndef RM_EMU
  // ---- the last block to get here runs the sorted path for whatever was put on the second list, and re-arms
  __shared__ int s_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = atomicAdd(&t.ctr->split_done, 1u) == gridDim.x - 1 ? 1 : 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int nbig_now2 = *(volatile int*)&t.ctr->n_big;
  const int nbig = min(nbig_now2, Bmax);
  if (nbig > 0) {
    const int w = rm_sort_block(mkey, mval, rkey, rval, nbig, passes);
    rm_split_apply_body<G>(t, w ? rkey : mkey, w ? rval : mval, Bmax, len_bits, seq_base, out_node, sufmax, krank, uniq_pos,
                           uniq_min, (int)(threadIdx.x >> 5), (int)(blockDim.x >> 5));
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    t.ctr->n_req = 0;
    t.ctr->n_big = 0;
    t.ctr->split_done = 0;
  }
#endif
}

// K2: values along the matched path, written back to front by walking parent links from the last
// node; depth_end gives each node's place in the output without a prefix sum.
template <int G>
__global__ void __launch_bounds__(RM_MATCH_TPB)
rm_gather_kernel(RmTree t, const int* __restrict__ nodes, const int* __restrict__ lens, const i64* __restrict__ out_off,
                 int B, i64* __restrict__ out) {
  RmGroup<G> g;
  const int q = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G);
  if (q >= B) return;
  const int mlen = lens[q];
  int node = nodes[q];
  i64* __restrict__ dst = out + out_off[q];
  while (node > 0) {
    const int dend = t.depth_end[node];
    const int elen = t.edge_len[node];
    const i64 eoff = t.edge_off[node];
    const int start = dend - elen;
    const int cnt = min(elen, mlen - start);   // the last node may be matched only in part (no-split modes)
    for (int x = g.lane; x < cnt; x += G) dst[start + x] = t.val[eoff + x];
    node = t.parent[node];
  }
}

static int rm_match_group() {
#ifdef RM_EMU
  return 1;
#else
  static int g = -1;
  if (g < 0) {
    const char* e = getenv("RM_MATCH_G");
    g = e ? atoi(e) : 8;
    if (g != 2 && g != 4 && g != 8 && g != 16 && g != 32) g = 8;
  }
  return g;
#endif
}

template <bool SPLIT, bool ROUTER>
static void rm_match_dispatch(rm_index* idx, const int* tokens, const i64* offsets, int B, u64 seq_base, int* out_len,
                              int* out_node, RmRouteOut ro, cudaStream_t s) {
#ifdef RM_EMU
  // the emulation build runs the run-following kernel with one lane per query unless RM_EMU_GROUP asks for the
  // hop-by-hop form
  static const bool emu_group = getenv("RM_EMU_GROUP") != nullptr;
  const int grid = (B + RM_MATCH_TPB - 1) / RM_MATCH_TPB;
  if (!emu_group) {
    auto kr = rm_match_rung_kernel<1, SPLIT, ROUTER, 2>;
    RM_LAUNCH(kr, grid, RM_MATCH_TPB, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,
              idx->sc.key[0], idx->sc.val[0], ro, 0);
    return;
  }
  auto k = rm_match_kernel<1, SPLIT, ROUTER>;
  RM_LAUNCH(k, grid, RM_MATCH_TPB, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,
            idx->sc.key[0], idx->sc.val[0], ro);
#else
  // RM_WALK selects the walk kernel: "run" (default) = run-following, RM_RUN_G lanes per query (default 32),
  // RM_RUN_DEPTH steps in flight (default 3); "table" = hop by hop through the child table, RM_MATCH_G lanes per
  // query.  The non-default forms are kept for A/B measurements and tests.
  static const char* walk = getenv("RM_WALK") ? getenv("RM_WALK") : "run";
  // RM_EXP_FLAGS: measurement experiments only (bit 0 = skip the stamp atomics; results then differ from the
  // reference in eviction order, never use it outside a profiling run)
  static const int run_flags = getenv("RM_EXP_FLAGS") ? (atoi(getenv("RM_EXP_FLAGS")) & 1) : 0;
  if (strcmp(walk, "table") != 0) {
    static const int rg = getenv("RM_RUN_G") ? atoi(getenv("RM_RUN_G")) : 32;
    static const int depth_env = getenv("RM_RUN_DEPTH") ? atoi(getenv("RM_RUN_DEPTH")) : 3;
    static const int rtpb_env = getenv("RM_RUN_TPB") ? atoi(getenv("RM_RUN_TPB")) : 0;
    const long long rthreads = (long long)B * rg;
    // 4 queries per block: a block's warps end at different times and its slots are only handed on when the last
    // one is done (128 threads: +2 % over 256 on both bench shapes)
    const int rtpb = rtpb_env ? rtpb_env : (rthreads >= 148LL * 8 * 128 ? 128 : 64);
    const int rgrid = (int)((rthreads + rtpb - 1) / rtpb);
#define RM_RLAUNCH(GG, DD)                                                                                        \
  do {                                                                                                            \
    auto kr = rm_match_rung_kernel<GG, SPLIT, ROUTER, DD>;                                                        \
    RM_LAUNCH(kr, rgrid, rtpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,        \
              idx->sc.key[0], idx->sc.val[0], ro, run_flags);                                                     \
  } while (0)
    switch (rg) {
      case 4: RM_RLAUNCH(4, 2); break;
      case 8: RM_RLAUNCH(8, 2); break;
      case 16: RM_RLAUNCH(16, 2); break;
      default:
        if (depth_env == 2) RM_RLAUNCH(32, 2);
        else if (depth_env == 4) RM_RLAUNCH(32, 4);
        else RM_RLAUNCH(32, 3);
        break;
    }
#undef RM_RLAUNCH
    return;
  }
  const int G = rm_match_group();
  const long long threads = (long long)B * G;
  // a grid of B*G threads must still cover all 148 SMs several times: narrow groups get small blocks
  const int tpb = threads >= 148LL * 8 * RM_MATCH_TPB ? RM_MATCH_TPB : (threads >= 148LL * 8 * 128 ? 128 : 64);
  const int grid = (int)((threads + tpb - 1) / tpb);
#define RM_CASE(GG)                                                                                              \
  case GG: {                                                                                                     \
    auto k = rm_match_kernel<GG, SPLIT, ROUTER>;                                                                 \
    RM_LAUNCH(k, grid, tpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,          \
              idx->sc.key[0], idx->sc.val[0], ro);                                                               \
  } break;
  switch (G) {
    RM_CASE(2)
    RM_CASE(4)
    RM_CASE(16)
    RM_CASE(32)
    default:
    RM_CASE(8)
  }
#undef RM_CASE
#endif
}

#ifdef RM_EMU
#define RM_WIDE_G 1
#else
#define RM_WIDE_G 32
#endif

int rm_launch_match(rm_index* idx, const int* tokens, const i64* offsets, int B, int* out_len, int* out_node,
                    cudaStream_t s) {
  if (B <= 0) return RM_OK;
  const u64 seq_base = idx->op_seq + 1;
  idx->op_seq += (u64)B;
  RmRouteOut ro = {nullptr, nullptr, nullptr};
  const bool split = idx->cfg.mode == RM_MODE_BASE || idx->cfg.mode == RM_MODE_PREFILL;
  if (!split) {
    rm_match_dispatch<false, false>(idx, tokens, offsets, B, seq_base, out_len, out_node, ro, s);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
  }
#ifndef RM_EMU
  const bool prof = idx->prof_on && idx->prof_n < RM_PROF_EVENTS / RM_PROF_STRIDE;
  cudaEvent_t* pev = idx->prof_ev + RM_PROF_STRIDE * idx->prof_n;
  if (prof) cudaEventRecord(pev[0], s);
#else
  RM_CUDA(cudaMemsetAsync(&idx->t.ctr->n_req, 0, 2 * sizeof(int), s));   // n_req and n_big (adjacent); on the GPU the split kernel re-arms them
#endif
  idx->t.req_link = idx->sc.link;
  rm_match_dispatch<true, false>(idx, tokens, offsets, B, seq_base, out_len, out_node, ro, s);
#ifndef RM_EMU
  if (prof) cudaEventRecord(pev[1], s);
#endif
  RM_CUDA(cudaGetLastError());
  // The split stage: one launch.  All counts stay on the device.
  RmScratch& sc = idx->sc;
  const int bits = idx->node_bits + idx->len_bits;
#ifndef RM_EMU
  static const int small_env = getenv("RM_SPLIT_SMALL_MAX") ? atoi(getenv("RM_SPLIT_SMALL_MAX")) : RM_SMALL_MAX;
  static const int thread_env = getenv("RM_SPLIT_THREAD_MAX") ? atoi(getenv("RM_SPLIT_THREAD_MAX")) : RM_THREAD_MAX;
  const int small_max = small_env < 1 ? 1 : (small_env > RM_SMALL_MAX ? RM_SMALL_MAX : small_env);
  const int thread_max = thread_env < 1 ? 1 : (thread_env > RM_THREAD_MAX ? RM_THREAD_MAX : thread_env);
  static const int roles_env = getenv("RM_SPLIT_ROLES") ? atoi(getenv("RM_SPLIT_ROLES")) : 1;
  int grid = 2 * ((B + RM_MATCH_TPB - 1) / RM_MATCH_TPB);
  if (grid > 148) grid = 148;            // grid-stride inside; one block per SM keeps the walk kernel's tail unobstructed
  if (!roles_env) grid |= 1;             // an odd grid runs without the role split (A/B)
  if (grid > 148) grid = 147;
  auto k3 = rm_split_kernel<RM_WIDE_G>;
  RM_LAUNCH_PDL(k3, grid, RM_MATCH_TPB, s, idx->t, sc.key[0], sc.val[0], B, idx->len_bits, (bits + 7) / 8, small_max, thread_max, seq_base,
                out_node, sc.key[1], sc.val[1], sc.sufmax, sc.krank, sc.uniq_pos, sc.uniq_min);
  if (prof) {
    cudaEventRecord(pev[2], s);
    ++idx->prof_n;
  }
#else
  static const int small_env = getenv("RM_SPLIT_SMALL_MAX") ? atoi(getenv("RM_SPLIT_SMALL_MAX")) : RM_SMALL_MAX;
  const int small_max = small_env < 1 ? 1 : (small_env > RM_SMALL_MAX ? RM_SMALL_MAX : small_env);
  auto k3 = rm_split_kernel<RM_WIDE_G>;
  RM_LAUNCH(k3, (B + RM_MATCH_TPB - 1) / RM_MATCH_TPB, RM_MATCH_TPB, s, idx->t, sc.key[0], sc.val[0], B, idx->len_bits,
            (bits + 7) / 8, small_max, 1, seq_base, out_node, sc.key[1], sc.val[1], sc.sufmax, sc.krank, sc.uniq_pos, sc.uniq_min);
  const int w = 1 ^ rm_sort_pairs(sc.key[1], sc.val[1], sc.key[0], sc.val[0], B, bits, sc.hist, s, &idx->t.ctr->n_big);
  auto k4 = rm_split_apply_kernel<RM_WIDE_G>;
  RM_LAUNCH(k4, (B + RM_MATCH_TPB - 1) / RM_MATCH_TPB, RM_MATCH_TPB, s, idx->t, sc.key[w], sc.val[w], B, idx->len_bits, seq_base,
            out_node, sc.sufmax, sc.krank, sc.uniq_pos, sc.uniq_min);
#endif
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_launch_gather(rm_index* idx, const int* nodes, const int* lens, const i64* out_offsets, int B, i64* out_values,
                     cudaStream_t s) {
  if (B <= 0) return RM_OK;
  const int G = RM_WIDE_G == 1 ? 1 : 8;
  const long long threads = (long long)B * G;
  const int grid = (int)((threads + RM_MATCH_TPB - 1) / RM_MATCH_TPB);
#ifdef RM_EMU
  auto k = rm_gather_kernel<1>;
#else
  auto k = rm_gather_kernel<8>;
#endif
  RM_LAUNCH(k, grid, RM_MATCH_TPB, s, idx->t, nodes, lens, out_offsets, B, out_values);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

// ROUTER-mode walk (never splits) that also evaluates the rank pick; used by rm_route.cu.
int rm_launch_router_match(rm_index* idx, const int* tokens, const i64* offsets, int B, int* out_len, int* out_node,
                           int* out_prefill, int* out_decode, unsigned char* out_err, cudaStream_t s) {
  if (B <= 0) return RM_OK;
  const u64 seq_base = idx->op_seq + 1;
  idx->op_seq += (u64)B;
  RmRouteOut ro = {out_prefill, out_decode, out_err};
  rm_match_dispatch<false, true>(idx, tokens, offsets, B, seq_base, out_len, out_node, ro, s);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

// Read-only pre-pass of the ordered insert (rm_insert.cu): walk every record of the batch against the
// tree as it stands, stamp the nodes it matches in full, and report the deepest such node.  Same kernel
// as the splitting match, with the split requests switched off.
static int rm_launch_walk_flags(rm_index* idx, const int* tokens, const i64* offsets, int N, u64 seq_base, int* out_len,
                                int* out_node, int walk_flags, cudaStream_t s) {
  if (N <= 0) return RM_OK;
  RmRouteOut ro = {nullptr, nullptr, nullptr};
#ifdef RM_EMU
  const int tpb = RM_MATCH_TPB;
  auto k = rm_match_rung_kernel<1, true, false, 2>;
  const long long wthreads = N;
#else
  // A batch that a warp per record still fits the GPU with (148 SMs x 64 warps) walks with 32 lanes and three steps
  // in flight per record - a 512-token record is 6 rounds instead of 32, and with a few thousand records the walk is
  // one record's latency; larger batches keep 8 lanes per record.  (RM_PREPASS_G=8 / 32 forces one form.)
  static const int g_env = getenv("RM_PREPASS_G") ? atoi(getenv("RM_PREPASS_G")) : 0;
  const bool wide = g_env ? g_env == 32 : N <= 148 * 64;
  if (wide) {
    auto kw = rm_match_rung_kernel<32, true, false, 3>;
    const long long wt = (long long)N * 32;
    RM_LAUNCH(kw, (int)((wt + 127) / 128), 128, s, idx->t, tokens, offsets, N, seq_base, idx->len_bits, out_len, out_node,
              idx->sc.key[0], idx->sc.val[0], ro, walk_flags);
    RM_CUDA(cudaGetLastError());
    return RM_OK;
  }
  const int tpb = N >= 148 * 8 * RM_MATCH_TPB ? RM_MATCH_TPB : (N >= 148 * 8 * 128 ? 128 : 64);
  auto k = rm_match_rung_kernel<8, true, false, 2>;
  const long long wthreads = (long long)N * 8;
#endif
  RM_LAUNCH(k, (int)((wthreads + tpb - 1) / tpb), tpb, s, idx->t, tokens, offsets, N, seq_base, idx->len_bits, out_len,
            out_node, idx->sc.key[0], idx->sc.val[0], ro, walk_flags);
  RM_CUDA(cudaGetLastError());
  return RM_OK;
}

int rm_launch_walk_records(rm_index* idx, const int* tokens, const i64* offsets, int N, u64 seq_base, int* out_len,
                           int* out_node, cudaStream_t s) {
  return rm_launch_walk_flags(idx, tokens, offsets, N, seq_base, out_len, out_node, 4, s);
}

// the same walk with nothing written to the tree at all (no stamp, no split request): locating GC keys
int rm_launch_walk_readonly(rm_index* idx, const int* tokens, const i64* offsets, int N, int* out_len, int* out_node,
                            cudaStream_t s) {
  return rm_launch_walk_flags(idx, tokens, offsets, N, 0, out_len, out_node, 4 | 1, s);
}
