// K1 batched longest-prefix match, K3 deferred split apply, K2 value gather (sm_100a).
//
// Replaces RadixCache._match_prefix_helper / _key_match_page_size1 / _split_node
// (radix_cache.py:252-294, 14-20) and RadixMesh._match_prefix_helper (radix_mesh.py:247-271).
//
// K1 is read-only on the tree structure: one group of G lanes walks one query.  Each hop is ONE
// dependent 32-byte load (the child-table slot, which carries child id, edge length, edge offset and
// owner rank); the probe for the NEXT hop is issued before the current edge's tokens are compared,
// so edge compares sit off the critical path and the chain per query is `hops` slot latencies.
// Partial edge matches in the splitting modes are recorded as (node, split point) requests; K3 sorts
// them, creates each distinct upper node once and derives every segment's last-access stamp in closed
// form so that the result equals the reference processing the batch query by query.
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
    while (k != want && k != RM_EMPTY_KEY) {
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
      const int* __restrict__ e = t.tok + eoff;
      const int* __restrict__ kq = key + pos;
      for (int base = 1; base < n; base += 4 * G) {
        int i0 = base + g.lane, i1 = i0 + G, i2 = i1 + G, i3 = i2 + G;
        int a0 = 0, b0 = 0, a1 = 0, b1 = 0, a2 = 0, b2 = 0, a3 = 0, b3 = 0;
        if (i0 < n) { a0 = __ldg(e + i0); b0 = __ldg(kq + i0); }
        if (i1 < n) { a1 = __ldg(e + i1); b1 = __ldg(kq + i1); }
        if (i2 < n) { a2 = __ldg(e + i2); b2 = __ldg(kq + i2); }
        if (i3 < n) { a3 = __ldg(e + i3); b3 = __ldg(kq + i3); }
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
      const int slot = atomicAdd(&t.ctr->n_req, 1);   // compact list of split requests
      req_key[slot] = req;
      req_val[slot] = (unsigned)q;
      // per-node chain of this batch's requests: lets K3 tell at once whether the node has a single request
      t.req_link[q] = atomicExch(&t.req_head[(int)(req >> len_bits)], q);
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

// K1, lane-per-query form (the default).  What the measurements said (profiles/README.md):
//   * the walk is a chain of dependent loads; a lone warp of the group form needs ~490 ns per hop while a
//     32-byte slot read from L2 costs ~200 ns (tools/chase.cu), the rest being instructions executed
//     between a slot's arrival and the next probe's issue;
//   * 89 % of the edges a Zipf workload walks are one token long and 96 % at most three, because every
//     first-sight match splits at its divergence point.
// So: one lane per query; the slot carries the edge's 2nd and 3rd token inline, and the lane keeps the
// next three query tokens in registers, so an edge of <= 3 tokens is verified, committed and followed by
// the next probe without touching the arena or the query again.  Longer edges are rare: they are
// verified RM_CH tokens per loop iteration, non-speculatively, while the other lanes keep hopping.
#define RM_CH 4
template <bool SPLIT, bool ROUTER>
__global__ void __launch_bounds__(RM_MATCH_TPB)
rm_match_lpq_kernel(RmTree t, const int* __restrict__ qtok, const i64* __restrict__ qoff, int B, u64 seq_base,
                    int len_bits, int* __restrict__ out_len, int* __restrict__ out_node, u64* __restrict__ req_key,
                    unsigned* __restrict__ req_val, RmRouteOut ro, int flags) {
  // flags >> 4 = log2(lanes left idle per query lane), a measurement knob (see tools/)
  const int spread = flags >> 4;
  const unsigned gt = blockIdx.x * (unsigned)blockDim.x + threadIdx.x;
  if (spread && (gt & ((1u << spread) - 1u))) return;
  const int q = (int)(gt >> spread);
  if (q >= B) return;
  const i64 o0 = qoff[q];
  const int L = (int)(qoff[q + 1] - o0);
  const int* __restrict__ key = qtok + o0;
  const u64 seq = ((seq_base + (u64)q) << 1) | 1ull;
  const RmSlot* __restrict__ table = t.table;
  const unsigned tmask = t.table_mask, tshift = t.table_shift;
  u64* __restrict__ stamp = t.stamp;
  const u64 qpol = rm_policy_evict_first();

  int node = 0, pos = 0;               // `node` is matched in full through `pos`
  u64 req = RM_NO_REQ;
  int cnt = 0, deepest = -1, last_prefill = -1;
  bool bad_below = false;

  RmSlotRegs S;                        // probe in flight, keyed (want_par, key[pos])
  unsigned want_par = 0, want_tok = 0, si = 0;
  int g1 = 0, g2 = 0, g3 = 0;          // key[pos+1], key[pos+2], key[pos+3] (0 past the end; lengths guard their use)
  // long edge being verified: found by its first three tokens, the rest not compared yet
  bool pend = false;
  int pc = 0, pelen = 0, pn = 0, pdone = 0, prank = 0;
  const int* __restrict__ pe = nullptr;
  int ca[RM_CH], cb[RM_CH];

  bool have_probe = L > 0;
  if (have_probe) {
    want_tok = (unsigned)rm_ld_query(key, qpol);
    si = rm_hash2(0u, want_tok, tshift);
    S = rm_ld_slot(table + si);
    if (1 < L) g1 = rm_ld_query(key + 1, qpol);
    if (2 < L) g2 = rm_ld_query(key + 2, qpol);
    if (3 < L) g3 = rm_ld_query(key + 3, qpol);
  }
  while (true) {
    // ---- fast path first: the probe hit a one-token edge (89 % of the hops) and the query goes on.
    // Nothing but the key compare, the hash and the address sits between the slot's arrival and the
    // next probe's issue; stamp, router state and look-ahead refill overlap with the new load.
    if (!pend && have_probe && (unsigned)S.a.x == want_tok && (unsigned)S.a.y == want_par &&
        ((unsigned)S.a.w & 0xFFFFFFu) == 1u && pos + 1 < L) {
      const int c1 = S.a.z;
      const unsigned w1 = (unsigned)S.a.w;
      want_par = (unsigned)c1;
      want_tok = (unsigned)g1;
      si = rm_hash2((unsigned)c1, want_tok, tshift);
      S = rm_ld_slot(table + si);
      if (ROUTER) {
        const int r1 = rm_slot_rank(w1);
        ++cnt;
        deepest = r1;
        if (r1 < t.n_prefill) {
          last_prefill = r1;
          bad_below = false;
        } else if (r1 >= t.n_prefill + t.n_decode) {
          bad_below = true;
        }
      }
      node = c1;
      pos += 1;
      g1 = g2;
      g2 = g3;
      g3 = (pos + 3 < L) ? rm_ld_query(key + pos + 3, qpol) : 0;
      continue;
    }
    int c, elen, m, erank;              // the edge decided in this iteration: m of its elen tokens match
    if (pend) {
      // ---- long edge: verify the chunk requested in the previous iteration
      const int lim = min(pn - pdone, RM_CH);
      int mm = lim;
#pragma unroll
      for (int x = RM_CH - 1; x >= 0; --x)
        if (x < lim && ca[x] != cb[x]) mm = x;
      pdone += mm;
      if (mm == lim && pdone < pn) {
        const int* __restrict__ ee = pe + pdone;
        const int* __restrict__ kk = key + pos + pdone;
        const int l2 = pn - pdone;
#pragma unroll
        for (int x = 0; x < RM_CH; ++x) {
          if (x < l2) {
            ca[x] = __ldg(ee + x);
            cb[x] = rm_ld_query(kk + x, qpol);
          }
        }
        continue;
      }
      pend = false;
      c = pc;
      elen = pelen;
      m = pdone;
      erank = prank;
    } else {
      if (!have_probe) break;           // the query ended exactly at a node boundary
      // ---- resolve the probe in flight (one step of linear probing per iteration)
      if ((unsigned)S.a.x != want_tok || (unsigned)S.a.y != want_par) {
        if ((S.a.x & S.a.y) == -1) break;   // empty slot: no child starts with this token
        si = (si + 1) & tmask;
        S = rm_ld_slot(table + si);
        continue;
      }
      c = S.a.z;
      elen = rm_slot_len((unsigned)S.a.w);
      erank = rm_slot_rank((unsigned)S.a.w);
      const int n = min(elen, L - pos);
      // tokens 1 and 2 of the edge are in the slot, tokens 1 and 2 after the key token are in g1, g2
      m = 1;
      if (n >= 2 && S.b.z == g1) {
        m = 2;
        if (n >= 3 && S.b.w == g2) m = 3;
      }
      if (elen <= 3 && m == elen) {
        // ---- short edge matched in full: issue the next probe first, commit while it is in flight
        const int tk = (elen == 1) ? g1 : (elen == 2) ? g2 : g3;
        const int nxt = pos + elen;
        have_probe = nxt < L;
        if (have_probe) {
          want_par = (unsigned)c;
          want_tok = (unsigned)tk;
          si = rm_hash2((unsigned)c, want_tok, tshift);
          S = rm_ld_slot(table + si);
        }
        if (ROUTER) {
          ++cnt;
          deepest = erank;
          if (erank < t.n_prefill) {
            last_prefill = erank;
            bad_below = false;
          } else if (erank >= t.n_prefill + t.n_decode) {
            bad_below = true;
          }
        }
        node = c;
        pos = nxt;
        // slide the three look-ahead tokens by elen
        if (elen == 1) {
          g1 = g2;
          g2 = g3;
        } else if (elen == 2) {
          g1 = g3;
          g2 = (nxt + 2 < L) ? rm_ld_query(key + nxt + 2, qpol) : 0;
        } else {
          g1 = (nxt + 1 < L) ? rm_ld_query(key + nxt + 1, qpol) : 0;
          g2 = (nxt + 2 < L) ? rm_ld_query(key + nxt + 2, qpol) : 0;
        }
        g3 = (nxt + 3 < L) ? rm_ld_query(key + nxt + 3, qpol) : 0;
        continue;
      }
      if (m == 3 && n > 3) {
        // ---- long edge whose first three tokens match: compare the rest chunk by chunk
        pend = true;
        pc = c;
        pelen = elen;
        pn = n;
        pdone = 3;
        prank = erank;
        pe = t.tok + (i64)(((u64)(unsigned)S.b.y << 32) | (unsigned)S.b.x);
        const int* __restrict__ ee = pe + 3;
        const int* __restrict__ kk = key + pos + 3;
        const int l2 = n - 3;
#pragma unroll
        for (int x = 0; x < RM_CH; ++x) {
          if (x < l2) {
            ca[x] = __ldg(ee + x);
            cb[x] = rm_ld_query(kk + x, qpol);
          }
        }
        continue;
      }
      // otherwise the edge is decided right here with m < elen (mismatch or query exhausted)
    }
    // ---- an edge has been decided with m of its elen tokens matching, and it is not a fully matched short edge
    const bool partial = m < elen;
    if (SPLIT && partial) {
      // the reference splits c here and returns the new upper node (radix_cache.py:262-266); K3 does it
      pos += m;
      req = ((u64)(unsigned)c << len_bits) | (u64)(unsigned)m;
      break;
    }
    if (ROUTER) {
      ++cnt;
      deepest = erank;
      if (erank < t.n_prefill) {
        last_prefill = erank;
        bad_below = false;
      } else if (erank >= t.n_prefill + t.n_decode) {
        bad_below = true;
      }
    }
    pos += m;
    node = c;
    if (partial) break;                 // no-split modes: the partially matched child is the last node
    // a long edge matched in full: only now look up the next hop
    have_probe = pos < L;
    if (!have_probe) break;
    want_par = (unsigned)c;
    want_tok = (unsigned)rm_ld_query(key + pos, qpol);
    si = rm_hash2((unsigned)c, want_tok, tshift);
    S = rm_ld_slot(table + si);
    g1 = (pos + 1 < L) ? rm_ld_query(key + pos + 1, qpol) : 0;
    g2 = (pos + 2 < L) ? rm_ld_query(key + pos + 2, qpol) : 0;
    g3 = (pos + 3 < L) ? rm_ld_query(key + pos + 3, qpol) : 0;
  }
  // Lazy stamps: an operation stamps only the deepest node it reached (radix_cache.py:253-260 stamps the whole
  // path; every node on it is an ancestor of this one, and eviction takes a node's stamp as the maximum over
  // its subtree, rm_maint.cu).  One atomic per query instead of one per hop.
  if (node > 0 && !(flags & 1)) atomicMax(stamp + node, seq);
  out_len[q] = pos;
  out_node[q] = node;
  if (SPLIT && req != RM_NO_REQ && !(flags & 4)) {   // flags bit 2: walk only (insert pre-pass)
    const int slot = atomicAdd(&t.ctr->n_req, 1);     // compact list of split requests
    req_key[slot] = req;
    req_val[slot] = (unsigned)q;
    // per-node chain of this batch's requests: lets K3 tell at once whether the node has a single request
    t.req_link[q] = atomicExch(&t.req_head[(int)(req >> len_bits)], q);
  }
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

// K1, run-following form (the default).  First-sight splits turn a hot prefix into a chain of one-token
// nodes: ~71 hops per 256-token Zipf query on average and ~225 at the tail, each hop a dependent ~0.2 us
// child-table read, which is what bounded the lane-per-query kernel above at ~0.18 ms per 64K batch.  But a
// split is zero-copy, so such a chain is still ONE contiguous run of the token arena.  The owner map
// seg[] (rm_common.cuh) says, per arena position, which node covers it and whether a node is the
// contiguous child of its predecessor.  A lane therefore probes the child table only where the path
// really branches (the root, a leaf appended by an insert, the final miss) and otherwise streams along
// the run, comparing arena tokens with query tokens RM_RUN positions per step and reading node
// boundaries from seg[] as it goes.  Every node passed is still stamped (its id comes from seg[]), so
// results and stamps are those of the hop-by-hop walk.
#define RM_RUN 8
template <bool SPLIT, bool ROUTER>
__global__ void __launch_bounds__(RM_MATCH_TPB)
rm_match_run_kernel(RmTree t, const int* __restrict__ qtok, const i64* __restrict__ qoff, int B, u64 seq_base,
                    int len_bits, int* __restrict__ out_len, int* __restrict__ out_node, u64* __restrict__ req_key,
                    unsigned* __restrict__ req_val, RmRouteOut ro, int flags) {
  const int q = (int)(blockIdx.x * (unsigned)blockDim.x + threadIdx.x);
  if (q >= B) return;
  const i64 o0 = qoff[q];
  const int L = (int)(qoff[q + 1] - o0);
  const int* __restrict__ key = qtok + o0;
  const u64 seq = ((seq_base + (u64)q) << 1) | 1ull;
  const RmSlot* __restrict__ table = t.table;
  const unsigned tmask = t.table_mask, tshift = t.table_shift;
  const int* __restrict__ tok = t.tok;
  const int* __restrict__ seg = t.seg;
  u64* __restrict__ stamp = t.stamp;
  const u64 qpol = rm_policy_evict_first();

  int node = 0, pos = 0;               // `node` is matched in full through `pos`
  u64 req = RM_NO_REQ;
  int cnt = 0, deepest = -1, last_prefill = -1;
  bool bad_below = false;

#define RM_ROUTER_SEE(rk)                          \
  do {                                             \
    ++cnt;                                         \
    deepest = (rk);                                \
    if ((rk) < t.n_prefill) {                      \
      last_prefill = (rk);                         \
      bad_below = false;                           \
    } else if ((rk) >= t.n_prefill + t.n_decode) { \
      bad_below = true;                            \
    }                                              \
  } while (0)

  bool done = L == 0;
  while (!done) {
    // ---- enter a child of `node` that starts with key[pos]
    const int tk = rm_ld_query(key + pos, qpol);
    int cur = -1;
    i64 cs = 0;                        // arena position where cur's run starts
    {
      // (the contiguous child, if any, was already ruled out when the previous run ended: table probe)
      unsigned si = rm_hash2((unsigned)node, (unsigned)tk, tshift);
      while (true) {
        const RmSlotRegs S = rm_ld_slot(table + si);
        if ((unsigned)S.a.x == (unsigned)tk && (unsigned)S.a.y == (unsigned)node) {
          cur = S.a.z;
          cs = (i64)(((u64)(unsigned)S.b.y << 32) | (unsigned)S.b.x);
          break;
        }
        if ((S.a.x & S.a.y) == -1) break;   // empty slot: no child starts with this token
        si = (si + 1) & tmask;
      }
      if (cur < 0) break;               // the walk ends at `node`
    }
    // ---- stream along the arena run from cs: token cs matched key[pos]
    i64 a = cs + 1;
    int qi = pos + 1;
    bool leave = false;                 // back to the outer loop for a table probe
    while (!leave && !done) {
      int ta[RM_RUN], sa[RM_RUN], ka[RM_RUN];
#pragma unroll
      for (int j = 0; j < RM_RUN; ++j) {
        ta[j] = __ldg(tok + a + j);
        sa[j] = __ldg(seg + a + j);
        ka[j] = (qi + j < L) ? rm_ld_query(key + qi + j, qpol) : 0;
      }
#pragma unroll
      for (int j = 0; j < RM_RUN; ++j) {
        if (leave || done) break;
        const int sg = sa[j];
        const int owner = sg < 0 ? -1 : (sg & 0x3FFFFFFF);
        if (owner != cur) {
          // cur's run ends here: cur is matched in full
          if (ROUTER) {
            const int rk = t.rank[cur];
            RM_ROUTER_SEE(rk);
          }
          node = cur;
          pos = qi;
          if (qi >= L) {
            done = true;
          } else if (sg >= 0 && (sg & RM_SEG_CONT) && ta[j] == ka[j]) {
            cur = owner;                // the contiguous child continues the match
            cs = a;
            ++a;
            ++qi;
          } else {
            leave = true;               // not the contiguous child: back to the child table
          }
        } else if (qi >= L || ta[j] != ka[j]) {
          // the query ends, or differs, inside cur: partial match of p = a - cs tokens (>= 1)
          const int p = (int)(a - cs);
          if (SPLIT) {
            // the reference splits cur here and returns the new upper node (radix_cache.py:262-266); K3 does it
            req = ((u64)(unsigned)cur << len_bits) | (u64)(unsigned)p;
            pos = qi;
          } else {
              if (ROUTER) {
              const int rk = t.rank[cur];
              RM_ROUTER_SEE(rk);
            }
            node = cur;                 // no-split modes return the partially matched child
            pos = qi;
          }
          done = true;
        } else {
          ++a;
          ++qi;
        }
      }
    }
  }
#undef RM_ROUTER_SEE
  // Lazy stamps: an operation stamps only the deepest node it reached (radix_cache.py:253-260 stamps the whole
  // path; every node on it is an ancestor of this one, and eviction takes a node's stamp as the maximum over
  // its subtree, rm_maint.cu).  One atomic per query instead of one per hop.
  if (node > 0 && !(flags & 1)) atomicMax(stamp + node, seq);
  out_len[q] = pos;
  out_node[q] = node;
  if (SPLIT && req != RM_NO_REQ && !(flags & 4)) {
    const int slot = atomicAdd(&t.ctr->n_req, 1);
    req_key[slot] = req;
    req_val[slot] = (unsigned)q;
    // per-node chain of this batch's requests: lets K3 tell at once whether the node has a single request
    t.req_link[q] = atomicExch(&t.req_head[(int)(req >> len_bits)], q);
  }
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

// K1, run-following GROUP form: G lanes per query, one arena position per lane and step, so the three
// streams (arena tokens, owner map, query tokens) are read as coalesced 32..128-byte pieces and a step
// costs one memory round trip for G positions.  (The one-lane form above issues the same loads as 3*8
// scattered 4-byte requests per lane and is bound by the load/store unit.)  All lanes of a group carry the
// same walk state; node boundaries inside a step are resolved with ballots: the first position where the
// match stops, the nodes that ended before it (each stamped by the lane that sees its end).
#ifndef RM_RUN_DEPTH
#define RM_RUN_DEPTH 2   // steps whose loads are in flight together; RM_RUN_DEPTH * 32 <= RM_ARENA_PAD
#endif
template <int G, bool SPLIT, bool ROUTER>
__global__ void __launch_bounds__(RM_MATCH_TPB, 8)
rm_match_rung_kernel(RmTree t, const int* __restrict__ qtok, const i64* __restrict__ qoff, int B, u64 seq_base,
                     int len_bits, int* __restrict__ out_len, int* __restrict__ out_node, u64* __restrict__ req_key,
                     unsigned* __restrict__ req_val, RmRouteOut ro, int flags) {
  RmGroup<G> g;
  rm_grid_dependency_trigger();        // the split kernels that follow may be scheduled now; they wait for this grid
  const int q = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G);
  if (q >= B) return;
  const i64 o0 = qoff[q];
  const int L = (int)(qoff[q + 1] - o0);
  const int* __restrict__ key = qtok + o0;
  const u64 seq = ((seq_base + (u64)q) << 1) | 1ull;
  const RmSlot* __restrict__ table = t.table;
  const unsigned tmask = t.table_mask, tshift = t.table_shift;
  const int* __restrict__ tok = t.tok;
  const int* __restrict__ seg = t.seg;
  u64* __restrict__ stamp = t.stamp;
  const u64 qpol = rm_policy_evict_first();
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
    const int tk = have_tk ? tk_next : rm_ld_query(key + pos, qpol);
    have_tk = false;
    int cur = -1;
    i64 cs = 0;
    {
      unsigned si = rm_hash2((unsigned)node, (unsigned)tk, tshift);
      while (true) {
        const RmSlotRegs S = rm_ld_slot(table + si);
        if ((unsigned)S.a.x == (unsigned)tk && (unsigned)S.a.y == (unsigned)node) {
          cur = S.a.z;
          cs = (i64)(((u64)(unsigned)S.b.y << 32) | (unsigned)S.b.x);
          break;
        }
        if ((S.a.x & S.a.y) == -1) break;
        si = (si + 1) & tmask;
      }
    }
    if (cur < 0) break;                 // the walk ends at `node`
    // ---- stream along the arena run: position a + lane against query token qi + lane
    i64 a = cs + 1;
    int qi = pos + 1;
    bool leave = false;
    while (!leave && !done) {
      // the loads of RM_RUN_DEPTH steps are issued together: a run is ~68 positions on the C2 tree, and every
      // step that waits for its own loads adds one memory round trip to the query's dependent chain
      int tjv[RM_RUN_DEPTH], sjv[RM_RUN_DEPTH], kjv[RM_RUN_DEPTH];
#pragma unroll
      for (int u = 0; u < RM_RUN_DEPTH; ++u) {
        const i64 ia = a + u * G + lane;
        const int iq = qi + u * G + lane;
        tjv[u] = __ldg(tok + ia);
        sjv[u] = __ldg(seg + ia);
        kjv[u] = iq < L ? rm_ld_query(key + iq, qpol) : 0;
      }
#pragma unroll
      for (int u = 0; u < RM_RUN_DEPTH; ++u) {
      const int tj = tjv[u], sj = sjv[u], kj = kjv[u];
      const bool inq = qi + lane < L;
      const int own = sj < 0 ? -1 : (sj & 0x3FFFFFFF);
      int prev = g.shfl_up(own, 1);
      if (lane == 0) prev = cur;
      const bool bnd = own != prev;                                   // node `prev` ends at this position
      const bool badb = bnd && !(sj >= 0 && (sj & RM_SEG_CONT));      // ... and the chain does not go on
      const bool stop = badb || !inq || tj != kj;
      const unsigned sm = g.ballot(stop);
      const int s = sm ? (__ffs(sm) - 1) : G;                         // first position where the match stops
      const bool commit = bnd && lane <= s;                           // `prev` is matched in full
      const unsigned cm = g.ballot(commit);
      if (ROUTER) {
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
      int jb = -1;                       // last boundary committed in this step
      if (cm) {
        jb = 31 - __clz((int)cm);
        node = g.bcast(prev, jb);
        pos = qi + jb;
        cs = a + jb;                     // the node that starts there is the current one from now on
        cur = g.bcast(own, jb);
      }
      if (s == G) {                      // everything matched: next G positions
        a += G;
        qi += G;
        continue;
      }
      // the match stops at relative position s
      const bool bnd_s = g.bcast(bnd ? 1 : 0, s) != 0;
      if (bnd_s) {
        // a node ended exactly there (it is committed); either the query is exhausted or the next node
        // is not a contiguous match: back to the child table
        if (qi + s >= L) {
          done = true;
        } else {
          leave = true;
          tk_next = g.bcast(kj, s);      // key[qi + s]: the token the next probe hashes
          have_tk = true;
        }
      } else {
        // the query ends, or differs, inside the node that covers position s: partial match
        const int pn = g.bcast(own, s);
        const int p = (int)(a + s - cs);
        if (SPLIT) {
          // the reference splits pn here and returns the new upper node (radix_cache.py:262-266); K3 does it
          req = ((u64)(unsigned)pn << len_bits) | (u64)(unsigned)p;
        } else {
          if (ROUTER) {
            const int rk = t.rank[pn];
            ++cnt;
            deepest = rk;
            if (rk < t.n_prefill) {
              last_prefill = rk;
              bad_below = false;
            } else if (rk >= t.n_prefill + t.n_decode) {
              bad_below = true;
            }
          }
          node = pn;                     // no-split modes return the partially matched child
        }
        pos = qi + s;
        done = true;
      }
      break;
      }
    }
  }
  if (lane == 0) {
    if (node > 0 && !(flags & 1)) atomicMax(stamp + node, seq);   // lazy stamps: the endpoint only
    out_len[q] = pos;
    out_node[q] = node;
    if (SPLIT && req != RM_NO_REQ && !(flags & 4)) {
      const int slot = atomicAdd(&t.ctr->n_req, 1);
      req_key[slot] = req;
      req_val[slot] = (unsigned)q;
      // per-node chain of this batch's requests: lets K3 tell at once whether the node has a single request
      t.req_link[q] = atomicExch(&t.req_head[(int)(req >> len_bits)], q);
    }
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

// K3: one group per edge that received split requests.  `skey` is sorted by (node, split point) and,
// inside equal keys, by query index (the sort is stable), `sval` holds the query index.
// (a device function: the stand-alone kernel and the one-launch form rm_split_fused_kernel share it.  The sorted
// list may have been written earlier in the SAME launch, by other blocks: it is read with ld.cg.)
template <int G>
__device__ __forceinline__ void rm_split_apply_body(const RmTree& t, const u64* skey_, const unsigned* sval_, int Bmax, int len_bits,
                                                    u64 seq_base, int* __restrict__ out_node, u64* __restrict__ sufmax,
                                                    int* __restrict__ krank, int* __restrict__ uniq_pos,
                                                    unsigned* __restrict__ uniq_min) {
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
  const int gstride = (int)((gridDim.x * (unsigned)blockDim.x) / G);
  for (int i = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) / G); i < B; i += gstride) {
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
    atomicMax(&t.ctr->node_hwm, u + 1);
    // owner map of segment [a, p): the first position keeps the edge's own flag (kk == 0) or marks u as the
    // contiguous child of the previous segment
    {
      const int first = kk == 0 ? ((t.seg[off] & RM_SEG_CONT) | u) : (u | RM_SEG_CONT);
      t.seg[off + a] = first;
      fill_off = off + a;          // positions a+1 .. p-1: filled by the whole group below
      fill_n = p - a;
      fill_id = u;
      if (kk == U - 1) t.seg[off + p] = c | RM_SEG_CONT;
    }
    if (kk == 0) {
      RmSlot s;
      int sidx = rm_find<true>(t, rm_key(par, t.tok[off]), s);
      if (sidx < 0) {
        atomicMax(&t.ctr->err, (int)RM_E_INVALID);
      } else {
        t.table[sidx].child = u;
        rm_slot_set_len(&t.table[sidx], p);
      }
    } else {
      rm_table_insert(t, rm_key(uprev, t.tok[off + a]), u, p - a, off + a, rk);
    }
    if (kk == U - 1) {
      rm_table_insert(t, rm_key(u, t.tok[off + p]), c, len - p, off + p, rk);
      t.parent[c] = u;
      t.edge_off[c] = off + p;
      t.edge_len[c] = len - p;
      u64 smin = ((seq_base + (u64)uniq_min[i + kk]) << 1) | 1ull;
      t.stamp[c] = smin > st_c ? smin : st_c;
    }
    }
    if (fill_n <= RM_FILL_COOP) {
      for (int x = 1; x < fill_n; ++x) t.seg[fill_off + x] = fill_id;
      fill_n = 0;
    }
    unsigned todo = g.ballot(fill_n > 1);
    while (todo) {
      const int l = __ffs(todo) - 1;
      todo &= todo - 1;
      const i64 o = g.bcast(fill_off, l);
      const int n = g.bcast(fill_n, l);
      const int id = g.bcast(fill_id, l);
      for (int x = 1 + g.lane; x < n; x += G) t.seg[o + x] = id;
    }
  }
  if (g.lane == 0) rm_add_i64(&t.ctr->n_nodes, (i64)U);
  // pass 4: every request's last node is the upper node that ends at its split point
  for (int j = i + g.lane; j < e; j += G) out_node[sval[j]] = rm_free_id_at(t, apos + (u64)krank[j]);
  }
}

template <int G>
__global__ void __launch_bounds__(RM_MATCH_TPB)
rm_split_apply_kernel(RmTree t, const u64* skey, const unsigned* sval, int Bmax, int len_bits, u64 seq_base,
                      int* __restrict__ out_node, u64* __restrict__ sufmax, int* __restrict__ krank,
                      int* __restrict__ uniq_pos, unsigned* __restrict__ uniq_min) {
  rm_grid_dependency_wait();               // launched while the sort may still be running (RM_LAUNCH_PDL)
  rm_split_apply_body<G>(t, skey, sval, Bmax, len_bits, seq_base, out_node, sufmax, krank, uniq_pos, uniq_min);
}

// K3, common case: an edge that received exactly ONE split request this batch (in steady state almost
// all of them: ~10K single requests per 64K-query batch against a few hundred shared ones).  One thread per
// request of K1's compact, unsorted list, so 32 independent splits share every memory round trip of a
// warp and nothing has to be sorted for them: a request is alone on its node iff it is the head of the
// node's request chain and has no predecessor.  Requests on shared edges are copied to a second list,
// which alone goes through the radix sort and rm_split_apply_kernel (a warp per edge).
// (device function, one call per request index i < B; every thread of a warp must make the call, also for i >= B)
__device__ __forceinline__ void rm_split_single_body(const RmTree& t, const u64* rkey, const unsigned* rval, int B, int i,
                                                     int len_bits, u64 seq_base, int* __restrict__ out_node, u64* mkey,
                                                     unsigned* mval) {
  // owner-map fill of the new upper node, done by the whole warp after the per-thread part (a thread writing
  // its own p-1 words one after another made this kernel 57 us on 4096-token keys)
  i64 fill_off = 0;
  int fill_n = 0, fill_id = 0;
  if (i < B) {
    const u64 ki = rkey[i];
    const int c = (int)(ki >> len_bits);
    const unsigned q = rval[i];
    // everything that depends only on (c, q) is loaded at once, before any of it is tested: one memory round trip
    // instead of four in a row (chain head, chain link, node record, edge tokens)
    const int head = t.req_head[c];
    const int link = t.req_link[q];
    const int par = t.parent[c];
    const i64 off = t.edge_off[c];
    const int len = t.edge_len[c];
    const int dend = t.depth_end[c];
    const int lk = t.lock_ref[c];
    const int rk = t.rank[c];
    const u64 st_c = t.stamp[c];
    const bool is_head = head == (int)q;
    const bool alone = is_head && link < 0;
    if (is_head) t.req_head[c] = -1;                                      // ready for the next batch
    if (!alone) {                                                         // shared edge: sorted, then a warp does it
      const int slot = atomicAdd(&t.ctr->n_big, 1);
      mkey[slot] = ki;
      mval[slot] = q;
    } else {
      const int p = (int)(ki & ((1ull << len_bits) - 1ull));
      const u64 seq = ((seq_base + (u64)q) << 1) | 1ull;
      // second round trip: the id allocation, the edge's tokens and its owner word, together
      const int tok0 = t.tok[off], tokp = t.tok[off + p];
      const int seg0 = t.seg[off];
      const u64 apos = rm_alloc_nodes(t, 1);
      if (apos != ~0ull) {
        // third: the new id and the parent's slot
        RmSlot s;
        const int sidx = rm_find<true>(t, rm_key(par, tok0), s);
        const int u = rm_free_id_at(t, apos);
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
        atomicMax(&t.ctr->node_hwm, u + 1);
        t.seg[off] = (seg0 & RM_SEG_CONT) | u;           // owner map: [off, off+p) is u's, c starts at off+p
        t.seg[off + p] = c | RM_SEG_CONT;
        fill_off = off;
        fill_n = p;
        fill_id = u;
        if (sidx < 0) {
          atomicMax(&t.ctr->err, (int)RM_E_INVALID);
        } else {
          t.table[sidx].child = u;
          rm_slot_set_len(&t.table[sidx], p);
        }
        rm_table_insert(t, rm_key(u, tokp), c, len - p, off + p, rk);
        t.parent[c] = u;
        t.edge_off[c] = off + p;
        t.edge_len[c] = len - p;
        t.stamp[c] = st;
        rm_add_i64(&t.ctr->n_nodes, (i64)1);
        out_node[q] = u;
      }
    }
  }
#ifdef RM_EMU
  for (int x = 1; x < fill_n; ++x) t.seg[fill_off + x] = fill_id;
#else
  {
    // short fills stay with their thread (32 of them proceed side by side); only long ones are worth
    // serialising the warp for
    if (fill_n <= RM_FILL_COOP) {
      for (int x = 1; x < fill_n; ++x) t.seg[fill_off + x] = fill_id;
      fill_n = 0;
    }
    const int lane = threadIdx.x & 31;
    unsigned todo = __ballot_sync(0xffffffffu, fill_n > 1);
    while (todo) {
      const int l = __ffs(todo) - 1;
      todo &= todo - 1;
      const i64 o = __shfl_sync(0xffffffffu, fill_off, l);
      const int n = __shfl_sync(0xffffffffu, fill_n, l);
      const int id = __shfl_sync(0xffffffffu, fill_id, l);
      for (int x = 1 + lane; x < n; x += 32) t.seg[o + x] = id;
    }
  }
#endif
}

__global__ void __launch_bounds__(RM_MATCH_TPB)
rm_split_single_kernel(RmTree t, const u64* rkey, const unsigned* rval, int Bmax, int len_bits, u64 seq_base,
                       int* __restrict__ out_node, u64* mkey, unsigned* mval) {
  rm_grid_dependency_trigger();
  rm_grid_dependency_wait();               // launched while the walk kernel may still be running (RM_LAUNCH_PDL)
  const int B = min(t.ctr->n_req, Bmax);
  rm_split_single_body(t, rkey, rval, B, (int)(blockIdx.x * (unsigned)blockDim.x + threadIdx.x), len_bits, seq_base, out_node,
                       mkey, mval);
}

#ifndef RM_EMU
// K3 in ONE launch: single-request splits, grid barrier, sort of the requests on shared edges, grid barrier,
// shared-edge splits.  Three dependent launches of ~10 us each (most of it launch and drain latency: each stage is
// a few microseconds of work on a few thousand requests) become one launch with two barriers of ~2 us.  The grid
// is sized to be resident as a whole (the barrier is a counter the blocks wait on, rm_sort.cuh).
// Measured (C2): 0.105 ms per step against 0.101-0.104 for the three launches with programmatic dependent launch:
// what a stage costs is its own chain of dependent memory round trips, not the launch.  Kept behind RM_FUSED_SPLIT=1.
static_assert(RM_MATCH_TPB == RM_SORT_TPB, "rm_sort_body is written for blocks of RM_SORT_TPB threads");
template <int G>
__global__ void __launch_bounds__(RM_MATCH_TPB)
rm_split_fused_kernel(RmTree t, u64* rkey, unsigned* rval, u64* mkey, unsigned* mval, int Bmax, int len_bits, int passes,
                      u64 seq_base, int* __restrict__ out_node, u64* __restrict__ sufmax, int* __restrict__ krank,
                      int* __restrict__ uniq_pos, unsigned* __restrict__ uniq_min, unsigned* hist, int tile) {
  rm_grid_dependency_wait();               // launched while the walk kernel may still be running (RM_LAUNCH_PDL)
  unsigned arrivals = 0;
  const int n_req = min(t.ctr->n_req, Bmax);
  const int gthreads = (int)(gridDim.x * blockDim.x);
  for (int base = (int)(blockIdx.x * blockDim.x); base < n_req; base += gthreads)
    rm_split_single_body(t, rkey, rval, n_req, base + (int)threadIdx.x, len_bits, seq_base, out_node, mkey, mval);
  arrivals += gridDim.x;
  rm_sort_grid_barrier(hist, arrivals);
  const int nbig_now = *(volatile int*)&t.ctr->n_big;
  const int nbig = min(nbig_now, Bmax);
  if (nbig > 0) rm_sort_body(mkey, mval, rkey, rval, nbig, passes, hist, tile, (int)blockIdx.x, (int)gridDim.x, arrivals);
  arrivals += gridDim.x;
  rm_sort_grid_barrier(hist, arrivals);
  const bool in_r = (passes & 1) != 0;     // where rm_sort_body left the result
  rm_split_apply_body<G>(t, in_r ? rkey : mkey, in_r ? rval : mval, Bmax, len_bits, seq_base, out_node, sufmax, krank,
                         uniq_pos, uniq_min);
  rm_sort_barrier_rearm(hist, (int)gridDim.x);
}

// blocks of rm_split_fused_kernel that are resident together on this device (0: do not use it)
static int rm_split_fused_grid() {
  static int grid = -1;
  if (grid < 0) {
    const char* e = getenv("RM_FUSED_SPLIT");
    int per_sm = 0, dev = 0, sms = 0;
    auto k = rm_split_fused_kernel<32>;
    if (!(e && e[0] == '1') || cudaGetDevice(&dev) != cudaSuccess ||   // opt-in: measured no faster than the three PDL launches
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, RM_MATCH_TPB, 0) != cudaSuccess)
      grid = 0;
    else
      grid = sms * (per_sm > 2 ? 2 : per_sm);
  }
  return grid;
}
#endif

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
    g = e ? atoi(e) : 1;
    if (g != 1 && g != 2 && g != 4 && g != 8 && g != 16 && g != 32) g = 1;
  }
  return g;
#endif
}

template <bool SPLIT, bool ROUTER>
static void rm_match_dispatch(rm_index* idx, const int* tokens, const i64* offsets, int B, u64 seq_base, int* out_len,
                              int* out_node, RmRouteOut ro, cudaStream_t s) {
  const int G = rm_match_group();
  const long long threads = (long long)B * G;
#ifndef RM_EMU
  // a grid of B*G threads must still cover all 148 SMs several times: narrow groups get small blocks
  const int tpb = threads >= 148LL * 8 * RM_MATCH_TPB ? RM_MATCH_TPB : (threads >= 148LL * 8 * 128 ? 128 : 64);
#else
  const int tpb = RM_MATCH_TPB;
#endif
  const int grid = (int)((threads + tpb - 1) / tpb);
#ifdef RM_EMU
  // the emulation build runs the lane-per-query kernel (the default on the GPU) unless RM_EMU_GROUP is set
  static const bool emu_group = getenv("RM_EMU_GROUP") != nullptr;
  static const bool emu_lpq = getenv("RM_EMU_LPQ") != nullptr;
  static const bool emu_run1 = getenv("RM_EMU_RUN1") != nullptr;
  if (!emu_group && !emu_lpq) {
    if (emu_run1) {
      auto kr = rm_match_run_kernel<SPLIT, ROUTER>;
      RM_LAUNCH(kr, grid, tpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,
                idx->sc.key[0], idx->sc.val[0], ro, 0);
    } else {
      auto kr = rm_match_rung_kernel<1, SPLIT, ROUTER>;
      RM_LAUNCH(kr, grid, tpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,
                idx->sc.key[0], idx->sc.val[0], ro, 0);
    }
    return;
  }
  if (!emu_group) {
    auto kq = rm_match_lpq_kernel<SPLIT, ROUTER>;
    RM_LAUNCH(kq, grid, tpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,
              idx->sc.key[0], idx->sc.val[0], ro, 0);
    return;
  }
  auto k = rm_match_kernel<1, SPLIT, ROUTER>;
  RM_LAUNCH(k, grid, tpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,
            idx->sc.key[0], idx->sc.val[0], ro);
#else
#define RM_CASE(GG)                                                                                              \
  case GG: {                                                                                                     \
    auto k = rm_match_kernel<GG, SPLIT, ROUTER>;                                                                 \
    RM_LAUNCH(k, grid, tpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,          \
              idx->sc.key[0], idx->sc.val[0], ro);                                                               \
  } break;
  // RM_WALK selects the walk kernel: "run" (default) = run-following, RM_RUN_G lanes per query (default 32);
  // "lane" = run-following, one lane per query; "table" = hop-by-hop through the child table, RM_MATCH_G lanes
  // per query (1 = the lane-per-query kernel).  The non-default forms are kept for A/B measurements and tests.
  static const char* walk = getenv("RM_WALK") ? getenv("RM_WALK") : "run";
  static const int run_flags = getenv("RM_EXP_FLAGS") ? (atoi(getenv("RM_EXP_FLAGS")) & 1) : 0;
  if (!strcmp(walk, "run")) {
    static const int rg = getenv("RM_RUN_G") ? atoi(getenv("RM_RUN_G")) : 32;
    const long long rthreads = (long long)B * rg;
    static const int rtpb_env = getenv("RM_RUN_TPB") ? atoi(getenv("RM_RUN_TPB")) : 0;
    // 4 queries per block: a block's warps end at different times and its slots are only handed on when the last
    // one is done (128 threads: +2 % over 256 on both bench shapes)
    const int rtpb = rtpb_env ? rtpb_env : (rthreads >= 148LL * 8 * 128 ? 128 : 64);
    const int rgrid = (int)((rthreads + rtpb - 1) / rtpb);
#define RM_RCASE(GG)                                                                                              \
  case GG: {                                                                                                      \
    auto kr = rm_match_rung_kernel<GG, SPLIT, ROUTER>;                                                            \
    RM_LAUNCH(kr, rgrid, rtpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,        \
              idx->sc.key[0], idx->sc.val[0], ro, run_flags);                                                     \
  } break;
    switch (rg) {
      RM_RCASE(4)
      RM_RCASE(8)
      RM_RCASE(16)
      default:
      RM_RCASE(32)
    }
#undef RM_RCASE
    return;
  }
  if (!strcmp(walk, "lane")) {
    const int ltpb = B >= 148 * 8 * RM_MATCH_TPB ? RM_MATCH_TPB : (B >= 148 * 8 * 128 ? 128 : 64);
    auto kr = rm_match_run_kernel<SPLIT, ROUTER>;
    RM_LAUNCH(kr, (B + ltpb - 1) / ltpb, ltpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,
              idx->sc.key[0], idx->sc.val[0], ro, run_flags);
    return;
  }
  switch (G) {
    case 1: {
      auto k = rm_match_lpq_kernel<SPLIT, ROUTER>;
      // RM_EXP_FLAGS: measurement experiments only (bit 0 = skip the stamp atomics; results then differ
      // from the reference in eviction order, never use it outside a profiling run)
      static const int exp_flags = getenv("RM_EXP_FLAGS") ? atoi(getenv("RM_EXP_FLAGS")) : 0;
      const int grid2 = (int)(((threads << (exp_flags >> 4)) + tpb - 1) / tpb);
      RM_LAUNCH(k, grid2, tpb, s, idx->t, tokens, offsets, B, seq_base, idx->len_bits, out_len, out_node,
                idx->sc.key[0], idx->sc.val[0], ro, exp_flags);
    } break;
    RM_CASE(2)
    RM_CASE(4)
    RM_CASE(8)
    RM_CASE(16)
    RM_CASE(32)
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
#else
  const bool prof = false;
#endif
  RM_CUDA(cudaMemsetAsync(&idx->t.ctr->n_req, 0, 2 * sizeof(int), s));   // n_req and n_big (adjacent)
  idx->t.req_link = idx->sc.link;
#ifndef RM_EMU
  if (prof) cudaEventRecord(pev[0], s);
#endif
  rm_match_dispatch<true, false>(idx, tokens, offsets, B, seq_base, out_len, out_node, ro, s);
#ifndef RM_EMU
  if (prof) cudaEventRecord(pev[1], s);
#endif
  RM_CUDA(cudaGetLastError());
  // requests that are alone on their edge are applied straight from K1's list; the rest (requests on
  // shared edges) are sorted by (node, split point) and applied a warp per edge.  All counts stay on the device.
  RmScratch& sc = idx->sc;
  const int bits = idx->node_bits + idx->len_bits;
#ifndef RM_EMU
  {
    const int fgrid = rm_split_fused_grid();
    const int tile = rm_sort_tile(B);
    if (B > 1 && fgrid > 0 && (B + tile - 1) / tile <= min(fgrid, RM_SORT_FUSED_TILES)) {
      auto kf = rm_split_fused_kernel<RM_WIDE_G>;
      RM_LAUNCH_PDL(kf, fgrid, RM_MATCH_TPB, s, idx->t, sc.key[0], sc.val[0], sc.key[1], sc.val[1], B, idx->len_bits, (bits + 7) / 8,
                    seq_base, out_node, sc.sufmax, sc.krank, sc.uniq_pos, sc.uniq_min, sc.hist, tile);
      if (prof) {
        if (idx->prof_on > 1) {
          cudaEventRecord(pev[2], s);
          cudaEventRecord(pev[3], s);
          cudaEventRecord(pev[4], s);
        }
        ++idx->prof_n;
      }
      RM_CUDA(cudaGetLastError());
      return RM_OK;
    }
  }
#endif
  RM_LAUNCH_PDL(rm_split_single_kernel, (B + RM_MATCH_TPB - 1) / RM_MATCH_TPB, RM_MATCH_TPB, s, idx->t, sc.key[0], sc.val[0], B,
                idx->len_bits, seq_base, out_node, sc.key[1], sc.val[1]);
#ifndef RM_EMU
  if (prof && idx->prof_on > 1) cudaEventRecord(pev[2], s);
#endif
  const int w = 1 ^ rm_sort_pairs(sc.key[1], sc.val[1], sc.key[0], sc.val[0], B, bits, sc.hist, s, &idx->t.ctr->n_big);
  RM_CUDA(cudaGetLastError());
#ifndef RM_EMU
  if (prof && idx->prof_on > 1) cudaEventRecord(pev[3], s);
#endif
  const long long threads = (long long)B * RM_WIDE_G;
  int grid = (int)((threads + RM_MATCH_TPB - 1) / RM_MATCH_TPB);
#ifndef RM_EMU
  if (grid > 148 * 8) grid = 148 * 8;    // persistent: 8 blocks of 8 warps per SM, grid-stride inside
#endif
  auto k3 = rm_split_apply_kernel<RM_WIDE_G>;
  RM_LAUNCH_PDL(k3, grid, RM_MATCH_TPB, s, idx->t, sc.key[w], sc.val[w], B, idx->len_bits, seq_base, out_node, sc.sufmax,
                sc.krank, sc.uniq_pos, sc.uniq_min);
#ifndef RM_EMU
  if (prof) {
    if (idx->prof_on > 1) cudaEventRecord(pev[4], s);
    ++idx->prof_n;
  }
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
#else
  const int tpb = N >= 148 * 8 * RM_MATCH_TPB ? RM_MATCH_TPB : (N >= 148 * 8 * 128 ? 128 : 64);
#endif
#ifdef RM_EMU
  auto k = rm_match_rung_kernel<1, true, false>;
  const long long wthreads = N;
#else
  auto k = rm_match_rung_kernel<8, true, false>;
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
