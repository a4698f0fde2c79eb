"""Full-size matched-length oracle (TEST INFRASTRUCTURE, not product code; same import rule as radix_port.py).

What it states.  `RadixCache.match_prefix(key)` (radix/sglang/srt/mem_cache/radix_cache.py:127-162, 252-275) returns
the values of the longest prefix of `key` that is a prefix of some key inserted so far: the walk follows children
while tokens agree (`_key_match_page_size1`, :14-20) and a partial edge match still counts the tokens that agreed
(`_split_node`, :262-266, 277-294).  Splits change node boundaries, never content, so the matched length depends
only on the SET of inserted keys:

    matched_len(q) = max over inserted keys k of  LCP(q, k)

and in lexicographic order that maximum is attained at one of q's two neighbours among the inserted keys.  This
module computes exactly that with numpy (a sort of the keys as big-endian byte strings + two vectorised LCPs per
query), which makes the 1 M-node / 512-token and 4096-token shapes checkable in seconds, where the pure-Python
restatement (`PortRadixCache`) needs an hour to build the tree.

Parity status: PINNED through `PortRadixCache` — `tests/test_lcp_oracle.py` compares the two on seeded Zipf
workloads and on small-vocabulary random keys (where chance collisions between documents are common); the port
itself is pinned against the unmodified reference (see radix_port.py).

Restrictions: all rows have the same length L (the SURVEY.md 8d generator's shape); tokens are non-negative int32.
"""
from __future__ import annotations

import numpy as np


def _as_sortable(rows: np.ndarray) -> np.ndarray:
    """[n, L] non-negative int32 -> n byte strings whose memcmp order is the lexicographic token order."""
    n, L = rows.shape
    be = np.ascontiguousarray(rows.astype(">u4"))
    return be.view(f"S{4 * L}").reshape(n)


def _lcp_rows(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    neq = a != b
    first = neq.argmax(axis=1)
    first[~neq.any(axis=1)] = a.shape[1]
    return first.astype(np.int32)


def longest_prefix_lengths(inserted: np.ndarray, queries: np.ndarray, chunk: int = 1 << 14) -> np.ndarray:
    """matched_len for every row of `queries` against the set of rows of `inserted` (both [*, L] int32)."""
    N, L = inserted.shape
    Q = queries.shape[0]
    assert queries.shape[1] == L
    out = np.zeros(Q, dtype=np.int32)
    if N == 0 or Q == 0:
        return out
    ins_sorted = inserted[np.argsort(_as_sortable(inserted), kind="stable")]
    keys_sorted = _as_sortable(ins_sorted)
    qk = _as_sortable(queries)
    pos = np.searchsorted(keys_sorted, qk, side="left")     # inserted[pos-1] < q <= inserted[pos]
    for a in range(0, Q, chunk):
        b = min(a + chunk, Q)
        p = pos[a:b]
        q = queries[a:b]
        lo = _lcp_rows(q, ins_sorted[np.maximum(p - 1, 0)])
        lo[p == 0] = 0
        hi = _lcp_rows(q, ins_sorted[np.minimum(p, N - 1)])
        hi[p == N] = 0
        out[a:b] = np.maximum(lo, hi)
    return out
