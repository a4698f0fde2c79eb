"""CPU oracle for the RadixMesh hot path (TEST INFRASTRUCTURE, not product code).

This file is a from-scratch CPU restatement of the reference's algorithm for the
one hot path this repo accelerates.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it; the
product package ``radixmesh_b200`` never does (it fails loudly without its CUDA
library).

Parity status: PINNED.  The reference is pure Python and imports in the build
container, so this restatement is validated there by differential tests against
the unmodified reference (``tests/test_oracle_vs_reference.py``, skipped where
``/root/reference`` is absent) and by the golden vectors generated from the
reference and committed under ``tests/golden/`` (``tools/make_golden.py``).

It keeps the reference's data structures on purpose (a dict of children per node
keyed by the first token, Python lists for keys that are re-sliced per hop, a
binary heap for eviction) so that, timed as the CPU baseline, it costs what the
reference costs.

What each piece follows (paths relative to /root/reference/python/src):
  * ``common_prefix_len``        radix/sglang/srt/mem_cache/radix_cache.py:14-20
  * ``common_prefix_len_paged``  radix_cache.py:23-32
  * ``Node``                     radix_cache.py:35-64  (TreeNode)
  * ``PortRadixCache``           radix_cache.py:87-377 (RadixCache)
  * ``OwnedValue / RankValue``   radix/radix_mesh.py:21-63
  * ``PortRadixMesh``            radix/radix_mesh.py:72-495 (tree ops only; the
                                 TCP ring, ticker and GC threads are out of scope)
  * ``keep_existing``            policy/conflict_resolve.py:1-6
  * ``PortConsistentHash``       router/cache_aware_router.py:42-118
  * ``port_route``               router/cache_aware_router.py:23-39

Clock.  The reference stamps nodes with ``time.monotonic()``; only the order of
stamps matters (eviction).  The oracle uses a logical clock: every public tree
operation gets the next *op sequence number* and every stamp taken inside it is
``(op_seq << SUB_BITS) + k`` with k counting the stamps of that operation, which
preserves the reference's "later call => larger stamp" order exactly.  The CUDA
path stores just ``op_seq`` per node (one stamp per operation), so
``stamp >> SUB_BITS`` is what parity tests compare.
"""
from __future__ import annotations

import bisect
import hashlib
import heapq
from typing import Dict, List, Optional, Tuple

import numpy as np

SUB_BITS = 20

MODE_BASE = 0      # plain RadixCache semantics (splitting match, no ranks)
MODE_PREFILL = 1   # RadixMesh, mode PREFILL (splitting match, owned values)
MODE_DECODE = 2    # RadixMesh, mode DECODE  (non-splitting match, owned values)
MODE_ROUTER = 3    # RadixMesh, mode ROUTER  (non-splitting match, rank-only values)


class LogicalClock:
    """Deterministic stand-in for time.monotonic(); see module docstring."""

    def __init__(self) -> None:
        self.op_seq = 0
        self._sub = 0

    def begin_op(self) -> int:
        self.op_seq += 1
        self._sub = 0
        return self.op_seq

    def now(self) -> int:
        self._sub += 1
        return (self.op_seq << SUB_BITS) + self._sub


def common_prefix_len(a, b) -> int:
    """radix_cache.py:14-20 — token-by-token scan."""
    n = 0
    for x, y in zip(a, b):
        if x != y:
            break
        n += 1
    return n


def common_prefix_len_paged(a, b, page: int) -> int:
    """radix_cache.py:23-32 — compare whole pages."""
    lim = min(len(a), len(b))
    n = 0
    while n < lim:
        if a[n:n + page] != b[n:n + page]:
            break
        n += page
    return n


class Node:
    """radix_cache.py:35-64.  ``uid`` is a creation counter like TreeNode.id; parity
    never compares it (replicas disagree on it), identity is the full key."""

    __slots__ = ("children", "parent", "key", "value", "lock_ref", "stamp", "uid")
    _next_uid = 0

    def __init__(self, clock: LogicalClock) -> None:
        self.children: Dict = {}
        self.parent: Optional["Node"] = None
        self.key: Optional[list] = None
        self.value = None
        self.lock_ref = 0
        self.stamp = clock.now()
        self.uid = Node._next_uid
        Node._next_uid += 1

    def __lt__(self, other: "Node") -> bool:
        return self.stamp < other.stamp


class MatchOut:
    """radix_cache.py:67-84 — (device_indices, last node)."""

    __slots__ = ("indices", "last_node")

    def __init__(self, indices: np.ndarray, last_node: Node) -> None:
        self.indices = indices
        self.last_node = last_node


_EMPTY = np.empty((0,), dtype=np.int64)


class PortRadixCache:
    """Restatement of RadixCache (radix_cache.py:87-377)."""

    def __init__(self, page_size: int = 1, clock: Optional[LogicalClock] = None,
                 free_hook=None) -> None:
        self.page = page_size
        self.clock = clock or LogicalClock()
        self.free_hook = free_hook          # allocator.free stand-in (radix_cache.py:195)
        self.freed: List[list] = []
        self.hops = 0
        self.reset()

    # -- helpers -----------------------------------------------------------
    def _ckey(self, key):
        # radix_cache.py:109-114
        return key[0] if self.page == 1 else tuple(key[:self.page])

    def _lcp(self, a, b) -> int:
        if self.page == 1:
            return common_prefix_len(a, b)
        return common_prefix_len_paged(a, b, self.page)

    # -- public surface ----------------------------------------------------
    def reset(self) -> None:
        # radix_cache.py:118-125
        self.clock.begin_op()
        self.root = Node(self.clock)
        self.root.key = []
        self.root.value = []
        self.root.lock_ref = 1
        self.evictable = 0
        self.protected = 0

    def match_prefix(self, key) -> MatchOut:
        # radix_cache.py:127-162
        if len(key) == 0:
            return MatchOut(_EMPTY, self.root)
        if self.page != 1:
            key = key[:len(key) // self.page * self.page]
        self.clock.begin_op()
        vals, last = self._match_walk(self.root, key)
        return MatchOut(np.concatenate(vals) if vals else _EMPTY, last)

    def insert(self, key, value=None) -> int:
        # radix_cache.py:164-170
        if value is None:
            value = np.asarray(key, dtype=np.int64)
        self.clock.begin_op()
        return self._insert_walk(self.root, key, value)

    def evict(self, num_tokens: int) -> None:
        # radix_cache.py:179-202
        self.clock.begin_op()
        heap = self._leaves()
        heapq.heapify(heap)
        done = 0
        while done < num_tokens and heap:
            x = heapq.heappop(heap)
            if x is self.root:
                break
            if x.lock_ref > 0:
                continue
            self._free(x.value)
            done += len(x.value)
            self._drop_leaf(x)
            if len(x.parent.children) == 0:
                heapq.heappush(heap, x.parent)

    def inc_lock_ref(self, node: Node) -> int:
        # radix_cache.py:204-216
        delta = 0
        while node is not self.root:
            if node.lock_ref == 0:
                self.evictable -= len(node.value)
                self.protected += len(node.value)
                delta -= len(node.value)
            node.lock_ref += 1
            node = node.parent
        return delta

    def dec_lock_ref(self, node: Node) -> int:
        # radix_cache.py:218-230
        delta = 0
        while node is not self.root:
            if node.lock_ref == 1:
                self.evictable += len(node.value)
                self.protected -= len(node.value)
                delta += len(node.value)
            node.lock_ref -= 1
            node = node.parent
        return delta

    def evictable_size(self) -> int:
        return self.evictable

    def protected_size(self) -> int:
        return self.protected

    def total_size(self) -> int:
        # radix_cache.py:354-364
        tot, stack = 0, [self.root]
        while stack:
            n = stack.pop()
            tot += len(n.value)
            stack.extend(c for c in n.children.values() if c.value is not None)
        return tot

    def all_values_flatten(self) -> np.ndarray:
        # radix_cache.py:239-248 (pre-order DFS in dict insertion order)
        out: List[np.ndarray] = []

        def rec(n: Node) -> None:
            for c in n.children.values():
                out.append(self._raw(c.value))
                rec(c)

        rec(self.root)
        return np.concatenate(out) if out else _EMPTY

    # -- internals ---------------------------------------------------------
    def _raw(self, v):
        return v

    def _free(self, v) -> None:
        rec = list(np.asarray(self._raw(v)).tolist())
        self.freed.append(rec)
        if self.free_hook is not None:
            self.free_hook(rec)

    def _match_walk(self, node: Node, key):
        # radix_cache.py:252-275 — splitting match
        node.stamp = self.clock.now()
        ck = self._ckey(key)
        vals = []
        while len(key) > 0 and ck in node.children:
            child = node.children[ck]
            child.stamp = self.clock.now()
            self.hops += 1          # instrumentation only (roofline bytes per lookup)
            p = self._lcp(child.key, key)
            if p < len(child.key):
                upper = self._split(child, p)
                vals.append(upper.value)
                node = upper
                break
            vals.append(child.value)
            node = child
            key = key[p:]
            if len(key):
                ck = self._ckey(key)
        return vals, node

    def _split(self, child: Node, at: int) -> Node:
        # radix_cache.py:277-294 — new upper node, the lower half keeps the object
        upper = Node(self.clock)
        upper.children = {self._ckey(child.key[at:]): child}
        upper.parent = child.parent
        upper.lock_ref = child.lock_ref
        upper.key = child.key[:at]
        upper.value = child.value[:at]
        full_first = self._ckey(child.key)
        child.parent = upper
        child.key = child.key[at:]
        child.value = child.value[at:]
        upper.parent.children[full_first] = upper
        return upper

    def _insert_walk(self, node: Node, key, value) -> int:
        # radix_cache.py:296-327
        node.stamp = self.clock.now()
        if len(key) == 0:
            return 0
        ck = self._ckey(key)
        total = 0
        while len(key) > 0 and ck in node.children:
            node = node.children[ck]
            node.stamp = self.clock.now()
            p = self._lcp(node.key, key)
            total += p
            key = key[p:]
            value = value[p:]
            if p < len(node.key):
                node = self._split(node, p)
            if len(key):
                ck = self._ckey(key)
        if len(key):
            leaf = Node(self.clock)
            leaf.parent = node
            leaf.key = key
            leaf.value = value
            node.children[ck] = leaf
            self.evictable += len(value)
        return total

    def _drop_leaf(self, node: Node) -> None:
        # radix_cache.py:347-352
        for k, v in node.parent.children.items():
            if v is node:
                break
        del node.parent.children[k]
        self.evictable -= len(node.key)

    def _leaves(self) -> List[Node]:
        # radix_cache.py:366-377
        out, stack = [], [self.root]
        while stack:
            n = stack.pop()
            if len(n.children) == 0:
                out.append(n)
            else:
                stack.extend(n.children.values())
        return out

    # -- parity helpers (not in the reference) -------------------------------
    def full_key(self, node: Node) -> list:
        # radix_mesh.py:459-464
        key = node.key
        while node is not self.root:
            node = node.parent
            key = node.key + key
        return key

    def iter_nodes(self):
        stack = [self.root]
        while stack:
            n = stack.pop()
            if n is not self.root:
                yield n
            stack.extend(n.children.values())

    def canonical(self) -> dict:
        """Sorted (full_key, value, rank, lock_ref, op_seq-of-stamp) per non-root node
        plus the two size counters (SURVEY.md 7.2 'canonical tree dump')."""
        rows = []
        for n in self.iter_nodes():
            v = n.value
            rank = getattr(v, "rank", -1)
            raw = self._raw(v)
            rows.append((tuple(self.full_key(n)), len(n.key),
                         tuple(int(x) for x in raw) if raw is not None else (),
                         rank, n.lock_ref, n.stamp >> SUB_BITS))
        rows.sort()
        return {"nodes": rows, "evictable": self.evictable, "protected": self.protected}


# ---------------------------------------------------------------------------
# Mesh semantics
# ---------------------------------------------------------------------------

class OwnedValue:
    """radix_mesh.py:21-44 — KV indices plus the rank that owns them."""

    __slots__ = ("arr", "rank")

    def __init__(self, arr, rank: int) -> None:
        self.arr = arr
        self.rank = rank

    def __getitem__(self, s: slice) -> "OwnedValue":
        return OwnedValue(self.arr[s], self.rank)

    def same(self, other: "OwnedValue") -> bool:
        # torch.equal: same shape and same elements (radix_mesh.py:33-36)
        return self.arr.shape == other.arr.shape and bool(np.array_equal(self.arr, other.arr))

    def __len__(self) -> int:
        return len(self.arr)


class RankValue:
    """radix_mesh.py:47-63 — router trees store only the owner rank; len() is 1."""

    __slots__ = ("rank",)

    def __init__(self, rank: int) -> None:
        self.rank = rank

    def __getitem__(self, s: slice) -> "RankValue":
        return RankValue(self.rank)

    def same(self, other: "RankValue") -> bool:
        return self.rank == other.rank

    def __len__(self) -> int:
        return 1


def keep_existing(now_rank: int, new_rank: int) -> bool:
    """policy/conflict_resolve.py:1-6 — the lower rank wins a multi-write."""
    return now_rank <= new_rank


class RouterPickError(AssertionError):
    """radix_mesh.py:233-237 assertion failures."""


class PortRadixMesh(PortRadixCache):
    """Tree operations of RadixMesh (radix_mesh.py:72-495) for one replica."""

    def __init__(self, mode: int, self_rank: int, n_prefill: int, n_decode: int,
                 clock: Optional[LogicalClock] = None) -> None:
        assert mode in (MODE_PREFILL, MODE_DECODE, MODE_ROUTER)
        self.mode = mode
        self.self_rank = self_rank
        self.n_prefill = n_prefill
        self.n_decode = n_decode
        self.dup_nodes: Dict[Tuple[Tuple[int, ...], int], Optional[Node]] = {}
        super().__init__(page_size=1, clock=clock)

    # rank predicates, config/cache_config.py:20-35
    def is_prefill_rank(self, r: int) -> bool:
        return r < self.n_prefill

    def is_decode_rank(self, r: int) -> bool:
        return self.n_prefill <= r < self.n_prefill + self.n_decode

    def local_rank(self, r: int) -> int:
        if r < self.n_prefill:
            return r
        if r < self.n_prefill + self.n_decode:
            return r - self.n_prefill
        return r - self.n_prefill - self.n_decode

    def reset(self) -> None:
        # radix_mesh.py:240-245 (master rank is 0, policy/sync_algo.py:7)
        super().reset()
        if self.mode == MODE_ROUTER:
            self.root.value = RankValue(0)
        else:
            self.root.value = OwnedValue(None, 0)

    def _raw(self, v):
        return v.arr if isinstance(v, OwnedValue) else None

    # -- match -------------------------------------------------------------
    def match_prefix(self, key):
        # radix_mesh.py:203-238; no empty-key guard -> IndexError like the reference
        self.clock.begin_op()
        vals, last = self._mesh_match_walk(self.root, key)
        if self.mode != MODE_ROUTER:
            arrs = [v.arr for v in vals]
            return MatchOut(np.concatenate(arrs) if arrs else _EMPTY, last)
        return self.pick_ranks([v.rank for v in vals])

    def pick_ranks(self, ranks: List[int]) -> Tuple[int, int]:
        # radix_mesh.py:220-238
        prefill = decode = -1
        if len(ranks) == 1:
            prefill = ranks[0]
        elif len(ranks) > 1:
            for r in reversed(ranks):
                if self.is_prefill_rank(r):
                    prefill = r
                    break
                if not self.is_decode_rank(r):
                    raise RouterPickError(f"rank {r} is invalid")
                if decode == -1:
                    decode = r
            if prefill == -1:
                raise RouterPickError("no prefill rank on the matched path")
        return prefill, decode

    def _mesh_match_walk(self, node: Node, key):
        # radix_mesh.py:247-271
        if self.mode == MODE_PREFILL:
            if len(key) == 0:
                raise IndexError("list index out of range")
            return self._match_walk(node, key)
        node.stamp = self.clock.now()
        ck = key[0]
        vals = []
        while len(key) > 0 and ck in node.children:
            child = node.children[ck]
            child.stamp = self.clock.now()
            p = common_prefix_len(child.key, key)
            if p < len(child.key):
                vals.append(child.value[:p])
                node = child
                break
            vals.append(child.value)
            node = child
            key = key[p:]
            if len(key):
                ck = key[0]
        return vals, node

    # -- insert ------------------------------------------------------------
    def insert(self, key, value=None) -> int:
        # radix_mesh.py:193-201 (local insert: owner = this node)
        assert self.mode in (MODE_PREFILL, MODE_DECODE), "only prefill/decode nodes insert"
        return self.apply_insert(key, value, self.self_rank)

    def apply_insert(self, key, value, owner_rank: int) -> int:
        """One INSERT oplog applied (radix_mesh.py:411-416 -> :197-201 -> :273-323)."""
        if self.mode == MODE_ROUTER:
            v = RankValue(owner_rank)
        else:
            v = OwnedValue(np.asarray(value, dtype=np.int64), owner_rank)
        self.clock.begin_op()
        return self._mesh_insert_walk(self.root, key, v)

    def _mesh_insert_walk(self, node: Node, key, value) -> int:
        # radix_mesh.py:273-323
        node.stamp = self.clock.now()
        if len(key) == 0:
            return 0
        ck = key[0]
        total = 0
        while len(key) > 0 and ck in node.children:
            node = node.children[ck]
            node.stamp = self.clock.now()
            p = common_prefix_len(node.key, key)
            total += p
            incoming = None
            lost = False
            if not value[:p].same(node.value[:p]):
                if keep_existing(node.value.rank, value.rank):
                    lost = True
                else:
                    incoming = value[:p]
            key = key[p:]
            value = value[p:]
            if p < len(node.key):
                node = self._split(node, p)
                if incoming is not None:
                    node = self._swap(incoming, node, ck)
            elif incoming is not None:
                node = self._swap(incoming, node, ck)
            if lost:
                self.dup_nodes[(tuple(self.full_key(node)), value.rank)] = None
            if len(key):
                ck = key[0]
        if len(key):
            leaf = Node(self.clock)
            leaf.parent = node
            leaf.key = key
            leaf.value = value
            node.children[ck] = leaf
            self.evictable += len(value)
        return total

    def _swap(self, incoming, node: Node, ck) -> Node:
        # radix_mesh.py:466-495.  Divergence (documented in DESIGN.md): the locked
        # branch of the reference iterates dict keys and raises AttributeError when
        # the node has children; here the children are re-parented (the evident intent).
        if self.mode == MODE_ROUTER:
            node.value = incoming
            return node
        old_rank = node.value.rank
        if node.lock_ref == 0:
            node.value = incoming
            self.dup_nodes[(tuple(self.full_key(node)), old_rank)] = None
            return node
        parent = node.parent
        repl = Node(self.clock)
        repl.value = incoming
        repl.key = node.key
        repl.parent = parent
        repl.children = node.children
        for c in node.children.values():
            c.parent = repl
        repl.lock_ref = 0
        parent.children[ck] = repl
        node.children = {}
        self.dup_nodes[(tuple(self.full_key(node)), old_rank)] = node
        return repl

    # -- lock refs / evict ---------------------------------------------------
    def dec_lock_ref(self, node: Node) -> int:
        # radix_mesh.py:168-179 — never below zero
        delta = 0
        while node is not self.root:
            if node.lock_ref == 1:
                self.evictable += len(node.value)
                self.protected -= len(node.value)
                delta += len(node.value)
            if node.lock_ref > 0:
                node.lock_ref -= 1
            node = node.parent
        return delta

    def evict(self, *_a, **_k) -> None:
        # radix_mesh.py:349-351 — no-op in the mesh
        return None

    # -- GC-dup candidate scan ----------------------------------------------
    def gc_candidates(self) -> List[Tuple[Tuple[int, ...], int]]:
        # radix_mesh.py:151-156
        return [k for k, n in self.dup_nodes.items() if n is not None and n.lock_ref == 0]

    def gc_agrees(self, k) -> bool:
        # radix_mesh.py:386-388
        return k in self.dup_nodes and (self.dup_nodes[k] is None or self.dup_nodes[k].lock_ref == 0)

    def gc_exec(self, keys, origin: bool) -> list:
        # radix_mesh.py:363-377: GC_EXEC pops the keys everywhere; the node that asked first frees the KV indices
        # of the parked nodes it holds for them (token_to_kv_pool_allocator.free(value.value.tolist()))
        freed = []
        for k in keys:
            n = self.dup_nodes.get(k)
            if origin and n is not None:
                freed.append([int(x) for x in self._raw(n.value)])
            self.dup_nodes.pop(k, None)
        return freed

    def dup_canonical(self) -> dict:
        return {k: (None if n is None else n.lock_ref) for k, n in self.dup_nodes.items()}


# ---------------------------------------------------------------------------
# Router
# ---------------------------------------------------------------------------

def md5_u32(text: str) -> int:
    """cache_aware_router.py:110-118 — first 4 digest bytes, big-endian."""
    return int.from_bytes(hashlib.md5(text.encode("utf-8")).digest()[:4], "big")


class PortConsistentHash:
    """cache_aware_router.py:42-108 — 3 virtual nodes per address, bisect-right."""

    def __init__(self, nodes, replicas: int = 3) -> None:
        self.ring: List[int] = []
        self.owner: Dict[int, str] = {}
        seen = set()
        for n in nodes or []:
            if n in seen:
                continue
            seen.add(n)
            for i in range(replicas):
                h = md5_u32(f"{n}#{i}")
                self.owner[h] = n
                bisect.insort(self.ring, h)

    def get_node(self, key):
        if not self.ring:
            return None
        h = md5_u32(str(key))
        return self.owner[self.ring[bisect.bisect(self.ring, h) % len(self.ring)]]


def port_route(mesh: PortRadixMesh, key, prefill_addrs, decode_addrs, warm: bool = True):
    """cache_aware_router.py:23-39 — returns (prefill_addr, decode_addr)."""
    if warm:
        prefill, decode = mesh.match_prefix(key)
    else:
        prefill, decode = -1, -1
    if prefill != -1:
        p_addr = prefill_addrs[prefill]                       # radix_mesh.py:447-451
    else:
        p_addr = PortConsistentHash(prefill_addrs).get_node(key)
    if decode != -1:
        d_addr = decode_addrs[mesh.local_rank(decode)]        # radix_mesh.py:453-457
    else:
        d_addr = PortConsistentHash(decode_addrs).get_node(key)
    return p_addr, d_addr


class PortMesh:
    """P prefill + D decode + R router replicas fed the same INSERT stream in the order
    the reference's ring delivers it (origin first, then around the ring, routers via
    rank 0; radix_mesh.py:339-347, policy/sync_algo.py:57-75)."""

    def __init__(self, n_prefill: int = 3, n_decode: int = 2, n_router: int = 1) -> None:
        self.P, self.D, self.R = n_prefill, n_decode, n_router
        self.nodes: List[PortRadixMesh] = []
        for r in range(n_prefill + n_decode + n_router):
            mode = MODE_PREFILL if r < n_prefill else MODE_DECODE if r < n_prefill + n_decode else MODE_ROUTER
            self.nodes.append(PortRadixMesh(mode, r, n_prefill, n_decode))
        self.p_addrs = [f"localhost:{50000 + i}" for i in range(n_prefill)]
        self.d_addrs = [f"localhost:{50000 + n_prefill + i}" for i in range(n_decode)]

    def insert(self, rank: int, key, value) -> int:
        ring = self.P + self.D
        assert rank < ring
        ret = self.nodes[rank].apply_insert(list(key), value, rank)
        for hop in range(1, ring):
            self.nodes[(rank + hop) % ring].apply_insert(list(key), value, rank)
        for r in range(ring, ring + self.R):
            self.nodes[r].apply_insert(list(key), value, rank)
        return ret

    def match(self, rank: int, key):
        return self.nodes[rank].match_prefix(list(key))

    def route(self, key, router: int = 0, warm: bool = True):
        return port_route(self.nodes[self.P + self.D + router], list(key), self.p_addrs, self.d_addrs, warm)

    def canonical(self, rank: int) -> dict:
        t = self.nodes[rank]
        rows = []
        for n in t.iter_nodes():
            raw = t._raw(n.value)
            rows.append((tuple(t.full_key(n)), len(n.key),
                         tuple(int(x) for x in raw) if raw is not None else (), n.value.rank, n.lock_ref))
        rows.sort()
        return {"nodes": rows, "evictable": t.evictable, "protected": t.protected, "dup": t.dup_canonical()}
