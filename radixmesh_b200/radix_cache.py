"""Drop-in for the reference's `RadixCache` (radix/sglang/srt/mem_cache/radix_cache.py:87-436) on the
device index.  Same constructor, same method names, same argument meaning and return values; the
tree itself lives in HBM behind the C ABI (`capi.DeviceIndex`).  `*_batch` methods are additions.

Single-key calls are thin wrappers over the batch entry points (a batch of one); the throughput
path is `match_prefix_batch` / `insert_batch` with tensors.
"""
from __future__ import annotations

import dataclasses
from typing import Any, List, Optional

import numpy as np

from . import capi

PAGE_FILL = -(1 << 31)     # fills up a short last page of a paged key (RadixCache.__init__); no key may contain it


class TreeNode:
    """Handle to a node of the device tree (reference: radix_cache.py:35-64).  Attributes are read from
    the device on access, a few dozen bytes per node through rm_node_info / rm_node_children whatever the
    size of the tree.  `id` is the node's pool id; like the reference's creation counter it is not
    comparable across replicas — identity for parity is the full key."""

    __slots__ = ("_cache", "id", "_gen")

    def __init__(self, cache: "RadixCache", node_id: int, gen: int = -1):
        self._cache = cache
        self.id = int(node_id)
        self._gen = int(gen)

    def _resolved(self) -> int:
        return self._cache._resolve(self.id, self._gen)

    @property
    def key(self) -> List[int]:
        return self._cache._node_tokens(self._resolved())

    @property
    def value(self):
        return self._cache._node_value(self._resolved())

    @property
    def lock_ref(self) -> int:
        return int(self._cache._info(self._resolved()).lock_ref)

    @property
    def last_access_time(self) -> int:
        return int(self._cache._eff_stamps()[self._resolved()])

    @property
    def parent(self) -> Optional["TreeNode"]:
        x = self._resolved()
        if x == 0:
            return None
        return self._cache._handle(int(self._cache._info(x).parent))

    @property
    def children(self) -> dict:
        """{child key: TreeNode} in the reference's dict order (one scan of the pool on the device, the children back)."""
        return self._cache._children(self._resolved())

    hit_count = 0          # never incremented in the reference either
    loading = False
    host_value = None

    @property
    def evicted(self) -> bool:
        return False

    @property
    def backuped(self) -> bool:
        return False

    def __eq__(self, other) -> bool:
        return isinstance(other, TreeNode) and other._cache is self._cache and other._resolved() == self._resolved()

    def __hash__(self) -> int:
        return hash((id(self._cache), self.id, self._gen))

    def __lt__(self, other: "TreeNode") -> bool:
        return self.last_access_time < other.last_access_time

    def __repr__(self) -> str:
        return f"TreeNode(id={self.id})"


@dataclasses.dataclass
class MatchResult:
    """radix_cache.py:67-84."""
    device_indices: Any
    last_device_node: Any
    last_host_node: Any
    host_hit_length: int = 0


class RadixCache:
    MODE = capi.MODE_BASE

    def __init__(self, req_to_token_pool, token_to_kv_pool_allocator, page_size: int, disable: bool = False,
                 enable_kv_cache_events: bool = False, *, max_nodes: int = 1 << 20, max_tokens: int = 1 << 26,
                 max_batch: int = 1 << 16, max_key_len: int = 1 << 15, max_dups: int = 1 << 12, device: int = 0,
                 max_page_symbols: Optional[int] = None, _index_kwargs: Optional[dict] = None, _backend=None, _library=None):
        if page_size < 1 or (page_size != 1 and self.MODE != capi.MODE_BASE):
            raise ValueError("page_size > 1 is a RadixCache feature; RadixMesh hard-codes 1 (radix_mesh.py:87)")
        # Paged keys (radix_cache.py:23-32, 109-114: child key = first page, keys compared a page at a time, splits at
        # page multiples).  A paged radix tree over tokens IS a radix tree over page symbols, so the device index runs
        # unchanged on symbols; pages are interned on the device (rm_pages_intern: hash table, every hit verified
        # against the stored page) and the KV indices of a page live in a device row store, the index's value of a
        # symbol position being the row id (capi.DevicePages, csrc/rm_pages.cu).  Lengths, sizes, lock deltas and the
        # eviction budget scale by page_size.  max_tokens stays a capacity in TOKENS.
        self._P = int(page_size)
        self._pg = None
        self.req_to_token_pool = req_to_token_pool
        self.token_to_kv_pool_allocator = token_to_kv_pool_allocator
        self.page_size = page_size
        self.disable = disable
        if enable_kv_cache_events:
            # the reference's constructor ends in reset() -> _record_all_cleared_event() -> AllBlocksCleared(), a class
            # it never imports (radix_cache.py:117-125, 422-424): same exception, said in full
            raise NameError("name 'AllBlocksCleared' is not defined (the reference's KV-event classes are never imported: "
                            "enable_kv_cache_events=True cannot construct a RadixCache there either; events are not produced)")
        self.enable_kv_cache_events = False
        self.kv_event_queue = []
        P = self._P
        kw = dict(mode=self.MODE, max_nodes=max_nodes, max_tokens=max(max_tokens // P, 1), max_batch=max_batch,
                  max_key_len=max(max_key_len // P, 1), max_dups=max_dups, device=device, backend=_backend, library=_library)
        kw.update(_index_kwargs or {})
        self._ix = capi.DeviceIndex(**kw)
        self._be = self._ix.be
        if P > 1:
            # distinct pages the interning table can hold at a time (those of evicted nodes are dropped when it fills)
            n_sym = int(max_page_symbols) if max_page_symbols else max(max_tokens // P, 1) + (1 << 12)
            self._pg_syms_total = n_sym
            self._pg_syms_after_retain = 0
            self._pg_rows_total = n_sym + max_batch * max(max_key_len // P, 1)
            self._pg = capi.DevicePages(P, n_sym, self._pg_rows_total, device=device, backend=self._be, library=self._ix.lib)
        # A key that is not a whole number of pages: the reference stores the short last page in the leaf, keyed by the
        # short tuple (radix_cache.py:109-114, 296-326).  Here the short page is filled up with a token no key may
        # contain, which makes it a page (and a symbol) of its own - equal only to the same short page, as the
        # reference's tuple compare has it; the host remembers which value rows are short, and by how much.
        self._ragged = {}              # value row id -> tokens the short page really holds
        self._deficit = 0              # sum over live short rows of (page_size - tokens held)
        self._deficit_protected = 0    # ... the part of it that sits in locked leaves
        self._arena_est = 0            # upper bound on the arena positions handed out (see _arena_room)
        self._arena_floor = 0          # arena positions in use right after the last compaction
        self.device = getattr(self._be, "device", "cpu")
        self._dump_cache = None
        self._info_cache = {}
        self.root_node = TreeNode(self, 0, 0)

    # ------------------------------------------------------------------ helpers
    def _invalidate(self) -> None:
        self._dump_cache = None
        self._info_cache = {}

    def _info(self, x: int):
        """rm_node_info row of node x (cached until the next mutation)."""
        self._flush()
        r = self._info_cache.get(x)
        if r is None:
            r = self._info_cache[x] = self._ix.node_info([x])[0]
        return r

    def _children(self, x: int) -> dict:
        self._flush()
        out = {}
        for cid, tok, gen in self._ix.node_children(x):
            key = tok if self._P == 1 else tuple(t for t in self._pages_of([tok]) if t != PAGE_FILL)   # a short page keys by the short tuple
            out[key] = self.root_node if cid == 0 else TreeNode(self, cid, gen)
        return out

    def _dump(self) -> dict:
        self._flush()          # staged remote inserts (RadixMesh) must be visible to every read
        if self._dump_cache is None:
            self._dump_cache = self._ix.dump()
        return self._dump_cache

    def _eff_stamps(self) -> np.ndarray:
        """Last-access stamp per node as the reference holds it.  The device stamps only the deepest node an
        operation reaches; a node's stamp is the maximum over its subtree (DESIGN.md 'lazy stamps')."""
        d = self._dump()
        if "eff" not in d:
            n = d["n"]
            eff = d["stamp"][:n].astype(np.uint64).copy()
            live = [x for x in range(1, n) if d["parent"][x] > 0 and d["gen"][x] >= 0]
            for x in sorted(live, key=lambda z: -int(d["depth"][z])):
                p = int(d["parent"][x])
                if eff[x] > eff[p]:
                    eff[p] = eff[x]
            d["eff"] = eff
        return d["eff"]

    def _flush(self) -> None:       # RadixMesh stages remote inserts; the base class has nothing pending
        return None

    def _handle(self, node_id: int) -> TreeNode:
        if node_id == 0:
            return self.root_node
        return TreeNode(self, node_id, int(self._info(node_id).gen))

    def _resolve(self, node_id: int, gen: int) -> int:
        if gen < 0 or node_id == 0:
            return node_id
        r = self._info(node_id)
        if int(r.gen) == gen:
            return node_id
        x = int(r.parked)
        while x >= 0:
            r = self._info(x)
            if int(r.gen) == -(gen + 1):
                return x
            x = int(r.parked)
        return node_id

    def _node_tokens(self, x: int) -> List[int]:
        if x == 0:
            return []
        r = self._info(x)
        tok, _ = self._ix.read_arena(int(r.edge_off), int(r.edge_len), values=False)
        if self._P > 1:
            toks = self._pages_of(tok)
            while toks and toks[-1] == PAGE_FILL:       # a short last page (see __init__)
                toks.pop()
            return toks
        return tok.tolist()

    def _node_value(self, x: int):
        import torch
        if x == 0:
            return []
        r = self._info(x)
        _, val = self._ix.read_arena(int(r.edge_off), int(r.edge_len))
        if self._P > 1:
            val = self._run_values(val)
        return torch.from_numpy(val)

    def _tensor(self, a):
        import torch
        return a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))

    # ------------------------------------------------------------------ paged keys: symbols and value rows on the device
    def _xp(self, x):
        """torch for device tensors, numpy for the host-emulation backend."""
        if hasattr(x, "data_ptr"):
            import torch
            return torch
        return np

    def _pages_of(self, syms) -> List[int]:
        """Page tokens of a list of symbols (TreeNode.key, child keys)."""
        syms = np.ascontiguousarray(syms, dtype=np.int32)
        n = len(syms)
        if n == 0:
            return []
        out = self._be.empty(n * self._P, np.int32)
        self._pg.tokens(self._be.to_device(syms), n, out)
        return self._be.to_host(out).tolist()

    def _rows_values(self, rows) -> np.ndarray:
        """KV indices of value rows (host array of row ids in, host array out)."""
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        n = len(rows)
        if n == 0:
            return np.zeros(0, dtype=np.int64)
        out = self._be.empty(n * self._P, np.int64)
        self._pg.rows_get(self._be.to_device(rows), n, out)
        return self._be.to_host(out)

    def _arena_room(self, n_tokens: int) -> None:
        """Before an insert of up to n_tokens arena positions.  The reference's tree gives memory back when a node is
        evicted; here an evicted run's arena positions come back only through rm_compact (arena space is handed out by
        a bump pointer), so a cache that streams inserts and evictions compacts by itself: when three quarters of the
        arena are handed out and at least a quarter of it belongs to evicted runs.  The device is asked only when the
        host-side bound passes the threshold."""
        if self.MODE == capi.MODE_ROUTER:      # a router replica holds no values and counts nodes, not tokens
            return
        self._arena_est += int(n_tokens)
        cap = int(self._ix.cfg.max_tokens)
        if self._arena_est * 4 <= cap * 3:
            return
        sz = self._ix.sizes()
        used = int(sz.arena_used) + int(n_tokens)
        # (a mesh replica never evicts - RadixMesh.evict is a no-op like the reference's, radix_mesh.py:349-351 - so
        # there only the runs that duplicate-node GC released come back; its parked nodes hold arena runs that neither
        # size counter shows, hence: after a compaction the next one waits until an eighth of the arena has been handed
        # out again)
        if (used * 4 > cap * 3 and (used - int(sz.evictable) - int(sz.protected_)) * 4 > cap
                and used - self._arena_floor > cap // 8):
            self._ix.compact()
            self._invalidate()
            sz = self._ix.sizes()
            self._arena_floor = int(sz.arena_used)
        self._arena_est = int(sz.arena_used) + int(n_tokens)

    def _run_values(self, rows) -> np.ndarray:
        """KV indices of one node's rows: only the last row of a run can be a short page."""
        vals = self._rows_values(rows)
        if len(rows) and self._ragged:
            held = self._ragged.get(int(rows[-1]))
            if held is not None:
                vals = vals[:len(vals) - (self._P - held)]
        return vals

    def _intern(self, tokens, offsets, B: int, create: bool):
        """Device tokens / offsets -> (symbols, symbol offsets [B+1], pages resolved per key), all on the device."""
        xp = self._xp(offsets)
        lens = offsets[1:] - offsets[:-1]
        npg = lens // self._P if xp is np else xp.div(lens, self._P, rounding_mode="floor")
        soffs = xp.zeros(B + 1, dtype=xp.int64) if xp is np else xp.zeros(B + 1, dtype=xp.int64, device=offsets.device)
        soffs[1:] = xp.cumsum(npg, 0)
        total = int(soffs[-1])
        syms = self._be.empty(total + 1, np.int32)
        npages = self._be.empty(B, np.int32)
        if total:
            self._pg.intern(tokens, offsets, B, create, soffs, total, syms, npages, self._be.empty(total + 1, np.int32))
        return syms, soffs, npages, total

    def _rows_of_prefixes(self, rows, soffs, prefix, B: int, total: int):
        """Rows of the first prefix[r] pages of every record r (the pages an insert found already present)."""
        xp = self._xp(rows)
        if xp is np:
            rec = np.repeat(np.arange(B), np.diff(soffs))
            pos = np.arange(total) - soffs[:-1][rec]
            return rows[:total][pos < prefix[rec]]
        rec = xp.repeat_interleave(xp.arange(B, device=rows.device), soffs[1:] - soffs[:-1])
        pos = xp.arange(total, device=rows.device) - soffs[:-1][rec]
        return rows[:total][pos < prefix.to(xp.int64)[rec]].contiguous()

    def _paged_insert(self, tokens, values, offsets, B: int):
        """Ordered batch of paged inserts on device arrays; returns matched prefix lengths in PAGES (device int32)."""
        syms, soffs, _, total = self._intern(tokens, offsets, B, create=True)
        out = self._be.empty(B, np.int32)
        if total == 0:
            return out, None
        rows = self._be.empty(total + 1, np.int64)
        self._pg.rows_put(values, total, rows)        # values are page aligned: row i = values[i*P : (i+1)*P]
        self._arena_room(total)
        self._ix.insert(syms, rows, soffs, None, B, out)
        dead = self._rows_of_prefixes(rows, soffs, out, B, total)
        if len(dead):
            self._pg.rows_free(dead, len(dead))
        self._invalidate()
        self._ix.check_device_error()
        in_use = self._pg.status()[0]
        # pages of evicted nodes keep their symbols until somebody asks for them back: when three quarters of the
        # symbol space are taken - and an eighth of it has been handed out since the last time - the pages no edge holds
        # are dropped (the reference's dict keys die with their nodes, radix_cache.py:347-352)
        if in_use > 0.75 * self._pg_syms_total and in_use - self._pg_syms_after_retain > self._pg_syms_total // 8:
            self._pg.retain(self._ix)
            self._pg_syms_after_retain = self._pg.status()[0]
        return out, rows

    # ------------------------------------------------------------------ public API (radix_cache.py:117-248)
    def reset(self) -> None:
        self._ix.reset()
        self._invalidate()
        if self._pg is not None:
            self._pg.reset()
            self._pg_syms_after_retain = 0
        self._ragged, self._deficit, self._deficit_protected = {}, 0, 0
        self._arena_est, self._arena_floor = 0, 0

    def match_prefix(self, key: List[int], **kwargs) -> MatchResult:
        import torch
        if self.disable or len(key) == 0:
            return MatchResult(torch.empty((0,), dtype=torch.int64), self.root_node, self.root_node)
        if self._P > 1:
            # page aligned (radix_cache.py:148-150); the pages become symbols on the device, an unseen page ends the key
            be = self._be
            n = len(key) // self._P * self._P
            if n == 0:
                return MatchResult(torch.empty((0,), dtype=torch.int64), self.root_node, self.root_node)
            if self._ragged and PAGE_FILL in key[:n]:
                raise ValueError(f"paged match: token {PAGE_FILL} is reserved (it fills up a short last page)")
            lens, nodes = self.match_prefix_batch(be.to_device(np.asarray(key[:n], dtype=np.int32)),
                                                  be.to_device(np.asarray([0, n], dtype=np.int64)))
            m = int(be.to_host(lens)[0])
            vals = be.empty(m + 1, np.int64)
            self.gather_values_batch(nodes, lens, be.to_device(np.asarray([0, m], dtype=np.int64)), vals)
            node = self._handle(int(be.to_host(nodes)[0]))
            return MatchResult(self._tensor(be.to_host(vals)[:m].copy()), node, node)
        vals, node = self._match_one(key)
        return MatchResult(self._tensor(vals), node, node)

    def _match_one(self, key):
        """One key end to end through rm_match_prefix_h (one synchronisation): (values of the matched prefix, last node)."""
        self._flush()
        ln, nd, vals, info = self._ix.match_prefix_host(np.asarray(key, dtype=np.int32))
        self._invalidate()
        if nd == 0:
            return vals, self.root_node
        self._info_cache[nd] = info
        return vals, TreeNode(self, nd, int(info.gen))

    def _match_keys(self, keys):
        """Host lists -> (lens, nodes, values, value offsets) as host arrays."""
        self._flush()
        flat = np.fromiter((t for k in keys for t in k), dtype=np.int32, count=sum(len(k) for k in keys))
        offs = np.zeros(len(keys) + 1, dtype=np.int64)
        np.cumsum([len(k) for k in keys], out=offs[1:])
        lens = np.zeros(len(keys), np.int32)
        nodes = np.zeros(len(keys), np.int32)
        self._ix.match_host(flat, offs, lens, nodes)
        self._invalidate()
        voffs = np.zeros(len(keys) + 1, dtype=np.int64)
        np.cumsum(lens, out=voffs[1:])
        be = self._be
        d_out = be.empty(int(voffs[-1]) + 1, np.int64)
        self._ix.gather(be.to_device(nodes), be.to_device(lens), be.to_device(voffs), len(keys), d_out)
        return lens, nodes, d_out[:int(voffs[-1])], voffs

    def match_prefix_batch(self, tokens, offsets, out_len=None, out_node=None):
        """Addition: device tensors in, device tensors out (int32 matched lengths in tokens, int32 last-node ids), equal
        to match_prefix on each query in order.  With page_size > 1 every query is matched in whole pages."""
        self._flush()
        B = int(offsets.shape[0]) - 1
        out_len = out_len if out_len is not None else self._be.empty(B, np.int32)
        out_node = out_node if out_node is not None else self._be.empty(B, np.int32)
        if self._P > 1:
            syms, soffs, _, total = self._intern(tokens, offsets, B, create=False)
            self._ix.match(syms, soffs, B, out_len, out_node)
            out_len *= self._P
        else:
            self._ix.match(tokens, offsets, B, out_len, out_node)
        self._invalidate()
        return out_len, out_node

    def gather_values_batch(self, nodes, lens, out_offsets, out_values):
        """Addition: the torch.cat of match_prefix for a whole batch (ragged, caller-provided offsets; lens and
        out_offsets in tokens, as match_prefix_batch returns them)."""
        B = int(nodes.shape[0])
        if self._P > 1:
            P = self._P
            xp = self._xp(lens)
            plens = lens // P if xp is np else xp.div(lens, P, rounding_mode="floor")
            poffs = out_offsets // P if xp is np else xp.div(out_offsets, P, rounding_mode="floor")
            n_rows = int(out_values.shape[0]) // P
            rows = self._be.empty(n_rows + 1, np.int64)
            rows[:] = -1
            self._ix.gather(nodes, plens, poffs, B, rows)
            self._pg.rows_get(rows, n_rows, out_values)
            return out_values
        self._ix.gather(nodes, lens, out_offsets, B, out_values)
        return out_values

    def insert(self, key: List, value=None) -> int:
        if self.disable:
            return 0
        if value is None:
            value = [x for x in key]
        if self._P > 1:
            P = self._P
            if len(key) == 0:
                return 0
            be = self._be
            k = np.asarray(key, dtype=np.int32)
            v = (value.cpu().numpy() if hasattr(value, "cpu") else np.asarray(value)).astype(np.int64, copy=False)
            held = len(k) % P                     # tokens of a short last page (0: whole pages)
            if held:
                if (k == PAGE_FILL).any():
                    raise ValueError(f"paged insert: token {PAGE_FILL} is reserved (it fills up a short last page)")
                k = np.concatenate([k, np.full(P - held, PAGE_FILL, dtype=np.int32)])
                v = np.concatenate([v, np.zeros(P - held, dtype=np.int64)])
            out, rows = self._paged_insert(be.to_device(k), be.to_device(v), be.to_device(np.asarray([0, len(k)], dtype=np.int64)), 1)
            matched = int(be.to_host(out)[0])
            if held and matched < len(k) // P:    # the short page went into a new leaf: its row stays
                row = int(be.to_host(rows[len(k) // P - 1:len(k) // P])[0])
                self._ragged[row] = held
                self._deficit += P - held
            # (a matched short page counts as a whole one, exactly like _key_match_paged, radix_cache.py:23-32)
            return matched * P
        return int(self._insert_records([(list(key), value, -1)])[0])

    def _insert_records(self, recs) -> np.ndarray:
        flat = np.fromiter((t for k, _, _ in recs for t in k), dtype=np.int32, count=sum(len(k) for k, _, _ in recs))
        offs = np.zeros(len(recs) + 1, dtype=np.int64)
        np.cumsum([len(k) for k, _, _ in recs], out=offs[1:])
        vals = None
        if self.MODE != capi.MODE_ROUTER:
            vals = np.zeros(int(offs[-1]), dtype=np.int64)
            for i, (_, v, _) in enumerate(recs):
                a = v.cpu().numpy() if hasattr(v, "cpu") else np.asarray(v)
                vals[offs[i]:offs[i + 1]] = a
        ranks = np.asarray([r for _, _, r in recs], dtype=np.int32) if self.MODE != capi.MODE_BASE else None
        self._arena_room(int(offs[-1]))
        out = self._ix.insert_host(flat, vals, offs, ranks)
        self._invalidate()
        self._ix.check_device_error()
        return out

    def insert_batch(self, tokens, values, offsets, ranks=None, out_prefix_len=None):
        """Addition: ordered batch of inserts on device tensors (rm_insert_batch); returns the matched prefix lengths in
        tokens.  With page_size > 1 every key (and its values) must be a whole number of pages."""
        self._flush()
        N = int(offsets.shape[0]) - 1
        if self._P > 1:
            out, _ = self._paged_insert(tokens, values, offsets, N)
            out *= self._P
            if out_prefix_len is not None:
                out_prefix_len[:] = out
                return out_prefix_len
            return out
        out = out_prefix_len if out_prefix_len is not None else self._be.empty(N, np.int32)
        self._arena_room(int(tokens.shape[0]))
        self._ix.insert(tokens, values, offsets, ranks, N, out)
        self._invalidate()
        return out

    def evict(self, num_tokens: int) -> None:
        if self.disable:
            return
        self._flush()
        P = self._P
        if P > 1:
            # sum(len * P) >= n  <=>  sum(len) >= ceil(n / P) while every page is whole.  A short last page frees
            # fewer tokens than its row counts for, so the call may stop early: it then goes on with what is missing
            # (the reference's heap loop continued is the same as collecting the leaves again, radix_cache.py:183-202)
            remaining = int(num_tokens)
            while remaining > 0:
                nodes, offs, lens = self._ix.evict((remaining + P - 1) // P)
                self._invalidate()
                if len(nodes) == 0:
                    break
                rows = self._ix.read_runs(offs, lens)           # the value rows of all evicted nodes in one copy
                ends = np.cumsum(lens)
                runs = [rows[b - n:b] for b, n in zip(ends.tolist(), lens.tolist())]
                vals = self._rows_values(rows)
                self._pg.rows_free(self._be.to_device(rows), len(rows))
                at, short = 0, False
                for r in runs:
                    held = self._ragged.pop(int(r[-1]), None) if self._ragged else None
                    n_tok = len(r) * P - (P - held if held is not None else 0)
                    if held is not None:
                        self._deficit -= P - held
                        short = True
                    if self.token_to_kv_pool_allocator is not None:
                        self.token_to_kv_pool_allocator.free(self._tensor(vals[at * P:at * P + n_tok].copy()))
                    at += len(r)
                    remaining -= n_tok
                if not short:
                    break
            return
        nodes, offs, lens = self._ix.evict(int(num_tokens))
        self._invalidate()
        if self.token_to_kv_pool_allocator is not None and len(nodes):
            # the KV indices of all evicted nodes in one gather and one copy (a read per node is a synchronisation per
            # node: 4 K nodes, 60 ms), then handed back node by node in eviction order, as radix_cache.py:195 does
            vals = self._ix.read_runs(offs, lens)
            at = 0
            for n in lens.tolist():
                self.token_to_kv_pool_allocator.free(self._tensor(vals[at:at + n].copy()))
                at += n

    def _lock(self, node: TreeNode, direction: int) -> int:
        self._flush()
        be = self._be
        d_out = be.empty(1, np.int64)
        short = 0
        if self._ragged and node.id != 0:
            # the handle's own node may end in a short page: its length counts for page_size - held tokens less when
            # its lock count crosses 0 <-> 1 (a short page is always the end of a leaf)
            r = self._info(node._resolved())
            crossing = (direction > 0 and int(r.lock_ref) == 0) or (direction < 0 and int(r.lock_ref) == 1)
            if crossing and int(r.edge_len) > 0:
                _, last = self._ix.read_arena(int(r.edge_off) + int(r.edge_len) - 1, 1)
                held = self._ragged.get(int(last[0]))
                if held is not None:
                    short = self._P - held
        self._ix.lock(be.to_device(np.asarray([node.id], np.int32)), be.to_device(np.asarray([node._gen], np.int32)),
                      1, direction, d_out)
        self._invalidate()
        self._deficit_protected += short if direction > 0 else -short
        # inc returns minus the tokens that became protected, dec the tokens that became evictable again
        return int(be.to_host(d_out)[0]) * self._P + short * (1 if direction > 0 else -1)

    def inc_lock_ref(self, node: TreeNode) -> int:
        return 0 if self.disable else self._lock(node, +1)

    def dec_lock_ref(self, node: TreeNode) -> int:
        return 0 if self.disable else self._lock(node, -1)

    def evictable_size(self) -> int:
        self._flush()
        return int(self._ix.sizes().evictable) * self._P - (self._deficit - self._deficit_protected)

    def protected_size(self) -> int:
        self._flush()
        return int(self._ix.sizes().protected_) * self._P - self._deficit_protected

    def total_size(self) -> int:
        """radix_cache.py:239-241, 354-364: reduced on the device (rm_total_size), eight bytes back."""
        self._flush()
        return self._ix.total_size() * self._P - self._deficit

    def all_values_flatten(self):
        """radix_cache.py:243-248: the values of the tree in the reference's order - pre-order, the children of a node
        in dict insertion order.  The device keeps that order as a key per node (the op that added the child to its
        parent; a split's upper node inherits the place of the node it splits, radix_cache.py:292)."""
        import torch
        d = self._dump()
        n = d["n"]
        order = self._ix.dump_order(n)
        live = np.nonzero((d["parent"][:n] >= 0) & (d["gen"][:n] >= 0))[0]
        live = live[live != 0]
        if len(live) == 0:
            return torch.empty((0,), dtype=torch.int64)
        # children lists in dict order: sort by (parent, order key), then cut at parent changes
        srt = live[np.lexsort((live, order[live], d["parent"][live]))]
        par = d["parent"][srt]
        starts = np.nonzero(np.r_[True, par[1:] != par[:-1]])[0]
        kids = {int(par[a]): srt[a:b] for a, b in zip(starts, np.r_[starts[1:], len(srt)])}
        used = int(self._ix.sizes().arena_used)
        _, val = self._ix.read_arena(0, used)
        out, stack = [], [iter(kids.get(0, ()))]
        while stack:
            x = next(stack[-1], None)
            if x is None:
                stack.pop()
                continue
            x = int(x)
            o, ln = int(d["edge_off"][x]), int(d["edge_len"][x])
            out.append(self._run_values(val[o:o + ln]) if self._P > 1 else val[o:o + ln])
            stack.append(iter(kids.get(x, ())))
        return torch.from_numpy(np.concatenate(out))

    def pretty_print(self) -> None:
        d = self._dump()
        kids = {}
        for x in range(1, d["n"]):
            if d["parent"][x] >= 0 and d["gen"][x] >= 0:
                kids.setdefault(int(d["parent"][x]), []).append(x)
        stack = [(0, 0)]
        while stack:
            x, ind = stack.pop()
            print(" " * ind, int(d["edge_len"][x]), self._node_tokens(x)[:10], f"r={int(d['lock_ref'][x])}")
            for c in kids.get(x, []):
                stack.append((c, ind + 2))
        print(f"#tokens: {self.total_size()}")

    def take_events(self):
        return []

    # parity helper shared with the oracle's canonical()
    def full_key(self, node: TreeNode) -> List[int]:
        x = node._resolved()
        parts = []
        while x != 0:
            parts.append(self._node_tokens(x))
            x = int(self._info(x).parent)
        return [t for p in reversed(parts) for t in p]
