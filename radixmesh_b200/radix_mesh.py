"""Drop-in for the reference's `RadixMesh` (radix/radix_mesh.py:72-495) on the device index.

What stays host Python, exactly as in the reference: who forwards what to whom (ring + routers fed by
rank 0), the oplog objects, the tick counting.  What moves to the GPU: the tree.  Remote INSERT oplogs
are staged on the host in arrival order and applied as ONE ordered `rm_insert_batch_h` before the next
operation that reads the tree, which is equivalent to the reference applying them one by one under its
`_tree_lock` (radix_mesh.py:198-199).
"""
from __future__ import annotations

import dataclasses
import threading
from typing import Dict, List

import numpy as np

from . import capi
from .host import CacheOplog, CacheOplogType, GCQuery, NodeKey, RadixMode, RingPolicy
from .oplog_codec import IngestBuffer, deserialize_any, with_header
from .radix_cache import MatchResult, RadixCache, TreeNode


@dataclasses.dataclass
class RouterMatchResult:              # radix_mesh.py:66-69
    prefill_node_rank: int = -1
    decode_node_rank: int = -1


class OwnedValue:
    """What `TreeNode.value` is on a prefill/decode replica (radix_mesh.py:21-44): `.value` tensor + `.node_rank`."""

    def __init__(self, value, node_rank: int):
        self.value = value
        self.node_rank = node_rank

    def __len__(self):
        return len(self.value)


class RankValue:
    """Router replicas keep only the owner rank (radix_mesh.py:47-63)."""

    def __init__(self, node_rank: int):
        self.node_rank = node_rank

    def __len__(self):
        return 1


_MODES = {RadixMode.PREFILL: capi.MODE_PREFILL, RadixMode.DECODE: capi.MODE_DECODE, RadixMode.ROUTER: capi.MODE_ROUTER}


class RadixMesh(RadixCache):
    def __init__(self, args, communicator=None, routers=None, token_to_kv_pool_allocator=None, **index_kwargs):
        if args.prefill_node_rank != -1:                       # radix_mesh.py:76-84
            self._global_node_rank, self.mode = args.prefill_node_rank, RadixMode.PREFILL
        elif args.decode_node_rank != -1:
            self._global_node_rank, self.mode = args.decode_node_rank, RadixMode.DECODE
        else:
            self._global_node_rank, self.mode = args.router_node_rank, RadixMode.ROUTER
        self.MODE = _MODES[self.mode]
        self.args = args
        self.sync_algo = RingPolicy()
        self.communicator = communicator
        self.router_communicators = routers
        self.logic_op_counter = 0
        # One lock for staging, taking and applying records: the reference holds _tree_lock across the whole
        # _insert_helper (radix_mesh.py:198-199).  Re-entrant, because _flush is reached from code that holds it.
        self._pending_lock = threading.RLock()
        self._ingest = IngestBuffer(with_values=self.mode != RadixMode.ROUTER)
        self.tick_received: Dict[int, int] = {}
        backend = index_kwargs.pop("_backend", None)
        library = index_kwargs.pop("_library", None)
        ikw = dict(self_rank=self._global_node_rank, n_prefill=len(args.prefill_cache_nodes),
                   n_decode=len(args.decode_cache_nodes))
        super().__init__(None, token_to_kv_pool_allocator, 1, _index_kwargs=ikw, _backend=backend, _library=library,
                         **index_kwargs)
        if self.communicator is not None:
            self.communicator.register_rcv_callback(self.oplog_received)

    # ------------------------------------------------------------------ reference surface
    def global_node_rank(self) -> int:
        return self._global_node_rank

    def prefill_cache_nodes(self, prefill_node_rank=None):
        if prefill_node_rank is None:
            return self.args.prefill_cache_nodes
        return self.args.prefill_cache_nodes[prefill_node_rank]

    def decode_cache_nodes(self, decode_node_rank=None):
        if decode_node_rank is None:
            return self.args.decode_cache_nodes
        return self.args.decode_cache_nodes[self.args.local_node_rank(decode_node_rank)]

    def match_prefix(self, key: List[int], **kwargs):
        if len(key) == 0:
            raise IndexError("list index out of range")          # radix_mesh.py:253 has no empty-key guard
        if self.mode == RadixMode.ROUTER:
            pf, dc, _, _, err = self._route_keys([key], warm=True)
            if err[0] == 1:
                raise AssertionError("[RadixMesh] Prefix match node_rank is invalid")   # radix_mesh.py:233-237
            return RouterMatchResult(int(pf[0]), int(dc[0]))
        vals, node = self._match_one(key)
        return MatchResult(self._tensor(vals), node, node)

    def insert(self, key: List, value=None) -> int:
        assert self.mode in (RadixMode.PREFILL, RadixMode.DECODE), "only support prefill/decode node to insert"
        return self._insert(list(key), value, self._global_node_rank)

    def _insert(self, key, value, owner_rank: int) -> int:
        # staged and applied under ONE hold of the lock: nobody can add or take records in between, so the
        # caller's record is the last one of the batch and batches reach the device in arrival order
        with self._pending_lock:
            self._ingest.add(key, value, owner_rank)
            ret = self._flush()
        self.send_insert_event(key, value, owner_rank)
        return ret

    def _flush(self):
        """Apply the staged INSERTs, in arrival order, as ordered device batches (rm_insert_batch_h).  Returns the
        prefix length of the LAST staged record (what _insert needs: it holds the lock across add + flush)."""
        with self._pending_lock:
            if len(self._ingest) == 0:
                return 0
            tok, val, offs, ranks = self._ingest.take()
            out = None
            n, step = len(ranks), int(self._ix.cfg.max_batch)
            for i in range(0, n, step):
                j = min(i + step, n)
                a, b = int(offs[i]), int(offs[j])
                self._arena_room(b - a)
                out = self._ix.insert_host(tok[a:b], val[a:b] if val is not None else None, offs[i:j + 1] - offs[i], ranks[i:j])
            self._invalidate()
            self._ix.check_device_error()
            return int(out[-1])

    def reset(self) -> None:
        """radix_mesh.py:240-245.  INSERTs that arrived before the reset and are still staged must not outlive it:
        the reference has applied them already when it resets, so they are gone with the tree."""
        with self._pending_lock:
            self._ingest.take()
            RadixCache.reset(self)

    def send_insert_event(self, key, value, owner_rank: int) -> None:
        # radix_mesh.py:325-347: routers never send; rank 0 also feeds the routers
        if self.mode == RadixMode.ROUTER or self.communicator is None:
            return
        vals = value.tolist() if hasattr(value, "tolist") else list(value)
        self.logic_op_counter += 1
        oplog = CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=owner_rank, local_logic_id=self.logic_op_counter,
                           key=list(key), value=vals, ttl=self.sync_algo.ttl(self.args))
        self.communicator.send(oplog)
        if self._global_node_rank == self.sync_algo.master_node_rank() and self.router_communicators:
            for r in self.router_communicators:
                r.send(oplog)

    def oplog_received(self, oplog) -> None:
        # radix_mesh.py:391-423
        oplog.ttl -= 1
        t = int(oplog.oplog_type)
        if t == CacheOplogType.TICK:                              # _tick_handle, radix_mesh.py:356-360
            self.tick_received[oplog.node_rank] = self.tick_received.get(oplog.node_rank, 0) + 1
            if oplog.ttl <= 0:
                return None
            self._send(oplog)
            return None
        if t in (CacheOplogType.GC_QUERY, CacheOplogType.GC_EXEC):
            return self._gc_handle(oplog)
        if oplog.node_rank == self._global_node_rank or oplog.ttl <= 0:
            return None
        if t == CacheOplogType.INSERT:
            with self._pending_lock:
                self._ingest.add(oplog.key, oplog.value, oplog.node_rank)
            self.send_insert_event(oplog.key, oplog.value, oplog.node_rank)   # re-forward with a fresh ttl
        elif t == CacheOplogType.RESET:
            self.reset()                                          # also drops INSERTs staged before the reset
        elif t != CacheOplogType.DELETE:
            assert False, "unknown oplog type"
        return None

    def oplog_received_wire(self, buf, size=None) -> None:
        """Addition: an inbound message as it came off the socket.  An INSERT - packed
        (oplog_codec.PackedSerializer) or the reference's JSON - is gated on its header fields and its payload
        copied straight into the staging arrays of the next ordered insert batch; anything else (TICK, RESET, GC,
        unusual JSON) is decoded and goes through oplog_received.  Re-forwarding re-uses the same bytes with a
        fresh ttl and logic id."""
        with self._pending_lock:
            what, oplog, _ttl = self._ingest.feed_wire(buf, size, self._global_node_rank)
        if what == "other":
            return self.oplog_received(oplog)
        if what == "insert" and self.mode != RadixMode.ROUTER and self.communicator is not None:
            raw = bytes(memoryview(buf)[:size] if size is not None else buf)
            self.logic_op_counter += 1
            fwd = with_header(raw, self.sync_algo.ttl(self.args), self.logic_op_counter)
            send_raw = getattr(self.communicator, "send_raw", None)
            if send_raw is not None:
                send_raw(fwd)
            else:
                self.communicator.send(deserialize_any(fwd))
            if self._global_node_rank == self.sync_algo.master_node_rank() and self.router_communicators:
                for r in self.router_communicators:
                    (getattr(r, "send_raw", None) or (lambda b, r=r: r.send(deserialize_any(b))))(fwd)
        return None

    def dec_lock_ref(self, node: TreeNode) -> int:
        return self._lock(node, -1)           # the device clamps at zero in mesh modes (radix_mesh.py:175-177)

    def inc_lock_ref(self, node: TreeNode) -> int:
        return self._lock(node, +1)

    def evict(self, node=None) -> None:       # radix_mesh.py:349-351
        return None

    def _node_value(self, x: int):
        rank = int(self._info(x).rank)
        if self.mode == RadixMode.ROUTER:
            return RankValue(rank)
        return OwnedValue(super()._node_value(x) if x else None, rank)

    @property
    def dup_nodes(self) -> dict:
        """{NodeKey(full key, rank): parked TreeNode or None} (radix_mesh.py:97, 310, 476, 493)."""
        self._flush()
        out = {}
        for node, rank, parked, _lk in self._ix.dup_entries().tolist():
            k = NodeKey(tuple(self.full_key(TreeNode(self, node, -1))), rank)
            out[k] = None if parked < 0 else TreeNode(self, parked, -1)
        return out

    def gc_candidates(self) -> List[NodeKey]:
        """The scan of RadixMesh._gc (radix_mesh.py:151-156)."""
        self._flush()
        return [NodeKey(tuple(self.full_key(TreeNode(self, n, -1))), r)
                for n, r, _, _ in self._ix.dup_entries(only_candidates=True).tolist()]

    # ------------------------------------------------------------------ duplicate-node GC (radix_mesh.py:148-166, 362-389)
    def _send(self, oplog) -> None:
        # radix_mesh.py:339-347
        if self.mode == RadixMode.ROUTER or self.communicator is None:
            return
        self.communicator.send(oplog)
        if self._global_node_rank == self.sync_algo.master_node_rank() and self.router_communicators:
            for r in self.router_communicators:
                r.send(oplog)

    def gc_scan_send(self) -> int:
        """One iteration of the reference's `_gc` loop (radix_mesh.py:148-166; the thread and its sleep stay with the
        caller): ask the ring about every parked, unlocked duplicate.  Returns the number of keys asked about."""
        assert self.mode != RadixMode.ROUTER, "only mode ref node need gc"
        query = [GCQuery(key=k) for k in self.gc_candidates()]
        if query:
            self._send(CacheOplog(oplog_type=CacheOplogType.GC_QUERY, node_rank=self._global_node_rank,
                                  ttl=self.sync_algo.ttl(self.args), gc_query=query))
        return len(query)

    def _gc_handle(self, oplog) -> None:
        """radix_mesh.py:362-389 on the device duplicate table (rm_gc_agree_h / rm_gc_exec_h).  Two departures from
        the letter of the reference, both its evident intent (its own version cannot run: CacheOplog.to_dict drops the
        payload): the node that asked also pops its own entries when it frees them (the reference would
        ask about them again on the next scan and free the same KV indices twice), and GC_EXEC is forwarded until its
        ttl runs out (the reference stops it at the first hop, so only one peer would ever pop)."""
        self._flush()
        if int(oplog.oplog_type) == CacheOplogType.GC_EXEC:
            keys = list(oplog.gc_exec)
            if keys:
                self._ix.gc_exec([list(k.key) for k in keys], [k.node_rank for k in keys], free_parked=False)
                self._invalidate()
            if oplog.node_rank != self._global_node_rank and oplog.ttl > 0:
                self._send(oplog)
            return None
        if oplog.node_rank == self._global_node_rank:
            full = self.sync_algo.ttl(self.args)
            keys = [q.key for q in oplog.gc_query if q.agree == full]
            if keys:
                off, ln = self._ix.gc_exec([list(k.key) for k in keys], [k.node_rank for k in keys], free_parked=True)
                if self.token_to_kv_pool_allocator is not None:
                    keep = (off >= 0) & (ln > 0)
                    vals = self._ix.read_runs(off[keep], ln[keep])      # the released runs in one copy
                    at = 0
                    for n in ln[keep].tolist():
                        self.token_to_kv_pool_allocator.free(self._tensor(vals[at:at + n].copy()))
                        at += n
                self._invalidate()
                self._send(CacheOplog(oplog_type=CacheOplogType.GC_EXEC, node_rank=self._global_node_rank,
                                      ttl=self.sync_algo.ttl(self.args), gc_exec=keys))
            return None
        if oplog.gc_query:
            agree = self._ix.gc_agree([list(q.key.key) for q in oplog.gc_query], [q.key.node_rank for q in oplog.gc_query])
            for q, a in zip(oplog.gc_query, agree.tolist()):
                if a:
                    q.agree += 1
        self._send(oplog)
        return None

    # ------------------------------------------------------------------ router side (used by CacheAwareRouter)
    def _route_keys(self, keys, warm: bool):
        self._flush()
        flat = np.fromiter((t for k in keys for t in k), dtype=np.int32, count=sum(len(k) for k in keys))
        offs = np.zeros(len(keys) + 1, dtype=np.int64)
        np.cumsum([len(k) for k in keys], out=offs[1:])
        be = self._be
        n = len(keys)
        outs = [be.empty(n, np.int32) for _ in range(4)]
        err = be.empty(n, np.uint8)
        self._ix.route(be.to_device(flat), be.to_device(offs), n, warm, outs[0], outs[1], outs[2], outs[3], err)
        self._invalidate()
        return [be.to_host(o) for o in outs] + [be.to_host(err)]
