"""Harness around the UNMODIFIED reference (TEST INFRASTRUCTURE, build container only).

Imports the reference from /root/reference/python and adds only the three shims
SURVEY.md 7.2 lists: a logical clock in place of ``time.monotonic``, a fake KV
allocator, and an in-process fake ``Communicator`` so a 3P+2D+1R mesh runs without
sockets.  Used to (1) validate ``oracle/radix_port.py`` differentially and
(2) generate the golden vectors under ``tests/golden/`` (``tools/make_golden.py``).
Nothing here runs on the GPU box: ``available()`` is False there.
"""
from __future__ import annotations

import os
import sys
import types
from typing import List, Optional

REF_ROOT = "/root/reference/python"
SUB_BITS = 20


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "src", "radix"))


class _ShimTime(types.SimpleNamespace):
    """Drop-in for the ``time`` module inside the reference's two tree modules."""

    def __init__(self) -> None:
        super().__init__()
        self.op_seq = 0
        self._sub = 0

    def begin_op(self) -> int:
        self.op_seq += 1
        self._sub = 0
        return self.op_seq

    def monotonic(self) -> int:
        self._sub += 1
        return (self.op_seq << SUB_BITS) + self._sub

    def sleep(self, _s) -> None:
        return None


_state = {}


def load():
    """Import the reference once, install the clock shim, return a namespace."""
    if _state:
        return _state["ns"]
    assert available(), "reference tree not present"
    sys.dont_write_bytecode = True
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import torch  # noqa: F401
    from src.radix.sglang.srt.mem_cache import radix_cache as rc
    from src.radix import radix_mesh as rm
    from src.radix.cache_oplog import CacheOplog, CacheOplogType, ImmutableNodeKey
    from src.radix.core_enum import RadixMode
    from src.config.cache_config import ServerArgs
    from src.router import cache_aware_router as car
    from src.communication.serializer import JsonSerializer
    from src.communication.communicator import Communicator

    clock = _ShimTime()
    rc.time = clock
    rm.time = clock
    ns = types.SimpleNamespace(rc=rc, rm=rm, car=car, clock=clock, CacheOplog=CacheOplog,
                               CacheOplogType=CacheOplogType, ImmutableNodeKey=ImmutableNodeKey,
                               RadixMode=RadixMode, ServerArgs=ServerArgs,
                               JsonSerializer=JsonSerializer, Communicator=Communicator, torch=torch)
    _state["ns"] = ns
    return ns


class FakeAllocator:
    """`.device` and `.free(tensor)` are all RadixCache needs (radix_cache.py:104-107, 195)."""

    def __init__(self) -> None:
        import torch
        self.device = torch.device("cpu")
        self.freed: List[list] = []

    def free(self, t) -> None:
        self.freed.append([int(x) for x in (t.tolist() if hasattr(t, "tolist") else t)])


class RefCache:
    """Reference RadixCache driven with the logical clock; ops mirror PortRadixCache."""

    def __init__(self, page_size: int = 1) -> None:
        ns = load()
        self.ns = ns
        self.alloc = FakeAllocator()
        ns.clock.op_seq = 0          # one logical clock per tree under test
        ns.clock.begin_op()
        self.tree = ns.rc.RadixCache(None, self.alloc, page_size)

    def match_prefix(self, key):
        if len(key):
            self.ns.clock.begin_op()
        return self.tree.match_prefix(list(key))

    def insert(self, key, value):
        self.ns.clock.begin_op()
        return self.tree.insert(list(key), self.ns.torch.tensor(list(value), dtype=self.ns.torch.int64))

    def evict(self, n):
        self.ns.clock.begin_op()
        self.tree.evict(n)

    def full_key(self, node):
        key = node.key
        while node is not self.tree.root_node:
            node = node.parent
            key = node.key + key
        return key

    def canonical(self) -> dict:
        rows = []
        stack = [self.tree.root_node]
        while stack:
            n = stack.pop()
            if n is not self.tree.root_node:
                rows.append((tuple(self.full_key(n)), len(n.key), tuple(int(x) for x in n.value.tolist()),
                             -1, n.lock_ref, n.last_access_time >> SUB_BITS))
            stack.extend(n.children.values())
        rows.sort()
        return {"nodes": rows, "evictable": self.tree.evictable_size_, "protected": self.tree.protected_size_}


class _LoopComm:
    """In-process stand-in for TcpCommunicator: JSON round trip through the reference's
    own serializer, then a synchronous call of the next hop's receive callback."""

    def __init__(self, ns, name: str) -> None:
        self.ns = ns
        self.name = name
        self.peer_cb = None
        self.cb = None
        self.ser = ns.JsonSerializer()

    def send(self, oplog) -> int:
        raw = self.ser.serialize(oplog)
        if self.peer_cb is not None:
            self.peer_cb(self.ser.deserialize(bytearray(raw), len(raw)))
        return len(raw)

    def register_rcv_callback(self, fn) -> None:
        self.cb = fn
        # satisfy _wait_all_nodes_ready (radix_mesh.py:435-443): two ticks from one origin
        for _ in range(2):
            fn(self.ns.CacheOplog(oplog_type=self.ns.CacheOplogType.TICK, node_rank=10_000, ttl=1))

    def is_ordered(self) -> bool:
        return True

    def target_address(self) -> str:
        return self.name


class RefMesh:
    """P prefill + D decode replicas in a ring, R routers fed by rank 0, all in-process."""

    def __init__(self, n_prefill: int = 3, n_decode: int = 2, n_router: int = 1) -> None:
        ns = load()
        self.ns = ns
        self.P, self.D, self.R = n_prefill, n_decode, n_router
        p_addrs = [f"localhost:{50000 + i}" for i in range(n_prefill)]
        d_addrs = [f"localhost:{50000 + n_prefill + i}" for i in range(n_decode)]
        r_addrs = [f"localhost:{50010 + i}" for i in range(n_router)]
        self.p_addrs, self.d_addrs, self.r_addrs = p_addrs, d_addrs, r_addrs
        ring = n_prefill + n_decode
        self.comms = [_LoopComm(ns, f"ring{i}") for i in range(ring)]
        self.router_in = [_LoopComm(ns, f"router{i}") for i in range(n_router)]
        self.to_router = [_LoopComm(ns, f"to_router{i}") for i in range(n_router)]
        self.nodes = []
        import threading
        real_thread = threading.Thread

        class _NoThread:  # the ticker / GC threads are out of scope; do not start them
            def __init__(self, *a, **k):
                pass

            def start(self):
                pass

        ns.rm.threading.Thread = _NoThread
        try:
            for rank in range(ring + n_router):
                args = ns.ServerArgs(prefill_cache_nodes=p_addrs, decode_cache_nodes=d_addrs,
                                     router_cache_nodes=r_addrs, protocol="test")
                if rank < n_prefill:
                    args.prefill_node_rank = rank
                    args.local_cache_addr = p_addrs[rank]
                elif rank < ring:
                    args.decode_node_rank = rank
                    args.local_cache_addr = d_addrs[rank - n_prefill]
                else:
                    args.router_node_rank = rank
                    args.local_cache_addr = r_addrs[rank - ring]
                ns.clock.begin_op()
                if rank < ring:
                    routers = self.to_router if rank == 0 else None
                    node = ns.rm.RadixMesh(args, communicator=self.comms[rank], routers=routers)
                else:
                    node = ns.rm.RadixMesh(args, communicator=self.router_in[rank - ring])
                self.nodes.append(node)
        finally:
            ns.rm.threading.Thread = real_thread
        for i in range(ring):
            self.comms[i].peer_cb = self.comms[(i + 1) % ring].cb
        for i in range(n_router):
            self.to_router[i].peer_cb = self.router_in[i].cb

    def insert(self, rank: int, key, value) -> int:
        self.ns.clock.begin_op()
        return self.nodes[rank].insert(list(key), self.ns.torch.tensor(list(value), dtype=self.ns.torch.int64))

    def match(self, rank: int, key):
        self.ns.clock.begin_op()
        return self.nodes[rank].match_prefix(list(key))

    def router(self, i: int = 0, warm: bool = True):
        return self.ns.car.CacheAwareRouter(self.nodes[self.P + self.D + i], skip_warm_up=warm)

    def full_key(self, rank: int, node):
        return self.nodes[rank]._full_key(node)

    def canonical(self, rank: int) -> dict:
        t = self.nodes[rank]
        rows = []
        stack = [t.root_node]
        while stack:
            n = stack.pop()
            if n is not t.root_node:
                v = n.value
                raw = tuple(int(x) for x in v.value.tolist()) if hasattr(v, "value") and v.value is not None else ()
                rows.append((tuple(t._full_key(n)), len(n.key), raw, v.node_rank, n.lock_ref))
            stack.extend(n.children.values())
        rows.sort()
        dup = {(tuple(k.key), k.node_rank): (None if n is None else n.lock_ref) for k, n in t.dup_nodes.items()}
        return {"nodes": rows, "evictable": t.evictable_size_, "protected": t.protected_size_, "dup": dup}
