"""Small host-side types the mesh drop-in needs when it runs WITHOUT the reference package.

Inside the reference, `RadixMesh` keeps using the reference's own `ServerArgs`, `CacheOplog`,
`RingSyncAlgo` and communicators (all duck-typed here); these stand-ins exist so the drop-in and its
tests run on a box where /root/reference is absent.  Control plane only — no arithmetic."""
from __future__ import annotations

import dataclasses
import enum
from typing import List


class RadixMode(str, enum.Enum):      # radix/core_enum.py:4-7
    PREFILL = "prefill"
    DECODE = "decode"
    ROUTER = "router"


class CacheOplogType(enum.IntEnum):   # radix/cache_oplog.py:13-22
    INSERT = 1
    DELETE = 2
    RESET = 3
    GC_QUERY = 4
    GC_EXEC = 5
    TICK = 10


@dataclasses.dataclass
class CacheOplog:                     # radix/cache_oplog.py:48-66 (fields the tree path reads)
    oplog_type: int
    node_rank: int
    ttl: int
    local_logic_id: int = 0
    key: list = dataclasses.field(default_factory=list)
    value: list = dataclasses.field(default_factory=list)
    gc_query: list = dataclasses.field(default_factory=list)   # of GCQuery; like the reference, NOT part of to_dict
    gc_exec: list = dataclasses.field(default_factory=list)    # of NodeKey

    def to_dict(self) -> dict:
        return {"oplog_type": int(self.oplog_type), "node_rank": self.node_rank, "local_logic_id": self.local_logic_id,
                "key": list(self.key), "value": list(self.value), "ttl": self.ttl}


@dataclasses.dataclass
class ServerArgs:                     # config/cache_config.py:6-35
    prefill_cache_nodes: List[str] = dataclasses.field(default_factory=list)
    router_cache_nodes: List[str] = dataclasses.field(default_factory=list)
    decode_cache_nodes: List[str] = dataclasses.field(default_factory=list)
    local_cache_addr: str = ""
    prefill_node_rank: int = -1
    decode_node_rank: int = -1
    router_node_rank: int = -1

    def is_decode_node_rank(self, r: int) -> bool:
        return len(self.prefill_cache_nodes) <= r < len(self.prefill_cache_nodes) + len(self.decode_cache_nodes)

    def is_prefill_node_rank(self, r: int) -> bool:
        return r < len(self.prefill_cache_nodes)

    def local_node_rank(self, r: int) -> int:
        p, d = len(self.prefill_cache_nodes), len(self.decode_cache_nodes)
        if r < p:
            return r
        if r < p + d:
            return r - p
        return r - p - d


class RingPolicy:                     # policy/sync_algo.py:50-110 (the few facts the tree path uses)
    MASTER = 0

    def master_node_rank(self) -> int:
        return self.MASTER

    def can_send(self, mode) -> bool:
        return mode != RadixMode.ROUTER

    def ttl(self, args) -> int:
        return len(args.prefill_cache_nodes) + len(args.decode_cache_nodes)


@dataclasses.dataclass(frozen=True)
class NodeKey:                        # radix/cache_oplog.py:25-40 (ImmutableNodeKey)
    key: tuple
    node_rank: int


@dataclasses.dataclass
class GCQuery:                        # radix/cache_oplog.py:43-45
    key: NodeKey
    agree: int = 1
