"""Drop-in for `CacheAwareRouter` / `ConsistentHash` (router/cache_aware_router.py:8-118)."""
from __future__ import annotations

import bisect
import dataclasses
import hashlib
from typing import List

import numpy as np


@dataclasses.dataclass
class RouteResult:                    # cache_aware_router.py:8-11
    prefill_addr: str
    decode_addr: str


def _md5_u32(text: str) -> int:       # cache_aware_router.py:110-118
    return int.from_bytes(hashlib.md5(text.encode("utf-8")).digest()[:4], "big")


class ConsistentHash:
    """Host-side ring builder (the ring is static per node list; the per-request hashing runs on the GPU).
    Same construction as cache_aware_router.py:42-73: `replicas` virtual nodes "addr#i" per address."""

    def __init__(self, nodes=None, replicas: int = 3):
        self.replicas = replicas
        self.ring: List[int] = []
        self.ring_nodes = {}
        self.nodes = set()
        for n in nodes or []:
            self.add_node(n)

    def add_node(self, node) -> None:
        if node in self.nodes:
            return
        self.nodes.add(node)
        for i in range(self.replicas):
            h = _md5_u32(f"{node}#{i}")
            self.ring_nodes[h] = node
            bisect.insort(self.ring, h)

    def get_node(self, key):
        if not self.ring:
            return None
        return self.ring_nodes[self.ring[bisect.bisect(self.ring, _md5_u32(str(key))) % len(self.ring)]]

    def device_arrays(self, addrs: List[str]):
        pts = np.asarray(self.ring, dtype=np.uint32)
        own = np.asarray([addrs.index(self.ring_nodes[p]) for p in self.ring], dtype=np.int32)
        return pts, own


class CacheAwareRouter:
    def __init__(self, radix_mesh, skip_warm_up: bool = False):
        self.radix_mesh = radix_mesh
        self.warm_up_done = skip_warm_up
        p_addrs, d_addrs = list(radix_mesh.prefill_cache_nodes()), list(radix_mesh.decode_cache_nodes())
        pp, po = ConsistentHash(p_addrs).device_arrays(p_addrs)
        dp, do = ConsistentHash(d_addrs).device_arrays(d_addrs)
        radix_mesh._ix.set_rings(pp, po, dp, do)

    def finish_warm_up(self) -> None:
        self.warm_up_done = True

    def cache_aware_route(self, key: List[int]) -> RouteResult:
        mesh = self.radix_mesh
        if self.warm_up_done and len(key) == 0:
            raise IndexError("list index out of range")          # radix_mesh.py:253
        _pf, _dc, pa, da, err = mesh._route_keys([key], warm=self.warm_up_done)
        if err[0] == 1:
            raise AssertionError("[RadixMesh] Prefix match node_rank is invalid")      # radix_mesh.py:233-237
        if err[0] == 2:
            raise IndexError("list index out of range")          # radix_mesh.py:451
        p = mesh.prefill_cache_nodes()[int(pa[0])] if pa[0] >= 0 else None
        d = mesh.decode_cache_nodes()[int(da[0])] if da[0] >= 0 else None
        return RouteResult(p, d)

    def route_batch(self, tokens, offsets, out_prefill, out_decode, out_paddr, out_daddr, out_err):
        """Addition: device tensors in/out; address indices refer to prefill_cache_nodes() / decode_cache_nodes()."""
        mesh = self.radix_mesh
        mesh._flush()
        mesh._ix.route(tokens, offsets, int(offsets.shape[0]) - 1, self.warm_up_done, out_prefill, out_decode,
                       out_paddr, out_daddr, out_err)
        mesh._invalidate()
