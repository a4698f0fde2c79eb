"""Mesh routing across the GPUs of one box (north_star item 4).

Two arrangements, both one process per GPU over `torch.distributed`:

* `ShardedMeshRouter`  — GPU g hosts mesh node g's tree with the keys rank g inserted.  A request
  batch is broadcast, every GPU walks its own tree (non-splitting match), the matched lengths are
  all-gathered over NVLink and `rm_pick_batch` picks the prefill / decode node (consistent hash on
  a miss).  No oplog replication is needed for routing.
* `ReplicatedMeshRouter` — every GPU holds a replica of the ROUTER-mode tree and routes its shard
  of the request batch with `rm_route_batch`; results are all-gathered.  Bit-exact with the
  reference by construction, error cases included; the parity anchor for the sharded form.
"""
from __future__ import annotations

import numpy as np

from . import capi


def _as_torch(x):
    import torch
    return x if isinstance(x, torch.Tensor) else torch.from_numpy(x)


class ShardedMeshRouter:
    def __init__(self, index: capi.DeviceIndex, rank: int, world: int, group=None):
        self.ix, self.rank, self.world, self.group = index, rank, world, group

    def route(self, tokens, offsets, B: int, src: int = 0):
        """tokens / offsets are device arrays; valid on `src`, overwritten elsewhere by the broadcast.
        Returns (prefill_rank, decode_rank, prefill_addr_idx, decode_addr_idx, err) device arrays."""
        import torch.distributed as dist
        be = self.ix.be
        if self.world > 1:
            dist.broadcast(_as_torch(tokens), src=src, group=self.group)
            dist.broadcast(_as_torch(offsets), src=src, group=self.group)
        lens = be.empty(B, np.int32)
        nodes = be.empty(B, np.int32)
        self.ix.match(tokens, offsets, B, lens, nodes)              # K1, non-splitting (DECODE-mode index)
        depths = be.empty(self.world * B, np.int32)
        if self.world > 1:
            dist.all_gather_into_tensor(_as_torch(depths), _as_torch(lens), group=self.group)
        else:
            _as_torch(depths).copy_(_as_torch(lens))
        outs = [be.empty(B, np.int32) for _ in range(4)]
        err = be.empty(B, np.uint8)
        self.ix.pick(depths, self.world, tokens, offsets, B, outs[0], outs[1], outs[2], outs[3], err)
        return outs[0], outs[1], outs[2], outs[3], err


class ReplicatedMeshRouter:
    def __init__(self, index: capi.DeviceIndex, rank: int, world: int, group=None):
        self.ix, self.rank, self.world, self.group = index, rank, world, group

    def route_shard(self, tokens, offsets, B_local: int, warm: bool = True):
        """Route this rank's shard; returns the all-gathered (world*B_local) result arrays."""
        import torch.distributed as dist
        be = self.ix.be
        local = [be.empty(B_local, np.int32) for _ in range(4)]
        err = be.empty(B_local, np.int32)
        err8 = be.empty(B_local, np.uint8)
        self.ix.route(tokens, offsets, B_local, warm, local[0], local[1], local[2], local[3], err8)
        _as_torch(err).copy_(_as_torch(err8))
        outs = []
        for a in local + [err]:
            g = be.empty(self.world * B_local, np.int32)
            if self.world > 1:
                dist.all_gather_into_tensor(_as_torch(g), _as_torch(a), group=self.group)
            else:
                _as_torch(g).copy_(_as_torch(a))
            outs.append(g)
        return outs
