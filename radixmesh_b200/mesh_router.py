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
    """`shard_pick=True` (default): the pick — and with it the MD5 of the consistent-hash fallback, the
    expensive part for long requests — is itself sharded: rank r picks for requests [r*B/W, (r+1)*B/W), so the
    matched lengths travel by all-to-all (every rank receives only its slice from everybody) and the five
    result arrays by one all-gather.  `shard_pick=False`: every rank gathers all lengths and picks for the
    whole batch (the first version; 2.3x slower at 8 GPUs on 2K-token requests)."""

    def __init__(self, index: capi.DeviceIndex, rank: int, world: int, group=None, shard_pick: bool = True,
                 bcast_group=None):
        self.ix, self.rank, self.world, self.group = index, rank, world, group
        self.shard_pick = shard_pick and world > 1
        # route_pipelined sends the request tokens on a communicator of their own, so that chunk k+1 travels while
        # chunk k is walked, exchanged and picked (pass a second group; collectives of one group run in issue order)
        self.bcast_group = bcast_group if bcast_group is not None else group

    def route_pipelined(self, tokens, offsets, B: int, src: int = 0, chunks: int = 8):
        """route() with the token broadcast cut into `chunks` pieces.  What limits the sharded form is the broadcast of
        the request tokens (every GPU must see every request; NVLink carries 4 bytes per token per GPU), so all the
        broadcasts are queued at once on their own communicator and each chunk is WALKED as soon as its tokens have
        landed - the walk hides under the transfer.  The exchange of the matched lengths, the pick and the result gather
        then run once for the whole batch: the pick's consistent-hash fallback is a dependent chain per request
        (MD5), whose latency a small chunk pays in full, so it is not chunked.  Same results as route()."""
        import torch.distributed as dist
        if self.world == 1 or chunks <= 1 or B < chunks:
            return self.route(tokens, offsets, B, src)
        tok_t, offs_t = _as_torch(tokens), _as_torch(offsets)
        dist.broadcast(offs_t, src=src, group=self.group)
        offs_h = offs_t.cpu()                                           # chunk boundaries in tokens: one host read per batch
        step = (B + chunks - 1) // chunks
        cuts = [min(k * step, B) for k in range(chunks + 1)]
        works = [dist.broadcast(tok_t[int(offs_h[cuts[k]]):int(offs_h[cuts[k + 1]])], src=src, group=self.bcast_group, async_op=True)
                 for k in range(chunks) if cuts[k + 1] > cuts[k]]
        lens = self.ix.be.empty(B, np.int32)
        nodes = self.ix.be.empty(B, np.int32)
        for k, w in enumerate(works):
            a, b = cuts[k], cuts[k + 1]
            w.wait()
            base = int(offs_h[a])
            sub_offs = offs_t[a:b + 1] - base
            self.ix.match(self._arr(tok_t[base:int(offs_h[b])]), self._arr(sub_offs), b - a, lens[a:b], nodes[a:b])
        return self.route(tokens, offsets, B, src, _lens=lens)

    def _arr(self, t):
        """torch tensor -> what the index's backend takes (numpy view for the host-emulation backend)."""
        return t if getattr(self.ix.be, "name", "") == "cuda" else t.numpy()

    def route(self, tokens, offsets, B: int, src: int = 0, _lens=None):
        """tokens / offsets are device arrays; valid on `src`, overwritten elsewhere by the broadcast.
        Returns (prefill_rank, decode_rank, prefill_addr_idx, decode_addr_idx, err) device arrays."""
        import torch.distributed as dist
        be = self.ix.be
        W = self.world
        if _lens is not None:                                        # route_pipelined: broadcast and walk already done
            lens = _lens
        else:
            if W > 1:
                dist.broadcast(_as_torch(tokens), src=src, group=self.group)
                dist.broadcast(_as_torch(offsets), src=src, group=self.group)
            lens = be.empty(B, np.int32)
            nodes = be.empty(B, np.int32)
            self.ix.match(tokens, offsets, B, lens, nodes)          # K1, non-splitting (DECODE-mode index)
        if not self.shard_pick:
            depths = be.empty(W * B, np.int32)
            if W > 1:
                dist.all_gather_into_tensor(_as_torch(depths), _as_torch(lens), group=self.group)
            else:
                _as_torch(depths).copy_(_as_torch(lens))
            outs = [be.empty(B, np.int32) for _ in range(4)]
            err = be.empty(B, np.uint8)
            self.ix.pick(depths, W, tokens, offsets, B, outs[0], outs[1], outs[2], outs[3], err)
            return outs[0], outs[1], outs[2], outs[3], err
        Bs = (B + W - 1) // W                                        # slice of rank j: [j*Bs, min((j+1)*Bs, B))
        bounds = [min(j * Bs, B) for j in range(W + 1)]
        lo, n = bounds[self.rank], bounds[self.rank + 1] - bounds[self.rank]
        depths = be.empty(max(W * n, 1), np.int32)                   # [g][i] = matched length of request lo+i on rank g
        dist.all_to_all_single(_as_torch(depths)[:W * n], _as_torch(lens), [n] * W,
                               [bounds[j + 1] - bounds[j] for j in range(W)], group=self.group)
        loc = be.empty(4 * Bs, np.int32)
        err = be.empty(Bs, np.uint8)
        _as_torch(loc).zero_()
        _as_torch(err).zero_()
        if n > 0:
            self.ix.pick(depths, W, tokens, offsets[lo:], n, loc[0:Bs], loc[Bs:2 * Bs], loc[2 * Bs:3 * Bs], loc[3 * Bs:4 * Bs], err)
        g = be.empty(W * 4 * Bs, np.int32)
        ge = be.empty(W * Bs, np.uint8)
        dist.all_gather_into_tensor(_as_torch(g), _as_torch(loc), group=self.group)
        dist.all_gather_into_tensor(_as_torch(ge), _as_torch(err), group=self.group)
        r = _as_torch(g).view(W, 4, Bs).permute(1, 0, 2).reshape(4, W * Bs)[:, :B].contiguous()
        return r[0], r[1], r[2], r[3], _as_torch(ge)[:B]


class ReplicatedMeshRouter:
    def __init__(self, index: capi.DeviceIndex, rank: int, world: int, group=None):
        self.ix, self.rank, self.world, self.group = index, rank, world, group

    def route_shard(self, tokens, offsets, B_local: int, warm: bool = True):
        """Route this rank's shard; returns the all-gathered (world*B_local) result arrays (prefill rank, decode rank,
        prefill address index, decode address index, error), gathered in ONE collective."""
        import torch
        import torch.distributed as dist
        be = self.ix.be
        W = self.world
        loc = be.empty(5 * B_local, np.int32)
        err8 = be.empty(B_local, np.uint8)
        self.ix.route(tokens, offsets, B_local, warm, loc[0:B_local], loc[B_local:2 * B_local], loc[2 * B_local:3 * B_local],
                      loc[3 * B_local:4 * B_local], err8)
        _as_torch(loc)[4 * B_local:].copy_(_as_torch(err8))
        if W == 1:
            t = _as_torch(loc)
            return [t[i * B_local:(i + 1) * B_local] for i in range(5)]
        g = be.empty(W * 5 * B_local, np.int32)
        dist.all_gather_into_tensor(_as_torch(g), _as_torch(loc), group=self.group)
        r = _as_torch(g).view(W, 5, B_local).permute(1, 0, 2).reshape(5, W * B_local)
        return [r[i] for i in range(5)]
