"""Intra-box oplog transport over NVLink.

The reference forwards every INSERT hop by hop around a TCP ring (`policy/sync_algo.py:57-75`,
`radix/radix_mesh.py:339-347`, `communication/communicator.py:183-210`): P + D sequential hops of JSON per insert,
one tree insert per hop.  Inside one box the mesh nodes are GPUs on NVSwitch, so the records of a round travel as
packed arrays in ONE exchange and every replica applies the same ordered batches:

    stage()               local inserts of this rank (owner rank = this rank), packed on the host (IngestBuffer)
    exchange_and_apply()  all-gather of (records, tokens) counts, then of the padded token / value / length arrays;
                          every rank applies the records of rank 0, 1, ..., W-1 in that order, each rank's records in
                          their local order, as ONE ordered batch through rm_insert_batch (device pointers, owner rank
                          per record, the conflict rule of radix_mesh.py:273-323 on the device).

All replicas see one total order, so they end identical (the ring gives every node a different interleaving and
relies on the lowest-rank-wins rule to converge; the same rule runs here, the order is just no longer a variable).
Cross-box traffic stays on the reference's TCP path: feed what arrives there through RadixMesh.oplog_received_wire.

One process per GPU, `torch.distributed` (NCCL on the GPUs; gloo with the host-emulation backend in the CPU tests).
"""
from __future__ import annotations

from typing import List

import numpy as np

from . import capi
from .oplog_codec import IngestBuffer


def _as_torch(x):
    import torch
    return x if isinstance(x, torch.Tensor) else torch.from_numpy(x)


class NvlinkOplogExchange:
    """`indexes`: the device indexes of THIS rank that replicate the mesh tree (its P/D tree, optionally a ROUTER
    tree as well); every one of them receives every record."""

    def __init__(self, indexes: List[capi.DeviceIndex], rank: int, world: int, group=None):
        self.ixs, self.rank, self.world, self.group = list(indexes), rank, world, group
        self.be = self.ixs[0].be
        self._stage = IngestBuffer(with_values=True, alloc=self._pinned if getattr(self.be, "name", "") == "cuda" else None)
        self.applied = 0

    @staticmethod
    def _pinned(n, dtype):
        """Staging arrays in pinned host memory: their copy to the device does not block the host."""
        import torch
        return torch.empty(int(n), dtype={np.int32: torch.int32, np.int64: torch.int64}[dtype], pin_memory=True).numpy()

    def stage(self, key, value) -> None:
        """A local insert (RadixMesh.insert on this node): applied, here and everywhere, at the next exchange."""
        self._stage.add(key, value, self.rank)

    def exchange_and_apply(self) -> int:
        """Collective.  Returns the number of records applied (the same on every rank)."""
        import torch
        import torch.distributed as dist
        be, W = self.be, self.world
        tok, val, offs, _ = self._stage.take()
        lens = np.diff(offs).astype(np.int32)
        mine = be.to_device(np.asarray([len(lens), int(offs[-1])], dtype=np.int64))
        counts = be.empty(2 * W, np.int64)
        if W > 1:
            dist.all_gather_into_tensor(_as_torch(counts), _as_torch(mine), group=self.group)
        else:
            _as_torch(counts).copy_(_as_torch(mine))
        cnt = be.to_host(counts).reshape(W, 2)
        max_rec, max_tok = int(cnt[:, 0].max()), int(cnt[:, 1].max())
        total = int(cnt[:, 0].sum())
        if total == 0:
            return 0
        # padded send buffers on the device, filled straight from the staged arrays (the pad is never written nor read:
        # every rank's true counts are known)
        n_tok_pad, n_rec_pad = max(max_tok, 1), max(max_rec, 1)
        d_tok, d_val, d_len = be.empty(n_tok_pad, np.int32), be.empty(n_tok_pad, np.int64), be.empty(n_rec_pad, np.int32)
        if len(tok):
            _as_torch(d_tok)[:len(tok)].copy_(torch.from_numpy(tok), non_blocking=True)
            _as_torch(d_val)[:len(val)].copy_(torch.from_numpy(val), non_blocking=True)
            _as_torch(d_len)[:len(lens)].copy_(torch.from_numpy(lens))
        g_tok = be.empty(W * n_tok_pad, np.int32)
        g_val = be.empty(W * n_tok_pad, np.int64)
        g_len = be.empty(W * n_rec_pad, np.int32)
        if W > 1:
            dist.all_gather_into_tensor(_as_torch(g_tok), _as_torch(d_tok), group=self.group)
            dist.all_gather_into_tensor(_as_torch(g_val), _as_torch(d_val), group=self.group)
            dist.all_gather_into_tensor(_as_torch(g_len), _as_torch(d_len), group=self.group)
        else:
            _as_torch(g_tok).copy_(_as_torch(d_tok))
            _as_torch(g_val).copy_(_as_torch(d_val))
            _as_torch(g_len).copy_(_as_torch(d_len))
        # One ordered batch out of the W padded slices (rank 0's records, then rank 1's, ...): the true counts select
        # the live part of every slice, the record offsets are one prefix sum, and the index takes the owner rank per
        # record - so a round is ceil(total / max_batch) insert calls instead of one (or more) per rank.
        dev = _as_torch(g_len).device
        n_rec = torch.as_tensor(cnt[:, 0], device=dev)
        n_tok = torch.as_tensor(cnt[:, 1], device=dev)
        live_rec = torch.arange(n_rec_pad, device=dev)[None, :] < n_rec[:, None]
        live_tok = torch.arange(n_tok_pad, device=dev)[None, :] < n_tok[:, None]
        lens_all = _as_torch(g_len).view(W, n_rec_pad)[live_rec].to(torch.int64)
        offs_all = torch.zeros(total + 1, dtype=torch.int64, device=dev)
        torch.cumsum(lens_all, 0, out=offs_all[1:])
        tok_all = _as_torch(g_tok).view(W, n_tok_pad)[live_tok]
        val_all = _as_torch(g_val).view(W, n_tok_pad)[live_tok]
        ranks_all = torch.repeat_interleave(torch.arange(W, dtype=torch.int32, device=dev), n_rec)
        out = torch.empty(total, dtype=torch.int32, device=dev)
        step = min(int(ix.cfg.max_batch) for ix in self.ixs)
        for a in range(0, total, step):
            b = min(a + step, total)
            # offsets are absolute positions in tok_all / val_all (rm_insert_batch takes a window of them)
            for ix in self.ixs:
                v = None if ix.mode == capi.MODE_ROUTER else self._arr(val_all)
                ix.insert(self._arr(tok_all), v, self._arr(offs_all[a:b + 1]), self._arr(ranks_all[a:b]), b - a, self._arr(out[a:b]))
        for ix in self.ixs:
            ix.check_device_error()
        self.applied += total
        return total

    def _arr(self, t):
        """torch tensor -> what the index's backend takes (numpy view for the host-emulation backend)."""
        return t if getattr(self.be, "name", "") == "cuda" else t.numpy()
