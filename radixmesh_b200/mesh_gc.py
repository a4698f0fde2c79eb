"""Duplicate-node GC across the replicas of one box (SURVEY.md 8e "GC-dup across GPUs", 8f-3).

The reference's protocol (radix_mesh.py:148-166, 362-389): a node sends GC_QUERY with its parked, unlocked dup_nodes
entries around the ring; every peer that also holds the key with an unlocked (or empty) entry adds one to `agree`;
back home, keys with full agreement are freed and a GC_EXEC makes everybody pop them.  Over the reference's real wire
this cannot complete (CacheOplog.to_dict drops gc_query / gc_exec), so the parity target is the
handler's logic, restated in oracle/radix_port.py (gc_candidates / gc_agrees / gc_exec).

Here one round is three collectives instead of two ring passes per node:
  1. every rank lists its candidates on the device (K7, rm_gc_candidates) and turns them into (full key, rank) pairs
     (rm_gc_keys_h); the lists are all-gathered;
  2. every rank answers every other rank's list (rm_gc_agree_h: read-only key lookup + duplicate-table probe); the
     answers are combined with an all-reduce(MIN) - a key survives only if every peer agreed;
  3. every rank pops the agreed keys (rm_gc_exec_h); the rank that asked releases its parked node and gets the KV
     indices back for token_to_kv_pool_allocator.free.
New behaviour relative to the reference, flagged as such: the MIN all-reduce replaces the agree counter.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

from . import capi


class GcRound:
    def __init__(self, index: capi.DeviceIndex, rank: int, world: int, group=None, device=None):
        self.ix, self.rank, self.world, self.group = index, rank, world, group
        self.device = device              # torch device of the collectives' tensors (None: CPU, gloo)

    def run(self) -> Tuple[List[Tuple[tuple, int]], List[List[int]]]:
        """Collective.  Returns (keys removed everywhere as (full key, rank), KV index runs this rank must free)."""
        import torch
        import torch.distributed as dist
        ix, W = self.ix, self.world
        cand = ix.dup_entries(only_candidates=True)                         # rows: node, rank, parked, lock_ref
        keys = ix.gc_keys(cand[:, 0]) if len(cand) else []
        mine = [(tuple(k), int(r)) for k, r in zip(keys, cand[:, 1].tolist())] if len(cand) else []
        mine.sort()
        lists: List[list] = [None] * W
        if W > 1:
            dist.all_gather_object(lists, mine, group=self.group)
        else:
            lists[0] = mine
        # answers for every origin's list, concatenated in rank order; an origin answers 1 for its own keys
        total = sum(len(l) for l in lists)
        if total == 0:
            return [], []
        answer = np.ones(total, dtype=np.int32)
        at = 0
        for o, l in enumerate(lists):
            if o != self.rank and l:
                answer[at:at + len(l)] = ix.gc_agree([list(k) for k, _ in l], [r for _, r in l]).astype(np.int32)
            at += len(l)
        t = torch.from_numpy(answer)
        if self.device is not None:
            t = t.to(self.device)
        if W > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        agreed = t.cpu().numpy().astype(bool)
        removed, freed = [], []
        # own list first: a key two ranks asked about must still release THIS rank's parked node
        order = [self.rank] + [o for o in range(W) if o != self.rank]
        starts = np.concatenate([[0], np.cumsum([len(l) for l in lists])])
        for o in order:
            l = [kv for kv, a in zip(lists[o], agreed[starts[o]:starts[o + 1]]) if a]
            if not l:
                continue
            off, ln = ix.gc_exec([list(k) for k, _ in l], [r for _, r in l], free_parked=(o == self.rank))
            if o == self.rank:
                keep = (off >= 0) & (ln > 0)
                vals = ix.read_runs(off[keep], ln[keep])        # the released runs in one copy
                at = 0
                for n in ln[keep].tolist():
                    freed.append(vals[at:at + n].tolist())
                    at += n
            removed.extend(l)
        ix.check_device_error()
        return sorted(set(removed)), freed
