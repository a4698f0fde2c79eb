"""Intra-box oplog exchange (radixmesh_b200/mesh_oplog.py): two gloo processes on CPU with the
host-emulation build of the kernels.  Every replica must end identical to the oracle's node fed the records in the
order the exchange defines (rank 0's records of the round, then rank 1's, ...)."""
import os
import random
import sys

import numpy as np

from oracle.radix_port import MODE_DECODE, MODE_PREFILL, MODE_ROUTER, PortRadixMesh
from tests.parity import device_canonical

ROUNDS = 4


def _records(rank, rnd_seed):
    rnd = random.Random(rnd_seed * 31 + rank)
    out = []
    for rd in range(ROUNDS):
        recs = []
        for _ in range(rnd.randrange(0, 7) if not (rank == 1 and rd == 2) else 0):   # one empty round on rank 1
            key = [rnd.randrange(3) for _ in range(rnd.randrange(1, 8))]
            variant = rnd.randrange(2)
            recs.append((key, [100 * (rank + 1) * variant + 10 * i + t for i, t in enumerate(key)]))
        out.append(recs)
    return out


def _canon(ix):
    c = device_canonical(ix, mesh=True)
    inv = {v: k for k, v in c["_ids"].items()}
    inv[0] = ()
    dup = sorted((inv[node], rank, parked >= 0) for node, rank, parked, _ in ix.dup_entries().tolist())
    return {"nodes": c["nodes"], "evictable": c["evictable"], "protected": c["protected"], "dup": dup}


def _worker(rank, world, port, seed, q):
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from radixmesh_b200.mesh_oplog import NvlinkOplogExchange
    from tests.emu.emu_backend import make_index
    # rank 0 = prefill node 0, rank 1 = decode node 1; both also keep a router replica; small max_batch: chunking
    mode = MODE_PREFILL if rank == 0 else MODE_DECODE
    kw = dict(n_prefill=1, n_decode=1, max_nodes=1024, max_tokens=1 << 14, max_batch=4, max_key_len=32, max_dups=256)
    tree = make_index(mode=mode, self_rank=rank, **kw)
    router = make_index(mode=MODE_ROUTER, self_rank=2, **kw)
    ex = NvlinkOplogExchange([tree, router], rank, world)
    applied = []
    for recs in _records(rank, seed):
        for key, val in recs:
            ex.stage(key, np.asarray(val, dtype=np.int64))
        applied.append(ex.exchange_and_apply())
    q.put((rank, applied, _canon(tree), _canon(router)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_process_exchange_matches_oracle():
    import torch.multiprocessing as mp
    seed = 5
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + random.randrange(2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, seed, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # the oracle: one node per replica, fed rank-major per round
    per_rank = [_records(r, seed) for r in range(2)]
    ports = [PortRadixMesh(MODE_PREFILL, 0, 1, 1), PortRadixMesh(MODE_DECODE, 1, 1, 1), PortRadixMesh(MODE_ROUTER, 2, 1, 1)]
    expected_counts = []
    for rd in range(ROUNDS):
        n = 0
        for r in range(2):
            for key, val in per_rank[r][rd]:
                for p in ports:
                    p.apply_insert(list(key), np.asarray(val, dtype=np.int64), r)
                n += 1
        expected_counts.append(n)

    def port_canon(t):
        rows = []
        for nd in t.iter_nodes():
            raw = t._raw(nd.value)
            rows.append((tuple(t.full_key(nd)), len(nd.key), tuple(int(x) for x in raw) if raw is not None else (),
                         nd.value.rank, nd.lock_ref))
        rows.sort()
        dup = sorted((tuple(k[0]), k[1], v is not None) for k, v in t.dup_canonical().items())
        return {"nodes": rows, "evictable": t.evictable, "protected": t.protected, "dup": dup}

    (_, a0, tree0, router0), (_, a1, tree1, router1) = res
    assert a0 == a1 == expected_counts and sum(expected_counts) > 10
    assert len(tree0["nodes"]) > 5 and len(tree0["dup"]) > 0, "the scenario must exercise conflicts"
    assert router0 == router1 == port_canon(ports[2])
    assert tree0 == port_canon(ports[0])
    assert tree1 == port_canon(ports[1])
