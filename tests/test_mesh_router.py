"""Mesh routing across GPUs: the pick rule on per-rank trees against the reference's router
semantics (oracle), and the torch.distributed orchestration with two gloo processes on CPU."""
import os
import random
import sys

import numpy as np
import pytest

from oracle.radix_port import MODE_DECODE, PortConsistentHash, PortMesh, RouterPickError
from tests.emu.emu_backend import make_index
from tests.parity import pack

P, D = 3, 2


def _rings(addrs):
    h = PortConsistentHash(addrs)
    return (np.asarray(h.ring, dtype=np.uint32), np.asarray([addrs.index(h.owner[p]) for p in h.ring], dtype=np.int32))


@pytest.mark.parametrize("seed", range(25))
def test_pick_from_per_rank_depths_matches_reference_router(seed):
    """Arrangement (A): tree g holds only what rank g inserted, the pick rule runs on the gathered depths.
    Lock-free meshes with distinct per-rank values; the reference's two error cases are excluded."""
    rnd = random.Random(seed)
    port = PortMesh(P, D, 1)
    trees = [make_index(mode=MODE_DECODE, self_rank=g, n_prefill=P, n_decode=D, max_nodes=2048, max_tokens=1 << 14,
                        max_batch=64, max_key_len=64) for g in range(P + D)]
    pp, po = _rings(port.p_addrs)
    dp, do = _rings(port.d_addrs)
    trees[0].set_rings(pp, po, dp, do)
    vc = 0
    for _ in range(rnd.randrange(3, 14)):
        rank = rnd.randrange(P + D)
        key = [rnd.randrange(3) for _ in range(rnd.randrange(1, 8))]
        val = [1000 * (rank + 1) + vc + i for i in range(len(key))]    # distinct values per insert
        vc += 10
        port.insert(rank, key, np.asarray(val, dtype=np.int64))
        tk, offs = pack([key])
        vv, _ = pack([val], np.int64)
        trees[rank].insert_host(tk, vv, offs, np.asarray([rank], np.int32))
    keys = [[rnd.randrange(3) for _ in range(rnd.randrange(1, 9))] for _ in range(40)]
    tk, offs = pack(keys)
    n = len(keys)
    depths = np.zeros((P + D, n), np.int32)
    for g, ix in enumerate(trees):
        on = np.zeros(n, np.int32)
        ix.match_host(tk, offs, depths[g], on)
    outs = [np.zeros(n, np.int32) for _ in range(4)]
    err = np.zeros(n, np.uint8)
    trees[0].pick(depths.reshape(-1).copy(), P + D, tk, offs, n, outs[0], outs[1], outs[2], outs[3], err)
    checked = 0
    for i, k in enumerate(keys):
        try:
            pf, dc = port.match(5, k)
        except RouterPickError:
            continue                       # all-decode path: the reference asserts
        if pf >= P:
            continue                       # single-node path owned by a decode rank: the reference raises later
        assert (int(outs[0][i]), int(outs[1][i])) == (pf, dc), f"query {k}"
        exp_p = port.p_addrs[pf] if pf != -1 else PortConsistentHash(port.p_addrs).get_node(k)
        exp_d = port.d_addrs[dc - P] if dc != -1 else PortConsistentHash(port.d_addrs).get_node(k)
        assert (port.p_addrs[outs[2][i]], port.d_addrs[outs[3][i]]) == (exp_p, exp_d)
        checked += 1
    assert checked > 0
    for ix in trees:
        ix.close()


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from radixmesh_b200.mesh_router import ReplicatedMeshRouter, ShardedMeshRouter
    from tests.emu.emu_backend import make_index as mk
    # sharded: rank 0 = prefill node 0, rank 1 = decode node 1 (P=1, D=1)
    ix = mk(mode=MODE_DECODE, self_rank=rank, n_prefill=1, n_decode=1, max_nodes=256, max_tokens=4096, max_batch=16, max_key_len=32)
    addrs_p, addrs_d = ["p0:1"], ["d0:2"]
    ix.set_rings(*_rings(addrs_p), *_rings(addrs_d))
    mine = [[1, 2, 3], [7, 7]] if rank == 0 else [[1, 2, 3, 4, 5], [9]]
    tk, offs = pack(mine)
    vv, _ = pack([[100 * (rank + 1) + i for i in range(len(k))] for k in mine], np.int64)
    ix.insert_host(tk, vv, offs, np.asarray([rank] * len(mine), np.int32))
    queries = [[1, 2, 3, 4, 5, 6], [1, 2, 3], [7, 7, 1], [9, 9], [5]]
    qt, qo = pack(queries)
    if rank != 0:
        qt[:] = 0                                   # only rank 0 holds the batch before the broadcast
    r = ShardedMeshRouter(ix, rank, world)
    pf, dc, pa, da, err = r.route(qt, qo, len(queries))
    sharded = (pf.tolist(), dc.tolist(), pa.tolist(), da.tolist(), err.tolist())
    # the same batch in pieces, token broadcasts queued ahead on a communicator of their own
    qt2, _ = pack(queries)
    if rank != 0:
        qt2[:] = 0
    rp = ShardedMeshRouter(ix, rank, world, bcast_group=dist.new_group(list(range(world))))
    piped = tuple(x.tolist() for x in rp.route_pipelined(qt2, qo.copy(), len(queries), chunks=3))
    assert piped == sharded, (piped, sharded)
    # replicated router tree, query batch sharded
    rt = mk(mode=3, self_rank=2, n_prefill=1, n_decode=1, max_nodes=256, max_tokens=4096, max_batch=16, max_key_len=32)
    rt.set_rings(*_rings(addrs_p), *_rings(addrs_d))
    allk = [[1, 2, 3], [7, 7], [1, 2, 3, 4, 5], [9]]
    owners = [0, 0, 1, 1]
    tk, offs = pack(allk)
    rt.insert_host(tk, None, offs, np.asarray(owners, np.int32))
    shard = queries[rank * 2:(rank + 1) * 2]
    st, so = pack(shard)
    rep = [a.tolist() for a in ReplicatedMeshRouter(rt, rank, world).route_shard(st, so, len(shard))]
    q.put((rank, sharded, rep))
    dist.barrier()
    dist.destroy_process_group()


def test_two_process_gloo_orchestration():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + random.randrange(2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    (_, s0, rep0), (_, s1, rep1) = res
    assert s0 == s1, "every rank computes the same pick from the gathered depths"
    pf, dc, pa, da, err = s0
    # [1..6]: prefill rank 0 matches 3, decode rank 1 matches 5 (deeper) -> (0, 1); [1,2,3] -> (0, -1) decode by hash;
    # [7,7,1] -> (0, -1); [9,9] -> only the decode rank matches -> prefill by hash, decode 1; [5] -> both by hash
    assert pf == [0, 0, 0, -1, -1] and dc == [1, -1, -1, 1, -1]
    assert pa == [0] * 5 and da == [0] * 5 and err == [0] * 5
    assert rep0 == rep1, "the gathered results of the replicated router are identical on every rank"
    # replicated router = the reference's own rule on its router tree (first four queries)
    assert rep0[0] == [0, 0, 0, 1] and rep0[1] == [1, -1, -1, -1]
