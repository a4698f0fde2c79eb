"""Duplicate-node GC across replicas (radixmesh_b200/mesh_gc.py; rm_gc_keys_h / rm_gc_agree_h / rm_gc_exec_h) on CPU:
two gloo processes with the host-emulation build, against the oracle's restatement of RadixMesh._gc_handle
(radix_mesh.py:362-389) run as the same protocol."""
import os
import random
import sys

import numpy as np

from oracle.radix_port import MODE_DECODE, MODE_PREFILL, PortRadixMesh
from tests.parity import pack

KEYS = [[1, 2, 3], [1, 2, 4, 4], [5, 5], [5, 6, 7]]


def _script(replica):
    """Ops every replica applies (the oplog stream is the same everywhere); locks differ per replica."""
    ops = []
    for i, k in enumerate(KEYS):                       # rank 1 owns everything first
        ops.append(("ins", 1, k, [100 + 10 * i + j for j in range(len(k))]))
    for i, k in enumerate(KEYS):                       # every key gets a reader: the node is locked
        ops.append(("lock", i))
    for i, k in enumerate(KEYS):                       # rank 0 re-inserts with other values: locked swap, old value parked
        ops.append(("ins", 0, k, [900 + 10 * i + j for j in range(len(k))]))
    unlock = [0, 1, 2] if replica == 0 else [0, 2, 3]  # key 1 stays locked on replica 1, key 3 on replica 0
    for i in unlock:
        ops.append(("unlock", i))
    return ops


def _apply_device(ix, ops):
    from tests.cases_mesh import mesh_canonical
    held = {}
    for op in ops:
        if op[0] == "ins":
            _, rank, k, v = op
            tk, offs = pack([k])
            vv, _ = pack([v], np.int64)
            ix.insert_host(tk, vv, offs, np.asarray([rank], np.int32))
        elif op[0] == "lock":
            k = KEYS[op[1]]
            tk, offs = pack([k])
            ol, on = np.zeros(1, np.int32), np.zeros(1, np.int32)
            ix.match_host(tk, offs, ol, on)
            assert int(ol[0]) == len(k)
            gen = int(mesh_canonical(ix)["_dump"]["gen"][int(on[0])])
            held[op[1]] = (int(on[0]), gen)
            d = ix.be.empty(1, np.int64)
            ix.lock(ix.be.to_device(np.asarray([on[0]], np.int32)), ix.be.to_device(np.asarray([gen], np.int32)), 1, +1, d)
        else:
            node, gen = held[op[1]]
            d = ix.be.empty(1, np.int64)
            ix.lock(ix.be.to_device(np.asarray([node], np.int32)), ix.be.to_device(np.asarray([gen], np.int32)), 1, -1, d)
    ix.check_device_error()


def _apply_oracle(port, ops):
    held = {}
    for op in ops:
        if op[0] == "ins":
            _, rank, k, v = op
            port.apply_insert(list(k), np.asarray(v, dtype=np.int64), rank)
        elif op[0] == "lock":
            m = port.match_prefix(list(KEYS[op[1]]))
            held[op[1]] = m.last_node
            port.inc_lock_ref(m.last_node)
        else:
            port.dec_lock_ref(held[op[1]])


def _dup_state(ix):
    from tests.cases_mesh import mesh_canonical
    return sorted((k, r, v) for (k, r), v in mesh_canonical(ix)["dup"].items())


def _worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from radixmesh_b200.mesh_gc import GcRound
    from tests.emu.emu_backend import make_index
    ix = make_index(mode=MODE_PREFILL if rank == 0 else MODE_DECODE, self_rank=rank, n_prefill=1, n_decode=1, max_nodes=256,
                    max_tokens=4096, max_batch=16, max_key_len=32, max_dups=64)
    _apply_device(ix, _script(rank))
    before = _dup_state(ix)
    free0 = ix.sizes().n_nodes
    removed, freed = GcRound(ix, rank, world).run()
    again = GcRound(ix, rank, world).run()              # nothing left that everybody agrees on
    q.put((rank, before, removed, sorted(freed), _dup_state(ix), again, int(ix.sizes().n_dups)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_replica_gc_round_matches_oracle():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 33500 + random.randrange(2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # the oracle: the same protocol on two PortRadixMesh replicas
    ports = [PortRadixMesh(MODE_PREFILL, 0, 1, 1), PortRadixMesh(MODE_DECODE, 1, 1, 1)]
    for r in range(2):
        _apply_oracle(ports[r], _script(r))
    exp_before = [sorted((k, r, v) for (k, r), v in p.dup_canonical().items()) for p in ports]
    queries = [sorted(p.gc_candidates()) for p in ports]
    agreed = [[k for k in queries[o] if all(ports[x].gc_agrees(k) for x in range(2) if x != o)] for o in range(2)]
    exp_freed = [[], []]
    for r in range(2):
        for o in [r] + [x for x in range(2) if x != r]:          # own list first, as GcRound does
            f = ports[r].gc_exec(agreed[o], origin=(o == r))
            if o == r:
                exp_freed[r] = sorted(f)
    exp_removed = sorted(set(agreed[0]) | set(agreed[1]))
    exp_after = [sorted((k, r, v) for (k, r), v in p.dup_canonical().items()) for p in ports]
    for r in range(2):
        _, before, removed, freed, after, again, n_dups = res[r]
        assert before == exp_before[r], f"replica {r}: duplicate table before the round"
        assert removed == exp_removed and len(exp_removed) == 2          # keys 0 and 2: unlocked on both replicas
        assert freed == exp_freed[r] and len(freed) == 2
        assert after == exp_after[r], f"replica {r}: duplicate table after the round"
        assert again == ([], []) and n_dups == len(exp_after[r])


def test_single_replica_round_and_slot_reuse():
    """World of one: every candidate is agreed and freed; the duplicate-table slots the round vacated are used again
    by later conflicts (same keys), and the table still equals the oracle's dup_nodes."""
    from radixmesh_b200.mesh_gc import GcRound
    from tests.emu.emu_backend import make_index
    ix = make_index(mode=MODE_DECODE, self_rank=1, n_prefill=1, n_decode=1, max_nodes=256, max_tokens=4096, max_batch=16,
                    max_key_len=32, max_dups=8)
    port = PortRadixMesh(MODE_DECODE, 1, 1, 1)
    ops = [op for op in _script(0) if op[0] != "unlock"] + [("unlock", i) for i in range(len(KEYS))]
    _apply_device(ix, ops)
    _apply_oracle(port, ops)
    removed, freed = GcRound(ix, 0, 1).run()
    cand = sorted(port.gc_candidates())
    assert removed == cand and len(cand) == len(KEYS)
    assert sorted(freed) == sorted(port.gc_exec(cand, origin=True))
    assert _dup_state(ix) == sorted((k, r, v) for (k, r), v in port.dup_canonical().items())
    assert int(ix.sizes().n_dups) == len(port.dup_nodes)
    for rnd in range(3):                                # rank 1 loses against rank 0's values again and again
        more = [("ins", 1, k, [7000 + 100 * rnd + 10 * i + j for j in range(len(k))]) for i, k in enumerate(KEYS)]
        _apply_device(ix, more)
        _apply_oracle(port, more)
        assert _dup_state(ix) == sorted((k, r, v) for (k, r), v in port.dup_canonical().items())
        assert int(ix.sizes().n_dups) == len(port.dup_nodes)
    ix.check_device_error()
