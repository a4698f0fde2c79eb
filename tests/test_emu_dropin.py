"""Drop-in classes on the host-emulation build: host logic of RadixCache / RadixMesh / CacheAwareRouter
(handles, staged oplog apply, ring forwarding, exceptions) against the reference's recorded results."""
import pytest

from tests import dropin_cases as dc
from tests.emu.emu_backend import NumpyBackend, emu_lib


def kw():
    return dict(_backend=NumpyBackend(), _library=emu_lib())


@pytest.mark.parametrize("block", dc.golden("correctness_blocks"), ids=lambda b: b["name"])
def test_reference_correctness_blocks(block):
    dc.replay_correctness_block(block, kw())


@pytest.mark.parametrize("case", dc.golden("mesh_sequences"), ids=lambda c: f"seed{c['seed']}")
def test_mesh_sequences(case):
    dc.replay_mesh_sequence(case, kw())


@pytest.mark.parametrize("case", dc.golden("base_sequences"), ids=lambda c: f"seed{c['seed']}")
def test_base_sequences(case):
    dc.replay_base_sequence(case, kw())


def test_gc_round_over_the_ring():
    """GC_QUERY / GC_EXEC through RadixMesh.oplog_received on a 3P+2D+1R mesh (radix_mesh.py:148-166, 362-389): a parked
    duplicate that every replica has unlocked is freed at the node that asked and popped everywhere; one that a
    single replica still locks survives everywhere."""
    import torch
    nodes = dc.build_mesh(kw(), with_allocators=True)
    keys = [[1, 2, 3], [4, 5]]
    held = {}
    for i, k in enumerate(keys):
        nodes[1].insert(k, torch.tensor([100 + 10 * i + j for j in range(len(k))]))
    for r in range(5):                      # every replica has a reader on both keys
        for i, k in enumerate(keys):
            held[(r, i)] = nodes[r].match_prefix(k).last_device_node
            nodes[r].inc_lock_ref(held[(r, i)])
    for i, k in enumerate(keys):            # rank 0 wins both: locked swap, the old value is parked on every replica
        nodes[0].insert(k, torch.tensor([900 + 10 * i + j for j in range(len(k))]))
    for r in range(5):
        nodes[r].dec_lock_ref(held[(r, 0)])
        if r != 3:
            nodes[r].dec_lock_ref(held[(r, 1)])   # replica 3 keeps its reader on the second key
    k0 = dc_key(keys[0], 1)
    k1 = dc_key(keys[1], 1)
    for r in range(5):
        d = nodes[r].dup_nodes
        assert d[k0] is not None and d[k1] is not None
    assert sorted(nodes[2].gc_candidates(), key=lambda k: k.key) == [k0, k1]
    assert nodes[2].gc_scan_send() == 2
    for r in range(5):
        d = nodes[r].dup_nodes
        assert k0 not in d, f"replica {r} still holds the agreed key"
        assert d[k1] is not None, f"replica {r} lost the key replica 3 still locks"
    assert nodes[2].token_to_kv_pool_allocator.freed == [[100, 101, 102]]
    for r in (0, 1, 3, 4):
        assert nodes[r].token_to_kv_pool_allocator.freed == []
    assert nodes[2].gc_candidates() == [k1]           # still asked about, still refused
    assert nodes[2].gc_scan_send() == 1
    assert nodes[2].token_to_kv_pool_allocator.freed == [[100, 101, 102]]
    nodes[3].dec_lock_ref(held[(3, 1)])
    assert nodes[4].gc_scan_send() == 1
    for r in range(5):
        assert k1 not in nodes[r].dup_nodes
    assert nodes[4].token_to_kv_pool_allocator.freed == [[110, 111]]
    # the trees still answer
    assert nodes[3].match_prefix(keys[1]).device_indices.tolist() == [910, 911]


def dc_key(key, rank):
    from radixmesh_b200.host import NodeKey
    return NodeKey(tuple(key), rank)
