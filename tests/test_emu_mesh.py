"""Mesh kernels (host emulation build) against the CPU oracle: conflict rule, duplicate table, locked
swap with handle generations, router pick, consistent-hash fallback (MD5 on the device)."""
import pytest

from tests import cases_mesh
from tests.emu.emu_backend import make_index


@pytest.mark.parametrize("seed", range(40))
def test_emu_mesh_sequences(seed):
    cases_mesh.run_mesh_sequence(make_index, seed)


@pytest.mark.parametrize("seed", range(10))
def test_emu_mesh_binary_vocab(seed):
    cases_mesh.run_mesh_sequence(make_index, 500 + seed, n_steps=30, vocab=2, maxlen=6)


@pytest.mark.parametrize("seed", range(30))
def test_emu_mesh_big_batches(seed):
    """Up to 48 records per insert batch: records of one batch lose against, replace and split the same nodes, in
    resumed groups and in groups that stay on the ordered path (rm_insert.cu, two-phase mesh apply)."""
    cases_mesh.run_mesh_sequence(make_index, 9000 + seed, n_steps=12, vocab=2 + seed % 2, maxlen=6 + seed % 4, max_recs=48)


@pytest.mark.parametrize("seed", range(4))
def test_emu_mesh_hundreds_of_records_on_shared_paths(seed):
    """Batches of up to 300 records over four 40-token documents, five owner ranks, two value variants: replacements,
    losses and splits of one batch land on the same long paths (grouping by the top replaced node, rm_insert.cu)."""
    cases_mesh.run_mesh_sequence(make_index, 4000 + seed, n_steps=6, vocab=2, maxlen=40, max_recs=300, max_batch=512, docs=4)


def test_emu_consistent_hash_vectors():
    """ConsistentHash golden vectors (recorded from the reference) through the MD5 fallback kernel."""
    import numpy as np
    from oracle.radix_port import PortConsistentHash
    from tests.golden_util import load
    from tests.parity import pack
    for case in load("consistent_hash"):
        addrs = case["addrs"]
        ix = make_index(mode=3, n_prefill=len(addrs), n_decode=len(addrs), max_nodes=64, max_tokens=64, max_batch=64,
                        max_key_len=512)
        h = PortConsistentHash(addrs)
        pts = np.asarray(case["ring"], dtype=np.uint32)
        own = np.asarray([addrs.index(h.owner[p]) for p in h.ring], dtype=np.int32)
        ix.set_rings(pts, own, pts, own)
        keys = [q["key"] for q in case["queries"]]
        tk, offs = pack(keys)
        n = len(keys)
        outs = [np.zeros(n, np.int32) for _ in range(4)]
        err = np.zeros(n, np.uint8)
        ix.route(tk, offs, n, False, outs[0], outs[1], outs[2], outs[3], err)
        assert [addrs[i] for i in outs[2].tolist()] == [q["node"] for q in case["queries"]]
        assert err.tolist() == [0] * n
        ix.close()


@pytest.mark.parametrize("scratch_mb", [None, "0", "1"], ids=["rendered", "streamed", "mixed"])
def test_emu_md5_awkward_inputs(scratch_mb, monkeypatch):
    """Device MD5 of str(list) against hashlib, through both forms of the hash (rm_route.cu)."""
    import numpy as np
    from oracle.radix_port import PortConsistentHash
    from tests.parity import pack
    if scratch_mb is not None:
        monkeypatch.setenv("RM_ROUTE_SCRATCH_MB", scratch_mb)
    addrs = [f"n{i}:1" for i in range(7)]
    h = PortConsistentHash(addrs)
    ix = make_index(mode=3, n_prefill=7, n_decode=7, max_nodes=64, max_tokens=64, max_batch=512, max_key_len=4096)
    pts = np.asarray(h.ring, dtype=np.uint32)
    own = np.asarray([addrs.index(h.owner[p]) for p in h.ring], dtype=np.int32)
    ix.set_rings(pts, own, pts, own)
    keys = cases_mesh.md5_awkward_keys()
    tk, offs = pack(keys)
    n = len(keys)
    outs = [np.zeros(n, np.int32) for _ in range(4)]
    err = np.zeros(n, np.uint8)
    ix.route(tk, offs, n, False, outs[0], outs[1], outs[2], outs[3], err)
    assert outs[2].tolist() == [addrs.index(h.get_node(k)) for k in keys]
    ix.close()
