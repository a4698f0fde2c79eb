"""The CPU oracle replayed against golden vectors recorded from the unmodified reference
(tools/make_golden.py).  Runs anywhere; this is what pins the oracle on the GPU box."""
import numpy as np
import pytest

from oracle.radix_port import PortConsistentHash, PortMesh, PortRadixCache, RouterPickError, md5_u32
from tests.golden_util import canon_json, load


def _replay_base(case, t):
    held = []
    for op in case["ops"]:
        if op["op"] == "ins":
            assert t.insert(list(op["key"]), np.asarray(op["val"], dtype=np.int64)) == op["ret"]
        elif op["op"] == "match":
            m = t.match_prefix(list(op["key"]))
            assert m.indices.tolist() == op["idx"]
            assert t.full_key(m.last_node) == op["last"]
            held.append(m.last_node)
        elif op["op"] == "lock":
            assert t.inc_lock_ref(held[op["h"]]) == op["ret"]
            held.append(held[op["h"]])
        elif op["op"] == "unlock":
            assert t.dec_lock_ref(held[op["h"]]) == op["ret"]
        elif op["op"] == "evict":
            before = len(t.freed)
            t.evict(op["n"])
            assert t.freed[before:] == op["freed"]
        assert canon_json(t.canonical()) == op["state"]


@pytest.mark.parametrize("case", load("base_sequences"), ids=lambda c: f"seed{c['seed']}")
def test_base_sequences(case):
    _replay_base(case, PortRadixCache(1))


@pytest.mark.parametrize("case", load("paged_sequences") + load("paged_ragged_sequences"),
                         ids=lambda c: f"seed{c['seed']}-page{c['page']}")
def test_paged_sequences(case):
    """page_size > 1, page-aligned inserts and inserts of any length (a short last page stays in the leaf)."""
    _replay_base(case, PortRadixCache(case["page"]))


@pytest.mark.parametrize("block", load("correctness_blocks"), ids=lambda b: b["name"])
def test_reference_correctness_blocks(block):
    mesh = PortMesh(3, 2, 1)
    for r, k, v in block["inserts"]:
        mesh.insert(r, k, np.asarray(v, dtype=np.int64))
    for row in block["expect"]:
        for r, idx in row["pd"].items():
            assert mesh.match(int(r), row["key"]).indices.tolist() == idx
        assert list(mesh.match(5, row["key"])) == row["router"]
        assert list(mesh.route(row["key"])) == row["route"]
    if "state" in block:
        for r, st in block["state"].items():
            assert canon_json(mesh.canonical(int(r))) == st


@pytest.mark.parametrize("case", load("mesh_sequences"), ids=lambda c: f"seed{c['seed']}")
def test_mesh_sequences(case):
    mesh = PortMesh(3, 2, 1)
    held = []
    for op in case["ops"]:
        if op["op"] == "ins":
            assert mesh.insert(op["rank"], op["key"], np.asarray(op["val"], dtype=np.int64)) == op["ret"]
        elif op["op"] == "match":
            m = mesh.match(op["rank"], op["key"])
            assert m.indices.tolist() == op["idx"]
            assert mesh.nodes[op["rank"]].full_key(m.last_node) == op["last"]
            held.append((op["rank"], m.last_node))
        elif op["op"] == "route":
            if op.get("error") == "AssertionError":
                with pytest.raises(RouterPickError):
                    mesh.match(5, op["key"])
                continue
            assert list(mesh.match(5, op["key"])) == op["ranks"]
            if op.get("error") == "IndexError":
                with pytest.raises(IndexError):
                    mesh.route(op["key"])
            else:
                assert list(mesh.route(op["key"])) == op["addrs"]
        elif op["op"] == "lock":
            r, n = held[op["h"]]
            assert mesh.nodes[r].inc_lock_ref(n) == op["ret"]
        elif op["op"] == "unlock":
            r, n = held[op["h"]]
            assert mesh.nodes[r].dec_lock_ref(n) == op["ret"]
    for r, st in case["state"].items():
        assert canon_json(mesh.canonical(int(r))) == st
    for r, cands in case["gc_candidates"].items():
        assert sorted([list(k[0]), k[1]] for k in mesh.nodes[int(r)].gc_candidates()) == cands


@pytest.mark.parametrize("case", load("consistent_hash"), ids=lambda c: f"n{len(c['addrs'])}")
def test_consistent_hash(case):
    h = PortConsistentHash(case["addrs"])
    assert h.ring == case["ring"]
    for q in case["queries"]:
        assert md5_u32(str(q["key"])) == q["hash"]
        assert h.get_node(q["key"]) == q["node"]
