"""Parity at the sizes bench.py measures: the benchmark's own generators, tree sizes and batch sizes.

C2 runs against the pure-Python restatement of RadixCache in full (matched lengths, last nodes, final tree with
stamp order, eviction order until the tree is empty).  The 1 M-node / 512-token target shape and the 4096-token
shape are too large for a Python tree, so they are checked in full through the size-independent property the
domain offers: matched_len(q) = max LCP(q, inserted keys), computed with numpy (oracle/lcp_oracle.py, itself
pinned against the restatement), plus last-node and tree-shape invariants read back from the device.
"""
import numpy as np
import pytest

import bench
from tests.parity import assert_same_digest, device_digest, oracle_digest

pytestmark = pytest.mark.gpu


def _build_device_tree(wl, docs, ix, dev, keep_rows=True, orc=None):
    """bench.py's build loop; returns the inserted rows (and feeds the oracle the same records when given)."""
    import torch
    L = wl["L"]
    nb = bench.INSERT_BATCH
    offs = torch.arange(0, (nb + 1) * L, L, dtype=torch.int64, device=dev)
    out = torch.empty(nb, dtype=torch.int32, device=dev)
    rows, vc = [], 0
    for batch in bench.insert_stream(wl, docs):
        tok = torch.from_numpy(batch.reshape(-1)).to(dev)
        val = torch.arange(vc, vc + nb * L, dtype=torch.int64, device=dev)
        ix.insert(tok, val, offs, None, nb, out)
        if orc is not None:
            exp = [orc.insert(r.tolist(), np.arange(vc + i * L, vc + (i + 1) * L, dtype=np.int64)) for i, r in enumerate(batch)]
            assert out.cpu().tolist() == exp, "insert prefix lengths differ from the oracle"
        if keep_rows:
            rows.append(batch)
        vc += nb * L
        if ix.sizes().n_nodes >= wl["tree_nodes"]:
            break
    ix.check_device_error()
    return np.concatenate(rows) if rows else None


def _match(ix, q, L, dev):
    import torch
    B = q.shape[0]
    tok = torch.from_numpy(q.reshape(-1)).to(dev)
    offs = torch.arange(0, (B + 1) * L, L, dtype=torch.int64, device=dev)
    ol = torch.empty(B, dtype=torch.int32, device=dev)
    on = torch.empty(B, dtype=torch.int32, device=dev)
    ix.match(tok, offs, B, ol, on)
    torch.cuda.synchronize()
    return ol.cpu().numpy(), on.cpu().numpy()


def test_gpu_c2_full_shape_against_the_oracle():
    """BASELINE.json configs[1] as bench.py runs it: tree grown to 100 K nodes with bench.insert_stream, then two
    65,536 x 256 batches through rm_match_batch, against PortRadixCache processing the same operations in order."""
    import torch
    from oracle.radix_port import PortRadixCache
    from radixmesh_b200 import capi
    wl = bench.WORKLOADS["c2"]
    L, B = wl["L"], wl["B"]
    dev = torch.device("cuda")
    docs = bench.gen_docs(wl)
    ix = capi.DeviceIndex(mode=capi.MODE_BASE, max_nodes=wl["tree_nodes"] * 2 + 4 * B, max_tokens=int(wl["tree_nodes"] * L * 0.9) + (1 << 22),
                          max_batch=B, max_key_len=1024)
    orc = PortRadixCache(1)
    _build_device_tree(wl, docs, ix, dev, keep_rows=False, orc=orc)
    assert_same_digest(device_digest(ix), oracle_digest(orc))
    last = []
    for step in range(2):
        q = bench.gen_batch(bench.query_seed(wl, 0, 0, step), B, L, docs)
        ol, on = _match(ix, q, L, dev)
        exp = [orc.match_prefix(r.tolist()) for r in q]
        assert ol.tolist() == [len(m.indices) for m in exp], f"matched lengths differ in batch {step}"
        last.append((on, [m.last_node for m in exp]))
    ix.check_device_error()
    dd, od = device_digest(ix), oracle_digest(orc)
    assert_same_digest(dd, od)                        # nodes, boundaries, keys, values, stamp order
    for on, nodes in last:                            # last_device_node of every query, by identity (full key)
        assert [dd["by_id"][x] for x in on.tolist()] == [od["by_node"][n] for n in nodes]
    # evict everything: same nodes, same order, same value runs handed back (radix_cache.py:179-202)
    val = dd["val"]
    _, offs, lens = ix.evict(1 << 40)
    orc.evict(1 << 40)
    assert lens.tolist() == [len(r) for r in orc.freed], "eviction order differs"
    got = np.concatenate([val[o:o + n] for o, n in zip(offs.tolist(), lens.tolist())])
    assert np.array_equal(got, np.fromiter((v for r in orc.freed for v in r), dtype=np.int64, count=len(got)))
    assert ix.sizes().n_nodes == 0 and ix.sizes().evictable == 0
    ix.check_device_error()
    ix.close()


def _property_check(name, steps, sample_nodes=4096, tree_nodes=None):
    import torch
    from oracle.lcp_oracle import longest_prefix_lengths
    from radixmesh_b200 import capi
    wl = dict(bench.WORKLOADS[name])
    if tree_nodes:
        wl["tree_nodes"] = tree_nodes
    L, B = wl["L"], wl["B"]
    dev = torch.device("cuda")
    docs = bench.gen_docs(wl)
    ix = capi.DeviceIndex(mode=capi.MODE_BASE, max_nodes=int(wl["tree_nodes"] * 1.3) + (steps + 1) * B, max_key_len=max(L, 1024),
                          max_tokens=int(wl["tree_nodes"] * L * 0.9) + (1 << 22), max_batch=B)
    ins = _build_device_tree(wl, docs, ix, dev)
    nodes_before = int(ix.sizes().n_nodes)
    evictable = int(ix.sizes().evictable)
    for step in range(steps):
        q = bench.gen_batch(bench.query_seed(wl, 0, 0, step), B, L, docs)
        ol, on = _match(ix, q, L, dev)
        exp = longest_prefix_lengths(ins, q)
        bad = np.nonzero(ol != exp)[0]
        assert len(bad) == 0, f"{name}: {len(bad)} of {B} matched lengths differ in batch {step}, first q={bad[:5]}"
        # a splitting match returns a node whose full key IS the matched prefix (radix_cache.py:262-266)
        d = ix.dump()
        assert np.array_equal(d["depth"][on], ol), "last node does not end at the matched length"
        rng = np.random.default_rng(step)
        for qi in rng.choice(B, size=sample_nodes // 16, replace=False).tolist():
            x, parts = int(on[qi]), []
            while x != 0:
                parts.append(ix.read_arena(int(d["edge_off"][x]), int(d["edge_len"][x]), values=False)[0])
                x = int(d["parent"][x])
            key = np.concatenate(parts[::-1]) if parts else np.zeros(0, np.int32)
            assert np.array_equal(key, q[qi, :ol[qi]]), f"full key of the last node of query {qi} is not its matched prefix"
        # an idempotence property: the same batch again matches the same lengths and creates no node
        n1 = int(ix.sizes().n_nodes)
        ol2, on2 = _match(ix, q, L, dev)
        assert np.array_equal(ol2, ol) and np.array_equal(on2, on) and int(ix.sizes().n_nodes) == n1
    s = ix.sizes()
    assert s.evictable == evictable and s.protected_ == 0          # splits move no tokens
    assert s.n_nodes >= nodes_before
    ix.check_device_error()
    ix.close()


def test_gpu_target_shape_matched_lengths():
    """north_star target shape (1 M-node tree, 65,536 x 512-token queries), every query of two batches."""
    _property_check("target", steps=2)


def test_gpu_c5r_shape_matched_lengths():
    """4096-token queries (BASELINE.json configs[4] generator), every query of one 16,384-query batch.  The tree is
    grown to 100 K nodes here, not bench.py's 500 K: the numpy oracle needs every inserted 16 KB row on the host."""
    _property_check("c5r", steps=1, tree_nodes=100_000)
