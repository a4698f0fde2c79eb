"""GPU parity: mesh kernels (conflict rule, duplicate table, locked swap, router pick, device MD5
fallback) and the drop-in classes against the oracle / the reference's recorded results."""
import numpy as np
import pytest

from tests import cases_mesh
from tests import dropin_cases as dc

pytestmark = pytest.mark.gpu


def make_index(**kw):
    from radixmesh_b200 import capi
    return capi.DeviceIndex(**kw)


@pytest.mark.parametrize("seed", range(8))
def test_gpu_mesh_sequences(seed):
    cases_mesh.run_mesh_sequence(make_index, seed)


@pytest.mark.parametrize("seed", range(3))
def test_gpu_mesh_binary_vocab(seed):
    cases_mesh.run_mesh_sequence(make_index, 500 + seed, n_steps=30, vocab=2, maxlen=6)


@pytest.mark.parametrize("seed", range(6))
def test_gpu_mesh_hundreds_of_records_on_shared_paths(seed):
    """Batches of up to 300 records over four 40-token documents, five owner ranks, two value variants: replacements,
    losses and splits of one batch land on the same long paths, with the groups running side by side on the GPU
    (grouping by the top replaced node, rm_insert.cu)."""
    cases_mesh.run_mesh_sequence(make_index, 4000 + seed, n_steps=6, vocab=2, maxlen=40, max_recs=300, max_batch=512, docs=4)


@pytest.mark.parametrize("block", dc.golden("correctness_blocks"), ids=lambda b: b["name"])
def test_gpu_reference_correctness_blocks(block):
    dc.replay_correctness_block(block, {})


@pytest.mark.parametrize("case", dc.golden("mesh_sequences")[:5], ids=lambda c: f"seed{c['seed']}")
def test_gpu_golden_mesh_sequences(case):
    dc.replay_mesh_sequence(case, {})


@pytest.mark.parametrize("case", dc.golden("base_sequences")[:6], ids=lambda c: f"seed{c['seed']}")
def test_gpu_golden_base_sequences(case):
    dc.replay_base_sequence(case, {})


def test_gpu_consistent_hash_vectors():
    """ConsistentHash golden vectors (recorded from the reference) through the device MD5 fallback."""
    import numpy as np
    import torch
    from radixmesh_b200 import capi
    from tests.parity import pack
    for case in dc.golden("consistent_hash"):
        addrs = case["addrs"]
        ix = capi.DeviceIndex(mode=capi.MODE_ROUTER, n_prefill=len(addrs), n_decode=len(addrs), max_nodes=64,
                              max_tokens=64, max_batch=64, max_key_len=512)
        pts = np.asarray(case["ring"], dtype=np.uint32)
        from oracle.radix_port import PortConsistentHash
        h = PortConsistentHash(addrs)
        own = np.asarray([addrs.index(h.owner[p]) for p in h.ring], dtype=np.int32)
        ix.set_rings(pts, own, pts, own)
        keys = [q["key"] for q in case["queries"]]
        tk, offs = pack(keys)
        n = len(keys)
        dev = torch.device("cuda")
        outs = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(4)]
        err = torch.empty(n, dtype=torch.uint8, device=dev)
        ix.route(torch.from_numpy(tk).to(dev), torch.from_numpy(offs).to(dev), n, False, outs[0], outs[1], outs[2], outs[3], err)
        pa = outs[2].cpu().tolist()
        assert [addrs[i] for i in pa] == [q["node"] for q in case["queries"]]
        assert outs[3].cpu().tolist() == pa and err.cpu().tolist() == [0] * n
        ix.close()


@pytest.mark.parametrize("scratch_mb", [None, "0", "1"], ids=["rendered", "streamed", "mixed"])
def test_gpu_md5_awkward_inputs(scratch_mb, monkeypatch):
    """Device MD5 of str(list) against hashlib through both forms of the hash: rendered by a block per request
    and compressed from scratch (default), streamed per thread (no scratch), and a batch that overflows a 1 MB
    scratch so that both forms run in one launch."""
    import numpy as np
    import torch
    from oracle.radix_port import PortConsistentHash
    from radixmesh_b200 import capi
    from tests.parity import pack
    if scratch_mb is not None:
        monkeypatch.setenv("RM_ROUTE_SCRATCH_MB", scratch_mb)
    addrs = [f"n{i}:1" for i in range(7)]
    h = PortConsistentHash(addrs)
    ix = capi.DeviceIndex(mode=capi.MODE_ROUTER, n_prefill=7, n_decode=7, max_nodes=64, max_tokens=64, max_batch=512,
                          max_key_len=4096)
    pts = np.asarray(h.ring, dtype=np.uint32)
    own = np.asarray([addrs.index(h.owner[p]) for p in h.ring], dtype=np.int32)
    ix.set_rings(pts, own, pts, own)
    keys = cases_mesh.md5_awkward_keys()
    tk, offs = pack(keys)
    n = len(keys)
    dev = torch.device("cuda")
    outs = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(4)]
    err = torch.empty(n, dtype=torch.uint8, device=dev)
    ix.route(torch.from_numpy(tk).to(dev), torch.from_numpy(offs).to(dev), n, False, outs[0], outs[1], outs[2], outs[3], err)
    exp = [addrs.index(h.get_node(k)) for k in keys]
    assert outs[2].cpu().tolist() == exp and outs[3].cpu().tolist() == exp
    assert err.cpu().tolist() == [0] * n
    ix.close()


def test_gpu_gc_keys_agree_exec():
    """rm_gc_keys_h / rm_gc_agree_h / rm_gc_exec_h on the GPU: two replicas in one process (the exchange itself is
    host code, tests/test_mesh_gc.py), against the oracle's restatement of _gc_handle (radix_mesh.py:362-389)."""
    from oracle.radix_port import MODE_DECODE, MODE_PREFILL, PortRadixMesh
    from radixmesh_b200 import capi
    from tests.test_mesh_gc import _apply_device, _apply_oracle, _dup_state, _script
    ixs, ports = [], []
    for r in range(2):
        ix = capi.DeviceIndex(mode=MODE_PREFILL if r == 0 else MODE_DECODE, self_rank=r, n_prefill=1, n_decode=1,
                              max_nodes=256, max_tokens=4096, max_batch=16, max_key_len=32, max_dups=64)
        _apply_device(ix, _script(r))
        ixs.append(ix)
        p = PortRadixMesh(MODE_PREFILL if r == 0 else MODE_DECODE, r, 1, 1)
        _apply_oracle(p, _script(r))
        ports.append(p)
    for o in range(2):
        cand = ixs[o].dup_entries(only_candidates=True)
        keys = ixs[o].gc_keys(cand[:, 0])
        q = sorted(zip(map(tuple, keys), cand[:, 1].tolist()))
        assert q == sorted(ports[o].gc_candidates())
        peer = 1 - o
        agree = ixs[peer].gc_agree([list(k) for k, _ in q], [r for _, r in q]).tolist()
        assert agree == [ports[peer].gc_agrees(k) for k in q]
        assert ixs[peer].gc_agree([[9, 9, 9], [1, 2]], [0, 1]).tolist() == [False, False]   # unknown key, inner prefix
        ok = [k for k, a in zip(q, agree) if a]
        off, ln = ixs[o].gc_exec([list(k) for k, _ in ok], [r for _, r in ok], free_parked=True)
        freed = sorted(ixs[o].read_arena(int(a), int(n))[1].tolist() for a, n in zip(off, ln) if a >= 0)
        assert freed == sorted(ports[o].gc_exec(ok, origin=True))
        ixs[peer].gc_exec([list(k) for k, _ in ok], [r for _, r in ok], free_parked=False)
        ports[peer].gc_exec(ok, origin=False)
        for r in range(2):
            assert _dup_state(ixs[r]) == sorted((k, rk, v) for (k, rk), v in ports[r].dup_canonical().items())
            ixs[r].check_device_error()
    for ix in ixs:
        ix.close()


def test_gpu_gc_round_over_the_ring():
    """GC_QUERY / GC_EXEC through the drop-in RadixMesh on a 3P+2D+1R mesh, device indexes on the GPU."""
    dc.run_gc_ring({})


def test_gpu_reset_tick_and_own_prefix():
    """Host-side oplog rules of the drop-in RadixMesh with the indexes on the GPU."""
    dc.run_reset_drops_staged({})
    dc.run_tick_ring({})
    dc.run_insert_returns_own_prefix({})


@pytest.mark.parametrize("case", dc.golden("paged_sequences") + dc.golden("paged_ragged_sequences"),
                         ids=lambda c: f"seed{c['seed']}-page{c['page']}")
def test_paged_sequences(case):
    """RadixCache(page_size > 1) recorded from the reference: the drop-in runs the device index on page symbols."""
    dc.replay_base_sequence(case, {})



@pytest.mark.parametrize("seed", range(3))
def test_gpu_oplog_wire_ingest(seed):
    """SURVEY.md 8f-1 on the GPU: packed and JSON wire bytes through RadixMesh.oplog_received_wire into a CUDA index
    (staged, gated on the header fields, applied as ordered rm_insert_batch_h calls) against the object path and the
    oracle's node."""
    from tests.test_oplog_codec import run_wire_case
    run_wire_case(seed, dict)


def test_gpu_oplog_exchange_on_one_rank():
    """SURVEY.md 8f-2, the apply half on the GPU (the all-gather needs one GPU per rank: bench.py --gpus N runs it over
    NCCL and compares the replicas): staged local inserts, packed, applied through rm_insert_batch on device pointers in
    max_batch chunks to a P/D tree and a ROUTER replica, against the oracle's nodes fed the same records."""
    import random
    from oracle.radix_port import MODE_PREFILL, MODE_ROUTER, PortRadixMesh
    from radixmesh_b200.mesh_oplog import NvlinkOplogExchange
    from tests.test_mesh_oplog import _canon
    kw = dict(n_prefill=1, n_decode=1, max_nodes=4096, max_tokens=1 << 16, max_batch=8, max_key_len=32, max_dups=512)
    tree = make_index(mode=1, self_rank=0, **kw)
    router = make_index(mode=3, self_rank=2, **kw)
    ex = NvlinkOplogExchange([tree, router], 0, 1)
    ports = [PortRadixMesh(MODE_PREFILL, 0, 1, 1), PortRadixMesh(MODE_ROUTER, 2, 1, 1)]
    rnd = random.Random(9)
    for rd in range(5):
        n = rnd.randrange(0, 30)
        for _ in range(n):
            key = [rnd.randrange(3) for _ in range(rnd.randrange(1, 9))]
            val = [10 * i + t for i, t in enumerate(key)]
            ex.stage(key, np.asarray(val, dtype=np.int64))
            for p in ports:
                p.apply_insert(list(key), np.asarray(val, dtype=np.int64), 0)
        assert ex.exchange_and_apply() == n

    def port_canon(t):
        rows = []
        for nd in t.iter_nodes():
            raw = t._raw(nd.value)
            rows.append((tuple(t.full_key(nd)), len(nd.key), tuple(int(x) for x in raw) if raw is not None else (), nd.value.rank, nd.lock_ref))
        rows.sort()
        dup = sorted((tuple(k[0]), k[1], v is not None) for k, v in t.dup_canonical().items())
        return {"nodes": rows, "evictable": t.evictable, "protected": t.protected, "dup": dup}

    assert _canon(tree) == port_canon(ports[0])
    assert _canon(router) == port_canon(ports[1])
    tree.close()
    router.close()


@pytest.mark.parametrize("seed", range(4))
def test_gpu_flatten_order_children_and_total_size(seed):
    """a12 on the GPU: all_values_flatten in the reference's pre-order / dict order, total_size reduced on the device,
    TreeNode attributes through rm_node_info / rm_node_children."""
    dc.run_flatten_order({}, seed)


@pytest.mark.parametrize("seed", range(6))
def test_gpu_paged_batch_entry_points(seed):
    """a2 on the GPU: page interning, value rows and the *_batch entry points with page_size 2..4 against the oracle."""
    dc.run_paged_batch({}, seed)


@pytest.mark.parametrize("seed", range(4))
def test_gpu_paged_keys_with_a_short_last_page(seed):
    """Keys that are not a whole number of pages through the single-key surface (the reference stores the short page in
    the leaf): insert / match / lock / evict / sizes / TreeNode reads against the oracle's paged RadixCache."""
    dc.run_paged_ragged({}, seed)


@pytest.mark.parametrize("seed,page", [(0, 1), (2, 2)])
def test_gpu_arena_of_evicted_runs_is_recycled(seed, page):
    """RadixCache compacts by itself when evicted runs hold a quarter of a nearly full arena (40 rounds of fresh keys
    through an arena of 2048 positions), against the oracle."""
    dc.run_arena_recycling({}, seed, page)


@pytest.mark.parametrize("seed", range(2))
def test_gpu_paged_short_pages_small_table_small_arena_together(seed):
    """Short last pages, a 160-symbol interning table, a 600-position arena, locks and evictions in one run."""
    dc.run_paged_mixed_pressure({}, seed, steps=150)


@pytest.mark.parametrize("seed", range(3))
def test_gpu_paged_random_traffic_through_a_small_symbol_table(seed):
    """Random paged inserts / matches / evictions with room for 256 page symbols: pages are dropped (rm_pages_retain)
    and interned again all the time (~17 reclaims per run), against the oracle."""
    dc.run_paged_batch({}, 300 + seed, vocab=12, max_page_symbols=256, rounds=40)


@pytest.mark.parametrize("seed", range(3))
def test_gpu_paged_symbols_of_evicted_pages_are_reused(seed):
    """rm_pages_retain on the GPU: a stream of fresh pages through a 192-symbol table, against the oracle."""
    dc.run_paged_symbol_reclaim({}, seed)


@pytest.mark.parametrize("seed", range(4))
def test_gpu_paged_random_against_oracle(seed):
    dc.run_flatten_order({}, 50 + seed, page=2)
