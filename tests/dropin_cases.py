"""Golden replays through the drop-in classes (RadixCache / RadixMesh / CacheAwareRouter): the same
scenarios the reference's own tests run (test/correctness.py) plus recorded op sequences, with the
results the unmodified reference produced (tests/golden/*.json)."""
import numpy as np
import pytest
import torch

from radixmesh_b200 import CacheAwareRouter, RadixCache, RadixMesh, ServerArgs
from tests.golden_util import load

P_ADDRS = [f"localhost:{50000 + i}" for i in range(3)]
D_ADDRS = [f"localhost:{50003 + i}" for i in range(2)]
R_ADDRS = ["localhost:50010"]


class LoopComm:
    """In-process stand-in for the reference's TcpCommunicator: delivers to the next hop synchronously."""

    def __init__(self):
        self.peer_cb = None
        self.cb = None

    def send(self, oplog):
        if self.peer_cb is not None:
            import copy
            self.peer_cb(copy.deepcopy(oplog))
        return 1

    def register_rcv_callback(self, fn):
        self.cb = fn

    def is_ordered(self):
        return True

    def target_address(self):
        return "loop"


class FakeAllocator:
    def __init__(self):
        self.device = torch.device("cpu")
        self.freed = []

    def free(self, t):
        self.freed.append([int(x) for x in t.tolist()])


def build_mesh(kw, with_allocators=False):
    """3 prefill + 2 decode in a ring, 1 router fed by rank 0 (policy/sync_algo.py:57-75)."""
    ring = [LoopComm() for _ in range(5)]
    to_router, router_in = LoopComm(), LoopComm()
    nodes = []
    for rank in range(6):
        args = ServerArgs(prefill_cache_nodes=P_ADDRS, decode_cache_nodes=D_ADDRS, router_cache_nodes=R_ADDRS)
        if rank < 3:
            args.prefill_node_rank = rank
        elif rank < 5:
            args.decode_node_rank = rank
        else:
            args.router_node_rank = rank
        small = dict(max_nodes=1 << 12, max_tokens=1 << 15, max_batch=64, max_key_len=64, max_dups=512)
        small.update(kw)
        if rank < 5:
            nodes.append(RadixMesh(args, communicator=ring[rank], routers=[to_router] if rank == 0 else None,
                                   token_to_kv_pool_allocator=FakeAllocator() if with_allocators else None, **small))
        else:
            nodes.append(RadixMesh(args, communicator=router_in, **small))
    for i in range(5):
        ring[i].peer_cb = ring[(i + 1) % 5].cb
    to_router.peer_cb = router_in.cb
    return nodes


def canon(mesh):
    d = mesh._dump()
    rows = []
    for x in range(1, d["n"]):
        if d["parent"][x] >= 0 and d["gen"][x] >= 0:
            from radixmesh_b200 import TreeNode
            h = TreeNode(mesh, x, -1)
            v = mesh._node_value(x)
            raw = [int(t) for t in v.value.tolist()] if hasattr(v, "value") and v.value is not None else []
            rows.append([mesh.full_key(h), int(d["edge_len"][x]), raw, int(d["rank"][x]), int(d["lock_ref"][x])])
    rows.sort()
    dup = sorted([[list(k.key), k.node_rank, None if n is None else n.lock_ref] for k, n in mesh.dup_nodes.items()])
    return {"nodes": rows, "evictable": mesh.evictable_size(), "protected": mesh.protected_size(), "dup": dup}


def replay_correctness_block(block, kw):
    nodes = build_mesh(kw)
    router = CacheAwareRouter(nodes[5], skip_warm_up=True)
    for r, k, v in block["inserts"]:
        nodes[r].insert(k, torch.tensor(v))
    for row in block["expect"]:
        for r, idx in row["pd"].items():
            res = nodes[int(r)].match_prefix(row["key"])
            assert res.device_indices.tolist() == idx            # test/correctness.py:52-55, 86-89, 158-162
        res = nodes[5].match_prefix(row["key"])
        assert [res.prefill_node_rank, res.decode_node_rank] == row["router"]   # :56-58, 90-93
        rr = router.cache_aware_route(row["key"])
        assert [rr.prefill_addr, rr.decode_addr] == row["route"]                # :66-70, 100-103
    if "state" in block:
        for r, st in block["state"].items():
            assert canon(nodes[int(r)]) == st, f"rank {r}"


def replay_mesh_sequence(case, kw):
    nodes = build_mesh(kw)
    router = CacheAwareRouter(nodes[5], skip_warm_up=True)
    held = []
    for op in case["ops"]:
        if op["op"] == "ins":
            assert nodes[op["rank"]].insert(op["key"], torch.tensor(op["val"])) == op["ret"]
        elif op["op"] == "match":
            m = nodes[op["rank"]].match_prefix(op["key"])
            assert m.device_indices.tolist() == op["idx"]
            assert nodes[op["rank"]].full_key(m.last_device_node) == op["last"]
            held.append((op["rank"], m.last_device_node))
        elif op["op"] == "route":
            if op.get("error") == "AssertionError":
                with pytest.raises(AssertionError):
                    nodes[5].match_prefix(op["key"])
                continue
            m = nodes[5].match_prefix(op["key"])
            assert [m.prefill_node_rank, m.decode_node_rank] == op["ranks"]
            if op.get("error") == "IndexError":
                with pytest.raises(IndexError):
                    router.cache_aware_route(op["key"])
            else:
                rr = router.cache_aware_route(op["key"])
                assert [rr.prefill_addr, rr.decode_addr] == op["addrs"]
        elif op["op"] == "lock":
            r, n = held[op["h"]]
            assert nodes[r].inc_lock_ref(n) == op["ret"]
        elif op["op"] == "unlock":
            r, n = held[op["h"]]
            assert nodes[r].dec_lock_ref(n) == op["ret"]
    for r, st in case["state"].items():
        assert canon(nodes[int(r)]) == st, f"rank {r}"
    for r, cands in case["gc_candidates"].items():
        assert sorted([list(k.key), k.node_rank] for k in nodes[int(r)].gc_candidates()) == cands


def replay_base_sequence(case, kw):
    alloc = FakeAllocator()
    small = dict(max_nodes=1 << 12, max_tokens=1 << 15, max_batch=64, max_key_len=64)
    small.update(kw)
    t = RadixCache(None, alloc, int(case.get("page", 1)), **small)
    held = []
    for op in case["ops"]:
        if op["op"] == "ins":
            assert t.insert(list(op["key"]), torch.tensor(op["val"], dtype=torch.int64)) == op["ret"]
        elif op["op"] == "match":
            m = t.match_prefix(list(op["key"]))
            assert m.device_indices.tolist() == op["idx"]
            assert t.full_key(m.last_device_node) == op["last"]
            held.append(m.last_device_node)
        elif op["op"] == "lock":
            assert t.inc_lock_ref(held[op["h"]]) == op["ret"]
            held.append(held[op["h"]])
        elif op["op"] == "unlock":
            assert t.dec_lock_ref(held[op["h"]]) == op["ret"]
        elif op["op"] == "evict":
            before = len(alloc.freed)
            t.evict(op["n"])
            assert alloc.freed[before:] == op["freed"]
        st = op["state"]
        assert t.evictable_size() == st["evictable"] and t.protected_size() == st["protected"]
        d = t._dump()
        from radixmesh_b200 import TreeNode
        rows = sorted([t.full_key(TreeNode(t, x, -1)), len(t._node_tokens(x)), t._node_value(x).tolist(), -1,
                       int(d["lock_ref"][x])] for x in range(1, d["n"]) if d["parent"][x] >= 0)
        assert rows == [r[:5] for r in st["nodes"]]


def golden(name):
    return load(name)


def run_gc_ring(kw):
    """GC_QUERY / GC_EXEC through RadixMesh.oplog_received on a 3P+2D+1R mesh (radix_mesh.py:148-166, 362-389): a parked
    duplicate that every replica has unlocked is freed at the node that asked and popped everywhere; one that a
    single replica still locks survives everywhere."""
    import torch
    nodes = build_mesh(kw, with_allocators=True)
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
    k0 = _key(keys[0], 1)
    k1 = _key(keys[1], 1)
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


def _key(key, rank):
    from radixmesh_b200.host import NodeKey
    return NodeKey(tuple(key), rank)


def run_reset_drops_staged(kw):
    """An INSERT oplog followed by a RESET with no read in between: the reference has applied the insert when the
    reset arrives, so nothing of it survives (radix_mesh.py:411-420); the staged record must go with the tree."""
    from radixmesh_b200.host import CacheOplog, CacheOplogType
    nodes = build_mesh(kw)
    n2 = nodes[2]
    n2.oplog_received(CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=1, key=[1, 2, 3], value=[10, 11, 12], ttl=5))
    n2.oplog_received(CacheOplog(oplog_type=CacheOplogType.RESET, node_rank=1, ttl=5))
    assert n2.match_prefix([1, 2, 3]).device_indices.tolist() == []
    assert n2.evictable_size() == 0
    # the same through a local reset()
    n2.oplog_received(CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=1, key=[4, 5], value=[40, 41], ttl=5))
    n2.reset()
    assert n2.match_prefix([4, 5]).device_indices.tolist() == []
    # and the tree still works afterwards
    assert n2.insert([4, 5], torch.tensor([70, 71])) == 0
    assert n2.match_prefix([4, 5, 6]).device_indices.tolist() == [70, 71]


def run_tick_ring(kw):
    """A TICK is counted and forwarded while its ttl lasts (_tick_handle, radix_mesh.py:356-360): with the ticker's
    two-round ttl 2*(P+D) every replica of the ring counts the origin twice (the _wait_all_nodes_ready condition)."""
    from radixmesh_b200.host import CacheOplog, CacheOplogType
    nodes = build_mesh(kw)
    ttl = 2 * 5
    nodes[3]._send(CacheOplog(oplog_type=CacheOplogType.TICK, node_rank=3, ttl=ttl))
    for r in range(5):
        assert nodes[r].tick_received.get(3, 0) == 2, f"replica {r} saw {nodes[r].tick_received}"
    # rank 0 also feeds the router, once per pass through rank 0's _send (radix_mesh.py:344-347)
    assert nodes[5].tick_received.get(3, 0) >= 1


def run_insert_returns_own_prefix(kw):
    """insert() returns the prefix length of the caller's own record although remote records were staged ahead of it
    (they are applied first, in arrival order: radix_mesh.py:197-201)."""
    from radixmesh_b200.host import CacheOplog, CacheOplogType
    nodes = build_mesh(kw)
    n1 = nodes[1]
    n1.oplog_received(CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=0, key=[1, 2, 3, 4], value=[1, 2, 3, 4], ttl=5))
    n1.oplog_received(CacheOplog(oplog_type=CacheOplogType.INSERT, node_rank=0, key=[9], value=[9], ttl=5))
    assert n1.insert([1, 2, 3, 4, 5, 6], torch.tensor([1, 2, 3, 4, 55, 66])) == 4
    assert n1.insert([7, 7], torch.tensor([5, 6])) == 0


def run_flatten_order(kw, seed, page=1):
    """all_values_flatten in the reference's order (pre-order, children in dict insertion order, radix_cache.py:239-248),
    total_size on the device, and TreeNode.children / .parent / .lock_ref through rm_node_info / rm_node_children,
    against the oracle under random insert / match (splitting) / lock / evict streams."""
    import random
    from oracle.radix_port import PortRadixCache
    rnd = random.Random(7000 + seed)
    alloc = FakeAllocator()
    small = dict(max_nodes=1 << 12, max_tokens=1 << 15, max_batch=64, max_key_len=64)
    small.update(kw)
    t = RadixCache(None, alloc, page, **small)
    orc = PortRadixCache(page)
    vc = 0
    held = []
    for step in range(70):
        kind = rnd.choice(["ins", "ins", "ins", "match", "match", "evict", "lock"])
        key = [rnd.randrange(4) for _ in range(rnd.randrange(0, 16))]
        if kind == "ins":
            key = key[:len(key) // page * page]
            val = list(range(vc, vc + len(key)))
            vc += len(key)
            assert t.insert(key, torch.tensor(val, dtype=torch.int64)) == orc.insert(list(key), np.asarray(val, dtype=np.int64))
        elif kind == "match":
            a, b = t.match_prefix(key), orc.match_prefix(list(key))
            assert a.device_indices.tolist() == b.indices.tolist()
            held.append((a.last_device_node, b.last_node))
        elif kind == "lock" and held:
            a, b = rnd.choice(held)
            if b is not orc.root and b.parent is not None and (b.key is not None):
                assert t.inc_lock_ref(a) == orc.inc_lock_ref(b)
                assert a.lock_ref == b.lock_ref
                assert t.dec_lock_ref(a) == orc.dec_lock_ref(b)
        elif kind == "evict":
            n = rnd.randrange(0, 12)
            t.evict(n)
            orc.evict(n)
            held = [(a, b) for a, b in held if b is orc.root or _alive(orc, b)]
        assert t.total_size() == orc.total_size()
        assert t.all_values_flatten().tolist() == orc.all_values_flatten().tolist(), f"step {step} ({kind})"
        # children of the root and of one random node on each side, in dict order
        assert [k if page == 1 else tuple(k) for k in t.root_node.children.keys()] == list(orc.root.children.keys())
        if held:
            a, b = held[-1]
            if b is orc.root or b.parent is not None:
                live = b is orc.root or _alive(orc, b)
                if live:
                    assert list(a.children.keys()) == list(b.children.keys())
                    if b is not orc.root:
                        assert t.full_key(a.parent) == orc.full_key(b.parent)


def _alive(orc, node):
    while node is not orc.root:
        p = node.parent
        if p is None or all(c is not node for c in p.children.values()):
            return False
        node = p
    return True


def run_paged_batch(kw, seed, vocab=3, max_page_symbols=None, rounds=12, evict_max=20):
    """The *_batch additions with page_size > 1: pages interned on the device (rm_pages_intern), values as rows of a
    device row store, the index running on page symbols - against the oracle's paged RadixCache, record by record."""
    import random
    from oracle.radix_port import PortRadixCache
    rnd = random.Random(8100 + seed)
    page = rnd.choice([2, 3, 4])
    alloc = FakeAllocator()
    small = dict(max_nodes=1 << 12, max_tokens=1 << 15, max_batch=64, max_key_len=64, max_page_symbols=max_page_symbols)
    small.update(kw)
    t = RadixCache(None, alloc, page, **small)
    be = t._be
    orc = PortRadixCache(page)
    vc = 0
    for rd in range(rounds):
        keys = []
        for _ in range(rnd.randrange(1, 12)):
            k = [rnd.randrange(vocab) for _ in range(rnd.randrange(0, 24))]
            keys.append(k[:len(k) // page * page])
        vals = []
        for k in keys:
            vals.append(list(range(vc, vc + len(k))))
            vc += len(k)
        offs = np.zeros(len(keys) + 1, dtype=np.int64)
        np.cumsum([len(k) for k in keys], out=offs[1:])
        flat = np.asarray([x for k in keys for x in k] or [0], dtype=np.int32)
        fv = np.asarray([x for v in vals for x in v] or [0], dtype=np.int64)
        got = be.to_host(t.insert_batch(be.to_device(flat), be.to_device(fv), be.to_device(offs))).tolist()
        assert got == [orc.insert(list(k), np.asarray(v, dtype=np.int64)) for k, v in zip(keys, vals)]
        qs = [[rnd.randrange(vocab) for _ in range(rnd.randrange(1, 26))] for _ in range(rnd.randrange(1, 16))]
        if vocab > 3:
            qs += [list(k) + [rnd.randrange(vocab)] * rnd.randrange(0, 5) for k in keys[:4]]   # queries that do match
        qo = np.zeros(len(qs) + 1, dtype=np.int64)
        np.cumsum([len(q) for q in qs], out=qo[1:])
        qf = np.asarray([x for q in qs for x in q], dtype=np.int32)
        lens, nodes = t.match_prefix_batch(be.to_device(qf), be.to_device(qo))
        exp = [orc.match_prefix(list(q)) for q in qs]
        hl = be.to_host(lens)
        assert hl.tolist() == [len(m.indices) for m in exp]
        vo = np.zeros(len(qs) + 1, dtype=np.int64)
        np.cumsum(hl, out=vo[1:])
        out = be.empty(int(vo[-1]) + 1, np.int64)
        t.gather_values_batch(nodes, lens, be.to_device(vo), out)
        hv = be.to_host(out)
        hn = be.to_host(nodes)
        for i, m in enumerate(exp):
            assert hv[vo[i]:vo[i + 1]].tolist() == m.indices.tolist()
            assert t.full_key(t._handle(int(hn[i]))) == orc.full_key(m.last_node)
        if rd % 3 == 2 or max_page_symbols:
            # (a small symbol table: evict down to a few dozen tokens, so that the live pages always fit it)
            n = max(orc.evictable - rnd.randrange(0, 40), 0) if max_page_symbols else rnd.randrange(0, evict_max)
            b0, b1 = len(alloc.freed), len(orc.freed)
            t.evict(n)
            orc.evict(n)
            assert alloc.freed[b0:] == orc.freed[b1:]
        assert (t.evictable_size(), t.total_size()) == (orc.evictable, orc.total_size())
    # every row the tree does not hold is back in the ring: rows in use = pages in the tree
    syms, free_rows, _, _ = t._pg.status()
    assert free_rows + t.total_size() // page == t._pg_rows_total


def run_paged_symbol_reclaim(kw, seed):
    """A stream of fresh pages through a small interning table: every round inserts keys made of pages nobody has seen,
    then evicts most of the tree.  The pages of evicted nodes must give their symbols back (rm_pages_retain), or the
    table (192 symbols here, ~60 new pages per round) is full after three rounds; results against the oracle's paged
    RadixCache throughout, including keys that survive several reclaims."""
    import random
    from oracle.radix_port import PortRadixCache
    rnd = random.Random(8300 + seed)
    page = rnd.choice([2, 4])
    alloc = FakeAllocator()
    small = dict(max_nodes=1 << 12, max_tokens=1 << 15, max_batch=64, max_key_len=64, max_page_symbols=192)
    small.update(kw)
    t = RadixCache(None, alloc, page, **small)
    be = t._be
    orc = PortRadixCache(page)
    vc = 0
    fresh = 10
    keep = [[1, 2, 3, 4, 5, 6, 7, 8][:2 * page]]      # a key re-inserted every round: it stays live
    retains = 0
    for rd in range(16):
        keys = [list(keep[0])]
        for _ in range(6):
            n_pages = rnd.randrange(4, 12)
            k = list(range(fresh, fresh + n_pages * page))       # tokens nobody has used: all pages new
            fresh += n_pages * page
            keys.append(k)
        if rd and rnd.random() < 0.5:
            keys.append(keys[rnd.randrange(1, len(keys))][:2 * page] + [7] * page)   # shares a prefix with a fresh key
        vals = []
        for k in keys:
            vals.append(list(range(vc, vc + len(k))))
            vc += len(k)
        offs = np.zeros(len(keys) + 1, dtype=np.int64)
        np.cumsum([len(k) for k in keys], out=offs[1:])
        flat = np.asarray([x for k in keys for x in k], dtype=np.int32)
        fv = np.asarray([x for v in vals for x in v], dtype=np.int64)
        before = t._pg_syms_after_retain
        got = be.to_host(t.insert_batch(be.to_device(flat), be.to_device(fv), be.to_device(offs))).tolist()
        retains += t._pg_syms_after_retain != before
        assert got == [orc.insert(list(k), np.asarray(v, dtype=np.int64)) for k, v in zip(keys, vals)]
        for q in keys[:3] + [keys[-1][:page] + [5] * page]:
            m = t.match_prefix(list(q))
            e = orc.match_prefix(list(q))
            assert m.device_indices.tolist() == e.indices.tolist()
        # evict everything but what the kept key needs (it was just stamped, so it goes last)
        n = max(orc.evictable - 2 * page, 0)
        b0, b1 = len(alloc.freed), len(orc.freed)
        t.evict(n)
        orc.evict(n)
        assert alloc.freed[b0:] == orc.freed[b1:]
        assert (t.evictable_size(), t.total_size()) == (orc.evictable, orc.total_size())
    assert retains >= 2, retains
    syms, free_rows, _, _ = t._pg.status()
    assert syms <= 192 and free_rows + t.total_size() // page == t._pg_rows_total


def run_paged_ragged(kw, seed):
    """RadixCache(page_size > 1) with keys of ANY length through the reference's own call shapes: the reference stores
    a short last page in the leaf (child key = the short tuple, a matched short page counted as a whole one by
    _key_match_paged).  Against the oracle's paged RadixCache, which tests/test_oracle_vs_reference.py pins on ragged
    keys: insert return values, match results and last nodes, lock deltas, freed indices of every evict, the three
    size counters, the values in tree order, and keys / child keys of the nodes."""
    import random
    from oracle.radix_port import PortRadixCache
    rnd = random.Random(8500 + seed)
    page = rnd.choice([2, 3, 4])
    alloc = FakeAllocator()
    small = dict(max_nodes=1 << 12, max_tokens=1 << 15, max_batch=64, max_key_len=64)
    small.update(kw)
    t = RadixCache(None, alloc, page, **small)
    orc = PortRadixCache(page)
    vc = 0
    held = []
    for step in range(70):
        op = rnd.choice(["ins", "ins", "ins", "match", "lock", "unlock", "evict"])
        if op == "ins":
            k = [rnd.randrange(2) for _ in range(rnd.randrange(0, 14))]
            v = list(range(vc, vc + len(k)))
            vc += len(k)
            assert t.insert(list(k), np.asarray(v, dtype=np.int64)) == orc.insert(list(k), np.asarray(v, dtype=np.int64)), (step, k)
        elif op == "match":
            q = [rnd.randrange(2) for _ in range(rnd.randrange(0, 14))]
            m, e = t.match_prefix(list(q)), orc.match_prefix(list(q))
            assert m.device_indices.tolist() == e.indices.tolist()
            assert t.full_key(m.last_device_node) == orc.full_key(e.last_node)
            held.append((m.last_device_node, e.last_node))
        elif op == "lock" and held:
            a, b = held[rnd.randrange(len(held))]
            assert t.inc_lock_ref(a) == orc.inc_lock_ref(b)
            held.append((a, b))
        elif op == "unlock" and held:
            a, b = held[rnd.randrange(len(held))]
            if b.lock_ref > 0 and b is not orc.root:
                assert t.dec_lock_ref(a) == orc.dec_lock_ref(b)
        elif op == "evict":
            n = rnd.randrange(0, 12)
            b0, b1 = len(alloc.freed), len(orc.freed)
            t.evict(n)
            orc.evict(n)
            assert alloc.freed[b0:] == orc.freed[b1:], (step, n)
            held = [(a, b) for a, b in held if b is orc.root or b.parent is not None and _alive(orc, b)]
        assert (t.evictable_size(), t.protected_size(), t.total_size()) == (orc.evictable, orc.protected, orc.total_size()), step
    if orc.total_size():
        assert t.all_values_flatten().tolist() == orc.all_values_flatten().tolist()
    # keys and child keys of the whole tree, short pages included
    def walk(a, b):
        assert a.key == list(b.key)
        ka, kb = a.children, b.children
        assert list(ka.keys()) == list(kb.keys())
        for key in kb:
            walk(ka[key], kb[key])
    walk(t.root_node, orc.root)


def _alive(orc, node):
    """node is still reachable from the oracle's root (evicted nodes keep their parent pointer)."""
    while node is not orc.root:
        p = node.parent
        if p is None or p.children.get(orc._ckey(node.key)) is not node:
            return False
        node = p
    return True


def run_arena_recycling(kw, seed, page=1):
    """A RadixCache that streams inserts and evictions through an arena far smaller than what it is handed in total:
    the reference gives memory back when a node is evicted; here evicted runs come back through rm_compact, which the
    drop-in calls by itself (RadixCache._arena_room).  40 rounds x ~600 fresh tokens through an arena of 2048
    positions, against the oracle (matches, freed indices, sizes); handles taken before a compaction stay valid."""
    import random
    from oracle.radix_port import PortRadixCache
    rnd = random.Random(8700 + seed)
    alloc = FakeAllocator()
    small = dict(max_nodes=1 << 12, max_tokens=2048 * page, max_batch=64, max_key_len=256)
    small.update(kw)
    t = RadixCache(None, alloc, page, **small)
    orc = PortRadixCache(page)
    vc, fresh = 0, 100
    keep = list(range(1, 1 + 8 * page))
    compactions = 0
    held = None
    for rd in range(40):
        keys = [list(keep)]
        for _ in range(5):
            n = rnd.randrange(40, 200) // page * page
            keys.append(list(range(fresh, fresh + n)))
            fresh += n
        before = t._arena_est
        for k in keys:
            v = list(range(vc, vc + len(k)))
            vc += len(k)
            assert t.insert(list(k), np.asarray(v, dtype=np.int64)) == orc.insert(list(k), np.asarray(v, dtype=np.int64))
        compactions += t._arena_est < before
        m, e = t.match_prefix(list(keep) + [7] * page), orc.match_prefix(list(keep) + [7] * page)
        assert m.device_indices.tolist() == e.indices.tolist()
        if held is None:
            held = (m.last_device_node, e.last_node)
        assert held[0].key == list(held[1].key) and held[0].value.tolist() == held[1].value.tolist()
        n = max(orc.evictable - 8 * page, 0)
        b0, b1 = len(alloc.freed), len(orc.freed)
        t.evict(n)
        orc.evict(n)
        assert alloc.freed[b0:] == orc.freed[b1:]
        assert (t.evictable_size(), t.total_size()) == (orc.evictable, orc.total_size())
    assert compactions >= 3, compactions



def run_paged_mixed_pressure(kw, seed, steps=250):
    """Everything the paged adapter keeps books on, at once and under pressure: keys of any length (short last pages),
    a 160-symbol interning table (pages dropped and interned again), an arena of 600 positions (the cache compacts by
    itself), locks, and evictions of up to 120 tokens - against the oracle's paged RadixCache after every step."""
    import random
    from oracle.radix_port import PortRadixCache
    rnd = random.Random(9100 + seed)
    page = rnd.choice([2, 3, 4])
    alloc = FakeAllocator()
    small = dict(max_nodes=1 << 12, max_tokens=600 * page, max_batch=64, max_key_len=64, max_page_symbols=160)
    small.update(kw)
    t = RadixCache(None, alloc, page, **small)
    orc = PortRadixCache(page)
    vc, held = 0, []
    for step in range(steps):
        op = rnd.choice(["ins", "ins", "ins", "match", "lock", "unlock", "evict", "evict"])
        if op == "ins":
            k = [rnd.randrange(9) for _ in range(rnd.randrange(0, 30))]
            v = list(range(vc, vc + len(k)))
            vc += len(k)
            assert t.insert(list(k), np.asarray(v, dtype=np.int64)) == orc.insert(list(k), np.asarray(v, dtype=np.int64)), (step, k)
        elif op == "match":
            q = [rnd.randrange(9) for _ in range(rnd.randrange(0, 30))]
            m, e = t.match_prefix(list(q)), orc.match_prefix(list(q))
            assert m.device_indices.tolist() == e.indices.tolist()
            held.append((m.last_device_node, e.last_node))
        elif op == "lock" and held:
            a, b = held[rnd.randrange(len(held))]
            if b is orc.root or _alive(orc, b):
                assert t.inc_lock_ref(a) == orc.inc_lock_ref(b)
                held.append((a, b))
        elif op == "unlock" and held:
            a, b = held[rnd.randrange(len(held))]
            if b.lock_ref > 0 and b is not orc.root and _alive(orc, b):
                assert t.dec_lock_ref(a) == orc.dec_lock_ref(b)
        elif op == "evict":
            n = rnd.randrange(0, 120)
            b0, b1 = len(alloc.freed), len(orc.freed)
            t.evict(n)
            orc.evict(n)
            assert alloc.freed[b0:] == orc.freed[b1:], (step, n)
            held = [(a, b) for a, b in held if b is orc.root or _alive(orc, b)]
        assert (t.evictable_size(), t.protected_size(), t.total_size()) == (orc.evictable, orc.protected, orc.total_size()), step
    if orc.total_size():
        assert t.all_values_flatten().tolist() == orc.all_values_flatten().tolist()
