"""Generate tests/golden/*.json by running the UNMODIFIED reference (build container only).

    python tools/make_golden.py

Every fixture is a recorded op sequence plus what the reference returned / ended
up with (canonical tree dump).  Tests replay the ops on the CPU
oracle (tests/test_oracle_golden.py, any box) and on the CUDA path
(tests/test_gpu_golden.py, B200 box) and require identical results.
"""
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def canon_json(c):
    d = {"nodes": [list(map(lambda x: list(x) if isinstance(x, tuple) else x, row)) for row in c["nodes"]],
         "evictable": c["evictable"], "protected": c["protected"]}
    if "dup" in c:
        d["dup"] = sorted([[list(k[0]), k[1], v] for k, v in c["dup"].items()])
    return d


def base_sequences():
    """Random op sequences on RadixCache (page_size 1): insert / match / lock / unlock / evict."""
    out = []
    for seed in range(12):
        rnd = random.Random(seed)
        vocab = rnd.choice([2, 3, 4, 50])
        maxlen = rnd.choice([6, 10, 24])
        ref = ref_harness.RefCache(1)
        held, ops, vc = [], [], 0
        for _ in range(rnd.randrange(20, 70)):
            kind = rnd.choice(["ins", "ins", "match", "match", "lock", "unlock", "evict"])
            key = [rnd.randrange(vocab) for _ in range(rnd.randrange(0, maxlen))]
            if kind == "ins":
                val = list(range(vc, vc + len(key)))
                vc += len(key)
                ops.append({"op": "ins", "key": key, "val": val, "ret": ref.insert(key, val)})
            elif kind == "match":
                r = ref.match_prefix(key)
                held.append(r.last_device_node)
                ops.append({"op": "match", "key": key, "idx": r.device_indices.tolist(),
                            "last": ref.full_key(r.last_device_node)})
            elif kind == "lock" and held:
                i = rnd.randrange(len(held))
                ops.append({"op": "lock", "h": i, "ret": ref.tree.inc_lock_ref(held[i])})
                held.append(held[i])
            elif kind == "unlock" and held:
                i = rnd.randrange(len(held))
                if held[i].lock_ref > 0 and held[i] is not ref.tree.root_node:
                    ops.append({"op": "unlock", "h": i, "ret": ref.tree.dec_lock_ref(held[i])})
            elif kind == "evict":
                n = rnd.randrange(0, 12)
                before = len(ref.alloc.freed)
                ref.evict(n)
                ops.append({"op": "evict", "n": n, "freed": ref.alloc.freed[before:]})
            else:
                continue
            ops[-1]["state"] = canon_json(ref.canonical())
        out.append({"seed": seed, "ops": ops})
    return out


def paged_sequences(ragged=False):
    """The same kind of sequences on RadixCache(page_size 2 / 4): keys matched a page at a time, page-aligned
    inserts (what the reference's callers do), radix_cache.py:23-32, 109-114, 148-150.  ragged: inserted keys of any
    length (page sizes 2 / 3 / 4) - the reference then keeps a short last page in the leaf."""
    out = []
    for seed in range(8):
        rnd = random.Random((900 if ragged else 700) + seed)
        page = rnd.choice([2, 3, 4] if ragged else [2, 4])
        vocab = rnd.choice([2, 3, 5])
        maxlen = rnd.choice([9, 17, 26])
        ref = ref_harness.RefCache(page)
        held, ops, vc = [], [], 0

        def alive(node):
            # a handle to a node that has been evicted since is outside the contract (the reference would still
            # count on the detached object; the device has recycled the id)
            while node is not ref.tree.root_node:
                par = node.parent
                if par is None or par.children.get(ref.tree.get_child_key_fn(node.key)) is not node:
                    return False
                node = par
            return True

        for _ in range(rnd.randrange(25, 70)):
            kind = rnd.choice(["ins", "ins", "match", "match", "lock", "unlock", "evict"])
            key = [rnd.randrange(vocab) for _ in range(rnd.randrange(0, maxlen))]
            if kind == "ins":
                if not ragged:
                    key = key[:len(key) // page * page]
                val = list(range(vc, vc + len(key)))
                vc += len(key)
                ops.append({"op": "ins", "key": key, "val": val, "ret": ref.insert(key, val)})
            elif kind == "match":
                r = ref.match_prefix(key)
                held.append(r.last_device_node)
                ops.append({"op": "match", "key": key, "idx": r.device_indices.tolist(),
                            "last": ref.full_key(r.last_device_node)})
            elif kind == "lock" and held:
                i = rnd.randrange(len(held))
                if not alive(held[i]):
                    continue
                ops.append({"op": "lock", "h": i, "ret": ref.tree.inc_lock_ref(held[i])})
                held.append(held[i])
            elif kind == "unlock" and held:
                i = rnd.randrange(len(held))
                if alive(held[i]) and held[i].lock_ref > 0 and held[i] is not ref.tree.root_node:
                    ops.append({"op": "unlock", "h": i, "ret": ref.tree.dec_lock_ref(held[i])})
            elif kind == "evict":
                n = rnd.randrange(0, 14)
                before = len(ref.alloc.freed)
                ref.evict(n)
                ops.append({"op": "evict", "n": n, "freed": ref.alloc.freed[before:]})
            else:
                continue
            ops[-1]["state"] = canon_json(ref.canonical())
        out.append({"seed": seed, "page": page, "ops": ops})
    return out


def correctness_blocks():
    """The four known-answer blocks of the reference's own test/correctness.py (:41-70, :77-103,
    :147-174, :178-211) replayed on the in-process 3P+2D+1R mesh."""
    blocks = []

    def snap(mesh, queries, ranks):
        res = []
        router = mesh.router()
        for q in queries:
            row = {"key": q, "pd": {}, "router": None, "route": None}
            for r in ranks:
                m = mesh.match(r, q)
                row["pd"][str(r)] = m.device_indices.tolist()
            m = mesh.match(5, q)
            row["router"] = [m.prefill_node_rank, m.decode_node_rank]
            rr = router.cache_aware_route(list(q))
            row["route"] = [rr.prefill_addr, rr.decode_addr]
            res.append(row)
        return res

    # sync_and_routing
    rnd = random.Random(1234)
    mesh = ref_harness.RefMesh(3, 2, 1)
    ins = [[1, [1, 2, 3], [10, 20, 30]]]
    for rank in range(3):
        for _ in range(10):
            n = rnd.randint(1, 20)
            ins.append([rank, [rnd.randint(100, 200) for _ in range(n)], [rnd.randint(1, 1000) for _ in range(n)]])
    for r, k, v in ins:
        mesh.insert(r, k, v)
    b1 = {"name": "sync_and_routing_prefill", "inserts": ins,
          "expect": snap(mesh, [[1, 2, 3], [1, 2, 3, 4]], [0, 1, 2])}
    ins2 = [[3, [1, 2, 3, 4, 5, 6], [10, 20, 30, 40, 50, 60]]]
    for r, k, v in ins2:
        mesh.insert(r, k, v)
    b2 = {"name": "sync_and_routing_decode", "inserts": ins + ins2,
          "expect": snap(mesh, [[1, 2, 3, 4, 5, 6], [1, 2, 3, 4, 5, 6, 7]], [0, 1, 2, 3, 4])}
    blocks += [b1, b2]
    assert b1["expect"][1]["router"] == [1, -1] and b2["expect"][1]["router"] == [1, 3]

    # multi_write
    for order in ([0, 1, 2], [2, 1, 0], [1, 2, 0]):
        mesh = ref_harness.RefMesh(3, 2, 1)
        ins = [[r, [1, 2, 3], [10, 20, 30 + r]] for r in order]
        for r, k, v in ins:
            mesh.insert(r, k, v)
        b3 = {"name": f"multi_write_1_order{order}", "inserts": list(ins),
              "expect": snap(mesh, [[1, 2, 3], [1, 2, 3, 4]], [0, 1, 2, 3, 4])}
        assert b3["expect"][0]["pd"]["4"] == [10, 20, 30] and b3["expect"][1]["router"] == [0, -1]
        k2, v2 = [1, 2, 3, 4, 5, 6, 7], [10, 20, 30, 40, 50, 60, 70]
        ins_b = [[2, k2, v2], [1, k2[:-1], v2[:-1]], [0, k2[:-2], v2[:-2]]]
        ins_b = [ins_b[i] for i in order]
        for r, k, v in ins_b:
            mesh.insert(r, k, v)
        b4 = {"name": f"multi_write_2_order{order}", "inserts": ins + ins_b,
              "expect": snap(mesh, [k2, k2[:-1], k2[:-2]], [0, 1, 2, 3, 4]),
              "state": {str(r): canon_json(mesh.canonical(r)) for r in range(6)}}
        blocks += [b3, b4]
    return blocks


def mesh_sequences():
    out = []
    seed = 0
    while len(out) < 10:
        seed += 1
        rnd = random.Random(1000 + seed)
        mesh = ref_harness.RefMesh(3, 2, 1)
        router = mesh.router()
        vocab = rnd.choice([2, 3, 5])
        ops, held, bad = [], [], False
        for _ in range(rnd.randrange(15, 45)):
            kind = rnd.choice(["ins", "ins", "ins", "match", "route", "lock", "unlock"])
            rank = rnd.randrange(5)
            key = [rnd.randrange(vocab) for _ in range(rnd.randrange(1, 8))]
            if kind == "ins":
                variant = rnd.randrange(2)
                val = [100 * (rank + 1) * variant + 10 * i + t for i, t in enumerate(key)]
                try:
                    ret = mesh.insert(rank, key, val)
                except AttributeError:
                    bad = True
                    break
                ops.append({"op": "ins", "rank": rank, "key": key, "val": val, "ret": ret})
            elif kind == "match":
                m = mesh.match(rank, key)
                held.append((rank, m.last_device_node))
                ops.append({"op": "match", "rank": rank, "key": key, "idx": m.device_indices.tolist(),
                            "last": mesh.full_key(rank, m.last_device_node)})
            elif kind == "route":
                try:
                    m = mesh.match(5, key)
                    picked = [m.prefill_node_rank, m.decode_node_rank]
                except AssertionError:
                    ops.append({"op": "route", "key": key, "error": "AssertionError"})
                    continue
                try:
                    rr = router.cache_aware_route(list(key))
                    ops.append({"op": "route", "key": key, "ranks": picked, "addrs": [rr.prefill_addr, rr.decode_addr]})
                except IndexError:
                    ops.append({"op": "route", "key": key, "ranks": picked, "error": "IndexError"})
            elif kind == "lock" and held:
                i = rnd.randrange(len(held))
                ops.append({"op": "lock", "h": i, "ret": mesh.nodes[held[i][0]].inc_lock_ref(held[i][1])})
            elif kind == "unlock" and held:
                i = rnd.randrange(len(held))
                ops.append({"op": "unlock", "h": i, "ret": mesh.nodes[held[i][0]].dec_lock_ref(held[i][1])})
        if bad:
            continue
        state = {str(r): canon_json(mesh.canonical(r)) for r in range(6)}
        gc = {str(r): sorted([list(k.key), k.node_rank] for k, n in mesh.nodes[r].dup_nodes.items()
                             if n is not None and n.lock_ref == 0) for r in range(5)}
        out.append({"seed": seed, "ops": ops, "state": state, "gc_candidates": gc})
    return out


def hash_vectors():
    ns = ref_harness.load()
    rnd = random.Random(99)
    out = []
    for n in (1, 2, 3, 5, 8):
        addrs = [f"10.1.{n}.{i}:{6000 + i}" for i in range(n)]
        h = ns.car.ConsistentHash(nodes=addrs)
        qs = []
        for _ in range(40):
            key = [rnd.randrange(0, 32000) for _ in range(rnd.choice([1, 2, 7, 13, 14, 15, 60, 300]))]
            qs.append({"key": key, "hash": h._hash(key), "node": h.get_node(key)})
        out.append({"addrs": addrs, "ring": h.ring, "queries": qs})
    return out


def oplog_wire():
    """Messages exactly as the reference's JsonSerializer puts them on the wire (communication/serializer.py:20-35),
    with the fields CacheOplog.model_validate_json gives back (radix/cache_oplog.py:48-66)."""
    ref_harness.load()
    from src.communication.serializer import JsonSerializer
    from src.radix.cache_oplog import CacheOplog, CacheOplogType
    rnd = random.Random(7)
    ser = JsonSerializer()
    out = []
    for i in range(60):
        typ = rnd.choice([CacheOplogType.INSERT] * 6 + [CacheOplogType.TICK, CacheOplogType.RESET, CacheOplogType.DELETE])
        n = rnd.choice([0, 1, 2, 7, 64, 300]) if typ == CacheOplogType.INSERT else 0
        key = [rnd.randrange(0, 32000) for _ in range(n)]
        val = [rnd.randrange(0, 1 << 40) for _ in range(n)]
        op = CacheOplog(oplog_type=typ, node_rank=rnd.randrange(0, 6), local_logic_id=rnd.randrange(0, 1 << 30), key=key,
                        value=val, ttl=rnd.randrange(1, 7))
        wire = ser.serialize(op)
        back = ser.deserialize(bytearray(wire), len(wire))
        out.append({"wire": wire.decode("utf-8"),
                    "fields": {"oplog_type": int(back.oplog_type), "node_rank": back.node_rank,
                               "local_logic_id": back.local_logic_id, "key": list(back.key), "value": list(back.value),
                               "ttl": back.ttl}})
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    only = set(sys.argv[1:])
    for name, fn in [("base_sequences", base_sequences), ("correctness_blocks", correctness_blocks),
                     ("mesh_sequences", mesh_sequences), ("consistent_hash", hash_vectors), ("oplog_wire", oplog_wire),
                     ("paged_sequences", paged_sequences),
                     ("paged_ragged_sequences", lambda: paged_sequences(ragged=True))]:
        if only and name not in only:
            continue
        data = fn()
        with open(os.path.join(OUT, name + ".json"), "w") as f:
            json.dump(data, f, separators=(",", ":"))
        print(name, os.path.getsize(os.path.join(OUT, name + ".json")), "bytes")


if __name__ == "__main__":
    main()
