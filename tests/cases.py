"""Op-sequence scenarios shared by the host-emulation tests (CPU) and the GPU parity tests: each one
drives a DeviceIndex through the C ABI and the CPU oracle side by side and requires identical results."""
import random

import numpy as np

from oracle.radix_port import PortRadixCache
from tests.parity import assert_same_tree, device_canonical, pack

def run_sequence(make_index, seed, n_steps=30, vocab=3, maxlen=8, batch=6):
    rnd = random.Random(seed)
    ix = make_index(mode=0, max_nodes=4096, max_tokens=1 << 16, max_batch=64, max_key_len=64)
    orc = PortRadixCache(1)
    held_dev, held_orc = [], []
    vc = 0
    for step in range(n_steps):
        kind = rnd.choice(["ins", "ins", "match", "match", "lock", "unlock", "evict"])
        if kind == "ins":
            keys = [[rnd.randrange(vocab) for _ in range(rnd.randrange(0, maxlen))] for _ in range(rnd.randrange(1, batch))]
            vals = []
            for k in keys:
                vals.append(list(range(vc, vc + len(k))))
                vc += len(k)
            tk, offs = pack(keys)
            vv, _ = pack(vals, np.int64)
            got = ix.insert_host(tk, vv, offs)
            exp = [orc.insert(list(k), np.asarray(v, dtype=np.int64)) for k, v in zip(keys, vals)]
            assert got.tolist() == exp
        elif kind == "match":
            keys = [[rnd.randrange(vocab) for _ in range(rnd.randrange(1, maxlen))] for _ in range(rnd.randrange(1, batch))]
            tk, offs = pack(keys)
            ol = np.zeros(len(keys), np.int32)
            on = np.zeros(len(keys), np.int32)
            ix.match_host(tk, offs, ol, on)
            exp = [orc.match_prefix(list(k)) for k in keys]
            assert ol.tolist() == [len(m.indices) for m in exp]
            # gather values
            oo = np.zeros(len(keys) + 1, np.int64)
            oo[1:] = np.cumsum(ol)
            d_out = ix.be.empty(int(oo[-1]) + 1, np.int64)
            ix.gather(ix.be.to_device(on), ix.be.to_device(ol), ix.be.to_device(oo), len(keys), d_out)
            outv = ix.be.to_host(d_out)
            for i, m in enumerate(exp):
                assert outv[oo[i]:oo[i + 1]].tolist() == m.indices.tolist()
            dev = device_canonical(ix)
            inv = {v: k for k, v in dev["_ids"].items()}
            inv[0] = ()
            for i, m in enumerate(exp):
                assert inv[int(on[i])] == tuple(orc.full_key(m.last_node)), f"last node of query {i}"
                held_dev.append(int(on[i]))
                held_orc.append(m.last_node)
        elif kind == "lock" and held_dev:
            sel = [rnd.randrange(len(held_dev)) for _ in range(rnd.randrange(1, 4))]
            nodes = np.asarray([held_dev[i] for i in sel], np.int32)
            d_out = ix.be.empty(len(sel), np.int64)
            ix.lock(ix.be.to_device(nodes), None, len(sel), +1, d_out)
            out = ix.be.to_host(d_out)
            exp = [orc.inc_lock_ref(held_orc[i]) for i in sel]
            assert out.tolist() == exp
            for i in sel:
                held_dev.append(held_dev[i])
                held_orc.append(held_orc[i])
        elif kind == "unlock" and held_dev:
            sel = [i for i in {rnd.randrange(len(held_dev)) for _ in range(3)}
                   if held_orc[i].lock_ref > 0 and held_orc[i] is not orc.root]
            # keep the oracle's lock_refs non-negative: at most lock_ref decrements per node path
            ok = []
            budget = {}
            for i in sel:
                n = held_orc[i]
                path = []
                while n is not orc.root:
                    path.append(n)
                    n = n.parent
                if all(budget.get(id(p), p.lock_ref) > 0 for p in path):
                    for p in path:
                        budget[id(p)] = budget.get(id(p), p.lock_ref) - 1
                    ok.append(i)
            if ok:
                nodes = np.asarray([held_dev[i] for i in ok], np.int32)
                d_out = ix.be.empty(len(ok), np.int64)
                ix.lock(ix.be.to_device(nodes), None, len(ok), -1, d_out)
                out = ix.be.to_host(d_out)
                exp = [orc.dec_lock_ref(held_orc[i]) for i in ok]
                assert out.tolist() == exp
        elif kind == "evict":
            n = rnd.randrange(0, 12)
            before = len(orc.freed)
            orc.evict(n)
            nodes, offs, lens = ix.evict(n)
            got = [ix.read_arena(int(o), int(l))[1].tolist() for o, l in zip(offs, lens)]
            assert got == orc.freed[before:]
            # evicted handles are dead on both sides
            dead = set(int(x) for x in nodes)
            keep = [i for i in range(len(held_dev)) if held_dev[i] not in dead]
            held_dev = [held_dev[i] for i in keep]
            held_orc = [held_orc[i] for i in keep]
            if rnd.random() < 0.5:
                used = ix.sizes().arena_used
                ix.compact()          # reclaims the evicted runs; nothing observable may change
                assert ix.sizes().arena_used <= used
        ix.check_device_error()
        assert_same_tree(device_canonical(ix), orc.canonical())
    ix.close()



def run_split_evict(make_index, seed):
    """Many queries of one batch split the same long edges; the stamps K3 derives in closed form must
    reproduce the reference's query-by-query stamps, which the full eviction order exposes."""
    rnd = random.Random(seed)
    ix = make_index(mode=0, max_nodes=4096, max_tokens=1 << 16, max_batch=128, max_key_len=64)
    orc = PortRadixCache(1)
    docs = [[rnd.randrange(2) for _ in range(rnd.randrange(10, 40))] for _ in range(rnd.randrange(1, 4))]
    vc = 0
    for rounds in range(rnd.randrange(1, 4)):
        keys = [list(d) for d in docs if rnd.random() < 0.8] or [list(docs[0])]
        vals = []
        for k in keys:
            vals.append(list(range(vc, vc + len(k))))
            vc += len(k)
        tk, offs = pack(keys)
        vv, _ = pack(vals, np.int64)
        got = ix.insert_host(tk, vv, offs)
        assert got.tolist() == [orc.insert(list(k), np.asarray(v, dtype=np.int64)) for k, v in zip(keys, vals)]
        qs = []
        for _ in range(rnd.randrange(5, 100)):
            d = rnd.choice(docs)
            cut = rnd.randrange(1, len(d) + 1)
            qs.append(d[:cut] + [rnd.randrange(2, 4) for _ in range(rnd.randrange(0, 3))])
        tk, offs = pack(qs)
        ol = np.zeros(len(qs), np.int32)
        on = np.zeros(len(qs), np.int32)
        ix.match_host(tk, offs, ol, on)
        exp = [orc.match_prefix(list(k)) for k in qs]
        assert ol.tolist() == [len(m.indices) for m in exp]
        dev = device_canonical(ix)
        inv = {v: k for k, v in dev["_ids"].items()}
        inv[0] = ()
        for i, m in enumerate(exp):
            assert inv[int(on[i])] == tuple(orc.full_key(m.last_node))
        assert_same_tree(dev, orc.canonical())
        # extend some docs so later rounds split fresh edges
        docs = [d + [rnd.randrange(2) for _ in range(rnd.randrange(0, 12))] for d in docs]
    while orc.canonical()["nodes"]:
        n = rnd.randrange(1, 9)
        before = len(orc.freed)
        orc.evict(n)
        _, offs, lens = ix.evict(n)
        got = [ix.read_arena(int(o), int(l))[1].tolist() for o, l in zip(offs, lens)]
        assert got == orc.freed[before:]
        assert_same_tree(device_canonical(ix), orc.canonical())
    ix.check_device_error()
    ix.close()


def run_stream_evict(make_index, seed, n_batches=12, batch=48, L=24, n_docs=12, capacity=1500):
    """BASELINE configs[2] in miniature: streaming insert batches into a pool kept at <= 90 % occupancy by
    evicting a batch's worth of tokens whenever it is exceeded (occupancy is a harness
    notion applied identically to both sides), with arena compaction when the bump arena fills up.
    Evicted runs (in order) and the final tree must equal the oracle's."""
    from tests.parity import zipf_prefix
    ix = make_index(mode=0, max_nodes=4096, max_tokens=2 * capacity, max_batch=max(batch, 64), max_key_len=4 * L)
    orc = PortRadixCache(1)
    docs = None
    vc = 0
    n_evicted = 0
    for b in range(n_batches):
        seqs, docs = zipf_prefix(seed * 100 + b, n_docs, batch, L, vocab=50, docs=docs)
        vals = np.arange(vc, vc + batch * L, dtype=np.int64)
        vc += batch * L
        offs = np.arange(0, (batch + 1) * L, L, dtype=np.int64)
        if ix.sizes().arena_used + batch * L > 2 * capacity:
            ix.compact()
        got = ix.insert_host(seqs.reshape(-1).copy(), vals, offs)
        exp = [orc.insert(r.tolist(), vals[i * L:(i + 1) * L]) for i, r in enumerate(seqs)]
        assert got.tolist() == exp
        ix.check_device_error()
        while orc.evictable_size() + orc.protected_size() > 0.9 * capacity:
            s = ix.sizes()
            assert (s.evictable, s.protected_) == (orc.evictable_size(), orc.protected_size())
            before = len(orc.freed)
            orc.evict(batch * L // 4)
            _, eo, el = ix.evict(batch * L // 4)
            freed = [ix.read_arena(int(o), int(n))[1].tolist() for o, n in zip(eo, el)]
            assert freed == orc.freed[before:]
            n_evicted += len(freed)
            if not freed:
                break
        assert_same_tree(device_canonical(ix), orc.canonical())
    assert n_evicted > 0
    ix.close()


def run_tombstone_pressure(make_index, seed, rounds=60):
    """A small index under a long insert / match (splitting) / evict stream: every eviction tombstones a child-table
    slot and every insert or split takes a free one, so without a rebuild the empty slots of the table run out (the
    256-slot table here sees well over 256 evictions) and every miss probe would walk the whole table.  rm_evict rebuilds
    the table when tombstones hold a quarter of it; results must stay identical to the oracle's throughout and the
    sticky RM_E_TABLE must never be raised."""
    rnd = random.Random(seed)
    ix = make_index(mode=0, max_nodes=128, max_tokens=1 << 14, max_batch=64, max_key_len=64)
    orc = PortRadixCache(1)
    vc = 0
    n_evicted = 0
    max_tombs = 0
    for r in range(rounds):
        keys = [[rnd.randrange(3) for _ in range(rnd.randrange(4, 14))] for _ in range(rnd.randrange(2, 6))]
        vals = []
        for k in keys:
            vals.append(list(range(vc, vc + len(k))))
            vc += len(k)
        if ix.sizes().arena_used + sum(len(k) for k in keys) > (1 << 14) - 256:
            ix.compact()
        tk, offs = pack(keys)
        vv, _ = pack(vals, np.int64)
        got = ix.insert_host(tk, vv, offs)
        assert got.tolist() == [orc.insert(list(k), np.asarray(v, dtype=np.int64)) for k, v in zip(keys, vals)]
        for _ in range(2):      # matches that split, then misses that probe for absent children
            qs = [[rnd.randrange(3) for _ in range(rnd.randrange(1, 12))] + [7] for _ in range(rnd.randrange(2, 10))]
            tk, offs = pack(qs)
            ol = np.zeros(len(qs), np.int32)
            on = np.zeros(len(qs), np.int32)
            ix.match_host(tk, offs, ol, on)
            assert ol.tolist() == [len(orc.match_prefix(list(k)).indices) for k in qs]
        n = orc.evictable_size() - rnd.randrange(0, 12)
        before = len(orc.freed)
        orc.evict(n)
        _, eo, el = ix.evict(n)
        assert [ix.read_arena(int(o), int(l))[1].tolist() for o, l in zip(eo, el)] == orc.freed[before:]
        n_evicted += len(eo)
        ix.check_device_error()
        s = ix.sizes()
        assert s.table_slots == 256 and 0 <= s.table_tombs <= s.table_slots // 4, (s.table_slots, s.table_tombs)
        max_tombs = max(max_tombs, s.table_tombs)
        if r % 10 == 9:
            assert_same_tree(device_canonical(ix), orc.canonical())
    assert n_evicted > 256, n_evicted          # more evictions than the table has slots ...
    assert max_tombs > 0                       # ... and tombstones were seen and kept in check
    assert_same_tree(device_canonical(ix), orc.canonical())
    ix.close()


def run_cold_batch(make_index, seed, n=700, L=72, n_docs=6):
    """The first batches into an empty index are applied as consecutive chunks of 256 records (rm_insert.cu: every record
    of a hot document would otherwise queue in one in-order group).  Same order, so the same prefix lengths, tree and
    eviction order as the oracle applying the records one by one; a third batch goes through the unchunked path."""
    from tests.parity import zipf_prefix
    ix = make_index(mode=0, max_nodes=1 << 16, max_tokens=max(1 << 18, 3 * n * L), max_batch=max(1024, n), max_key_len=4 * L)
    orc = PortRadixCache(1)
    docs = None
    vc = 0
    for b in range(3):
        seqs, docs = zipf_prefix(seed * 10 + b, n_docs, n, L, vocab=40, docs=docs)
        vals = np.arange(vc, vc + n * L, dtype=np.int64)
        vc += n * L
        offs = np.arange(0, (n + 1) * L, L, dtype=np.int64)
        got = ix.insert_host(seqs.reshape(-1).copy(), vals, offs)
        exp = [orc.insert(r.tolist(), vals[i * L:(i + 1) * L]) for i, r in enumerate(seqs)]
        assert got.tolist() == exp, f"batch {b}"
        ix.check_device_error()
        assert_same_tree(device_canonical(ix), orc.canonical())
    before = len(orc.freed)
    orc.evict(1 << 30)
    _, eo, el = ix.evict(1 << 30)
    assert el.tolist() == [len(r) for r in orc.freed[before:]]
    ix.close()
