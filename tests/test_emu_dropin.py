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
    dc.run_gc_ring(kw())


def test_reset_drops_staged_inserts():
    dc.run_reset_drops_staged(kw())


def test_tick_is_forwarded_around_the_ring():
    dc.run_tick_ring(kw())


def test_insert_returns_its_own_prefix_length():
    dc.run_insert_returns_own_prefix(kw())


@pytest.mark.parametrize("case", dc.golden("paged_sequences") + dc.golden("paged_ragged_sequences"),
                         ids=lambda c: f"seed{c['seed']}-page{c['page']}")
def test_paged_sequences(case):
    """RadixCache(page_size > 1) recorded from the reference: the drop-in runs the device index on page symbols."""
    dc.replay_base_sequence(case, kw())



def test_kv_events_flag_fails_as_in_the_reference():
    """radix_cache.py:117-125, 422-424: the constructor's reset() needs a class the module never imports."""
    from radixmesh_b200 import RadixCache
    with pytest.raises(NameError):
        RadixCache(None, None, 1, enable_kv_cache_events=True, max_nodes=64, max_tokens=256, max_batch=4, max_key_len=16, **kw())


def test_paged_restrictions():
    import torch
    from radixmesh_b200 import RadixCache, RadixMesh, ServerArgs
    t = RadixCache(None, None, 4, max_nodes=256, max_tokens=4096, max_batch=16, max_key_len=64, **kw())
    with pytest.raises(ValueError):
        t.insert([1, 2, -(1 << 31)], torch.arange(3))              # the token that fills up short pages is reserved
    assert t.insert([1, 2, 3, 4, 5, 6, 7, 8], torch.arange(8)) == 0
    assert t.match_prefix([1, 2, 3, 4, 5, 6, 7]).device_indices.tolist() == [0, 1, 2, 3]   # page aligned (radix_cache.py:148-150)
    assert t.match_prefix([1, 2, 3]).device_indices.tolist() == []
    assert (t.evictable_size(), t.total_size()) == (8, 8)
    t.reset()
    assert t.total_size() == 0 and t.match_prefix([1, 2, 3, 4]).device_indices.tolist() == []


@pytest.mark.parametrize("seed", range(12))
def test_paged_random_against_oracle(seed):
    """Random insert / match / evict on RadixCache(page_size 2..4) against the oracle's paged RadixCache."""
    import random
    import numpy as np
    import torch
    from oracle.radix_port import PortRadixCache
    from radixmesh_b200 import RadixCache
    rnd = random.Random(4200 + seed)
    page = rnd.choice([2, 3, 4])
    alloc = dc.FakeAllocator()
    t = RadixCache(None, alloc, page, max_nodes=1 << 12, max_tokens=1 << 15, max_batch=16, max_key_len=64, **kw())
    orc = PortRadixCache(page)
    vc = 0
    for _ in range(80):
        kind = rnd.choice(["ins", "ins", "match", "match", "evict"])
        key = [rnd.randrange(3) for _ in range(rnd.randrange(0, 20))]
        if kind == "ins":
            key = key[:len(key) // page * page]
            val = list(range(vc, vc + len(key)))
            vc += len(key)
            assert t.insert(key, torch.tensor(val, dtype=torch.int64)) == orc.insert(list(key), np.asarray(val, dtype=np.int64))
        elif kind == "match":
            a, b = t.match_prefix(key), orc.match_prefix(list(key))
            assert a.device_indices.tolist() == b.indices.tolist()
            assert t.full_key(a.last_device_node) == orc.full_key(b.last_node)
        else:
            n = rnd.randrange(0, 10)
            before_d, before_o = len(alloc.freed), len(orc.freed)
            t.evict(n)
            orc.evict(n)
            assert alloc.freed[before_d:] == orc.freed[before_o:]
        assert (t.evictable_size(), t.total_size()) == (orc.evictable, orc.total_size())


@pytest.mark.parametrize("seed", range(10))
def test_flatten_order_children_and_total_size(seed):
    dc.run_flatten_order(kw(), seed)


@pytest.mark.parametrize("seed", range(4))
def test_flatten_order_paged(seed):
    dc.run_flatten_order(kw(), 50 + seed, page=2)


@pytest.mark.parametrize("seed", range(10))
def test_paged_batch_entry_points(seed):
    dc.run_paged_batch(kw(), seed)


@pytest.mark.parametrize("seed", range(12))
def test_paged_keys_with_a_short_last_page(seed):
    dc.run_paged_ragged(kw(), seed)


@pytest.mark.parametrize("seed,page", [(0, 1), (1, 1), (2, 2), (3, 4)])
def test_arena_of_evicted_runs_is_recycled(seed, page):
    dc.run_arena_recycling(kw(), seed, page)


@pytest.mark.parametrize("seed", range(4))
def test_paged_symbols_of_evicted_pages_are_reused(seed):
    dc.run_paged_symbol_reclaim(kw(), seed)


@pytest.mark.parametrize("seed", range(8))
def test_paged_short_pages_small_table_small_arena_together(seed):
    dc.run_paged_mixed_pressure(kw(), seed)


@pytest.mark.parametrize("seed", range(8))
def test_paged_random_traffic_through_a_small_symbol_table(seed):
    """Random paged inserts / matches / evictions over a vocabulary of 12 tokens with room for 256 page symbols (a batch
    brings up to ~130 new pages, the tree is evicted down to a few dozen tokens every round): pages are dropped
    (rm_pages_retain) and interned again all the time, under the same checks as the other paged runs."""
    dc.run_paged_batch(kw(), 300 + seed, vocab=12, max_page_symbols=256, rounds=40)
