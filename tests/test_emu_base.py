"""Kernel logic (host emulation build, one lane per group) against the CPU oracle: RadixCache semantics.

This exercises the SAME .cu sources as the GPU library, compiled for the host (csrc/rm_emu.h), so
walk / split / in-order insert / lock / evict logic is checked in the GPU-less container.  The
GPU parity tests proper are tests/test_gpu_*.py.
"""
import pytest

from tests import cases
from tests.emu.emu_backend import make_index


@pytest.mark.parametrize("seed", range(40))
def test_emu_base_random_sequences(seed):
    cases.run_sequence(make_index, seed)


@pytest.mark.parametrize("seed", range(5))
def test_emu_base_longer_keys(seed):
    cases.run_sequence(make_index, 1000 + seed, n_steps=25, vocab=4, maxlen=40, batch=12)


@pytest.mark.parametrize("seed", range(30))
def test_emu_batched_match_splits_then_evict_order(seed):
    cases.run_split_evict(make_index, seed)


@pytest.mark.parametrize("seed", range(6))
def test_emu_streaming_insert_evict_compact(seed):
    cases.run_stream_evict(make_index, seed)


@pytest.mark.parametrize("seed", range(4))
def test_emu_evict_second_pass_with_exact_sizes(seed, monkeypatch):
    """rm_evict sorts at most RM_EVICT_M_OPT candidates without asking the device for the count; with more it must
    leave the tree alone and run again with exact sizes (forced here by a bound of 2 candidates)."""
    monkeypatch.setenv("RM_EVICT_M_OPT", "2")
    cases.run_split_evict(make_index, 40 + seed)
    cases.run_stream_evict(make_index, 40 + seed)


@pytest.mark.parametrize("seed", range(2))
def test_emu_cold_batches_are_applied_in_chunks(seed):
    cases.run_cold_batch(make_index, seed)


@pytest.mark.parametrize("seed", range(3))
def test_emu_tombstone_pressure_rebuilds_the_table(seed):
    cases.run_tombstone_pressure(make_index, seed)


@pytest.mark.parametrize("small_max", ["1", "3"])
def test_emu_split_stage_sorted_path(small_max):
    """The split stage orders up to 32 requests per edge in registers and sorts only beyond that; with the bound cut to
    1 or 3 the same scenarios push nearly every shared edge through the sorted path."""
    import os
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, '.'); from tests import cases; from tests.emu.emu_backend import make_index; "
            "[cases.run_split_evict(make_index, 200 + s) for s in range(10)]")
    subprocess.run([sys.executable, "-c", code], check=True, env=dict(os.environ, RM_SPLIT_SMALL_MAX=small_max),
                   cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
