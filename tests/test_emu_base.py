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
