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
