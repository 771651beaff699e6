"""The mbarrier protocols of the four kernels that wait on mbarriers, restated as concurrent programs and run under
random schedules (tools/protocol_model.py): stands in for compute-sanitizer's racecheck / synccheck, which this pool
does not allow.  The unmodified protocols must survive every schedule; every deliberately broken variant must be caught
(the checker's own check)."""
import importlib.util
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("protocol_model", os.path.join(ROOT, "tools", "protocol_model.py"))
pm = importlib.util.module_from_spec(spec)
spec.loader.exec_module(pm)


@pytest.mark.parametrize("name", sorted(pm.MODELS))
def test_protocol_survives_random_schedules(name):
    for seed in range(120):                      # even seeds: uniform choices, odd seeds: priority schedules
        assert pm.MODELS[name](seed) > 0


@pytest.mark.parametrize("name", sorted(pm.MUTANTS))
def test_broken_protocol_is_caught(name):
    caught = 0
    for seed in range(100):
        try:
            pm.MUTANTS[name](seed)
        except pm.ModelError:
            caught += 1
    assert caught >= 1, name


def test_mbarrier_model_semantics():
    b = pm.MBar("b", 1)
    assert b.passed(1) and not b.passed(0)       # a fresh barrier: "the phase before phase 0" counts as complete
    b.complete_tx(3)                              # data before the expectation: transaction count goes negative
    assert b.phase == 0
    b.expect_arrive(4)
    assert b.phase == 0 and b.tx == 1 and b.pending == 0
    b.complete_tx(1)
    assert b.phase == 1 and b.passed(0) and not b.passed(1) and b.pending == 1
    b.arrive()
    assert b.phase == 2 and not b.passed(0)      # two phases later the same parity blocks again: aliasing is modelled
