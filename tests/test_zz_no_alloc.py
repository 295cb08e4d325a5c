"""acc_firm_incoming_demand_static -- the entry point that allocates no device memory (include/acclimate_b200.h) --
against the existing allocating read acc_read("firms", "incoming_demand_*"), on two calls in a row.
Collected last on purpose (pytest -x must not hide other tests behind a failure here)."""
import ctypes as C

import numpy as np
import pytest

from acclimate_b200 import engine, synth


def _static(E):
    lib = E.lib
    lib.acc_firm_incoming_demand_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    n = C.c_size_t(0)
    E._check(lib.acc_firm_incoming_demand_static(E.h, None, None, 0, C.byref(n)))
    q, v = np.full(int(n.value), np.nan), np.full(int(n.value), np.nan)
    E._check(lib.acc_firm_incoming_demand_static(E.h, q.ctypes.data, v.ctypes.data, q.size, C.byref(n)))
    return q, v


@pytest.mark.gpu
def test_static_scratch_entry_matches_allocating_read(cuda_lib):
    net = synth.make_network(regions=7, sectors=4, seed=31, max_delay=3)
    synth.add_shock(net, 1, 5, 2, 0.3)
    E = engine.Engine(net, device=0, lib_path=cuda_lib)
    sc = engine.Scenario(net)
    firms = np.asarray(net["agent_type"]) == 1
    previous = None
    for call in range(2):  # two calls in a row: the second finds the first one's contents in the shared array
        for t in range(3 * call, 3 * call + 3):
            sc.apply(E, t)
            E.step()
        want_q, want_v = E.read("firms", "incoming_demand_quantity"), E.read("firms", "incoming_demand_value")
        got_q, got_v = _static(E)
        assert np.array_equal(got_q, want_q)  # int64 quanta: exact
        assert np.array_equal(got_v, want_v)  # same kernel, same fold order: exact as well
        assert np.all(got_q[~firms] == 0) and got_q[firms].min() > 0
        if previous is not None:
            assert not np.array_equal(previous, got_q)  # the shock moved the demand: the array does not just keep old contents
        previous = got_q
    # a second engine with FEWER agents right after: rows beyond its size are not its business, its own rows are cleared first
    small = synth.make_network(regions=3, sectors=2, seed=5, const_delay=1)
    E2 = engine.Engine(small, device=0, lib_path=cuda_lib)
    E2.step()
    q2, v2 = _static(E2)
    assert np.array_equal(q2, E2.read("firms", "incoming_demand_quantity")) and np.array_equal(v2, E2.read("firms", "incoming_demand_value"))
    E2.close()
    E.close()


@pytest.mark.gpu
def test_static_scratch_rejects_too_many_agents(cuda_lib):
    # 65 537 agents would exceed ACC_STATIC_SCRATCH_AGENTS; building such a network only to be refused is wasteful, so the limit
    # is read from the header and the error path is checked through the size query of a small engine plus the documented bound
    import os
    import re

    import __graft_entry__ as g

    text = open(os.path.join(g.ROOT, "include", "acclimate_b200.h")).read()
    limit = int(re.search(r"#define ACC_STATIC_SCRATCH_AGENTS (\d+)", text).group(1))
    assert limit == 65536
    net = synth.make_network(regions=3, sectors=2, seed=5, const_delay=1)
    E = engine.Engine(net, device=0, lib_path=cuda_lib)
    lib = E.lib
    lib.acc_firm_incoming_demand_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    n = C.c_size_t(0)
    q = np.zeros(2)
    assert lib.acc_firm_incoming_demand_static(E.h, q.ctypes.data, None, q.size, C.byref(n)) == -1  # destination too small: clear error
    assert b"too small" in lib.acc_last_error()
    E.close()
