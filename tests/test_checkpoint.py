"""State checkpoint (replaces the reference's DMTCP checkpoint, src/checkpointing.cpp:40-68): a run
that is saved, thrown away and resumed in a NEW engine continues bit for bit like the run that was never interrupted --
every readable state array and the event list, with a shock and a passage forcing active across the cut (so the
overflow arrays, which only exist once forcing was used, travel too)."""
import numpy as np
import pytest

from acclimate_b200 import engine, synth

NAMES = ["edge_D_q", "edge_D_p", "edge_ship_q", "sto_S_q", "sto_S_v", "sto_I_q", "sto_U_q", "sto_Ut_q", "sto_demand_q", "sto_demand_v",
         "sto_transit_q", "sto_transit_v", "sto_deficit_q", "sto_optimized_value", "firm_X_q", "firm_X_v", "firm_Xhat_q", "firm_Xexp_q", "firm_nbar",
         "firm_cutoff", "firm_costs_C", "firm_revenue_R", "region_import_q", "region_gdp_v", "consumer_utility"]


def _network():
    a = synth.make_network(regions=7, sectors=4, seed=21, max_delay=3)
    info = synth.add_routes(a, seed=21, hubs=3)
    synth.add_shock(a, 2, 9, 3, 0.3)
    synth.add_entity_passage(a, 3, 10, info["hub_loc"][0], 0.4)
    return a


def _advance(E, sc, t0, t1):
    for t in range(t0, t1):
        sc.apply(E, t)
        E.step()


def _snapshot(E):
    return {n: E.read_state(n).copy() for n in NAMES}


def _roundtrip(lib):
    net = _network()
    cut, end = 6, 14
    # uninterrupted run
    A = engine.Engine(net, lib_path=lib)
    scA = engine.Scenario(net)
    _advance(A, scA, 0, end)
    want, want_events = _snapshot(A), A.events()
    A.close()
    # run to the cut, save, drop the engine
    B = engine.Engine(net, lib_path=lib)
    scB = engine.Scenario(net)
    _advance(B, scB, 0, cut)
    blob = B.save_state().copy()
    B.close()
    # resume in a new engine
    Cn = engine.Engine(net, lib_path=lib)
    Cn.load_state(blob)
    assert Cn.steps_done == cut
    scC = engine.Scenario(net)
    _advance(Cn, scC, cut, end)
    got, got_events = _snapshot(Cn), Cn.events()
    for n in NAMES:
        assert np.array_equal(want[n], got[n], equal_nan=True), n
    assert sorted(want_events) == sorted(got_events)
    # a blob of another snapshot is refused
    other = engine.Engine(synth.make_network(regions=5, sectors=3, seed=2), lib_path=lib)
    with pytest.raises(engine.EngineError):
        other.load_state(blob)
    other.close()
    Cn.close()


def test_checkpoint_roundtrip_kernel_logic(emu_lib):
    _roundtrip(emu_lib)


@pytest.mark.gpu
def test_checkpoint_roundtrip_cuda(cuda_lib):
    _roundtrip(cuda_lib)
