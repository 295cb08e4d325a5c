"""`event_series` forcing (reference src/scenario/EventSeriesScenario.cpp:45-69): per-step grids of
agent forcing (time, sector, region; NaN = none) and of geo-entity passage (time, entity; negative = none) give the
same run as the equivalent list of shock / passage events of the reference's default scenario (Scenario.cpp:161-174)."""
import numpy as np
import pytest

from acclimate_b200 import engine, synth

NAMES = ["edge_D_q", "edge_D_p", "edge_ship_q", "sto_S_q", "sto_demand_q", "firm_X_q", "firm_nbar", "region_gdp_v"]


def _case():
    a = synth.make_network(regions=8, sectors=4, seed=8, max_delay=5)
    synth.add_shock(a, 2, 7, 5, 0.4)
    synth.add_passage(a, 3, 6, 4, 1, 0.25)
    steps = 12
    S, R = int(a["meta"][0]), int(a["meta"][1])
    beta = np.asarray(a["agent_beta"])
    at, ar, asec = np.asarray(a["agent_type"]), np.asarray(a["agent_region"]), np.asarray(a["agent_sector"])
    grid = np.full((steps, S + 1, R), np.nan)
    for f in np.nonzero((at == 1) & (ar == 5))[0]:
        grid[2:8, asec[f], 5] = 0.4 / beta[f]
    n_conn = R * (R - 1) // 2
    ent = np.full((steps, n_conn), -1.0)
    ent[3:7, 4 * 3 // 2 + 1] = 0.25
    return a, grid, ent, steps


def _run(lib, net, scenario, steps):
    E = engine.Engine(net, lib_path=lib)
    for t in range(steps):
        scenario.apply(E, t)
        E.step()
    out = {n: E.read_state(n).copy() for n in NAMES}
    E.close()
    return out


def _compare(lib):
    net, grid, ent, steps = _case()
    want = _run(lib, net, engine.Scenario(net), steps)
    got = _run(lib, net, engine.EventSeriesScenario(net, agent_grid=grid, entity_grid=ent), steps)
    for n in NAMES:
        assert np.array_equal(want[n], got[n]), n
    # the shock did something
    assert not np.array_equal(want["firm_X_q"], _run(lib, net, engine.EventSeriesScenario(net), steps)["firm_X_q"])


def test_event_series_kernel_logic(emu_lib):
    _compare(emu_lib)


@pytest.mark.gpu
def test_event_series_cuda(cuda_lib):
    _compare(cuda_lib)
