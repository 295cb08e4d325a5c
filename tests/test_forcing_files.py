"""Forcing files of the `event_series` scenario (the file layer): reference src/scenario/ExternalScenario.cpp:43-191,
src/scenario/ExternalForcing.cpp:29-51, src/scenario/EventSeriesScenario.cpp:45-119.

The settings-driven run on two forcing files (a shock record, then a file one reference year later whose record lifts it) must equal
the GOLDEN run of the reference under the equivalent `events` scenario (tests/golden/small_shock12, produced by the reference's own
Scenario::iterate), bit for bit: the file layer is anchored on the reference's run, not on this engine.  The stepping rule itself
(record times, seconds, templates, year roll-over, end of the last file, errors) is checked against hand-derived schedules."""
import os
import tempfile

import numpy as np
import pytest

import parity
import test_settings
from acclimate_b200 import forcing, modelrun, synth


class Recorder:
    """stands in for the engine: keeps what a step sent to the device"""

    def __init__(self):
        self.agent, self.entity = [], []

    def set_agent_forcing(self, lam):
        self.agent.append(np.array(lam))

    def set_entity_passage(self, nu):
        self.entity.append(np.array(nu))


def small():
    net = synth.make_config("small")
    sec, reg = test_settings.names(net)
    return net, sec, reg


def shock_files(tmp, net, sec, reg):
    nS, nR = len(sec), len(reg)
    at, ar, asec = (np.asarray(net[k]) for k in ("agent_type", "agent_region", "agent_sector"))
    beta = np.asarray(net["agent_beta"])
    shock = np.full((1, nS, nR), np.nan)
    for f in np.nonzero((at == 1) & (ar == 3))[0]:
        shock[0, asec[f], 3] = 0.5 / beta[f]
    # file 0: one record at day 2; file 1 counts from the next year: its day 3 is model day 3 + (3 + 1) = 7
    forcing.write_forcing_file(os.path.join(tmp, "f_0.nc"), "forcing", [2], shock, units="days since 2020-1-1", sectors=sec, regions=reg)
    forcing.write_forcing_file(os.path.join(tmp, "f_1.nc"), "forcing", [3], np.full((1, nS, nR), np.nan), units="days since 2021-01-01",
                               sectors=sec, regions=reg)


def run_settings(lib):
    net, sec, reg = small()
    with tempfile.TemporaryDirectory() as tmp:
        shock_files(tmp, net, sec, reg)
        head, tail = test_settings.DOC.split("scenario:")
        doc = head + f"""scenario:
  type: event_series
  start: 0
  stop: 11
  basedate: 2020-1-1
  parameters:
    stem: f
  forcing:
    variable: forcing
    file: {tmp}/[[stem]]_[[index]].nc
    index_from: 0
    index_to: 1
outputs:
  - format: array
    observables:
      firms:
        forcing: ~
"""
        mr, steps = modelrun.ModelRun.from_settings(doc, net, sec, reg, lib_path=lib)
        assert steps == 12 and isinstance(mr.scenario, forcing.EventSeriesFiles)
        _, gold, _ = parity.load_case("small_shock12")
        problems = []
        for t in range(steps):
            mr.step()
            problems += parity.compare_step("small_shock12", mr.engine, gold, t)
        lam = mr.sinks[0][1].arrays()[("firms", "forcing")]
        mr.close()
    assert not problems, "\n".join(problems[:10])
    hit = (np.asarray(net["agent_type"]) == 1) & (np.asarray(net["agent_region"]) == 3)
    assert np.allclose(lam[2:7][:, hit], 0.5 / 1.15) and np.all(lam[7:][:, hit] == 1.0) and np.all(lam[:2][:, hit] == 1.0)


def test_forcing_files_give_the_reference_run_emulation(emu_lib):
    run_settings(emu_lib)


@pytest.mark.gpu
def test_forcing_files_give_the_reference_run_cuda(cuda_lib):
    run_settings(cuda_lib)


def series(tmp, node, steps, net, sec, reg, entity_names=None, start=0.0, delta_t=1.0):
    sc = forcing.EventSeriesFiles(net, sec, reg, node, entity_names=entity_names, start_time=start, delta_t=delta_t)
    rec = Recorder()
    sent = []
    for t in range(steps):
        n = len(rec.agent) + len(rec.entity)
        sc.apply(rec, t)
        sent.append(len(rec.agent) + len(rec.entity) - n)
    sc.close()
    return rec, sent


def test_stepping_rule_seconds_and_end_of_files():
    net, sec, reg = small()
    nS, nR = len(sec), len(reg)
    day = 24 * 60 * 60
    with tempfile.TemporaryDirectory() as tmp:
        recs = np.full((3, nS, nR), np.nan)
        recs[0, 0, 0], recs[1, 0, 0], recs[2, 1, 2] = 0.25, 0.75, 0.5
        forcing.write_forcing_file(os.path.join(tmp, "a.nc"), "lam", [1 * day, 4 * day, 5 * day], recs, units="seconds since 2020-1-1 00:00:00",
                                   sectors=sec, regions=reg)
        node = {"forcing": {"variable": "lam", "file": os.path.join(tmp, "a.nc")}}
        rec, sent = series(tmp, node, 9, net, sec, reg)
    # records at days 1, 4, 5; day 6 finds the file exhausted and no further file: everything back to 1.0, then nothing more
    assert sent == [0, 1, 0, 0, 1, 1, 1, 0, 0]
    a00 = next(a for a in range(int(net["meta"][2])) if net["agent_sector"][a] == 0 and net["agent_region"][a] == 0)
    a12 = next(a for a in range(int(net["meta"][2])) if net["agent_sector"][a] == 1 and net["agent_region"][a] == 2)
    assert [x[a00] for x in rec.agent] == [0.25, 0.75, 1.0, 1.0]   # NaN in a later record lifts the forcing (EventSeriesScenario.cpp:50-51)
    assert [x[a12] for x in rec.agent] == [1.0, 1.0, 0.5, 1.0]
    assert all(np.all(np.delete(x, [a00, a12]) == 1.0) for x in rec.agent)


def test_start_time_and_delta_t():
    net, sec, reg = small()
    with tempfile.TemporaryDirectory() as tmp:
        recs = np.full((2, len(sec), len(reg)), 0.5)
        forcing.write_forcing_file(os.path.join(tmp, "a.nc"), "lam", [10, 13], recs, sectors=sec, regions=reg)
        node = {"forcing": {"variable": "lam", "file": os.path.join(tmp, "a.nc")}}
        # model time = 8 + 2 t: 8, 10, 12, 14, 16 -> records once time >= 10 and >= 13, then the reset
        _, sent = series(tmp, node, 5, net, sec, reg, start=8.0, delta_t=2.0)
    assert sent == [0, 1, 0, 1, 1]


def test_sea_route_records_and_unknown_names():
    net, sec, reg = small()
    with tempfile.TemporaryDirectory() as tmp:
        forcing.write_forcing_file(os.path.join(tmp, "s.nc"), "nu", [0, 1], [[0.5, -3.0, 0.1], [-1.0, 0.25, 0.9]], sea_routes=["suez", "nowhere", "panama"])
        node = {"forcing": {"variable": "nu", "file": os.path.join(tmp, "s.nc")}}
        with pytest.warns(UserWarning, match="Sea object 'nowhere' not found"):
            rec, sent = series(tmp, node, 4, net, sec, reg, entity_names=["panama", "suez", "malacca"])
    assert sent == [1, 1, 1, 0] and len(rec.entity) == 2
    assert rec.entity[0].tolist() == [0.1, 0.5, -1.0] and rec.entity[1].tolist() == [0.9, -1.0, -1.0]   # negative -> -1 (:60-61)


def test_template_and_errors():
    net, sec, reg = small()
    one = np.full((1, len(sec), len(reg)), np.nan)
    with tempfile.TemporaryDirectory() as tmp:
        w = lambda name, **kw: forcing.write_forcing_file(os.path.join(tmp, name), "lam", [0], one, sectors=sec, regions=reg, **kw)
        w("x_3.nc"), w("x_4.nc", units="days since 2002-1-1"), w("c_0.nc"), w("c_1.nc", calendar="noleap"), w("bad.nc", units="hours since 2000-1-1")
        node = {"parameters": {"who": "x"}, "forcing": {"variable": "lam", "file": tmp + "/[[who]]_[[index]].nc", "index_from": 3, "index_to": 4}}
        sc = forcing.EventSeriesFiles(net, sec, reg, node)
        assert sc.fill_template("[[who]]-[[index]]-[[who]]") == "x-4-x"   # the index has moved on to the next file
        r = Recorder()
        sc.apply(r, 0)
        with pytest.raises(forcing.ForcingError, match="differ by more than a year"):
            sc.apply(r, 1)
        sc = forcing.EventSeriesFiles(net, sec, reg, {"forcing": {"variable": "lam", "file": tmp + "/c_[[index]].nc", "index_to": 1}})
        sc.apply(r, 0)
        with pytest.raises(forcing.ForcingError, match="differ in calendar"):
            sc.apply(r, 1)
        with pytest.raises(forcing.ForcingError, match="not found"):
            forcing.EventSeriesFiles(net, sec, reg, {"forcing": {"variable": "lam", "file": tmp + "/none.nc"}})
        with pytest.raises(forcing.ForcingError, match="Variable 'other' not found"):
            forcing.EventSeriesFiles(net, sec, reg, {"forcing": {"variable": "other", "file": tmp + "/c_0.nc"}})
        with pytest.raises(forcing.ForcingError, match="unknown key scenario.forcing.path"):
            forcing.EventSeriesFiles(net, sec, reg, {"forcing": {"variable": "lam", "file": "x", "path": 1}})
    assert forcing.get_ref_year("f", "days since 2020-1-1") == 2020 and forcing.get_ref_year("f", "days since 1999-01-01") == 1999
    # the reference compares the characters from position 19 on, one past the '-' behind the year (ExternalScenario.cpp:76): no
    # "seconds since" string passes, so a units change between files in seconds is an error there -- and here
    with pytest.raises(forcing.ForcingError, match="invalid time units"):
        forcing.get_ref_year("f", "seconds since 1999-01-01 00:00:00")
    with pytest.raises(forcing.ForcingError, match="invalid time units"):
        forcing.get_ref_year("f", "days since 2020-3-1")
    with pytest.raises(forcing.ForcingError, match="invalid time units"):
        forcing.get_ref_year("f", "hours since 2020-1-1")
