"""The output surface: acc_observe against the reference's own observe<>() visitors.

tests/golden/observe10.shim.acc.gz holds, at three steps of a shocked run with multi-hop routes and passage forcing, EVERY
observable the reference offers -- collected by the oracle harness through the reference's observe<Observer, H>() templates
the way ArrayOutput does (src/output/ArrayOutput.cpp:115-158), in ArrayOutput's array shapes (firms / consumers [agents],
sectors, regions, storages [sectors][agents], flows [from][to], NaN where there is no object).  The engine must reproduce all of
them through acc_observe: NaN pattern identical, quantities / prices / counts exactly, values to 1e-12 (region sums and
in-transit values to 1e-9, as in tests/parity.py), with and without a selection."""
import numpy as np
import pytest

import parity
from acclimate_b200 import engine

# value sums the reference folds in lock-arrival / ring-position order (tests/parity.py LOOSE, TRANSIT)
LOOSE = ("regions:import_value", "regions:export_value", "regions:gdp_value", "flows:total_flow_value", "flows:flow_mean_value")
# not provided: Model::duration (wall clock) has no group here; everything else the oracle dumps must be reproduced
STEPS = [0, 4, 9]


def close(name, got, want, scale=None):
    if scale is not None:  # a difference of two nearly equal values: judged against their size
        bad = np.abs(got - want) > 1e-12 * np.abs(scale)
        bad &= ~(np.isnan(got) & np.isnan(want))
        return f"{name}: {int(bad.sum())} cells off by more than 1e-12 of the production value" if bad.any() else None
    if got.shape != want.shape:
        return f"{name}: shape {got.shape} != {want.shape}"
    nan_g, nan_w = np.isnan(got), np.isnan(want)
    if name.startswith("flows:flow_mean_"):
        # the reference divides by the connection's delay: a zero delay gives inf or garbage there, NaN here
        nan_w = nan_w | ~np.isfinite(want) | nan_g
        want = np.where(nan_w, np.nan, want)
    if not np.array_equal(nan_g, nan_w):
        return f"{name}: NaN pattern differs in {(nan_g != nan_w).sum()} cells"
    g, w = got[~nan_g], want[~nan_w]
    tol = 1e-9 if name.split("[")[0] in LOOSE else 1e-12
    err = np.abs(g - w) / np.maximum(np.abs(w), 1e-300)
    err[g == w] = 0.0
    if err.size and err.max() > tol:
        i = int(err.argmax())
        return f"{name}: rel err {err.max():.3e}: {g[i]!r} != {w[i]!r}"
    return None


def run(lib_path):
    net, gold, steps = parity.load_case("observe10")
    E = engine.Engine(net, device=0, lib_path=lib_path)
    sc = engine.Scenario(net)
    nA, nS = int(net["meta"][2]), int(net["meta"][0])
    problems, seen = [], set()
    for t in range(steps):
        sc.apply(E, t)
        E.step()
        E.events()
        if t not in STEPS:
            continue
        pre = "t%05d/obs:" % t
        names = sorted(k[len(pre):] for k in gold if k.startswith(pre))
        assert len(names) > 80, "the fixture holds no observables"
        for name in names:
            want = gold[pre + name]
            group, var = name.split(":")
            seen.add(name)
            try:
                got = E.observe(group, var)
            except engine.EngineError as ex:
                problems.append(f"step {t} {name}: {ex}")
                continue
            if name == "firms:total_value_loss":  # X*.value - X.value: zero up to the rounding of the two sums it subtracts
                msg = close(name, got, want, scale=gold[pre + "firms:production_value"])
                if msg:
                    problems.append(f"step {t} {msg}")
                continue
            msg = close(name, got, want)
            if msg:
                problems.append(f"step {t} {msg}")
            # the same variable for a selection (ArrayOutput's select_* of an outputs entry)
            if group in ("firms", "consumers"):
                sel = np.array([nA - 1, 2, 0, 7])
                msg = close(name + "[sel]", E.observe(group, var, select_from=sel), want[sel])
            elif group in ("sectors", "regions"):
                sel = np.array([1, 0])
                msg = close(name + "[sel]", E.observe(group, var, select_from=sel), want[sel])
            elif group == "storages":
                sel = np.array([5, 1, nA - 2])
                msg = close(name + "[sel]", E.observe(group, var, select_to=sel), want.reshape(nS, nA)[:, sel].reshape(-1))
            else:
                sel = np.array([3, 0, nA - 3])
                msg = close(name + "[from]", E.observe(group, var, select_from=sel), want.reshape(nA, nA)[sel, :].reshape(-1))
                if not msg:
                    msg = close(name + "[to]", E.observe(group, var, select_to=sel), want.reshape(nA, nA)[:, sel].reshape(-1))
            if msg:
                problems.append(f"step {t} {msg}")
    E.close()
    return problems, seen


def test_observables_match_the_references_visitors_emulation(emu_lib):
    problems, seen = run(emu_lib)
    assert not problems, "\n".join(problems[:20])
    # every group is covered
    assert {n.split(":")[0] for n in seen} == {"firms", "consumers", "sectors", "regions", "storages", "flows"}


@pytest.mark.gpu
def test_observables_match_the_references_visitors_cuda(cuda_lib):
    problems, _ = run(cuda_lib)
    assert not problems, "\n".join(problems[:20])


def test_observe_rejects_what_the_reference_rejects(emu_lib):
    net, gold, steps = parity.load_case("observe10")
    E = engine.Engine(net, device=0, lib_path=emu_lib)
    E.step()
    with pytest.raises(engine.EngineError, match="both source agent and target agent"):
        E.observe("flows", "sent_flow_quantity", select_from=[0], select_to=[1])
    with pytest.raises(engine.EngineError, match="unknown observable"):
        E.observe("firms", "no_such_variable")
    with pytest.raises(engine.EngineError, match="out of range"):
        E.observe("firms", "production_quantity", select_from=[10 ** 6])
    with pytest.raises(engine.EngineError, match="unknown group"):
        E.observe("locations", "total_flow_quantity")
    E.close()
