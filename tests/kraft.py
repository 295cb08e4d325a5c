"""How far the engine's run lies from the reference run under an INDEPENDENT optimiser.

tests/golden/<case>.kraft.acc.gz is the reference's own model code with oracle/nlopt_kraft (Kraft's SLSQP restated from
the publication, nothing of the product in it) behind its NLopt calls.  The engine runs its own solver (csrc/sqp_core.h,
csrc/sqp_dense.h).  Both minimise the same objective to the reference's stopping tolerance (xtol_abs = 1e-3 quantity units
per connection, src/model/PurchasingManager.cpp:640-644), so the two trajectories are two admissible answers of the SAME
model, not bit-identical ones; this module MEASURES the distance:

    python tests/kraft.py            writes tests/golden/kraft_deviation.json (host emulation of the kernels; int64 fields of
                                     the CUDA run are bit-identical to it, tests/test_parity_gpu.py)

per case and field: the first step at which anything differs, the largest aggregate deviation  sum|a - b| / sum|b|  of a
step, the largest per-entry relative deviation among entries of at least one unit (1000 quanta), and -- at the first step
at which the demand requests differ, where both solvers still saw IDENTICAL inputs -- how the optimised objective values
of the purchasing managers compare.  tests/test_kraft_deviation.py holds the engine to the recorded numbers.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import parity  # noqa: E402
from acclimate_b200 import engine, netfile  # noqa: E402

FIELDS = ["edge_D_q", "edge_D_p", "edge_ship_q", "firm_X_q", "firm_nbar", "sto_S_q", "sto_demand_q", "sector_production_q", "region_consumption_q", "region_gdp_v"]
RECORD = os.path.join(parity.GOLDEN_DIR, "kraft_deviation.json")
CASES = [c for c in parity.CASES if os.path.exists(parity.fixture_path(c, "kraft"))]


def measure(name, lib_path=None, device=0):
    net, shim_gold, steps = parity.load_case(name)
    gold = netfile.read(parity.fixture_path(name, "kraft"))
    parity.check_stamp(gold, "kraft", name)
    E = engine.Engine(net, device=device, lib_path=lib_path)
    sc = engine.Scenario(net)
    out = {f: {"first_diff_step": None, "max_rel_l1": 0.0, "max_rel_entry": 0.0, "worst_step": None} for f in FIELDS}
    onset = None
    for t in range(steps):
        sc.apply(E, t)
        E.step()
        E.events()
        full, _ = parity.step_arrays(gold, t)
        for f in FIELDS:
            if f not in full:
                continue
            got, want = E.read_state(f).astype(np.float64), full[f].astype(np.float64)
            diff = np.abs(got - want)
            if not diff.any():
                continue
            rec = out[f]
            if rec["first_diff_step"] is None:
                rec["first_diff_step"] = t
            l1 = float(diff.sum() / max(np.abs(want).sum(), 1e-300))
            if l1 > rec["max_rel_l1"]:
                rec["max_rel_l1"], rec["worst_step"] = l1, t
            big = np.abs(want) >= (1000.0 if f.endswith("_q") else 0.0)
            if f in ("edge_D_p", "firm_nbar"):
                big = want > 0
            if big.any():
                rec["max_rel_entry"] = max(rec["max_rel_entry"], float((diff[big] / np.abs(want[big])).max()))
        if onset is None and "edge_D_q" in full and np.any(E.read_state("edge_D_q") != full["edge_D_q"]) and "sto_optimized_value" in full:
            # identical inputs up to here: compare what the two optimisers made of the same sub-problems
            a, b = E.read_state("sto_optimized_value"), full["sto_optimized_value"]
            scale = np.maximum(np.abs(b), 1e-300)
            rel = (a - b) / scale  # > 0: the engine's optimum is the better one (the objective is maximised)
            onset = {"step": t, "managers": int(a.size), "engine_better": int((rel > 1e-9).sum()), "kraft_better": int((rel < -1e-9).sum()),
                     "equal_to_1e-9": int((np.abs(rel) <= 1e-9).sum()), "median_rel_gap": float(np.median(rel)),
                     "p95_kraft_lead": float(np.percentile(np.maximum(-rel, 0.0), 95))}
    E.close()
    return {"fields": out, "objective_at_onset": onset, "steps": steps}


def main():
    import __graft_entry__ as g

    lib = g.build_emu()
    rec = {}
    for name in CASES:
        rec[name] = measure(name, lib_path=lib)
        f = rec[name]["fields"]
        print("%-20s D_q l1 %.2e (first %s)  D_p l1 %.2e  ship l1 %.2e  X l1 %.2e  gdp l1 %.2e | onset %s" % (
            name, f["edge_D_q"]["max_rel_l1"], f["edge_D_q"]["first_diff_step"], f["edge_D_p"]["max_rel_l1"], f["edge_ship_q"]["max_rel_l1"],
            f["firm_X_q"]["max_rel_l1"], f["region_gdp_v"]["max_rel_l1"], rec[name]["objective_at_onset"]), flush=True)
    with open(RECORD, "w") as fp:
        json.dump(rec, fp, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
