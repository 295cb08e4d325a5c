"""BASELINE.json configs[2]: the EORA-shaped network with a regional production-loss shock, a 3-year run (1 095 daily steps),
tracking the recovery dynamics (`firms/production`, `regions/gdp`).

  python scripts/flood_recovery.py [--config eora|cut40] [--steps 1095] [--out profiles/r02_flood_recovery_eora.csv]

Runs on the GPU through the C ABI (ModelRun + acc_observe pulls every step) and writes one row per step: total GDP, GDP of the
shocked region, total production, production of the shocked region's firms, wall time of the step.  The 40 x 26 cut is the run
tests/test_parity_gpu.py::test_three_year_flood_matches_reference compares with the reference bit for bit."""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from acclimate_b200 import engine, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="eora", choices=["eora", "cut40"])
    ap.add_argument("--steps", type=int, default=1095)
    ap.add_argument("--region", type=int, default=17)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    net = synth.make_config("eora_flood") if args.config == "eora" else synth.make_network(regions=40, sectors=26, seed=20260101 + 3)
    if args.config != "eora":
        synth.add_shock(net, 10, 24, args.region, 0.5)
    E = engine.Engine(net)
    sc = engine.Scenario(net)
    at, ar = np.asarray(net["agent_type"]), np.asarray(net["agent_region"])
    hit = (at == 1) & (ar == args.region)
    rows = []
    t0 = time.perf_counter()
    for t in range(args.steps):
        sc.apply(E, t)
        E.step()
        gdp = E.observe("regions", "gdp_value")
        x = E.observe("firms", "production_quantity")
        E.events()
        rows.append((t, gdp.sum(), gdp[args.region], np.nansum(x), np.nansum(x[hit]), float(sc.lam[hit].min())))
    E.sync()
    secs = time.perf_counter() - t0
    out = args.out or os.path.join(ROOT, "profiles", "r02_flood_recovery_%s.csv" % args.config)
    with open(out, "w") as fp:
        fp.write("# %s, %d connections, shock: region %d at 50 %% capacity from step 10 to 24; %d steps in %.1f s (%.2f ms per step incl. two observable pulls)\n"
                 % (args.config, E.n_edges, args.region, args.steps, secs, 1e3 * secs / args.steps))
        fp.write("step,gdp_total,gdp_shocked_region,production_total,production_shocked_region,forcing\n")
        for r in rows:
            fp.write("%d,%.6f,%.6f,%.3f,%.3f,%.6f\n" % r)
    base = rows[0]
    worst = min(rows, key=lambda r: r[1])
    print("steps %d in %.1f s; GDP baseline %.1f, minimum %.1f at step %d (%.3f %%), end %.1f (%.5f %% of baseline)"
          % (args.steps, secs, base[1], worst[1], worst[0], 100 * worst[1] / base[1], rows[-1][1], 100 * rows[-1][1] / base[1]))
    E.close()


if __name__ == "__main__":
    main()
