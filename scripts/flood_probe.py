"""Where the time of a shocked step goes (configs[2]): solver statistics of the purchasing managers per step.

  python scripts/flood_probe.py [steps]     (GPU box)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from acclimate_b200 import engine, synth  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
net = synth.make_config("eora_flood")
e = engine.Engine(net)
sc = engine.Scenario(net)
e.profile(True)
for t in range(steps):
    sc.apply(e, t)
    e.step()
    e.sync()
    kt = e.kernel_times(reset=True)
    if t >= 8:
        it = e.read_state("sto_solver_iters")
        ev = e.read_state("sto_solver_evals")
        ms = {k: round(v[0], 2) for k, v in kt.items()} if isinstance(kt, dict) else kt
        print("t=%d iters max %d mean %.2f  evals max %d mean %.2f  PMs with evals>20: %d, >100: %d  ms %s" % (
            t, it.max(), it.mean(), ev.max(), ev.mean(), int((ev > 20).sum()), int((ev > 100).sum()), ms), flush=True)
ev_ = e.events()
print("events:", len(ev_))
e.close()
