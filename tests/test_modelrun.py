"""The ModelRun mirror drives the engine phase by phase like the reference's run loop (CPU tier: host emulation)."""
import numpy as np

import parity
from acclimate_b200 import modelrun


def test_phase_split_equals_reference(emu_lib):
    net, gold, steps = parity.load_case("outage25")
    run = modelrun.ModelRun(net, outputs=[("firms", "production_quantity"), ("regions", "gdp_value")], lib_path=emu_lib)
    series = run.run(steps)
    x = series[("firms", "production_quantity")]
    assert x.shape == (steps, int(net["meta"][2]))
    for t in range(steps):
        want = gold["t%05d/firm_X_q" % t] * float.fromhex("0x1.0624dd2f1a9fdp-10")
        assert np.array_equal(x[t], want)
    assert len(run.duration_ms) == steps
    run.close()
