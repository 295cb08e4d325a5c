"""Host-side mirror of the reference's run driver for this path.

`ModelRun.run()` follows the loop body of the reference's ModelRun::run (reference src/ModelRun.cpp:153-229):
scenario -> switch_registers + consumption_and_production -> expectation -> purchase -> investment (no-op) -> outputs ->
tick.  Model state lives on the GPU behind the C ABI (include/acclimate_b200.h); the forcing tables are applied with the
rule of Scenario::iterate (reference src/scenario/Scenario.cpp:161-174); observables are pulled by name after each step,
the way ArrayOutput's visitor sweep does (reference src/output/ArrayOutput.cpp:479-646).
"""
from __future__ import annotations

import time

import numpy as np

from . import engine

PHASE_CP, PHASE_EXP, PHASE_PUR, PHASE_TICK = 0, 1, 2, 3


class ModelRun:
    def __init__(self, network, outputs=(), device=0, lib_path=None):
        """network: snapshot arrays (synth.make_network / netfile.read); outputs: iterable of (group, variable) names."""
        self.engine = engine.Engine(network, device=device, lib_path=lib_path)
        self.scenario = engine.Scenario(network)
        self.outputs = list(outputs)
        self.series = {o: [] for o in self.outputs}
        self.events = []
        self.duration_ms = []  # the reference's model/duration observable (ModelRun.cpp:192-194)
        self.time = 0

    def step(self):
        e = self.engine
        self.scenario.apply(e, self.time)                 # scenario->iterate()
        t0 = time.perf_counter()
        e.phase(PHASE_CP)                                 # switch_registers + iterate_consumption_and_production
        e.phase(PHASE_EXP)                                # iterate_expectation
        e.phase(PHASE_PUR)                                # iterate_purchase (+ iterate_investment: no-op bodies)
        e.sync()
        self.duration_ms.append(1e3 * (time.perf_counter() - t0))
        for o in self.outputs:                            # outputs->iterate()
            self.series[o].append(e.read(*o))
        self.events += e.events()                         # ModelRun::event: drained every step (the device buffer is bounded)
        e.phase(PHASE_TICK)                               # model->tick(); ++time
        self.time += 1

    def run(self, steps):
        for _ in range(steps):
            self.step()
        return {o: np.stack(v) for o, v in self.series.items()}

    def close(self):
        self.engine.close()
