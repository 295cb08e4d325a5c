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
        self.network = network
        self.engine = engine.Engine(network, device=device, lib_path=lib_path)
        self.scenario = engine.Scenario(network)
        self.outputs = list(outputs)
        self.series = {o: [] for o in self.outputs}
        self.events = []
        self.duration_ms = []  # the reference's model/duration observable (ModelRun.cpp:192-194)
        self.time = 0
        self.sinks = []        # (pulls, sink) pairs of a settings document's `outputs:` section
        self.start_time = 0.0
        self.delta_t = 1.0

    @classmethod
    def from_settings(cls, doc, network, sector_names, region_names, device=0, lib_path=None, flows=None, entity_names=None):
        """The reference's entry point (src/main.cpp:58-124 -> ModelRun::ModelRun, src/ModelRun.cpp:77-150) on a settings document:
        `network` is the snapshot (or, with `flows`, the agents of a candidate MRIO table that is filtered by network.threshold and
        cleaned on the device first, acclimate_b200/ingest.py; with network = None the table, the agents and the names are read from
        the document's network.file, acclimate_b200/mrio.py); model / sectors / firms / consumers / scenario / outputs come from
        the document (acclimate_b200/settings.py).  Returns (run, number of timesteps)."""
        from . import ingest, output, settings

        st = doc if isinstance(doc, settings.Settings) else settings.Settings(doc)
        nw = st.doc.get("network", {}) or {}
        if network is None:                                # build_agent_network, ModelInitializer.cpp:822-976: the table file
            from . import mrio

            if nw.get("type") != "netcdf" or "file" not in nw or "threshold" not in nw:
                raise settings.SettingsError("without a snapshot the document needs network.type = netcdf, network.file and network.threshold")
            tb = mrio.read_table(nw["file"])
            network, flows = mrio.seed_network(tb, st.doc.get("transport")), tb["flows"]
            sector_names, region_names = tb["sector_names"], tb["region_names"]
        net = st.apply_to(network, sector_names, region_names) if flows is None else dict(network)
        if flows is not None:
            seed = dict(network)
            seed["model_f64"] = st.model_params()
            net, _ = ingest.ingest_table(seed, flows, float(nw.get("threshold", 0.0)), device=device, lib_path=lib_path)
            net = st.apply_to(net, sector_names, region_names)
        run = cls(net, device=device, lib_path=lib_path)
        start, _, basedate = st.run_window()
        run.start_time, run.delta_t = start, float(st.doc["model"]["delta_t"])
        if st.doc["scenario"]["type"] == "event_series" and "forcing" in st.doc["scenario"]:
            from . import forcing                          # ModelRun.cpp:124-126: the forcing files of scenario.forcing

            run.scenario = forcing.EventSeriesFiles(net, sector_names, region_names, st.doc["scenario"], entity_names=entity_names,
                                                    start_time=start, delta_t=run.delta_t)
        for _, fmt, path, pulls in st.output_plan(net, sector_names, region_names):
            if fmt == "progress":
                continue
            sink = output.ArraySink() if fmt == "array" else output.NetCDFSink(path, net, sector_names, region_names, basedate=basedate)
            run.sinks.append((pulls, sink))
        return run, st.steps()

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
        new_events = e.events()                           # ModelRun::event: drained every step (the device buffer is bounded)
        self.events += new_events
        if self.sinks:
            from . import output

            for pulls, sink in self.sinks:                # ArrayOutput::iterate / NetCDFOutput::iterate: gathered on the device
                sink.write(self.start_time + self.time * self.delta_t, output.pull(e, pulls), new_events)
        e.phase(PHASE_TICK)                               # model->tick(); ++time
        self.time += 1

    def run(self, steps):
        for _ in range(steps):
            self.step()
        return {o: np.stack(v) for o, v in self.series.items()}

    def close(self):
        for _, sink in self.sinks:
            sink.close()
        if hasattr(self.scenario, "close"):
            self.scenario.close()                          # ExternalScenario::end
        self.engine.close()
