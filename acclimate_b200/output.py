"""Output sinks behind the observable pull: what the reference's ArrayOutput / NetCDFOutput do with the values.

  ArraySink   ArrayOutput with only_current_timestep = false (reference src/output/ArrayOutput.cpp:160-183, :479-646): one array
              per variable, [time, ...group shape...], NaN where nothing was written; events as rows (time, type, index1, index2, value)
  NetCDFSink  the layout of NetCDFOutput::start / ::iterate (reference src/output/NetCDFOutput.cpp:90-283): dimensions time
              (unlimited), sector, region, agent, agent_from, agent_to, event; coordinate variables; one variable per
              observable over (time, ...); the time axis in "days since <basedate>"; events.

The reference writes NetCDF-4 (groups per observable group, compound types for agents and events) through libnetcdf, which is
not on this box (no netcdf.h, no HDF5).  NetCDFSink writes NetCDF-3 classic with scipy.io.netcdf_file instead and flattens
exactly those two things: a variable of group G is named "G.<variable>", and a compound becomes one variable per field
("agent_name", "agent_type_index", "agent_sector", "agent_region"; "event_time", "event_type_index", "event_index1",
"event_index2", "event_value").  Dimension names, shapes, units and the NaN fill are the reference's.

Both sinks are fed by `pull(engine, plan_entry)`, which runs the (group, variables, select_from, select_to) pulls of a
settings.Settings.output_plan through acc_observe -- the values are gathered on the device, only the selected cells travel.
"""
from __future__ import annotations

import numpy as np

EVENT_NAMES = ["NO_CONSUMPTION", "STORAGE_UNDERRUN", "STORAGE_OVERRUN", "NO_EXP_PRODUCTION_SUPPLY_SHORTAGE", "NO_PRODUCTION_SUPPLY_SHORTAGE",
               "NO_EXP_PRODUCTION_DEMAND_QUANTITY_SHORTAGE", "NO_PRODUCTION_DEMAND_QUANTITY_SHORTAGE", "NO_EXP_PRODUCTION_DEMAND_VALUE_SHORTAGE",
               "NO_PRODUCTION_DEMAND_VALUE_SHORTAGE", "NO_EXP_PRODUCTION_HIGH_COSTS", "NO_PRODUCTION_HIGH_COSTS", "NO_PRODUCTION_PRODUCTION_CAPACITY",
               "DEMAND_FULFILL_HISTORY_UNDERFLOW", "OPTIMIZER_MAXITER", "OPTIMIZER_ROUNDOFF_LIMITED", "OPTIMIZER_TIMEOUT"]  # include/ModelRun.h:39-73

# every observable of a group, in the order of the reference's observe<>() visitors (what "no variable listed" means)
ALL_VARIABLES = {
    "firms": ["business_connections", "consumption", "demand", "input_flow", "storage", "communicated_possible_production", "desired_production_capacity",
              "direct_loss", "expected_offer_price", "expected_production", "forcing", "incoming_demand", "offer_price", "production", "production_capacity",
              "possible_production_capacity", "tax", "total_loss", "total_production_costs", "total_revenue", "total_value_loss", "unit_production_costs"],
    "consumers": ["business_connections", "consumption", "demand", "input_flow", "storage", "utility", "local_optimal_utility"],
    "sectors": ["offer_price", "total_production", "total_demand"],
    "regions": ["import", "export", "consumption", "gdp", "total_flow", "total_outflow"],
    "storages": ["business_connections", "content", "demand", "desired_used_flow", "expected_costs", "input_flow", "optimized_value", "possible_use", "purchase",
                 "shipment", "storage_demand", "total_transport_penalty", "use", "used_flow"],
    "flows": ["initial_flow", "demand_request", "flow_deficit", "flow_mean", "received_flow", "sent_flow", "total_flow", "minimum_passage"],
}
# observables that are Flows / Stocks: split into <name>_quantity and <name>_value (ArrayOutput.cpp:58-66)
PRICED = {"consumption", "demand", "input_flow", "storage", "communicated_possible_production", "direct_loss", "expected_production", "incoming_demand",
          "production", "total_loss", "total_production", "total_demand", "import", "export", "gdp", "total_flow", "total_outflow", "content",
          "desired_used_flow", "possible_use", "purchase", "shipment", "storage_demand", "use", "used_flow", "initial_flow", "demand_request", "flow_mean",
          "received_flow", "sent_flow"}


def expand(group, variables):
    """Names as acc_observe takes them: priced observables split, an empty list = all of the group."""
    out = []
    for v in (variables or ALL_VARIABLES[group]):
        if v.endswith("_quantity") or v.endswith("_value"):
            out.append(v)
        elif v in PRICED:
            out += [v + "_quantity", v + "_value"]
        else:
            out.append(v)
    return out


def shape_of(group, n_sectors, n_regions, n_agents, select_from, select_to):
    n0 = n_agents if select_from is None else len(select_from)
    n1 = n_agents if select_to is None else len(select_to)
    if group in ("firms", "consumers"):
        return (n0,)
    if group == "sectors":
        return (n_sectors if select_from is None else len(select_from),)
    if group == "regions":
        return (n_regions if select_from is None else len(select_from),)
    if group == "storages":
        return (n_sectors, n1)
    return (n0, n1)


def pull(engine, pulls):
    """One timestep's values: {(group, variable): array shaped like ArrayOutput's data of one step}."""
    out = {}
    dims = (engine.n_sectors, engine.n_regions, engine.n_agents)
    for group, variables, sel_from, sel_to in pulls:
        if group in ("model", "locations"):
            continue  # model/duration is the caller's wall clock; a snapshot has no "other locations" (ports, sea areas) as objects
        for var in expand(group, variables):
            out[(group, var)] = engine.observe(group, var, select_from=sel_from, select_to=sel_to).reshape(shape_of(group, *dims, sel_from, sel_to)).copy()
    return out


class ArraySink:
    def __init__(self):
        self.series, self.events, self.times = {}, [], []

    def write(self, time, values, events=()):
        self.times.append(time)
        for k, v in values.items():
            self.series.setdefault(k, []).append(v)
        self.events += [(time, e[1], e[2], e[3], e[4]) for e in events]

    def arrays(self):
        return {k: np.stack(v) for k, v in self.series.items()}

    def close(self):
        pass


class NetCDFSink:
    DIMS = {"firms": ("agent",), "consumers": ("agent",), "sectors": ("sector",), "regions": ("region",), "storages": ("sector", "agent"),
            "flows": ("agent_from", "agent_to")}

    def __init__(self, path, net, sector_names, region_names, basedate="2000-1-1", calendar="standard"):
        from scipy.io import netcdf_file

        self.f = f = netcdf_file(path, "w", version=2)
        nS, nR, nA = int(net["meta"][0]), int(net["meta"][1]), int(net["meta"][2])
        f.createDimension("time", None)
        for name, n in (("sector", nS), ("region", nR), ("agent", nA), ("agent_from", nA), ("agent_to", nA), ("event_type", len(EVENT_NAMES)),
                        ("agent_type", 2), ("name_len", 25)):
            f.createDimension(name, n)
        self.time = f.createVariable("time", "i", ("time",))
        self.time.units = "days since " + basedate
        self.time.calendar = calendar

        def names(var, dim, values):
            v = f.createVariable(var, "c", (dim, "name_len"))
            for i, s in enumerate(values):
                v[i] = np.frombuffer(s.encode()[:24].ljust(25, b"\0"), dtype="S1")

        names("sector", "sector", sector_names)
        names("region", "region", region_names)
        names("event_type", "event_type", EVENT_NAMES)
        names("agent_type", "agent_type", ["firm", "consumer"])
        at, asec, areg = (np.asarray(net[k], dtype=np.int64) for k in ("agent_type", "agent_sector", "agent_region"))
        names("agent_name", "agent", [sector_names[s] + ":" + region_names[r] for s, r in zip(asec, areg)])
        for var, vals in (("agent_type_index", np.where(at == 1, 0, 1)), ("agent_sector", asec), ("agent_region", areg)):
            v = f.createVariable(var, "i", ("agent",))
            v[:] = vals.astype(np.int32)
        self.vars, self.step, self.events = {}, 0, []
        self.sel = {}

    def write(self, time, values, events=()):
        f = self.f
        self.time[self.step] = int(time)
        for (group, var), data in values.items():
            key = group + "." + var
            if key not in self.vars:
                dims = list(self.DIMS[group])
                for d, n in zip(range(len(dims)), data.shape):  # a selection shortens an axis: its own dimension, as NetCDFOutput does for select_*
                    full = f.dimensions[dims[d]]
                    if n != full:
                        dn = "%s_%s_%d" % (group, dims[d], n)
                        if dn not in f.dimensions:
                            f.createDimension(dn, n)
                        dims[d] = dn
                v = f.createVariable(key, "d", ("time",) + tuple(dims))
                v._FillValue = np.nan
                self.vars[key] = v
            self.vars[key][self.step] = data
        self.events += [(int(time), e[1], e[2], e[3], e[4]) for e in events]
        self.step += 1

    def close(self):
        f = self.f
        n = len(self.events)
        f.createDimension("event", max(n, 1))
        cols = list(zip(*self.events)) if n else [[0], [0], [0], [0], [np.nan]]
        for name, typ, col in (("event_time", "i", cols[0]), ("event_type_index", "i", cols[1]), ("event_index1", "i", cols[2]), ("event_index2", "i", cols[3]),
                               ("event_value", "d", cols[4])):
            v = f.createVariable(name, typ, ("event",))
            v[:] = np.asarray(col)
        f.n_events = n
        f.close()
