"""The reference's `settings.yml` surface in front of the engine ("keeping the ... settings.yml ... surface").

The reference reads ONE YAML document (reference src/main.cpp:58-124 -> ModelRun::ModelRun, src/ModelRun.cpp:77-150):

  model:      Parameters::ModelParameters           src/input/ModelInitializer.cpp:58-62, 1015-1079
  sectors:    <name> | ALL -> sector parameters     :145-185 (get_named_property: the sector's own node first, then ALL)
  firms:      <sector:region> | <sector> | <region> | ALL -> possible_overcapacity_ratio      :74-106
  consumers:  <name> | <region> | ALL -> consumer_baskets, inter_basket_substitution_coefficient, bool_utilitarian,
              consumption_price_elasticity           :108-143
  network:    where the MRIO table comes from (NetCDF in the reference); here a snapshot file or a seeded synthetic table
  scenario:   type events | event_series, start, stop, basedate, events: [{type: shock, from, to, targets: [...]}]
                                                     src/ModelRun.cpp:99-130, src/scenario/Scenario.cpp:40-174
  outputs:    [{format: netcdf | array | progress, file, observables: {<group>: {select_*: [...], ...variables}}}]
                                                     src/ModelRun.cpp:132-150, src/output/ArrayOutput.cpp:185-470

`Settings(...)` validates the document the way the reference's readers do -- a key the reference does not read is an error
here (the reference silently ignores it; a typo then changes nothing, which is the worse failure) and an option whose
implementation this engine does not provide is refused by name -- and turns it into what the C ABI takes: the
`model_f64` / `sector_f64` rows of a snapshot, per-agent `agent_beta`, the consumers' baskets, a forcing table with the
rule of Scenario::iterate, and an output plan of (group, variable, selection) pulls for acc_observe.
"""
from __future__ import annotations

import numpy as np

from . import netfile

try:  # PyYAML is on the box; the surface also accepts an already parsed dict
    import yaml
except ImportError:  # pragma: no cover
    yaml = None


class SettingsError(ValueError):
    pass


# model.* keys the reference reads (ModelInitializer.cpp:58-62, 1015-1079) -> (snapshot name or None, default or REQUIRED)
REQUIRED = object()
MODEL_KEYS = {
    "delta_t": ("delta_t", REQUIRED),
    "no_self_supply": ("no_self_supply", True),
    "transport_penalty_small": ("transport_penalty_small", REQUIRED),
    "transport_penalty_large": ("transport_penalty_large", REQUIRED),
    "optimization_maxiter": ("optimization_maxiter", REQUIRED),
    "optimization_timeout": (None, REQUIRED),  # wall-clock limit of one NLopt call: a device solve has no clock (maxiter bounds it)
    "quadratic_transport_penalty": ("quadratic_transport_penalty", True),
    "maximal_decrease_reservation_price_limited_by_markup": ("maximal_decrease_reservation_price_limited_by_markup", False),
    "always_extend_expected_demand_curve": ("always_extend_expected_demand_curve", False),
    "naive_expectations": ("naive_expectations", True),
    "deviation_penalty": ("deviation_penalty", False),
    "min_storage": ("min_storage", 0.0),
    "cheapest_price_range_preserve_seller_price": ("cheapest_price_range_preserve_seller_price", False),
    "cheapest_price_range_width": ("cheapest_price_range_width", REQUIRED),  # a Price or "auto"
    "relative_transport_penalty": ("relative_transport_penalty", True),
    "optimization_precision_adjustment": ("optimization_precision_adjustment", 1.0),
    "debug_purchasing_steps": (None, []),
}
# options whose implementation the engine does not provide: accepted only at the reference's default (value, why)
MODEL_FIXED = {
    "optimization_algorithm": ("slsqp", "the purchasing solver is the on-device SQP that replaces NLopt's SLSQP"),
    "utility_optimization_algorithm": ("slsqp", "the consumers' solver is the on-device dense SQP"),
    "cost_correction": (False, "the reference refuses it too (ModelInitializer.cpp:1036-1038)"),
    "global_purchasing_optimization": (False, "global (augmented-Lagrangian) purchasing optimisation is not provided"),
    "local_purchasing_optimization": (True, "the local SLSQP optimisation is the only purchasing mode"),
    "optimization_restart_baseline": (False, "restart at baseline demand is not provided"),
    "start_purchasing_at_baseline": (False, "not provided"),
    "purchasing_halfway_baseline": (False, "not provided"),
    "global_utility_optimization": (False, "the global MLSL utility search is not provided"),
    "budget_inequality_constrained": (False, "only the equality budget constraint is provided"),
    "elastic_budget": (False, "not provided"),
}
# read by the reference only for modes that are refused above: tolerated, unused
MODEL_IGNORED = {"global_sampling_points", "global_optimization_algorithm", "lagrangian_optimization_algorithm", "global_utility_optimization_algorithm",
                 "utility_optimization_maxiter", "utility_optimization_timeout", "global_utility_optimization_maxiter", "global_utility_optimization_timeout",
                 "utility_optimization_precision_adjustment", "global_optimization_precision_adjustment", "global_utility_optimization_precision_adjustment",
                 "global_optimization_maxiter", "global_optimization_timeout"}
SECTOR_KEYS = {  # sectors.<name>|ALL.<key> (ModelInitializer.cpp:159-180) -> snapshot name
    "upper_storage_limit": "upper_storage_limit", "initial_storage_fill_factor": "initial_storage_fill_factor", "transport": "transport_type",
    "supply_elasticity": "supply_elasticity", "price_increase_production_extension": "price_increase_production_extension",
    "estimated_price_increase_production_extension": "estimated_price_increase_production_extension", "initial_markup": "initial_markup",
    "target_storage_refill_time": "target_storage_refill_time", "target_storage_withdraw_time": "target_storage_withdraw_time",
}
TRANSPORT_TYPES = {"aviation": 0, "immediate": 1, "roadsea": 2}  # Sector::map_transport_type, src/model/Sector.cpp:65-78
FIRM_KEYS = {"possible_overcapacity_ratio"}
CONSUMER_KEYS = {"consumer_baskets", "inter_basket_substitution_coefficient", "bool_utilitarian", "consumption_price_elasticity"}
SCENARIO_KEYS = {"type", "start", "stop", "basedate", "baseyear", "events", "seed", "forcings", "forcing", "parameters"}
OUTPUT_KEYS = {"format", "file", "observables", "events", "flush", "only_current_timestep"}
GROUPS = ("model", "firms", "consumers", "sectors", "regions", "locations", "storages", "flows")
SELECTORS = {  # ArrayOutput.cpp:200-470: which select_* keys a group takes and whether they pick the first or second index
    "firms": {"select_sector": 0, "select_region": 0}, "consumers": {"select_region": 0}, "sectors": {"select_sector": 0},
    "regions": {"select_region": 0}, "storages": {"select_sector_to": 1, "select_region_to": 1, "select_agent_to": 1},
    "flows": {"select_sector_from": 0, "select_region_from": 0, "select_sector_to": 1, "select_region_to": 1},
    "model": {}, "locations": {},
}


def _load(doc):
    if isinstance(doc, dict):
        return doc
    if yaml is None:
        raise SettingsError("PyYAML is not available: pass the parsed settings as a dict")
    if hasattr(doc, "read"):
        return yaml.safe_load(doc)
    text = str(doc)
    if "\n" not in text and (text.endswith(".yml") or text.endswith(".yaml")):
        with open(text) as fp:
            return yaml.safe_load(fp)
    return yaml.safe_load(text)


def _named(node, name, key, what):
    """get_named_property (ModelInitializer.cpp:64-72): the node's own entry first, then ALL."""
    if name in node and isinstance(node[name], dict) and key in node[name]:
        return node[name][key]
    if "ALL" in node and key in node["ALL"]:
        return node["ALL"][key]
    raise SettingsError(f"{what}.{name}.{key} (or {what}.ALL.{key}) is required")


class Settings:
    """One parsed and validated settings document."""

    def __init__(self, doc):
        self.doc = _load(doc)
        if not isinstance(self.doc, dict):
            raise SettingsError("settings must be a mapping")
        known = {"model", "sectors", "firms", "consumers", "network", "transport", "scenario", "scenarios", "outputs", "output", "run", "checkpoint"}
        for k in self.doc:
            if k not in known:
                raise SettingsError(f"unknown top-level key '{k}'")
        for k in ("model", "sectors", "firms", "consumers", "scenario"):
            if k not in self.doc:
                raise SettingsError(f"settings lack the '{k}' section")
        self._check_model()

    # ---- model ----
    def _check_model(self):
        m = self.doc["model"]
        for k in m:
            if k not in MODEL_KEYS and k not in MODEL_FIXED and k not in MODEL_IGNORED:
                raise SettingsError(f"unknown key model.{k} (the reference reads no such parameter)")
        for k, (want, why) in MODEL_FIXED.items():
            if k in m and m[k] != want:
                raise SettingsError(f"model.{k} = {m[k]!r}: {why}")
        for k, (_, default) in MODEL_KEYS.items():
            if default is REQUIRED and k not in m:
                raise SettingsError(f"model.{k} is required")

    def model_params(self, utilitarian=True):
        """The `model_f64` row of a snapshot (netfile.MODEL_PARAMS order)."""
        m = self.doc["model"]
        out = {}
        for k, (name, default) in MODEL_KEYS.items():
            if name is None:
                continue
            v = m.get(k, default)
            if k == "cheapest_price_range_width":
                v = -1.0 if v == "auto" else float(v)  # ModelInitializer.cpp:1028-1031
            out[name] = float(v)
        out["utilitarian"] = 1.0 if utilitarian else 0.0
        return np.array([out[k] for k in netfile.MODEL_PARAMS], dtype=np.float64)

    # ---- sectors / firms / consumers ----
    def sector_params(self, sector_names):
        """`sector_f64` rows for the named sectors (consumer sector included: it has storages of its own on no one)."""
        node = self.doc["sectors"]
        for name, sub in node.items():
            if isinstance(sub, dict):
                for k in sub:
                    if k not in SECTOR_KEYS:
                        raise SettingsError(f"unknown key sectors.{name}.{k}")
        rows = []
        for s in sector_names:
            row = {}
            for key, snap in SECTOR_KEYS.items():
                if key == "estimated_price_increase_production_extension":
                    try:
                        v = _named(node, s, key, "sectors")
                    except SettingsError:  # defaults to price_increase_production_extension, ModelInitializer.cpp:166-169
                        v = _named(node, s, "price_increase_production_extension", "sectors")
                else:
                    v = _named(node, s, key, "sectors")
                if key == "transport":
                    if v not in TRANSPORT_TYPES:
                        raise SettingsError(f"sectors.{s}.transport = {v!r}: unknown transport type")
                    v = TRANSPORT_TYPES[v]
                row[snap] = float(v)
            rows.append([row[k] for k in netfile.SECTOR_PARAMS])
        return np.array(rows, dtype=np.float64)

    def firm_property(self, sector, region, key="possible_overcapacity_ratio"):
        """get_firm_property (ModelInitializer.cpp:74-93): name, sector:region, sector, region, ALL."""
        if key not in FIRM_KEYS:
            raise SettingsError(f"unknown firm property {key}")
        node = self.doc["firms"]
        for k in (f"{sector}:{region}", sector, region, "ALL"):
            if k in node and isinstance(node[k], dict) and key in node[k]:
                return node[k][key]
        raise SettingsError(f"firms.ALL.{key} is required")

    def consumer_property(self, name, region, key, default=REQUIRED):
        """get_consumer_property (ModelInitializer.cpp:108-119): name, region, ALL."""
        if key not in CONSUMER_KEYS:
            raise SettingsError(f"unknown consumer property {key}")
        node = self.doc["consumers"]
        for k in (name, region, "ALL"):
            if k in node and isinstance(node[k], dict) and key in node[k]:
                return node[k][key]
        if default is REQUIRED:
            raise SettingsError(f"consumers.ALL.{key} is required")
        return default

    def consumer_baskets(self, sector_names):
        """basket_of_sector / basket_coeff / inter_basket_coeff / utilitarian of a snapshot (ModelInitializer.cpp:121-139; the
        engine takes ONE set of baskets for all consumers, so consumers.ALL decides)."""
        util = bool(self.consumer_property("", "", "bool_utilitarian", True))
        if not util:
            return None
        baskets = self.consumer_property("", "", "consumer_baskets")
        of = np.full(len(sector_names), -1, dtype=np.int64)
        coeff = []
        for b, basket in enumerate(baskets):
            for s in basket["sectors"]:
                if s not in sector_names:
                    raise SettingsError(f"consumer basket names an unknown sector '{s}'")
                of[sector_names.index(s)] = b
            coeff.append(float(basket["substituition_coefficient"]))  # the reference's own spelling of the key (:130)
        inter = float(self.consumer_property("", "", "inter_basket_substitution_coefficient"))
        return dict(basket_of_sector=of, basket_coeff=np.array(coeff), inter_basket_coeff=np.array([inter]))

    # ---- applying the document to a snapshot ----
    def apply_to(self, net, sector_names, region_names):
        """Overwrites the parameter rows of the snapshot `net` (synth.make_network / netfile.read layout) with this document's
        and adds the scenario's forcing table; the topology (agents, connections, Z*) stays what the network source gave."""
        net = dict(net)
        nS, nR = int(net["meta"][0]), int(net["meta"][1])
        if len(sector_names) != nS or len(region_names) != nR:
            raise SettingsError("sector / region names do not match the snapshot")
        baskets = self.consumer_baskets(sector_names)
        net["model_f64"] = self.model_params(utilitarian=baskets is not None)
        net["sector_f64"] = self.sector_params(sector_names).reshape(-1)
        asec, areg, atype = (np.asarray(net[k], dtype=np.int64) for k in ("agent_sector", "agent_region", "agent_type"))
        beta = np.asarray(net["agent_beta"], dtype=np.float64).copy()
        for a in np.nonzero(atype == 1)[0]:
            beta[a] = float(self.firm_property(sector_names[asec[a]], region_names[areg[a]]))
        net["agent_beta"] = beta
        if baskets is not None:
            net.update(baskets)
        else:
            eps = float(self.consumer_property("", "", "consumption_price_elasticity", -0.5))
            net["consumer_elasticity"] = np.full(nS, eps)
        forcing, passage = self.scenario_tables(net, sector_names, region_names)
        net["forcing"], net["passage"] = forcing.reshape(-1), passage.reshape(-1)
        return net

    # ---- scenario ----
    def run_window(self):
        sc = self.doc["scenario"]
        for k in sc:
            if k not in SCENARIO_KEYS:
                raise SettingsError(f"unknown key scenario.{k}")
        if sc.get("type") not in ("events", "event_series"):
            raise SettingsError(f"Unknown scenario type '{sc.get('type')}'")  # src/ModelRun.cpp:127-129
        if "start" not in sc or "stop" not in sc:
            raise SettingsError("scenario.start and scenario.stop are required")
        basedate = sc.get("basedate", (str(sc["baseyear"]) + "-1-1") if "baseyear" in sc else "2000-1-1")  # :102-106
        return float(sc["start"]), float(sc["stop"]), basedate

    def steps(self):
        """Number of timesteps of the run: ModelRun::run iterates while time <= stop (src/ModelRun.cpp:170-228)."""
        start, stop, _ = self.run_window()
        dt = float(self.doc["model"]["delta_t"])
        return int(np.floor((stop - start) / dt)) + 1

    def scenario_tables(self, net, sector_names, region_names):
        """`forcing` rows (from, to, agent, lambda) and `passage` rows (from, to, geo entity, nu) for engine.Scenario, from
        scenario.events with the target grammar of Scenario::apply_target (src/scenario/Scenario.cpp:79-158)."""
        sc = self.doc["scenario"]
        self.run_window()
        asec, areg, atype = (np.asarray(net[k], dtype=np.int64) for k in ("agent_sector", "agent_region", "agent_type"))
        beta = np.asarray(net["agent_beta"], dtype=np.float64)
        nR = len(region_names)
        forcing, passage = [], []
        if sc["type"] == "event_series":
            return np.zeros((0, 4)), np.zeros((0, 4))  # records are streamed: forcing.EventSeriesFiles / engine.EventSeriesScenario
        for ev in sc.get("events", []) or []:
            if ev.get("type") != "shock":
                continue  # Scenario::iterate only acts on shocks (:163-172)
            t0, t1 = float(ev["from"]), float(ev["to"])
            for targets in ev["targets"]:
                for kind, it in targets.items():
                    if kind == "firm":
                        sel = atype == 1
                        if "id" in it:
                            s, r = it["id"].split(":")
                            sel &= (asec == self._index(sector_names, s, "Sector")) & (areg == self._index(region_names, r, "Region"))
                        else:
                            if "sector" in it:
                                sel &= asec == self._index(sector_names, it["sector"], "Sector")
                            if "region" in it:
                                sel &= areg == self._index(region_names, it["region"], "Region")
                        idx = np.nonzero(sel)[0]
                        if idx.size == 0:
                            raise SettingsError(f"scenario target {it} matches no firm")
                        if "remaining_capacity" in it:  # Scenario.cpp:49-50: forcing = remaining capacity / beta
                            forcing += [(t0, t1, a, float(it["remaining_capacity"]) / beta[a]) for a in idx]
                        elif "forcing" in it:
                            forcing += [(t0, t1, a, float(it["forcing"])) for a in idx]
                    elif kind == "consumer":
                        sel = atype == 0
                        if "region" in it:
                            sel &= areg == self._index(region_names, it["region"], "Region")
                        if "remaining_consumption_rate" in it:
                            forcing += [(t0, t1, a, float(it["remaining_consumption_rate"])) for a in np.nonzero(sel)[0]]
                    elif kind == "sea":
                        # a sea route of the reference's network transport; here: "<region>-<region>", the connection entity
                        if "sea_route" in it and "passage" in it:
                            ra, rb = (self._index(region_names, x, "Region") for x in it["sea_route"].split("-"))
                            hi, lo = max(ra, rb), min(ra, rb)
                            passage.append((t0, t1, hi * (hi - 1) // 2 + lo, float(it["passage"])))
                    else:
                        raise SettingsError(f"unknown scenario target type '{kind}'")
        del nR
        return np.array(forcing, dtype=np.float64).reshape(-1, 4), np.array(passage, dtype=np.float64).reshape(-1, 4)

    @staticmethod
    def _index(names, name, what):
        if name not in names:
            raise SettingsError(f"{what} {name} not found")
        return names.index(name)

    # ---- outputs ----
    def output_plan(self, net, sector_names, region_names):
        """[(output entry index, format, file, [(group, variable, select_from, select_to)])]: what to pull with acc_observe after every
        step, following ArrayOutput's constructor (src/output/ArrayOutput.cpp:185-470): per group the listed variables (all of the
        group when none is listed) and the agents picked by its select_* keys."""
        asec, areg, atype = (np.asarray(net[k], dtype=np.int64) for k in ("agent_sector", "agent_region", "agent_type"))
        plan = []
        for i, node in enumerate(self.doc.get("outputs", []) or []):
            for k in node:
                if k not in OUTPUT_KEYS:
                    raise SettingsError(f"unknown key outputs[{i}].{k}")
            fmt = node.get("format")
            if fmt not in ("netcdf", "array", "progress"):
                raise SettingsError(f"Unknown output format '{fmt}'")  # src/ModelRun.cpp:146-148
            pulls = []
            for group, obs in (node.get("observables") or {}).items():
                if group not in GROUPS:
                    raise SettingsError(f"unknown observable group '{group}'")
                obs = obs or {}
                sel = [None, None]
                variables = []
                for key, val in obs.items():
                    if key.startswith("select_"):
                        if key not in SELECTORS[group]:
                            raise SettingsError(f"outputs[{i}].observables.{group}.{key}: no such selection")
                        which = SELECTORS[group][key]
                        picked = []
                        for name in val:
                            if "sector" in key and group == "sectors":
                                picked.append(self._index(sector_names, name, "Sector"))
                            elif "region" in key and group == "regions":
                                picked.append(self._index(region_names, name, "Region"))
                            elif "sector" in key:  # all firms of the sector, ArrayOutput.cpp:436-447
                                picked += list(np.nonzero((asec == self._index(sector_names, name, "Sector")) & (atype == 1))[0])
                            elif "region" in key:  # all agents of the region, :449-459
                                picked += list(np.nonzero(areg == self._index(region_names, name, "Region"))[0])
                        # indices in the order the names are listed, appended to what an earlier select_* of the same axis picked
                        # (ArrayOutput.cpp:436-459 push_back, no sorting)
                        sel[which] = np.array((sel[which].tolist() if sel[which] is not None else []) + [int(x) for x in picked], dtype=np.int64)
                    elif key == "variables":
                        variables += list(val)
                    else:
                        variables.append(key)
                pulls.append((group, variables, sel[0], sel[1]))
            plan.append((i, fmt, node.get("file"), pulls))
        return plan
