"""Generates the golden fixtures of tests/golden/ by running the REFERENCE's own model code (built from /root/reference
by oracle/Makefile) on seeded synthetic networks.  Run here (where /root/reference exists):

    python tests/golden/make_golden.py [case ...]

Every case exists in two flavours, which differ only in the NLopt provider linked behind the reference's optimiser calls
(reference include/optimization.h:123-190):

  <case>.shim.acc.gz   oracle/_ref/acclimate_ref        provider oracle/nlopt_shim = the product's own solver source.  Pins
                                                        everything AROUND the optimiser bit for bit (tests/parity.py).
  <case>.kraft.acc.gz  oracle/_ref/acclimate_ref_kraft  provider oracle/nlopt_kraft = an independent restatement of Kraft's
                                                        SLSQP (includes nothing of the product).  The CUDA run is
                                                        measured against it (tests/kraft.py, tests/golden/kraft_deviation.json).

A fixture is a gzip-compressed ACCARR1 file (acclimate_b200/netfile.py): the network snapshot under "net/<name>", the
reference's per-step state under "tNNNNN/<name>" (every array the harness dumps, per-connection arrays included), the
event list, and "stamp/<file>": the git blob hash of every solver source the flavour depends on -- the tests refuse a
fixture whose stamp does not match the sources they run (a solver change must regenerate the fixtures, visibly).
Long runs are stored compactly: full arrays at a few steps, position-weighted checksums at all others.
"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from acclimate_b200 import netfile, synth  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
ORACLE = {"shim": os.path.join(ROOT, "oracle", "_ref", "acclimate_ref"), "kraft": os.path.join(ROOT, "oracle", "_ref", "acclimate_ref_kraft")}
# solver sources a flavour's numbers depend on (beyond the reference's own code)
SOLVER_SOURCES = {
    "shim": ["acclimate_b200/csrc/sqp_core.h", "acclimate_b200/csrc/sqp_dense.h", "acclimate_b200/csrc/coop.cuh", "acclimate_b200/csrc/purchase_core.cuh",
             "oracle/nlopt_shim/nlopt_shim.cpp", "oracle/ref_harness/harness.cpp"],
    "kraft": ["oracle/nlopt_kraft/slsqp_kraft.cpp", "oracle/nlopt_kraft/nlopt_kraft.cpp", "oracle/ref_harness/harness.cpp"],
}
# what the kraft flavour keeps (the fields north_star names: flows, prices, demand requests; plus production and storage)
KRAFT_KEEP = ("edge_D_q", "edge_D_p", "edge_ship_q", "firm_X_q", "firm_nbar", "sto_S_q", "sto_demand_q", "sto_optimized_value", "region_gdp_v",
              "region_consumption_q", "sector_production_q")


def blob_hash(path):
    """git's blob hash of a file (what `git hash-object` prints), as 20 small integers."""
    data = open(os.path.join(ROOT, path), "rb").read()
    h = hashlib.sha1(b"blob %d\0" % len(data) + data).digest()
    return np.frombuffer(h, dtype=np.uint8).astype(np.int64)


def stamps(flavour):
    return {"stamp/" + p: blob_hash(p) for p in SOLVER_SOURCES[flavour]}


def checksum(arr):
    """Position-weighted checksum of one array: int64 arrays exactly (mod 2^64), f64 arrays as two weighted sums."""
    a = np.asarray(arr)
    if a.dtype.kind == "i":
        w = (2 * np.arange(a.size, dtype=np.uint64) + 1)
        return np.array([np.sum(a.astype(np.uint64) * w, dtype=np.uint64)]).view(np.int64)
    w = 0.5 + (np.arange(a.size) * 0.6180339887498949) % 1.0
    return np.array([a.sum(), (a * w).sum()])


def all_firms(net, sector=None, region=None):
    at, asec, ar = np.asarray(net["agent_type"]), np.asarray(net["agent_sector"]), np.asarray(net["agent_region"])
    m = at == 1
    if sector is not None:
        m &= asec == sector
    if region is not None:
        m &= ar == region
    return np.nonzero(m)[0]


def add_forcing(net, t_from, t_to, agents, value):
    rows = np.stack([np.full(len(agents), float(t_from)), np.full(len(agents), float(t_to)), np.asarray(agents, dtype=np.float64), np.full(len(agents), float(value))], axis=1)
    net["forcing"] = np.concatenate([np.asarray(net["forcing"], dtype=np.float64).reshape(-1, 4), rows]).reshape(-1)


def cases():
    """(name, network, steps, options).  options: full_steps = list of steps kept in full (others as checksums);
    no_edges = per-connection arrays left out; no_kraft = no kraft flavour (dense SLSQP in 199 unknowns takes hours)."""
    a = synth.make_config("small")
    synth.add_shock(a, 2, 6, 3, 0.5)
    yield "small_shock12", a, 12, {}
    a = synth.make_network(regions=8, sectors=4, seed=7, max_delay=6, model=dict(utilitarian=0))  # price-elastic consumers
    synth.add_shock(a, 3, 12, 2, 0.05)
    yield "severe_shock30", a, 30, {}
    a = synth.make_network(regions=8, sectors=4, seed=8, max_delay=5)
    synth.add_passage(a, 2, 9, 5, 1, 0.3)
    synth.add_passage(a, 4, 6, 3, 2, 0.0)
    yield "passage25", a, 25, {}
    a = synth.make_network(regions=6, sectors=3, seed=9, const_delay=2, model=dict(utilitarian=0))
    synth.add_shock(a, 2, 5, 1, 0.0)
    yield "outage25", a, 25, {}
    a = synth.make_network(regions=6, sectors=4, seed=10, const_delay=1,
                           model=dict(quadratic_transport_penalty=0, relative_transport_penalty=0, cheapest_price_range_width=0.01, min_storage=0.2,
                                      deviation_penalty=1, maximal_decrease_reservation_price_limited_by_markup=1))
    synth.add_shock(a, 1, 8, 0, 0.3)
    yield "flags20", a, 20, {}
    # utilitarian consumers (the reference's default, ModelInitializer.cpp:136): nested-CES utility optimisation, one basket
    a = synth.make_network(regions=8, sectors=4, seed=11, max_delay=4, model=dict(utilitarian=1))
    synth.add_shock(a, 2, 9, 3, 0.3)
    yield "utilitarian20", a, 20, {}
    # three baskets with different substitution coefficients, one sector outside every basket, severe shock + blocked passage
    a = synth.make_network(regions=6, sectors=7, seed=12, max_delay=3, model=dict(utilitarian=1))
    a["basket_of_sector"] = np.array([0, 0, 1, 2, 2, -1, 1, -1])
    a["basket_coeff"] = np.array([2.0, 0.7, 4.0])
    a["inter_basket_coeff"] = np.array([1.3])
    synth.add_shock(a, 1, 7, 2, 0.1)
    synth.add_passage(a, 3, 8, 4, 1, 0.2)
    yield "baskets20", a, 20, {}
    # multi-hop routes ("network" transport): access connections, zero-delay hub locations, sea lanes; passage forcing
    # on a zero-delay hub, a blocked lane and a throttled access connection
    a = synth.make_network(regions=7, sectors=4, seed=21, max_delay=3)
    info = synth.add_routes(a, seed=21, hubs=3)
    synth.add_shock(a, 2, 8, 3, 0.3)
    synth.add_entity_passage(a, 3, 9, info["hub_loc"][0], 0.4)
    synth.add_entity_passage(a, 5, 7, list(info["lane"].values())[0], 0.0)
    synth.add_entity_passage(a, 4, 10, info["access"][2], 0.6)
    yield "routes25", a, 25, {}
    # naive_expectations = false: sellers extend their expected demand curve with the supply elasticity and look for
    # the production where marginal costs meet marginal revenue (bisection); second half: always_extend as well
    a = synth.make_network(regions=6, sectors=4, seed=14, max_delay=3, model=dict(naive_expectations=0),
                           sector=dict(supply_elasticity=0.4))
    synth.add_shock(a, 2, 10, 1, 0.2)
    yield "expectations20", a, 20, {}
    a = synth.make_network(regions=5, sectors=3, seed=15, const_delay=2, model=dict(naive_expectations=0, always_extend_expected_demand_curve=1))
    synth.add_shock(a, 3, 8, 0, 0.5)
    yield "extend15", a, 15, {}
    # 200 regions x 1 sector: every purchasing manager has 199 connections, more than two per thread of the base solver
    # group, so engine and oracle switch to the larger group (accsqp_group_for); per-connection arrays are left out
    a = synth.make_network(regions=200, sectors=1, seed=13, max_delay=3)
    synth.add_shock(a, 1, 4, 7, 0.4)
    yield "wide8", a, 8, {"no_edges": True, "no_kraft": True}
    # ---- round 2 ----
    # BASELINE.json configs[0] as specified: 10 x 5, 365 daily steps, regional shock from step 10 to 24 and the recovery
    a = synth.make_config("small")
    synth.add_shock(a, 10, 24, 3, 0.5)
    yield "config0_365", a, 365, {"full_steps": [0, 9, 10, 11, 24, 25, 40, 100, 364]}
    # the two transport-penalty variants no other case has (PurchasingManager.cpp:381-451): quadratic + absolute, linear + relative
    a = synth.make_network(regions=6, sectors=4, seed=16, const_delay=2, model=dict(quadratic_transport_penalty=1, relative_transport_penalty=0))
    synth.add_shock(a, 1, 7, 2, 0.3)
    yield "penalty_quad_abs15", a, 15, {}
    a = synth.make_network(regions=6, sectors=4, seed=17, const_delay=2, model=dict(quadratic_transport_penalty=0, relative_transport_penalty=1))
    synth.add_shock(a, 1, 7, 2, 0.3)
    yield "penalty_lin_rel15", a, 15, {}
    # optimization_maxiter = 3: OPTIMIZER_MAXITER events and the restart loop of run_optimizer (PurchasingManager.cpp:465-546)
    a = synth.make_network(regions=6, sectors=4, seed=18, const_delay=2, model=dict(optimization_maxiter=3))
    synth.add_shock(a, 1, 6, 1, 0.2)
    yield "maxiter12", a, 12, {}
    # events: one sector stops producing everywhere while its buyers hold little stock of it -> storages run empty
    # (STORAGE_UNDERRUN), firms cannot produce (NO_PRODUCTION_SUPPLY_SHORTAGE and its expectation twin), price-elastic
    # consumers find nothing to consume (NO_CONSUMPTION); a region throttled to 10 % keeps receiving what it ordered into
    # storages capped at 1.2 x baseline (STORAGE_OVERRUN)
    a = synth.make_network(regions=5, sectors=3, seed=19, const_delay=1, model=dict(utilitarian=0),
                           sector=dict(initial_storage_fill_factor=2.0, upper_storage_limit=1.2))
    add_forcing(a, 1, 14, all_firms(a, sector=1), 0.0)
    add_forcing(a, 3, 12, all_firms(a, region=2), 0.1 / 1.15)
    # no kraft flavour: with nothing to consume the bounds of the consumers' (discarded) utility optimisation collapse below
    # its start value, which NLopt -- and oracle/nlopt_kraft with it -- rejects as NLOPT_INVALID_ARGS: the reference aborts the
    # run there ("optimization failed, ... (for 3 consumption goods)", Consumer.cpp:497-502); the shim clips the start value
    yield "events18", a, 18, {"no_kraft": True}
    # the OUTPUT SURFACE: every observable of the reference's observe<>() visitors (firms, consumers, sectors, regions, storages,
    # flows), collected by the oracle the way ArrayOutput does, at three steps of a run with a shock, multi-hop routes and
    # passage forcing on them (overflow, minimum passage < 1) and utilitarian consumers
    a = synth.make_network(regions=5, sectors=3, seed=23, max_delay=3)
    info = synth.add_routes(a, seed=23, hubs=2)
    synth.add_shock(a, 1, 6, 2, 0.4)
    synth.add_entity_passage(a, 2, 7, info["hub_loc"][0], 0.5)
    synth.add_entity_passage(a, 3, 5, info["access"][1], 0.3)
    yield "observe10", a, 10, {"no_kraft": True, "observe": True, "full_steps": [0, 4, 9]}
    if "flood40_1095" in sys.argv[1:]:  # only on request: a quarter of an hour of oracle time
        # BASELINE.json configs[2] as specified -- regional production loss from step 10 to 24, THREE YEARS of daily steps -- on a
        # 40 x 26 cut of the EORA generator (1.1 M connections; the full network would keep the serial reference busy for a day):
        # every 15th step as checksums of all per-storage / per-firm / region / sector arrays, plus the recovery series in full
        yield "flood40_1095", synth_network("flood40_1095"), 1095, {"no_kraft": True, "full_steps": [], "synth": True, "no_edges": True,
                                                                    "oracle_args": ["--every", "15", "--no-edges"], "keep_full": ["region_gdp_v", "firm_X_q"]}
    if "eora_flood4" in sys.argv[1:]:  # only on request: minutes of oracle time, 20 GB of memory
        yield "eora_flood4", synth_network("eora_flood4"), 4, {"no_kraft": True, "full_steps": [], "synth": True}


def synth_network(name):
    """Networks too large to store in a fixture are rebuilt from the seeded generator by name."""
    if name == "eora_flood4":
        # BASELINE.json configs[2] at FULL size (187 x 26, 24.5 M connections) through the shock onset: region 17 at half capacity
        # from step 1 on.  Only checksums of the oracle's arrays are kept (the dump is 7 GB); the serial oracle needs ~6 minutes.
        a = synth.make_config("eora")
        synth.add_shock(a, 1, 3, 17, 0.5)
        return a
    if name == "flood40_1095":
        a = synth.make_network(regions=40, sectors=26, seed=20260101 + 3)
        synth.add_shock(a, 10, 24, 17, 0.5)
        return a
    raise KeyError(name)


def run_oracle(flavour, net, steps, tmp, observe=False, extra=()):
    npath, dpath = os.path.join(tmp, "n.acc"), os.path.join(tmp, "d_%s.acc" % flavour)
    if not os.path.exists(npath):
        netfile.write(npath, net)
    r = subprocess.run([ORACLE[flavour], npath, str(steps), dpath] + (["--observe"] if observe else []) + list(extra), capture_output=True, text=True,
                       env=dict(os.environ, ORACLE_SOLVER_STATS="1"))
    if r.returncode != 0:
        raise RuntimeError("oracle (%s) failed: %s" % (flavour, r.stderr[-2000:]))
    return netfile.read(dpath), r.stderr.strip().splitlines()[-1]


def build_fixture(flavour, net, steps, dump, opts):
    out = {}
    if flavour == "shim" and not opts.get("synth"):
        out.update({"net/" + k: v for k, v in net.items()})
    out["steps"] = [steps]
    out.update(stamps(flavour))
    full = opts.get("full_steps")
    for k, v in dump.items():
        if "/" not in k:
            out[k] = v  # events, event_values
            continue
        t, base = k.split("/")
        if base.startswith("edge_") and opts.get("no_edges"):
            continue
        if flavour == "kraft" and base not in KRAFT_KEEP:
            continue
        if full is None or int(t[1:]) in full or base in opts.get("keep_full", ()):
            out[k] = v
        elif base.startswith("obs:"):
            continue  # observables are kept at the full steps only
        else:
            out["chk/" + k] = checksum(v)
    if full is not None:
        out["full_steps"] = np.array(full, dtype=np.int64)
    return out


def main():
    want = set(sys.argv[1:])
    spath = os.path.join(GOLDEN, "SUMMARY.json")
    summary = json.load(open(spath)) if os.path.exists(spath) else {}
    for name, net, steps, opts in cases():
        if want and name not in want:
            continue
        with tempfile.TemporaryDirectory() as tmp:
            for flavour in ("shim", "kraft"):
                if flavour == "kraft" and opts.get("no_kraft"):
                    continue
                dump, stats = run_oracle(flavour, net, steps, tmp, observe=bool(opts.get("observe")), extra=opts.get("oracle_args", ()))
                path = os.path.join(GOLDEN, "%s.%s.acc.gz" % (name, flavour))
                netfile.write(path, build_fixture(flavour, net, steps, dump, opts))
                ev = dump["events"].reshape(-1, 4)
                types = sorted(set(int(x) for x in ev[:, 1]))
                summary["%s.%s" % (name, flavour)] = {"events": int(ev.shape[0]), "event_types": types, "solver": stats}
                print("%-22s %-5s %6d KiB  events %d types %s | %s" % (name, flavour, os.path.getsize(path) // 1024, ev.shape[0], types, stats), flush=True)
    with open(spath, "w") as fp:
        json.dump(summary, fp, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
