"""Shared checker: run an engine over a golden fixture and compare EVERY dumped array with the reference oracle.

Fixtures: tests/golden/<case>.shim.acc.gz (made by tests/golden/make_golden.py from the reference's own model code with
the NLopt provider oracle/nlopt_shim).  Bar (DESIGN.md "Parity"): int64 fields (quantities, prices) bit-exact, including
everything downstream of the optimiser; f64 values that the reference folds in a replicable order (storage / firm /
purchase values) to 1e-12 relative; region sums (folded in lock-arrival order by the reference) to 1e-9 relative.
An array the fixture holds and the engine cannot produce FAILS the case (no silent skips); so does a fixture whose solver
stamp does not match the sources in the tree.
"""
import os
import sys

import numpy as np

from acclimate_b200 import engine, netfile

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLDEN_DIR)
import make_golden  # noqa: E402  (blob_hash / checksum / SOLVER_SOURCES: the fixture's own definitions)

CASES = ["small_shock12", "severe_shock30", "passage25", "outage25", "flags20", "utilitarian20", "baskets20", "wide8", "routes25", "expectations20",
         "extend15", "penalty_quad_abs15", "penalty_lin_rel15", "maxiter12", "events18", "config0_365"]

LOOSE = {"region_import_v", "region_export_v", "region_gdp_v"}
# per-connection in-transit value: the reference folds a queue in ring-position order, the engine in travel order
TRANSIT = {"edge_transit_v"}
# (case, array) the engine does not keep: nothing may be missing
UNAVAILABLE = set()
EV_ROUNDOFF = 14


def fixture_path(name, flavour="shim"):
    return os.path.join(GOLDEN_DIR, "%s.%s.acc.gz" % (name, flavour))


def check_stamp(gold, flavour, name):
    for src in make_golden.SOLVER_SOURCES[flavour]:
        key = "stamp/" + src
        if key not in gold or not np.array_equal(gold[key], make_golden.blob_hash(src)):
            raise AssertionError(f"{name}.{flavour}: fixture was generated from another version of {src}; regenerate with tests/golden/make_golden.py")


def load_case(name):
    g = netfile.read(fixture_path(name))
    check_stamp(g, "shim", name)
    net = {k[4:]: v for k, v in g.items() if k.startswith("net/")}
    if not net:
        net = make_golden.synth_network(name)  # too large to store: rebuilt from the seeded generator
    steps = int(g["steps"][0])
    return net, g, steps


def step_arrays(gold, t):
    """(name -> full array) and (name -> checksum) the fixture holds for step t."""
    pre, cpre = "t%05d/" % t, "chk/t%05d/" % t
    full = {k[len(pre):]: v for k, v in gold.items() if k.startswith(pre)}
    chk = {k[len(cpre):]: v for k, v in gold.items() if k.startswith(cpre)}
    return full, chk


def differs(name, got, want):
    if got.shape != want.shape:
        return f"{name}: {got.size} entries, reference has {want.size}"
    if got.dtype == np.int64:
        bad = got != want
        if bad.any():
            i = int(np.nonzero(bad)[0][0])
            return f"{name}: {int(bad.sum())} of {got.size} differ, first at {i}: {got[i]} != {want[i]}"
        return None
    tol = 1e-9 if name in LOOSE or name in TRANSIT else 1e-12
    err = np.abs(got - want) / np.maximum(np.abs(want), 1e-300)
    err[got == want] = 0.0
    if err.size and err.max() > tol:
        i = int(err.argmax())
        return f"{name}: rel err {err.max():.3e} at {i}: {got[i]!r} != {want[i]!r}"
    return None


def expected_events(gold, net):
    """The reference's events.  Price-elastic consumers run (and discard) a utility optimisation in the reference
    (Consumer.cpp:270) whose OPTIMIZER_ROUNDOFF_LIMITED the engine cannot raise: it does not run that optimisation."""
    utilitarian = float(np.asarray(net["model_f64"])[netfile.MODEL_PARAMS.index("utilitarian")]) != 0.0
    rows = gold["events"].reshape(-1, 4)
    return sorted((int(r[0]), int(r[1]), int(r[2]), int(r[3])) for r in rows if utilitarian or int(r[1]) != EV_ROUNDOFF)


def compare_step(name, E, gold, t, select=None):
    """Mismatches of step t.  select(array_name, got, want) -> (got, want) restricts the comparison (partitioned runs)."""
    problems = []
    full, chk = step_arrays(gold, t)
    if not full and not chk:
        problems.append(f"step {t}: the fixture holds nothing for this step")
    for n, want in sorted(full.items()):
        if (name, n) in UNAVAILABLE:
            continue
        try:
            got = E.read_state(n)
        except engine.EngineError as ex:
            problems.append(f"step {t} {n}: the engine cannot produce this array ({ex})")
            continue
        if select is not None:
            got, want = select(n, got, want)
        msg = differs(n, got, want)
        if msg:
            problems.append(f"step {t} {msg}")
    for n, want in sorted(chk.items()):
        if (name, n) in UNAVAILABLE or select is not None:
            continue
        got = make_golden.checksum(E.read_state(n))
        if got.dtype == np.int64:
            if not np.array_equal(got, want):
                problems.append(f"step {t} {n}: checksum differs")
        elif np.abs(got - want).max() > (1e-8 if n in LOOSE or n in TRANSIT else 1e-11) * max(np.abs(want).max(), 1e-300):
            problems.append(f"step {t} {n}: checksum {got} != {want}")
    return problems


def run_case(name, lib_path=None, device=0):
    net, gold, steps = load_case(name)
    E = engine.Engine(net, device=device, lib_path=lib_path)
    sc = engine.Scenario(net)
    problems, events = [], []
    for t in range(steps):
        sc.apply(E, t)
        E.step()
        problems += compare_step(name, E, gold, t)
        events += E.events()  # drained every step: the device buffer is bounded
        if len(problems) > 10:
            break
    else:
        ev = sorted((e[0], e[1], e[2], e[3]) for e in events)
        oe = expected_events(gold, net)
        if ev != oe:
            problems.append(f"events differ: engine {len(ev)} reference {len(oe)}; first engine {ev[:3]} reference {oe[:3]}; "
                            f"only engine {sorted(set(ev) - set(oe))[:3]} only reference {sorted(set(oe) - set(ev))[:3]}")
    launches = E.launches
    E.close()
    return problems, launches


# ---- region-partitioned run: every rank checks what it owns against the same golden dump ----
def global_storage_ids(net):
    """(input sector, buying agent) -> creation index of the storage in the full network (first connection wins)."""
    sec = np.asarray(net["agent_sector"], dtype=np.int64)[np.asarray(net["edge_from"], dtype=np.int64)]
    key = sec * int(net["meta"][2]) + np.asarray(net["edge_to"], dtype=np.int64)
    uniq, first = np.unique(key, return_index=True)
    order = np.argsort(first, kind="stable")
    return {int(uniq[k]): i for i, k in enumerate(order)}


def run_case_partitioned(name, lib_path, device=0, tensor_device="cpu", region_owner=None, engine_factory=None):
    """Called on every rank of an initialised process group; returns this rank's list of mismatches."""
    import torch.distributed as dist

    from acclimate_b200 import partition

    net, gold, steps = load_case(name)
    make = engine_factory or partition.PartitionedEngine
    E = make(net, region_owner=region_owner, device=device, lib_path=lib_path, tensor_device=tensor_device)
    part = E.part
    sc = engine.Scenario(net)
    nA = int(net["meta"][2])
    gsid = global_storage_ids(net)
    l_agent, l_sector = E.read_state("sto_agent"), E.read_state("sto_sector")
    sto_owned = part.owned[l_agent].astype(bool)
    sto_gid = np.array([gsid[int(s) * nA + int(a)] for s, a in zip(l_sector[sto_owned], l_agent[sto_owned])], dtype=np.int64)
    agents = np.nonzero(part.owned)[0]
    regions = np.unique(np.asarray(net["agent_region"], dtype=np.int64)[agents])

    def select(n, got, want):
        if n.startswith("firm_") or n in ("consumer_utility", "agent_forcing"):
            return got[agents], want[agents]
        if n.startswith("sto_"):
            return got[sto_owned], want[sto_gid]
        if n.startswith("edge_"):
            # the buyer's part is authoritative for demand requests and everything that arrives, the seller's for shipments
            mine = part.buys if (n.startswith("edge_D") or n.startswith("edge_deliv") or n.startswith("edge_transit") or n.startswith("edge_deficit")) else part.sells
            return got[mine], want[part.edge_gid][mine]
        if n.startswith("region_"):
            return got[regions], want[regions]
        return got[:0], want[:0]  # sector sums span all parts: checked by the unpartitioned run

    problems, mine_ev = [], []
    for t in range(steps):
        sc.apply(E, t)
        E.step()
        problems += [f"rank {part.rank} {m}" for m in compare_step(name, E, gold, t, select=select)]
        mine_ev += [(e[0], e[1], e[2], e[3]) for e in E.events()]
        if len(problems) > 10:
            break
    everyone = [None] * dist.get_world_size()
    dist.all_gather_object(everyone, mine_ev)
    if dist.get_rank() == 0 and len(problems) <= 10:
        ev = sorted(x for part_events in everyone for x in part_events)
        oe = expected_events(gold, net)
        if ev != oe:
            problems.append(f"events differ: engines {len(ev)} reference {len(oe)}; first engines {ev[:3]} reference {oe[:3]}")
    halo = (int(part.send_edges.size), int(part.recv_edges.size))
    E.close()
    return problems, halo
