"""Shared checker: run an engine over a golden fixture and compare every dumped array with the reference oracle.

Bar (DESIGN.md "Parity"): int64 fields (quantities, prices) bit-exact, including everything downstream of the optimiser;
f64 values that the reference folds in a replicable order (storage / firm / purchase values) to 1e-12 relative; region sums
(folded in lock-arrival order by the reference) to 1e-9 relative.
"""
import os

import numpy as np

from acclimate_b200 import engine, netfile

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["small_shock12", "severe_shock30", "passage25", "outage25", "flags20", "utilitarian20", "baskets20", "wide8", "routes25", "expectations20", "extend15"]

EXACT_OR_TIGHT = [
    "edge_D_q", "edge_D_p", "edge_ship_q", "sto_S_q", "sto_S_v", "sto_I_q", "sto_I_v", "sto_U_q", "sto_U_v", "sto_Ut_q", "sto_Ut_v",
    "sto_demand_q", "sto_demand_v", "sto_expected_costs", "sto_optimized_value", "sto_transport_penalty", "firm_X_q", "firm_X_v", "firm_Xhat_q",
    "firm_Xhat_p", "firm_Xexp_q", "firm_Xexp_v", "firm_nbar", "firm_costs_C", "firm_revenue_R", "firm_desired_q", "firm_possible_q",
    "consumer_utility", "sector_production_q", "sector_production_v", "region_import_q", "region_export_q", "region_consumption_q", "region_consumption_v",
]
LOOSE = ["region_import_v", "region_export_v", "region_gdp_v"]
# the reference's discarded consumer utility optimisation (Consumer.cpp:270) raises these; the engine does not run it
IGNORED_EVENT_TYPES = {14}


def load_case(name):
    g = netfile.read(os.path.join(GOLDEN_DIR, name + ".acc"))
    net = {k[4:]: v for k, v in g.items() if k.startswith("net/")}
    steps = int(g["steps"][0])
    return net, g, steps


def run_case(name, lib_path=None, device=0):
    net, gold, steps = load_case(name)
    E = engine.Engine(net, device=device, lib_path=lib_path)
    sc = engine.Scenario(net)
    problems = []
    for t in range(steps):
        sc.apply(E, t)
        E.step()
        pre = "t%05d/" % t
        for n in EXACT_OR_TIGHT + LOOSE:
            if pre + n not in gold:
                continue
            got, want = E.read_state(n), gold[pre + n]
            if got.dtype == np.int64:
                if not np.array_equal(got, want):
                    i = int(np.nonzero(got != want)[0][0])
                    problems.append(f"step {t} {n}: {(got != want).sum()} of {got.size} differ, first at {i}: {got[i]} != {want[i]}")
            else:
                tol = 1e-9 if n in LOOSE else 1e-12
                err = np.abs(got - want) / np.maximum(np.abs(want), 1e-300)
                err[got == want] = 0.0
                if err.size and err.max() > tol:
                    i = int(err.argmax())
                    problems.append(f"step {t} {n}: rel err {err.max():.3e} at {i}: {got[i]!r} != {want[i]!r}")
        if len(problems) > 10:
            break
    ev = sorted((e[0], e[1], e[2], e[3]) for e in E.events())
    oe = sorted((int(r[0]), int(r[1]), int(r[2]), int(r[3])) for r in gold["events"].reshape(-1, 4) if int(r[1]) not in IGNORED_EVENT_TYPES)
    if ev != oe:
        problems.append(f"events differ: engine {len(ev)} reference {len(oe)}; first engine {ev[:3]} reference {oe[:3]}")
    launches = E.launches
    E.close()
    return problems, launches


# ---- region-partitioned run: every rank checks what it owns against the same golden dump ----
PER_AGENT = [n for n in EXACT_OR_TIGHT if n.startswith("firm_")] + ["consumer_utility"]
PER_STORAGE = [n for n in EXACT_OR_TIGHT if n.startswith("sto_")]
PER_EDGE = ["edge_D_q", "edge_D_p", "edge_ship_q"]
PER_REGION = ["region_import_q", "region_export_q", "region_consumption_q", "region_consumption_v", "region_import_v", "region_export_v", "region_gdp_v"]


def global_storage_ids(net):
    """(input sector, buying agent) -> creation index of the storage in the full network (first connection wins)."""
    sec = np.asarray(net["agent_sector"], dtype=np.int64)[np.asarray(net["edge_from"], dtype=np.int64)]
    key = sec * int(net["meta"][2]) + np.asarray(net["edge_to"], dtype=np.int64)
    uniq, first = np.unique(key, return_index=True)
    order = np.argsort(first, kind="stable")
    return {int(uniq[k]): i for i, k in enumerate(order)}


def _differs(name, got, want, loose):
    if got.dtype == np.int64:
        bad = got != want
        if bad.any():
            i = int(np.nonzero(bad)[0][0])
            return f"{name}: {int(bad.sum())} of {got.size} differ, first at {i}: {got[i]} != {want[i]}"
        return None
    tol = 1e-9 if loose else 1e-12
    err = np.abs(got - want) / np.maximum(np.abs(want), 1e-300)
    err[got == want] = 0.0
    if err.size and err.max() > tol:
        i = int(err.argmax())
        return f"{name}: rel err {err.max():.3e} at {i}: {got[i]!r} != {want[i]!r}"
    return None


def run_case_partitioned(name, lib_path, device=0, tensor_device="cpu", region_owner=None):
    """Called on every rank of an initialised process group; returns this rank's list of mismatches."""
    import torch.distributed as dist

    from acclimate_b200 import partition

    net, gold, steps = load_case(name)
    E = partition.PartitionedEngine(net, region_owner=region_owner, device=device, lib_path=lib_path, tensor_device=tensor_device)
    part = E.part
    sc = engine.Scenario(net)
    nA = int(net["meta"][2])
    gsid = global_storage_ids(net)
    l_agent, l_sector = E.read_state("sto_agent"), E.read_state("sto_sector")
    sto_owned = part.owned[l_agent].astype(bool)
    sto_gid = np.array([gsid[int(s) * nA + int(a)] for s, a in zip(l_sector[sto_owned], l_agent[sto_owned])], dtype=np.int64)
    agents = np.nonzero(part.owned)[0]
    regions = np.unique(np.asarray(net["agent_region"], dtype=np.int64)[agents])
    problems = []
    for t in range(steps):
        sc.apply(E, t)
        E.step()
        pre = "t%05d/" % t
        for n in PER_AGENT + PER_STORAGE + PER_EDGE + PER_REGION:
            if pre + n not in gold:
                continue
            got, want = E.read_state(n), gold[pre + n]
            if n in PER_AGENT:
                got, want = got[agents], want[agents]
            elif n in PER_STORAGE:
                got, want = got[sto_owned], want[sto_gid]
            elif n in PER_EDGE:
                want = want[part.edge_gid]
            else:
                got, want = got[regions], want[regions]
            msg = _differs(n, got, want, n in LOOSE)
            if msg:
                problems.append(f"rank {part.rank} step {t} {msg}")
        if len(problems) > 10:
            break
    mine = [(e[0], e[1], e[2], e[3]) for e in E.events()]
    everyone = [None] * dist.get_world_size()
    dist.all_gather_object(everyone, mine)
    if dist.get_rank() == 0:
        ev = sorted(x for part_events in everyone for x in part_events)
        oe = sorted((int(r[0]), int(r[1]), int(r[2]), int(r[3])) for r in gold["events"].reshape(-1, 4) if int(r[1]) not in IGNORED_EVENT_TYPES)
        if ev != oe:
            problems.append(f"events differ: engines {len(ev)} reference {len(oe)}; first engines {ev[:3]} reference {oe[:3]}")
    halo = (int(part.send_edges.size), int(part.recv_edges.size))
    E.close()
    return problems, halo
