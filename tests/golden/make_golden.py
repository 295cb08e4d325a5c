"""Generates the golden fixtures of tests/golden/ by running the REFERENCE's own model code
(oracle/_ref/acclimate_ref, built from /root/reference by oracle/Makefile) on seeded synthetic networks.

Run here (where /root/reference exists):  python tests/golden/make_golden.py
Each fixture is an ACCARR1 file holding the network snapshot under "net/<name>" and the reference's per-step state
under "tNNNNN/<name>"; per-connection arrays are kept for demand requests and shipments only, to keep the files small.
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from acclimate_b200 import netfile, synth  # noqa: E402

ORACLE = os.path.join(ROOT, "oracle", "_ref", "acclimate_ref")
KEEP_EDGE = ("edge_D_q", "edge_D_p", "edge_ship_q")


def cases():
    a = synth.make_config("small")
    synth.add_shock(a, 2, 6, 3, 0.5)
    yield "small_shock12", a, 12
    a = synth.make_network(regions=8, sectors=4, seed=7, max_delay=6, model=dict(utilitarian=0))  # price-elastic consumers
    synth.add_shock(a, 3, 12, 2, 0.05)
    yield "severe_shock30", a, 30
    a = synth.make_network(regions=8, sectors=4, seed=8, max_delay=5)
    synth.add_passage(a, 2, 9, 5, 1, 0.3)
    synth.add_passage(a, 4, 6, 3, 2, 0.0)
    yield "passage25", a, 25
    a = synth.make_network(regions=6, sectors=3, seed=9, const_delay=2, model=dict(utilitarian=0))
    synth.add_shock(a, 2, 5, 1, 0.0)
    yield "outage25", a, 25
    a = synth.make_network(regions=6, sectors=4, seed=10, const_delay=1,
                           model=dict(quadratic_transport_penalty=0, relative_transport_penalty=0, cheapest_price_range_width=0.01, min_storage=0.2,
                                      deviation_penalty=1, maximal_decrease_reservation_price_limited_by_markup=1))
    synth.add_shock(a, 1, 8, 0, 0.3)
    yield "flags20", a, 20
    # utilitarian consumers (the reference's default, ModelInitializer.cpp:136): nested-CES utility optimisation, one basket
    a = synth.make_network(regions=8, sectors=4, seed=11, max_delay=4, model=dict(utilitarian=1))
    synth.add_shock(a, 2, 9, 3, 0.3)
    yield "utilitarian20", a, 20
    # three baskets with different substitution coefficients, one sector outside every basket, severe shock + blocked passage
    a = synth.make_network(regions=6, sectors=7, seed=12, max_delay=3, model=dict(utilitarian=1))
    a["basket_of_sector"] = np.array([0, 0, 1, 2, 2, -1, 1, -1])
    a["basket_coeff"] = np.array([2.0, 0.7, 4.0])
    a["inter_basket_coeff"] = np.array([1.3])
    synth.add_shock(a, 1, 7, 2, 0.1)
    synth.add_passage(a, 3, 8, 4, 1, 0.2)
    yield "baskets20", a, 20
    # multi-hop routes ("network" transport): access connections, zero-delay hub locations, sea lanes; passage forcing
    # on a zero-delay hub, a blocked lane and a throttled access connection
    a = synth.make_network(regions=7, sectors=4, seed=21, max_delay=3)
    info = synth.add_routes(a, seed=21, hubs=3)
    synth.add_shock(a, 2, 8, 3, 0.3)
    synth.add_entity_passage(a, 3, 9, info["hub_loc"][0], 0.4)
    synth.add_entity_passage(a, 5, 7, list(info["lane"].values())[0], 0.0)
    synth.add_entity_passage(a, 4, 10, info["access"][2], 0.6)
    yield "routes25", a, 25
    # naive_expectations = false: sellers extend their expected demand curve with the supply elasticity and look for
    # the production where marginal costs meet marginal revenue (bisection); second half: always_extend as well
    a = synth.make_network(regions=6, sectors=4, seed=14, max_delay=3, model=dict(naive_expectations=0),
                           sector=dict(supply_elasticity=0.4))
    synth.add_shock(a, 2, 10, 1, 0.2)
    yield "expectations20", a, 20
    a = synth.make_network(regions=5, sectors=3, seed=15, const_delay=2, model=dict(naive_expectations=0, always_extend_expected_demand_curve=1))
    synth.add_shock(a, 3, 8, 0, 0.5)
    yield "extend15", a, 15
    # 200 regions x 1 sector: every purchasing manager has 199 connections, more than two per thread of the base solver
    # group, so engine and oracle switch to the larger group (accsqp_group_for); per-connection arrays are left out
    a = synth.make_network(regions=200, sectors=1, seed=13, max_delay=3)
    synth.add_shock(a, 1, 4, 7, 0.4)
    yield "wide8", a, 8


def main():
    for name, net, steps in cases():
        with tempfile.TemporaryDirectory() as tmp:
            npath, dpath = os.path.join(tmp, "n.acc"), os.path.join(tmp, "d.acc")
            netfile.write(npath, net)
            subprocess.run([ORACLE, npath, str(steps), dpath], check=True, stdout=subprocess.DEVNULL)
            dump = netfile.read(dpath)
        out = {"net/" + k: v for k, v in net.items()}
        out["steps"] = [steps]
        for k, v in dump.items():
            base = k.split("/")[-1]
            if base.startswith("edge_") and (base not in KEEP_EDGE or name == "wide8"):
                continue
            out[k] = v
        path = os.path.join(ROOT, "tests", "golden", name + ".acc")
        netfile.write(path, out)
        print(name, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
