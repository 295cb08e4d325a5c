"""Seeded synthetic multi-regional input-output (MRIO) networks of the shapes BASELINE.json names.

EORA data is not available offline, so every configuration runs on a generated network
(SURVEY.md section 8(d)): Dirichlet technical coefficients, log-normal region sizes, gravity-like
trade weights, a Leontief fixed point for gross output, continuous jitter so that a seller's
demand requests do not tie.  The result is the flat snapshot the reference's ModelInitializer
would hand to the model (agents, per-edge initial flows Z*, hop delays, sector parameters).

Agent order is region-major: agent = region * (n_firm_sectors + 1) + sector, the last slot of each
region being its consumer (sector "FCON").  Edges are listed seller-major (then by buyer), which
makes the file order the CSR-by-seller order.
"""
from __future__ import annotations

import os

import numpy as np

from . import netfile

DEFAULT_MODEL = dict(
    transport_penalty_small=0.1,
    transport_penalty_large=1.0,
    min_storage=0.0,
    cheapest_price_range_width=-1.0,
    optimization_maxiter=500000,
    optimization_precision_adjustment=1,
    quadratic_transport_penalty=1,
    relative_transport_penalty=1,
    maximal_decrease_reservation_price_limited_by_markup=0,
    deviation_penalty=0,
    cheapest_price_range_preserve_seller_price=0,
    no_self_supply=1,
    utilitarian=1,  # consumers.<..>.bool_utilitarian defaults to true (reference ModelInitializer.cpp:136)
    delta_t=1.0,
    naive_expectations=1,
    always_extend_expected_demand_curve=0,
)
DEFAULT_SECTOR = dict(
    upper_storage_limit=3.0,
    initial_storage_fill_factor=10.0,
    transport_type=2,
    supply_elasticity=0.1,
    price_increase_production_extension=5.0,
    estimated_price_increase_production_extension=5.0,
    initial_markup=0.05,
    target_storage_refill_time=2.0,
    target_storage_withdraw_time=2.0,
)


def _haversine_km(lon1, lat1, lon2, lat2):
    # same great-circle formula family as reference src/model/GeoPoint.cpp:29-37
    r = 6371.0
    p1, p2 = np.radians(lat1), np.radians(lat2)
    dphi = p2 - p1
    dl = np.radians(lon2 - lon1)
    a = np.sin(dphi / 2) ** 2 + np.cos(p1) * np.cos(p2) * np.sin(dl / 2) ** 2
    return 2 * r * np.arcsin(np.sqrt(np.clip(a, 0, 1)))


def make_network(regions, sectors, seed, const_delay=None, world_output=None, beta=1.15, model=None, sector=None, max_delay=30, part=None):
    """Returns an OrderedDict of arrays (see netfile.py) for a dense `regions` x `sectors` network.

    part = (region_owner, rank) emits only the connections a part of a region-partitioned run holds (seller or buyer in
    a region of `rank`), in the global order, plus "global_edges" = the size of the whole network: a part of a network
    that is too large for one host buffer can be generated without ever materialising the rest (partition.Part.from_local)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    R, S = int(regions), int(sectors)
    A = rng.dirichlet(np.full(S, 0.3), size=S).T  # A[s_from, s_to], columns sum to 1
    u = rng.uniform(0.35, 0.65, size=S)           # intermediate input share of each buying sector
    A = A * u[None, :]
    g = rng.lognormal(0.0, 1.2, size=R)
    lon = rng.uniform(-180, 180, size=R)
    lat = np.degrees(np.arcsin(rng.uniform(-1, 1, size=R)))
    dist = _haversine_km(lon[:, None], lat[:, None], lon[None, :], lat[None, :])
    w = g[:, None] * np.where(np.eye(R, dtype=bool), 8.0, np.exp(-dist / 4000.0))  # w[r_from, r_to]
    w = w / w.sum(axis=0, keepdims=True)
    csh = rng.dirichlet(np.full(S, 1.0))           # consumption shares by supplying sector
    fd = csh[:, None] * g[None, :]                 # final demand of region r for sector s goods
    f = fd @ w.T                                   # f[s_from, r_from]
    # Leontief fixed point  X = A (X w^T) + f   (X[s, r] gross output)
    X = f.copy()
    for _ in range(200):
        X = A @ X @ w.T + f
    if world_output is None:
        world_output = 2.0e4 * R * S               # daily units; keeps the smallest flows above one quantum mostly
    scale = world_output / X.sum()
    X *= scale
    fd *= scale

    n_per = S + 1
    n_agents = R * n_per
    sigma = 0.5
    to_sector = np.arange(n_per)
    sidx = np.arange(S)

    def block(rf):
        """Jittered flows sold by the firms of region rf: Z[s_from, r_to, s_to] (s_to == S: the consumer)."""
        brng = np.random.Generator(np.random.PCG64([seed, rf]))
        base = np.empty((S, R, n_per))
        base[:, :, :S] = A[:, None, :] * X.T[None, :, :] * w[rf, :][None, :, None]
        base[:, :, S] = fd * w[rf, :][None, :]
        z = base * brng.lognormal(0.0, sigma, size=(S, R, n_per))
        z[sidx, rf, sidx] = 0.0  # no self supply
        return z

    # pass 1: inputs per buyer and sales per seller, then lift each seller's final-demand sales so that
    # value added stays >= 20 % of output (clean_network would delete the firm otherwise)
    in_q = np.zeros((R, S))
    inter = np.zeros((R, S))
    fsale = np.zeros((R, S))
    for rf in range(R):
        z = block(rf)
        in_q += z[:, :, :S].sum(axis=0)
        inter[rf] = z[:, :, :S].sum(axis=(1, 2))
        fsale[rf] = z[:, :, S].sum(axis=1)
    lift = np.maximum(1.0, (1.25 * in_q - inter) / fsale)
    # pass 2: emit seller-major edges
    blocks_from, blocks_to, blocks_z = [], [], []
    buyers_all = (np.arange(R)[:, None] * n_per + to_sector[None, :]).reshape(-1)
    global_edges = 0
    out_q = np.zeros(n_agents)
    in_q2 = np.zeros(n_agents)
    owned_buyer = None
    if part is not None:
        owner, rank = np.asarray(part[0]), int(part[1])
        owned_buyer = np.repeat(owner == rank, n_per)
    for rf in range(R):
        z = block(rf)
        z[:, :, S] *= lift[rf][:, None]
        zq = np.maximum(1, np.rint(z / 1e-3)).astype(np.int64).reshape(S, -1)
        for sf in range(S):
            seller = rf * n_per + sf
            keep = buyers_all != seller
            global_edges += int(keep.sum())
            out_q[seller] = zq[sf][keep].sum()
            in_q2 += np.where(keep, zq[sf], 0)
            if owned_buyer is not None and owner[rf] != rank:
                keep = keep & owned_buyer
            blocks_from.append(np.full(int(keep.sum()), seller, dtype=np.int64))
            blocks_to.append(buyers_all[keep])
            blocks_z.append(zq[sf][keep])
    edge_from = np.concatenate(blocks_from)
    edge_to = np.concatenate(blocks_to)
    edge_zq = np.concatenate(blocks_z)

    agent_type = np.ones(n_agents, dtype=np.int64)
    agent_type[S::n_per] = 0
    agent_sector = np.tile(np.arange(n_per), R).astype(np.int64)
    agent_region = np.repeat(np.arange(R), n_per).astype(np.int64)
    agent_beta = np.full(n_agents, float(beta))

    mp = dict(DEFAULT_MODEL)
    mp.update(model or {})
    sp = dict(DEFAULT_SECTOR)
    sp.update(sector or {})
    sector_f64 = np.tile(np.array([sp[k] for k in netfile.SECTOR_PARAMS], dtype=np.float64), (S + 1, 1))

    if const_delay is not None:
        delay = np.full((R, R), int(const_delay), dtype=np.int64)
    else:
        # delay rule of reference ModelInitializer.cpp:646-660 ("centroids" transport), clamped to max_delay
        thr, road_speed, sea_speed = 3000.0, 35.0, 28.0
        d_road = np.rint(dist / road_speed / 24.0 / mp["delta_t"])
        d_sea = np.rint(dist / sea_speed / 24.0 / mp["delta_t"])
        delay = np.where(dist >= thr, d_sea, d_road).astype(np.int64)
        delay = np.where(delay > 0, delay - 1, delay)
        delay = np.clip(delay, 0, max_delay)
        delay = np.maximum(delay, delay.T)
    np.fill_diagonal(delay, 0)

    # invariants the reference's clean_network checks (ModelInitializer.cpp:230-236)
    firms = agent_type == 1
    if not np.all(out_q[firms] - in_q2[firms] > 0):
        raise ValueError("generated network has a firm without positive value added")

    from collections import OrderedDict

    arrays = OrderedDict()
    arrays["meta"] = np.array([S + 1, R, n_agents, edge_from.size], dtype=np.int64)
    arrays["model_f64"] = np.array([mp[k] for k in netfile.MODEL_PARAMS], dtype=np.float64)
    arrays["sector_f64"] = sector_f64.reshape(-1)
    arrays["agent_type"] = agent_type
    arrays["agent_sector"] = agent_sector
    arrays["agent_region"] = agent_region
    arrays["agent_beta"] = agent_beta
    arrays["edge_from"] = edge_from
    arrays["edge_to"] = edge_to
    arrays["edge_zq"] = edge_zq
    arrays["delay"] = delay.reshape(-1)
    arrays["consumer_elasticity"] = np.full(S + 1, -0.5)
    arrays["forcing"] = np.zeros(0)
    arrays["passage"] = np.zeros(0)
    if part is not None:
        arrays["global_edges"] = np.array([global_edges], dtype=np.int64)
    return arrays


def add_routes(arrays, seed, hubs=3):
    """Multi-hop transport routes (the reference's "network" transport, ModelInitializer.cpp:339-607): every region
    reaches a sea hub over an access connection, hubs are zero-delay locations joined by sea lanes; road/sea sectors
    ship region -> access -> hub [-> lane -> hub] -> access -> region, aviation sectors over one direct connection per
    region pair.  Adds entity_delay / route_off / route_entity (see oracle/ref_harness and include/acclimate_b200.h);
    passage forcing (add_entity_passage) then addresses these entities."""
    rng = np.random.Generator(np.random.PCG64([seed, 77]))
    R = int(arrays["meta"][1])
    hubs = max(1, min(int(hubs), R))
    hub_of = rng.integers(0, hubs, size=R)
    delays = []
    access = []
    for r in range(R):
        access.append(len(delays)); delays.append(int(rng.integers(0, 3)))
    hub_loc = []
    for h in range(hubs):
        hub_loc.append(len(delays)); delays.append(0)
    lane = {}
    for a in range(hubs):
        for b in range(a):
            lane[(a, b)] = lane[(b, a)] = len(delays); delays.append(int(rng.integers(1, 4)))
    direct = {}
    for i in range(R):
        for j in range(i):
            direct[(i, j)] = direct[(j, i)] = len(delays); delays.append(1)
    off, ent = [0], []
    for k in range(2):
        for i in range(R):
            for j in range(R):
                if i != j:
                    if k == 0:
                        ent += [direct[(i, j)]]
                    else:
                        hi, hj = int(hub_of[i]), int(hub_of[j])
                        path = [access[i], hub_loc[hi]]
                        if hi != hj:
                            path += [lane[(hi, hj)], hub_loc[hj]]
                        path += [access[j]]
                        ent += path
                off.append(len(ent))
    arrays["entity_delay"] = np.array(delays, dtype=np.int64)
    arrays["route_off"] = np.array(off, dtype=np.int64)
    arrays["route_entity"] = np.array(ent, dtype=np.int64)
    return dict(access=access, hub_loc=hub_loc, lane=lane, direct=direct, hub_of=hub_of)


def add_entity_passage(arrays, t_from, t_to, entity, nu):
    """Passage forcing of one geo entity of a routed network (index into entity_delay)."""
    rec = np.array([t_from, t_to, int(entity), nu], dtype=np.float64)
    arrays["passage"] = np.concatenate([np.asarray(arrays.get("passage", np.zeros(0)), dtype=np.float64).reshape(-1), rec])


def add_shock(arrays, t_from, t_to, region, remaining_capacity):
    """Production-loss shock on every firm of `region` (reference Scenario.cpp:45-56: forcing = capacity / beta)."""
    at, ar, ab = arrays["agent_type"], arrays["agent_region"], arrays["agent_beta"]
    idx = np.nonzero((at == 1) & (ar == region))[0]
    rows = np.stack([np.full(idx.size, float(t_from)), np.full(idx.size, float(t_to)), idx.astype(np.float64), remaining_capacity / ab[idx]], axis=1)
    arrays["forcing"] = np.concatenate([np.asarray(arrays["forcing"], dtype=np.float64).reshape(-1, 4), rows]).reshape(-1)
    return arrays


def add_passage(arrays, t_from, t_to, region_a, region_b, nu):
    """Passage forcing on the connection between two regions (reference Scenario.cpp:69-77, 'sea' targets)."""
    a, b = max(region_a, region_b), min(region_a, region_b)
    conn = a * (a - 1) // 2 + b  # index in the (i, j<i) enumeration used by the oracle harness and the engine
    row = np.array([float(t_from), float(t_to), float(conn), float(nu)])
    arrays["passage"] = np.concatenate([np.asarray(arrays["passage"], dtype=np.float64), row])
    return arrays


CONFIGS = {
    # BASELINE.json configs[0..3]
    "small": dict(regions=10, sectors=5, seed=20260101 + 1, const_delay=3),
    "eora": dict(regions=187, sectors=26, seed=20260101 + 2),
    "eora_flood": dict(regions=187, sectors=26, seed=20260101 + 3),
    "x4": dict(regions=748, sectors=26, seed=20260101 + 4),
}


def make_config(name, part=None):
    cfg = dict(CONFIGS[name])
    if name == "x4" and os.environ.get("ACC_X4_REGIONS"):  # dry runs of the partitioned generator path at a small size
        cfg["regions"] = int(os.environ["ACC_X4_REGIONS"])
    arrays = make_network(part=part, **cfg)
    if name == "eora_flood":
        add_shock(arrays, 10, 24, 17, 0.5)
    return arrays
