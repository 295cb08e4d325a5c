"""Network ingest in front of acc_create: candidate MRIO table -> threshold filter -> clean_network -> snapshot.

The reference does this on the host inside ModelInitializer (reference src/input/ModelInitializer.cpp:822-976 reads the table,
:187-213 creates the connections, :215-299 cleans the network); here the filter, the per-agent sums and the removal fix point run
on the device behind acc_ingest_table (csrc/ingest.cuh) and this module only renumbers the survivors into the snapshot layout
acc_create takes.  The MRIO file readers themselves (four NetCDF layouts) stay outside the path: the table arrives as an array.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import engine as _engine


def ingest_table(arrays, flows, threshold, time_factor=None, device=0, lib_path=None):
    """arrays: a snapshot WITHOUT connections (meta, agent_*, model_f64, sector_f64, delay, ...: what the settings and the index
    file of the MRIO give); flows: [n_agents, n_agents] annual flows, row = selling agent (float32, as the reference reads
    them); threshold: network.threshold.  Returns (snapshot with the surviving agents and connections, info)."""
    lib = _engine.load_library(lib_path)
    nS, nR, nA = (int(x) for x in np.asarray(arrays["meta"])[:3])
    mp = np.asarray(arrays["model_f64"], dtype=np.float64)
    from . import netfile

    dt = float(mp[netfile.MODEL_PARAMS.index("delta_t")])
    if time_factor is None:
        time_factor = dt / 365.0  # ModelInitializer.cpp:833: annual table -> flow per timestep
    no_self = int(mp[netfile.MODEL_PARAMS.index("no_self_supply")] != 0.0)
    table = np.ascontiguousarray(flows, dtype=np.float32)
    if table.shape != (nA, nA):
        raise ValueError("flows must be [n_agents, n_agents]")
    at = np.ascontiguousarray(arrays["agent_type"], dtype=np.int64)
    asec = np.ascontiguousarray(arrays["agent_sector"], dtype=np.int64)
    lib.acc_ingest_table.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    alive = np.zeros(nA, dtype=np.uint8)
    n = C.c_int64(0)
    sweeps, removed = C.c_int64(0), C.c_int64(0)

    def call(ef, et, ez, cap):
        rc = lib.acc_ingest_table(int(device), nA, nS, at.ctypes.data, asec.ctypes.data, table.ctypes.data, float(threshold), float(time_factor), no_self,
                                  alive.ctypes.data, C.byref(n), ef, et, ez, cap, C.byref(sweeps), C.byref(removed))
        if rc != 0:
            raise _engine.EngineError(lib.acc_last_error().decode())

    call(None, None, None, 0)
    ne = int(n.value)
    ef, et, ez = (np.empty(ne, dtype=np.int64) for _ in range(3))
    call(ef.ctypes.data, et.ctypes.data, ez.ctypes.data, ne)
    # renumber: the reference erases removed agents from model()->economic_agents, the others keep their order (:290-293)
    keep = alive.astype(bool)
    new_id = np.cumsum(keep) - 1
    out = {k: v for k, v in arrays.items()}
    for k in ("agent_type", "agent_sector", "agent_region", "agent_beta"):
        out[k] = np.asarray(arrays[k])[keep]
    out["edge_from"], out["edge_to"], out["edge_zq"] = new_id[ef], new_id[et], ez
    meta = np.array(arrays["meta"], dtype=np.int64).copy()
    meta[2], meta[3] = int(keep.sum()), ne
    out["meta"] = meta
    info = {"agent_alive": keep, "sweeps": int(sweeps.value), "removed": int(removed.value), "edge_from": ef, "edge_to": et, "edge_zq": ez}
    return out, info
