"""MRIO table files in front of the device ingest (the file layer): `network: {type: netcdf, file, threshold}`.

Reference: src/input/ModelInitializer.cpp
  :829-845   sectors and regions are the file's `sector` / `region` name variables, in file order; the table is `flows` (or `flow`)
  :847-855   (sector, region, sector, region): agents sector-major
  :857-863   (region, sector, region, sector): agents region-major
  :865-872   (index, index) with index_sector / index_region: agents in index order
  :874-915   (flow) -- a list of compound records over compound agents: NetCDF-4 user-defined types, which the classic format cannot
             hold; not read here (an agent list + table built from such a file by the caller goes to ingest.ingest_table directly)
  :144-151   add_standard_agent: sector "FCON" or "*income_quintile*" makes a consumer, any other a firm
  :947-965   the dense table is read as FLOAT and walked agent by agent, row = seller
  :978-993   transport.type = const: every pair of distinct regions `transport.value` days apart (csv / centroids / network read
             geographic files and stay outside)

libnetcdf / HDF5 are not on this box: files are NetCDF-3 classic (scipy.io.netcdf_file), names as fixed-width character arrays.
The table goes to the device as it is read (float32, [agents][agents]); filter, quantisation and clean_network happen there
(csrc/ingest.cuh).
"""
from __future__ import annotations

import numpy as np
from scipy.io import netcdf_file

from .forcing import _names


class TableError(RuntimeError):
    pass


def is_consumer_sector(name):
    return name == "FCON" or "income_quintile" in name


def read_table(path):
    """-> dict(sector_names, region_names, agent_sector, agent_region, agent_type (1 firm / 0 consumer), flows float32 [nA, nA])"""
    with netcdf_file(path, "r", mmap=False) as f:
        v, d = f.variables, f.dimensions
        for k in ("sector", "region"):
            if k not in d or k not in v:
                raise TableError(f"'{k}' not found in '{path}'")
        sectors, regions = _names(v["sector"]), _names(v["region"])
        if len(set(sectors)) != len(sectors) or len(set(regions)) != len(regions):
            raise TableError(f"Ambiguous sector or region name in '{path}'")     # add_sector / add_region, :153-170
        fv = v["flows"] if "flows" in v else v.get("flow")
        if fv is None:
            raise TableError(f"Variable 'flows' not found in '{path}'")
        dims = tuple(fv.dimensions)
        nS, nR = len(sectors), len(regions)
        if dims == ("sector", "region", "sector", "region"):
            asec, areg = np.repeat(np.arange(nS), nR), np.tile(np.arange(nR), nS)
        elif dims == ("region", "sector", "region", "sector"):
            areg, asec = np.repeat(np.arange(nR), nS), np.tile(np.arange(nS), nR)
        elif dims == ("index", "index"):
            for k in ("index_sector", "index_region"):
                if k not in v or tuple(v[k].dimensions) != ("index",):
                    raise TableError(f"'{k}' over (index) not found in '{path}'")
            asec, areg = np.array(v["index_sector"][:], dtype=np.int64), np.array(v["index_region"][:], dtype=np.int64)
            if asec.min(initial=0) < 0 or asec.max(initial=0) >= nS or areg.min(initial=0) < 0 or areg.max(initial=0) >= nR:
                raise TableError(f"index_sector / index_region out of range in '{path}'")
        elif dims == ("flow",):
            raise TableError(f"'{path}': the compound (flow) layout needs NetCDF-4 user-defined types; pass the table to ingest.ingest_table")
        else:
            raise TableError(f"Wrong dimensions for flows variable in '{path}'")
        nA = asec.size
        flows = np.array(fv[:], dtype=np.float32).reshape(-1)                     # "use float to save memory", :947
        if flows.size != nA * nA:
            raise TableError(f"flows variable of '{path}' has {flows.size} cells, {nA * nA} expected")
    atype = np.array([0 if is_consumer_sector(sectors[s]) else 1 for s in asec], dtype=np.int64)
    return {"sector_names": sectors, "region_names": regions, "agent_sector": asec.astype(np.int64), "agent_region": areg.astype(np.int64),
            "agent_type": atype, "flows": flows.reshape(nA, nA)}


def seed_network(table, transport):
    """The snapshot skeleton ingest.ingest_table takes (no connections yet), transport as the `transport:` settings node says."""
    if not isinstance(transport, dict) or transport.get("type") != "const":
        raise TableError(f"transport.type = {None if not isinstance(transport, dict) else transport.get('type')!r}: only 'const' is built "
                         "without geographic files (csv / centroids / network: give the snapshot's `delay` table yourself)")
    if "value" not in transport:
        raise TableError("transport.value is required")
    nS, nR = len(table["sector_names"]), len(table["region_names"])
    nA = int(table["agent_sector"].size)
    delay = np.full((nR, nR), int(transport["value"]), dtype=np.int64)
    np.fill_diagonal(delay, 0)
    return {"meta": np.array([nS, nR, nA, 0], dtype=np.int64), "agent_type": table["agent_type"], "agent_sector": table["agent_sector"],
            "agent_region": table["agent_region"], "agent_beta": np.ones(nA), "delay": delay.reshape(-1),
            "consumer_elasticity": np.full(nS, -0.5),  # Consumer's default; the document's consumers section overwrites it (settings.apply_to)
            "forcing": np.zeros(0), "passage": np.zeros(0)}


def write_table(path, layout, sector_names, region_names, flows, index_sector=None, index_region=None):
    """A classic NetCDF table in one of the three dense layouts ("srsr", "rsrs", "index"); flows is [agents, agents] in that layout's
    agent order (fixtures, converters)."""
    flows = np.asarray(flows, dtype=np.float32)
    nS, nR = len(sector_names), len(region_names)
    with netcdf_file(path, "w") as f:
        def names(dim, vals):
            w = max(len(s) for s in vals) + 1
            f.createDimension(dim, len(vals))
            f.createDimension(dim + "_len", w)
            v = f.createVariable(dim, "c", (dim, dim + "_len"))
            for i, s in enumerate(vals):
                v[i] = np.frombuffer(s.encode().ljust(w, b"\0"), dtype="S1")

        names("sector", sector_names)
        names("region", region_names)
        if layout == "srsr":
            f.createVariable("flows", "f4", ("sector", "region", "sector", "region"))[:] = flows.reshape(nS, nR, nS, nR)
        elif layout == "rsrs":
            f.createVariable("flows", "f4", ("region", "sector", "region", "sector"))[:] = flows.reshape(nR, nS, nR, nS)
        elif layout == "index":
            n = len(index_sector)
            f.createDimension("index", n)
            f.createVariable("index_sector", "i4", ("index",))[:] = np.asarray(index_sector, dtype=np.int32)
            f.createVariable("index_region", "i4", ("index",))[:] = np.asarray(index_region, dtype=np.int32)
            f.createVariable("flows", "f4", ("index", "index"))[:] = flows.reshape(n, n)
        else:
            raise TableError(f"unknown layout {layout!r}")
