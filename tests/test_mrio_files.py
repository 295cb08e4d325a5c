"""MRIO table files in front of the device ingest (the file layer): reference src/input/ModelInitializer.cpp:829-872
(the three dense layouts, agent order per layout, FCON -> consumer) and :947-965 (float table, row = seller).

A table written in the (region, sector, region, sector) and (index, index) layouts and read back must give the ingest the SAME agents and
cells as the array the oracle-made fixture was generated from (tests/golden/ingest_cases.acc.gz: the reference's own filter and
clean_network), so the surviving network out of a FILE equals the reference's; the (sector, region, sector, region) layout is the same
table with the agents sector-major."""
import os
import tempfile

import numpy as np
import pytest

import test_ingest
import test_settings
from acclimate_b200 import engine, ingest, modelrun, mrio, settings


def case0():
    arrays, table = test_ingest.candidate(*test_ingest.CASES[0][:5])
    sec, reg = test_settings.names(arrays)
    return arrays, table, sec, reg


def want0():
    g = test_ingest.golden()
    return {n.split("/", 1)[1]: v for n, v in g.items() if n.startswith("case0/")}


@pytest.mark.parametrize("layout", ["rsrs", "index"])
def test_table_file_gives_the_reference_network_emulation(layout, emu_lib):
    arrays, table, sec, reg = case0()
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "t.nc")
        mrio.write_table(path, layout, sec, reg, table, index_sector=arrays["agent_sector"], index_region=arrays["agent_region"])
        tb = mrio.read_table(path)
    assert tb["sector_names"] == sec and tb["region_names"] == reg
    for k in ("agent_sector", "agent_region", "agent_type"):
        assert np.array_equal(tb[k], np.asarray(arrays[k])), k
    assert tb["flows"].dtype == np.float32 and np.array_equal(tb["flows"], table)
    seed = dict(arrays)
    for k in ("agent_sector", "agent_region", "agent_type"):
        seed[k] = tb[k]
    _, info = ingest.ingest_table(seed, tb["flows"], test_ingest.CASES[0][5], time_factor=1.0 / 365.0 * 30.0, lib_path=emu_lib)
    want = want0()
    assert np.array_equal(info["agent_alive"].astype(np.int64), want["agent_alive"])
    assert np.array_equal(info["edge_from"], want["edge_from"]) and np.array_equal(info["edge_to"], want["edge_to"])
    assert np.array_equal(info["edge_zq"], want["edge_zq"])


def test_sector_major_layout():
    arrays, table, sec, reg = case0()
    nS, nR = len(sec), len(reg)
    asec, areg = np.asarray(arrays["agent_sector"]), np.asarray(arrays["agent_region"])
    # agent (s, r) of the sector-major order is agent r * nS + s of the region-major table
    perm = np.array([r * nS + s for s in range(nS) for r in range(nR)])
    assert np.array_equal(asec[perm], np.repeat(np.arange(nS), nR))
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "t.nc")
        mrio.write_table(path, "srsr", sec, reg, table[np.ix_(perm, perm)])
        tb = mrio.read_table(path)
    assert np.array_equal(tb["agent_sector"], asec[perm]) and np.array_equal(tb["agent_region"], areg[perm])
    assert np.array_equal(tb["agent_type"], np.asarray(arrays["agent_type"])[perm])
    assert np.array_equal(tb["flows"], table[np.ix_(perm, perm)])


def test_refusals():
    from scipy.io import netcdf_file

    arrays, table, sec, reg = case0()
    with tempfile.TemporaryDirectory() as tmp:
        p = os.path.join(tmp, "bad.nc")
        with netcdf_file(p, "w") as f:
            f.createDimension("sector", 1); f.createDimension("region", 1); f.createDimension("w", 2); f.createDimension("flow", 3)
            f.createVariable("sector", "c", ("sector", "w"))[0] = np.frombuffer(b"A\0", dtype="S1")
            f.createVariable("region", "c", ("region", "w"))[0] = np.frombuffer(b"B\0", dtype="S1")
            f.createVariable("flows", "f4", ("flow",))[:] = 0
        with pytest.raises(mrio.TableError, match="compound"):
            mrio.read_table(p)
        p2 = os.path.join(tmp, "bad2.nc")
        with netcdf_file(p2, "w") as f:
            f.createDimension("sector", 1); f.createDimension("region", 1); f.createDimension("w", 2)
            f.createVariable("sector", "c", ("sector", "w"))[0] = np.frombuffer(b"A\0", dtype="S1")
            f.createVariable("region", "c", ("region", "w"))[0] = np.frombuffer(b"B\0", dtype="S1")
            f.createVariable("flows", "f4", ("region", "sector"))[:] = 0
        with pytest.raises(mrio.TableError, match="Wrong dimensions"):
            mrio.read_table(p2)
    assert mrio.is_consumer_sector("FCON") and mrio.is_consumer_sector("income_quintile_3") and not mrio.is_consumer_sector("AGRI")
    with pytest.raises(mrio.TableError, match="only 'const'"):
        mrio.seed_network({"sector_names": sec, "region_names": reg, "agent_sector": np.zeros(1)}, {"type": "centroids", "file": "x"})


def settings_run(lib):
    arrays, table, sec, reg = case0()
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "t.nc")
        mrio.write_table(path, "rsrs", sec, reg, table)
        head = test_settings.DOC.split("network:")[0].replace("[SEC00, SEC01, SEC02, SEC03, SEC04]", "[" + ", ".join(sec[:-1]) + "]")
        doc = head + f"""network:
  type: netcdf
  file: {path}
  threshold: 20.0
transport:
  type: const
  value: 1
scenario:
  type: events
  start: 0
  stop: 2
outputs:
  - format: array
    observables:
      firms:
        production: ~
"""
        mr, steps = modelrun.ModelRun.from_settings(doc, None, None, None, lib_path=lib)
        st = settings.Settings(doc)
        seed = dict(arrays)
        seed["model_f64"] = st.model_params()
        seed["delay"] = np.asarray(mrio.seed_network(mrio.read_table(path), {"type": "const", "value": 1})["delay"])
        want, info = ingest.ingest_table(seed, table, 20.0, lib_path=lib)
    assert steps == 3 and info["removed"] >= 3
    E, got = mr.engine, mr.network
    assert int(got["meta"][2]) == int(want["meta"][2]) and int(got["meta"][3]) == int(want["meta"][3])
    for k in ("edge_from", "edge_to", "edge_zq", "agent_sector", "agent_region", "delay"):
        assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), k
    for _ in range(steps):
        mr.step()
    z = np.asarray(want["edge_zq"])
    assert np.all(np.abs(E.read_state("edge_D_q") - z) <= np.maximum(5, z // 100000))   # the cleaned network sits at its baseline
    mr.close()


def test_settings_with_a_table_file_emulation(emu_lib):
    settings_run(emu_lib)


@pytest.mark.gpu
def test_settings_with_a_table_file_cuda(cuda_lib):
    settings_run(cuda_lib)
