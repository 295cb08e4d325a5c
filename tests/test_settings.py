"""The settings.yml surface (north_star: "keeping the Model/ModelRun, settings.yml and NetCDF output surface").

A document with the reference's sections (model, sectors, firms, consumers, network, scenario, outputs) drives a run end to end:
parameters with the reference's look-up rules and defaults, the scenario's shock with Scenario::apply_target's grammar, outputs
with ArrayOutput's select_* keys pulled through acc_observe and written in NetCDFOutput's layout.  The run must equal the same
run driven through the snapshot API (tests/golden), and the surface must refuse what the reference would not understand."""
import os
import tempfile

import numpy as np
import pytest

import parity
from acclimate_b200 import modelrun, output, settings, synth

PREC3 = float.fromhex("0x1.0624dd2f1a9fdp-10")  # FlowQuantity::precision, reference include/types.h:313-315

DOC = """
model:
  delta_t: 1
  transport_penalty_small: 0.1
  transport_penalty_large: 1.0
  optimization_maxiter: 500000
  optimization_timeout: 60
  cheapest_price_range_width: auto
sectors:
  ALL:
    upper_storage_limit: 3.0
    initial_storage_fill_factor: 10
    transport: roadsea
    supply_elasticity: 0.1
    price_increase_production_extension: 5.0
    initial_markup: 0.05
    target_storage_refill_time: 2
    target_storage_withdraw_time: 2
firms:
  ALL:
    possible_overcapacity_ratio: 1.15
consumers:
  ALL:
    bool_utilitarian: true
    inter_basket_substitution_coefficient: 1.5
    consumer_baskets:
      - sectors: [SEC00, SEC01, SEC02, SEC03, SEC04]
        substituition_coefficient: 2.0
network:
  type: snapshot
scenario:
  type: events
  start: 0
  stop: 11
  basedate: 2020-1-1
  events:
    - type: shock
      from: 2
      to: 6
      targets:
        - firm:
            region: RG003
            remaining_capacity: 0.5
outputs:
  - format: netcdf
    file: OUT
    observables:
      firms:
        production: ~
        forcing: ~
        select_region: [RG003, RG001]
      regions:
        gdp: ~
      storages:
        content: ~
        select_region_to: [RG002]
      flows:
        sent_flow: ~
        select_sector_from: [SEC01]
  - format: array
    observables:
      sectors:
        total_production: ~
"""


def names(net):
    nS, nR = int(net["meta"][0]), int(net["meta"][1])
    return ["SEC%02d" % s for s in range(nS - 1)] + ["FCON"], ["RG%03d" % r for r in range(nR)]


def run(lib):
    net = synth.make_config("small")  # the network of tests/golden/small_shock12: its shock comes from the document here
    sec, reg = names(net)
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "out.nc")
        mr, steps = modelrun.ModelRun.from_settings(DOC.replace("file: OUT", "file: " + path), net, sec, reg, lib_path=lib)
        assert steps == 12
        gold_net, gold, gsteps = parity.load_case("small_shock12")
        problems = []
        for t in range(steps):
            mr.step()
            problems += parity.compare_step("small_shock12", mr.engine, gold, t)
        arrays = mr.sinks[1][1].arrays()
        mr.close()
        from scipy.io import netcdf_file

        f = netcdf_file(path, "r", mmap=False)
        out = {k: np.array(v[:]) for k, v in f.variables.items()}
        dims = dict(f.dimensions)
        units = f.variables["time"].units.decode()
        f.close()
    return problems, out, dims, units, arrays, gold, net


def check(lib):
    problems, out, dims, units, arrays, gold, net = run(lib)
    # the settings-driven run IS the golden run: same parameters, same shock (remaining_capacity / beta), bit for bit
    assert not problems, "\n".join(problems[:10])
    nA = int(net["meta"][2])
    assert units == "days since 2020-1-1" and dims["agent"] == nA and dims["sector"] == 6 and dims["region"] == 10
    assert list(out["time"]) == list(range(12))
    areg, atype = np.asarray(net["agent_region"]), np.asarray(net["agent_type"])
    sel = np.concatenate([np.nonzero(areg == 3)[0], np.nonzero(areg == 1)[0]])  # in the order the regions are listed
    prod = out["firms.production_quantity"]
    assert prod.shape == (12, sel.size)
    want = np.stack([gold["t%05d/firm_X_q" % t][sel] * PREC3 for t in range(12)])
    firm = atype[sel] == 1
    assert np.array_equal(prod[:, firm], want[:, firm]) and np.isnan(prod[:, ~firm]).all()
    forcing = out["firms.forcing"]
    hit = firm & (areg[sel] == 3)
    assert np.allclose(forcing[2:7][:, hit], 0.5 / 1.15) and np.all(forcing[7:][:, hit] == 1.0) and np.all(forcing[:2][:, hit] == 1.0)
    assert out["regions.gdp_value"].shape == (12, 10)
    assert np.allclose(out["regions.gdp_value"], np.stack([gold["t%05d/region_gdp_v" % t] for t in range(12)]), rtol=1e-9)
    assert out["storages.content_quantity"].shape == (12, 6, int((areg == 2).sum()))
    n_from = int(((np.asarray(net["agent_sector"]) == 1) & (atype == 1)).sum())
    assert out["flows.sent_flow_quantity"].shape == (12, n_from, nA)
    assert np.array_equal(arrays[("sectors", "total_production_quantity")][:, :5],
                          np.stack([gold["t%05d/sector_production_q" % t][:5] * PREC3 for t in range(12)]))


def test_settings_document_drives_the_golden_run_emulation(emu_lib):
    check(emu_lib)


@pytest.mark.gpu
def test_settings_document_drives_the_golden_run_cuda(cuda_lib):
    check(cuda_lib)


def test_settings_refuse_what_the_reference_would_not_read():
    import yaml

    doc = yaml.safe_load(DOC)
    bad = dict(doc, model=dict(doc["model"], transport_penalty_smal=0.2))
    with pytest.raises(settings.SettingsError, match="unknown key model.transport_penalty_smal"):
        settings.Settings(bad)
    bad = dict(doc, model=dict(doc["model"], global_purchasing_optimization=True))
    with pytest.raises(settings.SettingsError, match="not provided"):
        settings.Settings(bad)
    bad = dict(doc, model={k: v for k, v in doc["model"].items() if k != "delta_t"})
    with pytest.raises(settings.SettingsError, match="model.delta_t is required"):
        settings.Settings(bad)
    bad = dict(doc, scenario=dict(doc["scenario"], type="hazard"))
    net = synth.make_config("small")
    with pytest.raises(settings.SettingsError, match="Unknown scenario type"):
        settings.Settings(bad).apply_to(net, *names(net))
    bad = dict(doc, outputs=[dict(format="hdf5")])
    with pytest.raises(settings.SettingsError, match="Unknown output format"):
        settings.Settings(bad).output_plan(net, *names(net))
    st = settings.Settings(doc)
    # look-up order of firm properties: sector:region, sector, region, ALL (ModelInitializer.cpp:74-93)
    st.doc["firms"]["SEC01"] = {"possible_overcapacity_ratio": 1.3}
    st.doc["firms"]["SEC01:RG002"] = {"possible_overcapacity_ratio": 1.4}
    assert st.firm_property("SEC01", "RG002") == 1.4 and st.firm_property("SEC01", "RG000") == 1.3 and st.firm_property("SEC02", "RG000") == 1.15
    assert st.sector_params(["SEC00"])[0][5] == 5.0  # estimated_price_increase defaults to price_increase (:166-169)
    assert output.expand("firms", ["production", "forcing"]) == ["production_quantity", "production_value", "forcing"]
