"""The one analytic known answer the reference offers: without a shock the model sits at its baseline
fixed point -- every demand request and shipment equals Z*, every price is 1.000000, storages stay at S*, firms produce
X*, no event is raised -- at every step.  Checked on the kernel logic (CPU tier) and on the CUDA library (GPU tier) for
both consumer variants, a routed network, and at the headline size on the GPU."""
import numpy as np
import pytest

from acclimate_b200 import engine, synth


def _check(net, lib, steps, device=0):
    E = engine.Engine(net, device=device, lib_path=lib)
    Z = np.asarray(net["edge_zq"], dtype=np.int64)
    S0 = E.read_state("sto_S_q").copy()
    X0 = E.read_state("firm_X_q").copy()
    for _ in range(steps):
        E.step()
    assert np.array_equal(E.read_state("edge_D_q"), Z)
    assert np.array_equal(E.read_state("edge_ship_q"), Z)
    assert np.all(E.read_state("edge_D_p") == 1000000)
    assert np.array_equal(E.read_state("sto_S_q"), S0)
    assert np.array_equal(E.read_state("firm_X_q"), X0)
    firms = np.asarray(net["agent_type"]) == 1
    assert np.all(E.read_state("firm_nbar")[firms] == 1000000)
    assert E.events() == []
    E.close()


def _nets():
    yield synth.make_network(regions=8, sectors=4, seed=3, max_delay=5)                                # utilitarian consumers
    yield synth.make_network(regions=8, sectors=4, seed=3, max_delay=5, model=dict(utilitarian=0))     # price-elastic consumers
    a = synth.make_network(regions=6, sectors=3, seed=4, max_delay=3, model=dict(naive_expectations=0))
    synth.add_routes(a, seed=4, hubs=2)
    yield a


def test_baseline_fixed_point_kernel_logic(emu_lib):
    for net in _nets():
        _check(net, emu_lib, 40)


@pytest.mark.gpu
def test_baseline_fixed_point_cuda(cuda_lib):
    for net in _nets():
        _check(net, cuda_lib, 40)


@pytest.mark.gpu
def test_baseline_fixed_point_headline_size(cuda_lib):
    """BASELINE.json configs[1] (187 x 26, 24.5 M connections): a size-independent property at the full size."""
    _check(synth.make_config("eora"), cuda_lib, 5)
