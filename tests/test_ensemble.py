"""Ensembles of scenarios on one network (BASELINE.json configs[4]; SURVEY.md 8e: "replicate topology once per GPU, batch state
along a scenario axis"): acc_fork members share the parent's network tables and own their state.  Every member must run exactly
the trajectory a separate engine of its own would run -- here: the golden runs of the reference oracle -- whatever the others
do in between, and the shared tables must survive the parent."""
import numpy as np
import pytest

import parity
from acclimate_b200 import engine


def run(lib):
    net, gold, steps = parity.load_case("small_shock12")
    base = engine.Engine(net, device=0, lib_path=lib)
    members = [base.fork() for _ in range(3)]
    sc = engine.Scenario(net)
    # member 0 runs the golden scenario; member 1 another shock; member 2 and the parent stay unshocked
    other = np.ones(base.n_agents)
    other[np.asarray(net["agent_region"]) == 5] = 0.3
    problems = []
    for t in range(steps):
        sc.apply(members[0], t)
        if t == 1:
            members[1].set_agent_forcing(other)
        for e in (members[1], base, members[0], members[2]):  # interleaved on purpose
            if e.h:
                e.step()
        problems += parity.compare_step("small_shock12", members[0], gold, t)
        if t == 5:
            base.close()  # the parent goes first: its tables must stay until the last member is gone
    z = np.asarray(net["edge_zq"])
    assert np.array_equal(members[2].read_state("edge_D_q"), z), "an unshocked member left the baseline"
    assert not np.array_equal(members[1].read_state("firm_X_q"), members[2].read_state("firm_X_q")), "the members' states are not independent"
    # a member forked from a running member continues its trajectory
    late = members[0].fork()
    members[0].step(); late.step()
    for name in ("edge_D_q", "sto_S_q", "firm_X_q"):
        assert np.array_equal(members[0].read_state(name), late.read_state(name)), name
    for e in members + [late]:
        e.close()
    return problems


def test_members_share_the_network_and_run_independently_emulation(emu_lib):
    problems = run(emu_lib)
    assert not problems, "\n".join(problems[:10])


@pytest.mark.gpu
def test_members_share_the_network_and_run_independently_cuda(cuda_lib):
    problems = run(cuda_lib)
    assert not problems, "\n".join(problems[:10])
