"""A consumer's utility optimisation that cannot make progress any more must end, not kill the run.

Found at step 406 of the three-year flood run on the 40 x 26 network (BASELINE.json configs[2]; tests/golden/make_golden.py
flood40_1095): the objective sits at its optimum to 1e-12, the multiplier equals the penalty weight of the merit function, every line
search runs out of trials and the iteration crawls on in steps of three to four xtol.  NLopt ends such a solve as "roundoff limited" and
the reference carries on with the best point and an OPTIMIZER_ROUNDOFF_LIMITED event (Consumer.cpp:471-483); csrc/sqp_dense.h used to
run into its iteration cap, which the engine reports like the reference's "optimization failed" (Consumer.cpp:505-512): the end of the run.
The consumer's inputs (content, arrived flow and previous consumption of its 26 storages) were captured on the GPU before the failing
step (scripts/consumer_replay.py) and are replayed here on the host emulation of the same kernel code."""
import ctypes as C
import os

import numpy as np

from acclimate_b200 import engine, synth

HERE = os.path.dirname(os.path.abspath(__file__))
SQP_ROUNDOFF_LIMITED = -4   # csrc/sqp_core.h


def test_stalled_utility_optimisation_ends_roundoff_limited(emu_lib):
    net = synth.make_network(regions=40, sectors=26, seed=20260101 + 3)   # the network of flood40_1095
    agent = 323
    assert net["agent_type"][agent] == 0
    buf = np.ascontiguousarray(np.load(os.path.join(HERE, "golden", "consumer_stall_inputs.npy")))
    e = engine.Engine(net, lib_path=emu_lib)
    lib = e.lib
    lib.acc_debug_consumer_replay.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.c_void_p]
    ret = C.c_int(0)
    x = np.zeros(buf.size // 6)
    assert lib.acc_debug_consumer_replay(e.h, agent, buf.ctypes.data, buf.size, C.byref(ret), x.ctypes.data) == 0, lib.acc_last_error()
    e.close()
    assert ret.value == SQP_ROUNDOFF_LIMITED
    # the best point is the optimum the crawl was circling: the consumption the 200-iteration run had reached, to 1e-5
    prev = buf.reshape(-1, 6)[:, 4]
    assert np.all(x > 0) and np.allclose(x, prev, rtol=0.05)
    crawl = np.array([6.92786000e+07, 5.20552160e+07, 3.92292850e+07, 2.16003870e+07, 1.78222740e+07, 3.06653445e+08])
    assert np.allclose(x[:6], crawl, rtol=1e-5)
