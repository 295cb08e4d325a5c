"""Debugging aid: find a consumer whose utility optimisation fails in a long shocked run, and replay that one solve elsewhere.

  capture (GPU box, library built with -DACC_DEBUG_API):
      ACCLIMATE_B200_LIB=variants/libacc_dbg.so python scripts/consumer_replay.py capture [steps] [first_step_to_record]
      -> gpurun_out/consumer_fail.npz (inputs of every consumer before the failing step, failing agents, step)
  replay (anywhere, e.g. the host emulation built with -DACC_DEBUG_API [-DACCSQP_TRACE]):
      python scripts/consumer_replay.py replay <lib.so> [file.npz]
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from acclimate_b200 import engine, synth  # noqa: E402


def network():
    a = synth.make_network(regions=40, sectors=26, seed=20260101 + 3)  # tests/golden/make_golden.py: flood40_1095
    synth.add_shock(a, 10, 24, 17, 0.5)
    return a


def inputs(e, lib, agent):
    n = C.c_size_t(0)
    lib.acc_debug_consumer_inputs.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    assert lib.acc_debug_consumer_inputs(e.h, agent, None, 0, C.byref(n)) == 0
    buf = np.zeros(n.value)
    assert lib.acc_debug_consumer_inputs(e.h, agent, buf.ctypes.data, buf.size, C.byref(n)) == 0, lib.acc_last_error()
    return buf


def capture(steps, first):
    net = network()
    e = engine.Engine(net)
    lib = e.lib
    sc = engine.Scenario(net)
    cons = np.nonzero(np.asarray(net["agent_type"]) == 0)[0]
    last = None
    for t in range(steps):
        sc.apply(e, t)
        if t >= first:
            last = {int(a): inputs(e, lib, int(a)) for a in cons}
        try:
            e.step()
            e.sync()
        except engine.EngineError as ex:
            print("step", t, "failed:", ex, flush=True)
            ret = e.read_state("consumer_solver_ret")
            bad = [int(a) for a in cons if ret[a] == -1]
            print("failing consumers:", bad, "return codes:", {int(a): int(ret[a]) for a in cons if ret[a] < 0})
            if last is not None:
                os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
                np.savez(os.path.join(ROOT, "gpurun_out", "consumer_fail.npz"), step=t, bad=np.array(bad), agents=np.array(sorted(last)),
                         inputs=np.stack([last[a] for a in sorted(last)]))
            return
        if t % 100 == 0:
            print("step", t, "ok", flush=True)
    print("no failure in", steps, "steps")


def replay(lib_path, path):
    d = np.load(path)
    net = network()
    e = engine.Engine(net, lib_path=lib_path)
    lib = e.lib
    lib.acc_debug_consumer_replay.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.c_void_p]
    agents = list(d["agents"])
    only = [int(x) for x in os.environ.get("REPLAY_AGENTS", "").split(",") if x]
    for a in (only or list(d["bad"]) or agents):
        buf = np.ascontiguousarray(d["inputs"][agents.index(a)])
        ret = C.c_int(0)
        x = np.zeros(buf.size // 6)
        rc = lib.acc_debug_consumer_replay(e.h, int(a), buf.ctypes.data, buf.size, C.byref(ret), x.ctypes.data)
        assert rc == 0, lib.acc_last_error()
        print("agent", a, "ret", ret.value, "consumption", x[:6], "...")
    e.close()


if __name__ == "__main__":
    if sys.argv[1] == "capture":
        capture(int(sys.argv[2]) if len(sys.argv) > 2 else 1095, int(sys.argv[3]) if len(sys.argv) > 3 else 1000)
    else:
        replay(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "gpurun_out", "consumer_fail.npz"))
