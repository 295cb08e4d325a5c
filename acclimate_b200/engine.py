"""ctypes binding of the C ABI in include/acclimate_b200.h.

`Engine` is the thin host handle used by the tests, bench.py and the ModelRun mirror (modelrun.py).
It loads libacclimate_b200.so (CUDA only).  There is no CPU implementation behind it: a missing
library or a missing GPU raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import netfile

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libacclimate_b200.so")

MP_COUNT = len(netfile.MODEL_PARAMS)
SP_COUNT = len(netfile.SECTOR_PARAMS)

# reference include/ModelRun.h:39-73
EVENT_NAMES = [
    "NO_CONSUMPTION", "STORAGE_UNDERRUN", "STORAGE_OVERRUN", "NO_PRODUCTION_SUPPLY_SHORTAGE",
    "NO_PRODUCTION_DEMAND_QUANTITY_SHORTAGE", "NO_PRODUCTION_DEMAND_VALUE_SHORTAGE", "NO_PRODUCTION_HIGH_COSTS",
    "NO_EXP_PRODUCTION_SUPPLY_SHORTAGE", "NO_EXP_PRODUCTION_DEMAND_QUANTITY_SHORTAGE",
    "NO_EXP_PRODUCTION_DEMAND_VALUE_SHORTAGE", "NO_EXP_PRODUCTION_HIGH_COSTS", "DEMAND_FULFILL_HISTORY_UNDERFLOW",
    "OPTIMIZER_TIMEOUT", "OPTIMIZER_MAXITER", "OPTIMIZER_ROUNDOFF_LIMITED",
]


class _Network(C.Structure):
    _fields_ = [
        ("n_sectors", C.c_int64), ("n_regions", C.c_int64), ("n_agents", C.c_int64), ("n_edges", C.c_int64),
        ("agent_type", C.c_void_p), ("agent_sector", C.c_void_p), ("agent_region", C.c_void_p), ("agent_beta", C.c_void_p),
        ("edge_from", C.c_void_p), ("edge_to", C.c_void_p), ("edge_zq", C.c_void_p), ("delay", C.c_void_p),
        ("sector_params", C.c_void_p), ("consumer_elasticity", C.c_void_p),
        ("n_baskets", C.c_int64), ("basket_of_sector", C.c_void_p), ("basket_substitution", C.c_void_p), ("inter_basket_substitution", C.c_double),
        ("n_entities", C.c_int64), ("entity_delay", C.c_void_p), ("route_off", C.c_void_p), ("route_entity", C.c_void_p),
    ]


class _Params(C.Structure):
    _fields_ = [("model", C.c_double * MP_COUNT)]


class _Event(C.Structure):
    _fields_ = [("step", C.c_int64), ("type", C.c_int32), ("a", C.c_int32), ("b", C.c_int32), ("value", C.c_double)]


class EngineError(RuntimeError):
    pass


def load_library(path=None):
    path = path or os.environ.get("ACCLIMATE_B200_LIB") or LIB_PATH  # the override serves kernel-variant experiments
    if not os.path.exists(path):
        raise EngineError(f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` (CUDA library, no CPU fallback)")
    lib = C.CDLL(path)
    lib.acc_create.restype = C.c_void_p
    lib.acc_create.argtypes = [C.POINTER(_Network), C.POINTER(_Params), C.c_int]
    lib.acc_destroy.argtypes = [C.c_void_p]
    lib.acc_destroy.restype = None
    lib.acc_last_error.restype = C.c_char_p
    lib.acc_set_agent_forcing.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    lib.acc_set_entity_passage.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    lib.acc_step.argtypes = [C.c_void_p]
    lib.acc_phase.argtypes = [C.c_void_p, C.c_int]
    lib.acc_sync.argtypes = [C.c_void_p]
    lib.acc_drain_events.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    lib.acc_set_event_capacity.argtypes = [C.c_void_p, C.c_size_t]
    lib.acc_read.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    lib.acc_read_state.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_int)]
    for name in ("acc_edge_count", "acc_step_count", "acc_launch_count"):
        getattr(lib, name).restype = C.c_int64
        getattr(lib, name).argtypes = [C.c_void_p]
    lib.acc_stream.restype = C.c_void_p
    lib.acc_stream.argtypes = [C.c_void_p]
    lib.acc_profile.argtypes = [C.c_void_p, C.c_int]
    lib.acc_kernel_times.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    lib.acc_state_bytes.argtypes = [C.c_void_p, C.POINTER(C.c_size_t)]
    lib.acc_save_state.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    lib.acc_load_state.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    return lib


def consumer_baskets(arrays):
    """Consumer baskets of a snapshot: (basket_of_sector, basket_coeff, inter_basket_coeff).  A snapshot without them gets
    the oracle harness's default: one basket holding every supplying sector, substitution 2.0 inside, 1.5 between."""
    nS = int(np.asarray(arrays["meta"])[0])
    if "basket_of_sector" in arrays:
        bos = np.ascontiguousarray(arrays["basket_of_sector"], dtype=np.int64)
        bco = np.ascontiguousarray(arrays["basket_coeff"], dtype=np.float64)
        inter = float(np.asarray(arrays["inter_basket_coeff"]).reshape(-1)[0])
    else:
        bos = np.zeros(nS, dtype=np.int64)
        bos[nS - 1] = -1  # the consumer sector supplies nothing
        bco = np.array([2.0])
        inter = 1.5
    if bos.size != nS:
        raise EngineError("basket_of_sector needs one entry per sector")
    return bos, bco, inter


def network_struct(arrays):
    """acc_network / acc_params views of a snapshot dict; the third value keeps the arrays alive."""
    a = {k: np.ascontiguousarray(v) for k, v in arrays.items()}
    meta = a["meta"].astype(np.int64)
    keep = [
        a["agent_type"].astype(np.int64), a["agent_sector"].astype(np.int64), a["agent_region"].astype(np.int64),
        a["agent_beta"].astype(np.float64), a["edge_from"].astype(np.int64), a["edge_to"].astype(np.int64),
        a["edge_zq"].astype(np.int64), a["delay"].astype(np.int64), a["sector_f64"].astype(np.float64),
        a["consumer_elasticity"].astype(np.float64),
    ]
    bos, bco, inter = consumer_baskets(a)
    keep += [bos, bco]
    routes = [0, None, None, None]
    if "entity_delay" in a:
        rt = [a[k].astype(np.int64) for k in ("entity_delay", "route_off", "route_entity")]
        if rt[1].size != 2 * int(meta[1]) ** 2 + 1:
            raise EngineError("route_off needs 2 * n_regions^2 + 1 entries")
        keep += rt
        routes = [int(rt[0].size)] + [x.ctypes.data for x in rt]
    net = _Network(int(meta[0]), int(meta[1]), int(meta[2]), int(meta[3]), *[x.ctypes.data for x in keep[:10]], int(bco.size), bos.ctypes.data,
                   bco.ctypes.data, inter, *routes)
    par = _Params()
    mp = a["model_f64"].astype(np.float64)
    if mp.size != MP_COUNT:
        raise EngineError("model_f64 has the wrong length")
    for i in range(MP_COUNT):
        par.model[i] = float(mp[i])
    return net, par, keep


class Engine:
    """One device-resident model.  `arrays` is a network snapshot as produced by synth.make_network / netfile.read."""

    def __init__(self, arrays, device=0, lib_path=None):
        self.lib = load_library(lib_path)
        net, par, self._keep = network_struct(arrays)
        self.n_sectors, self.n_regions, self.n_agents, self.n_edges = (int(x) for x in np.asarray(arrays["meta"])[:4])
        self.h = self.lib.acc_create(C.byref(net), C.byref(par), int(device))
        if not self.h:
            raise EngineError(self.lib.acc_last_error().decode())

    def _check(self, rc):
        if rc != 0:
            raise EngineError(self.lib.acc_last_error().decode())

    def fork(self):
        """acc_fork: an ensemble member -- shares this engine's network tables on the device, owns a copy of its current state."""
        self.lib.acc_fork.restype = C.c_void_p
        self.lib.acc_fork.argtypes = [C.c_void_p]
        h = self.lib.acc_fork(self.h)
        if not h:
            raise EngineError(self.lib.acc_last_error().decode())
        member = object.__new__(type(self))
        member.lib, member._keep = self.lib, self._keep
        member.n_sectors, member.n_regions, member.n_agents, member.n_edges = self.n_sectors, self.n_regions, self.n_agents, self.n_edges
        member.h = h
        return member

    def close(self):
        if getattr(self, "h", None):
            self.lib.acc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_agent_forcing(self, lam):
        lam = np.ascontiguousarray(lam, dtype=np.float64)
        self._check(self.lib.acc_set_agent_forcing(self.h, lam.ctypes.data, lam.size))

    def set_entity_passage(self, nu):
        nu = np.ascontiguousarray(nu, dtype=np.float64)
        self._check(self.lib.acc_set_entity_passage(self.h, nu.ctypes.data, nu.size))

    def step(self):
        self._check(self.lib.acc_step(self.h))

    def phase(self, which):
        self._check(self.lib.acc_phase(self.h, int(which)))

    def sync(self):
        self._check(self.lib.acc_sync(self.h))

    @property
    def steps_done(self):
        return int(self.lib.acc_step_count(self.h))

    @property
    def launches(self):
        return int(self.lib.acc_launch_count(self.h))

    def stream(self):
        return self.lib.acc_stream(self.h)

    def pending_events(self):
        """Number of events raised since the last drain (nothing is taken)."""
        n = C.c_size_t(0)
        self._check(self.lib.acc_drain_events(self.h, None, 0, C.byref(n)))
        return int(n.value)

    def events(self):
        """Takes all pending events: (step, type, a, b, value), ordered by step, type and indices.  Raises when the engine's
        event buffer overflowed since the last drain (drain every step, or set_event_capacity)."""
        cap = max(self.pending_events(), 1)
        buf = (_Event * cap)()
        n = C.c_size_t(0)
        self._check(self.lib.acc_drain_events(self.h, buf, cap, C.byref(n)))
        m = min(cap, int(n.value))
        return [(int(buf[k].step), int(buf[k].type), int(buf[k].a), int(buf[k].b), float(buf[k].value)) for k in range(m)]

    def set_event_capacity(self, capacity):
        self._check(self.lib.acc_set_event_capacity(self.h, int(capacity)))

    def state_bytes(self):
        """Size of the state blob (acc_state_bytes): what one more ensemble member costs in device memory."""
        n = C.c_size_t(0)
        self._check(self.lib.acc_state_bytes(self.h, C.byref(n)))
        return int(n.value)

    def save_state(self):
        """Checkpoint: everything that changes from step to step as one blob (numpy uint8 array); see acc_save_state."""
        n = C.c_size_t(0)
        self._check(self.lib.acc_state_bytes(self.h, C.byref(n)))
        blob = np.empty(int(n.value), dtype=np.uint8)
        self._check(self.lib.acc_save_state(self.h, blob.ctypes.data, blob.nbytes, C.byref(n)))
        return blob[: int(n.value)]

    def load_state(self, blob):
        """Restart from a blob of save_state(), taken from an engine created from the same snapshot."""
        blob = np.ascontiguousarray(blob, dtype=np.uint8)
        self._check(self.lib.acc_load_state(self.h, blob.ctypes.data, blob.nbytes))

    def read_state(self, name):
        n = C.c_size_t(0)
        isf = C.c_int(0)
        self._check(self.lib.acc_read_state(self.h, name.encode(), None, 0, C.byref(n), C.byref(isf)))
        out = np.empty(int(n.value), dtype=np.float64 if isf.value else np.int64)
        self._check(self.lib.acc_read_state(self.h, name.encode(), out.ctypes.data, out.nbytes, C.byref(n), C.byref(isf)))
        return out

    def read(self, group, variable, out=None):
        n = C.c_size_t(0)
        if out is None:
            self._check(self.lib.acc_read(self.h, group.encode(), variable.encode(), None, 0, C.byref(n)))
            out = np.empty(int(n.value), dtype=np.float64)
        self._check(self.lib.acc_read(self.h, group.encode(), variable.encode(), out.ctypes.data, out.size, C.byref(n)))
        return out

    def observe(self, group, variable, select_from=None, select_to=None, out=None):
        """acc_observe: one observable of one group, shaped like ArrayOutput's data of one timestep (NaN where there is no
        object).  select_from / select_to: agent (sector, region) indices -- the reference's select_* of an `outputs:` entry."""
        lib = self.lib
        lib.acc_observe.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]
        i0 = None if select_from is None else np.ascontiguousarray(select_from, dtype=np.int64)
        i1 = None if select_to is None else np.ascontiguousarray(select_to, dtype=np.int64)
        a0 = (i0.ctypes.data, i0.size) if i0 is not None else (None, 0)
        a1 = (i1.ctypes.data, i1.size) if i1 is not None else (None, 0)
        n = C.c_size_t(0)
        if out is None:
            self._check(lib.acc_observe(self.h, group.encode(), variable.encode(), a0[0], a0[1], a1[0], a1[1], None, 0, C.byref(n)))
            out = np.empty(int(n.value), dtype=np.float64)
        self._check(lib.acc_observe(self.h, group.encode(), variable.encode(), a0[0], a0[1], a1[0], a1[1], out.ctypes.data, out.size, C.byref(n)))
        return out[: int(n.value)]

    def profile(self, enable=True):
        self._check(self.lib.acc_profile(self.h, 1 if enable else 0))

    def kernel_times(self, reset=True, cap=32):
        names = (C.c_char_p * cap)()
        ms = (C.c_double * cap)()
        launches = (C.c_int64 * cap)()
        cnt = self.lib.acc_kernel_times(self.h, names, ms, launches, cap, 1 if reset else 0)
        if cnt < 0:
            raise EngineError(self.lib.acc_last_error().decode())
        return {names[k].decode(): (float(ms[k]), int(launches[k])) for k in range(min(cnt, cap))}


class Scenario:
    """Forcing tables of a snapshot applied with the rule of the reference's Scenario::iterate
    (reference src/scenario/Scenario.cpp:161-174): value inside [from, to], reset at to + delta_t."""

    def __init__(self, arrays):
        self.n_agents = int(arrays["meta"][2])
        R = int(arrays["meta"][1])
        self.n_conn = int(np.asarray(arrays["entity_delay"]).size) if "entity_delay" in arrays else R * (R - 1) // 2
        self.forcing = np.asarray(arrays.get("forcing", np.zeros(0)), dtype=np.float64).reshape(-1, 4)
        self.passage = np.asarray(arrays.get("passage", np.zeros(0)), dtype=np.float64).reshape(-1, 4)
        self.dt = float(np.rint(arrays["model_f64"][netfile.MODEL_PARAMS.index("delta_t")]))
        self.lam = np.ones(self.n_agents)
        self.nu = np.full(self.n_conn, -1.0)

    def apply(self, engine, step):
        t = step * self.dt
        changed_l = changed_n = False
        for f, to, idx, val in self.forcing:
            f, to = np.rint(f), np.rint(to)
            if f <= t <= to:
                self.lam[int(idx)] = val
                changed_l = True
            elif t == to + self.dt:
                self.lam[int(idx)] = 1.0
                changed_l = True
        for f, to, idx, val in self.passage:
            f, to = np.rint(f), np.rint(to)
            if f <= t <= to:
                self.nu[int(idx)] = val
                changed_n = True
            elif t == to + self.dt:
                self.nu[int(idx)] = -1.0
                changed_n = True
        if changed_l:
            engine.set_agent_forcing(self.lam)
        if changed_n:
            engine.set_entity_passage(self.nu)


class EventSeriesScenario:
    """Per-step forcing grids streamed to the device, the rule of the reference's `event_series` scenario
    (reference src/scenario/EventSeriesScenario.cpp:45-69): at time index t every firm (sector s, region r) gets the
    forcing `agent_grid[t, s, r]` (NaN -> 1.0: no forcing) and every geo entity e the passage `entity_grid[t, e]`
    (negative -> -1: none).  The reference reads the grids from NetCDF files with the dimensions (time, sector, region)
    and (time, sea_route) (:71-77); here they are arrays -- the file layer is outside the path.  Time indices past the end
    of a grid leave the last forcing in place, as the reference's reader does when its file ends
    (src/scenario/ExternalScenario.cpp:83-110)."""

    def __init__(self, arrays, agent_grid=None, entity_grid=None):
        self.n_agents = int(arrays["meta"][2])
        at = np.asarray(arrays["agent_type"])
        self.firms = np.nonzero(at == 1)[0]
        self.f_sector = np.asarray(arrays["agent_sector"])[self.firms].astype(np.int64)
        self.f_region = np.asarray(arrays["agent_region"])[self.firms].astype(np.int64)
        self.agent_grid = None if agent_grid is None else np.asarray(agent_grid, dtype=np.float64)
        self.entity_grid = None if entity_grid is None else np.asarray(entity_grid, dtype=np.float64)
        if self.agent_grid is not None and self.agent_grid.ndim != 3:
            raise ValueError("agent_grid must have the dimensions (time, sector, region)")
        if self.entity_grid is not None and self.entity_grid.ndim != 2:
            raise ValueError("entity_grid must have the dimensions (time, geo entity)")
        self.lam = np.ones(self.n_agents)

    def apply(self, engine, step):
        if self.agent_grid is not None and step < self.agent_grid.shape[0]:
            v = self.agent_grid[step][self.f_sector, self.f_region]
            self.lam[self.firms] = np.where(np.isnan(v), 1.0, v)
            engine.set_agent_forcing(self.lam)
        if self.entity_grid is not None and step < self.entity_grid.shape[0]:
            v = self.entity_grid[step]
            engine.set_entity_passage(np.ascontiguousarray(np.where(v < 0.0, -1.0, v)))
