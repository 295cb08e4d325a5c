"""Region-partitioned run: one engine per GPU, agents split by region.

The reference runs its agent loops with OpenMP over shared memory (reference src/model/Model.cpp:59-118).  Across GPUs
the same loops are split by the REGION of the agent; what the reference passes between agents through pointers
becomes three exchanges per timestep:

  shipments       seller -> buyer   BusinessConnection::push_flow_Z            (reference src/model/BusinessConnection.cpp:87-94)
  supply params   seller -> all     SalesManager::communicated_parameters     (reference src/model/SalesManager.cpp:611-624)
  demand requests buyer  -> seller  BusinessConnection::send_demand_request_D  (reference src/model/BusinessConnection.cpp:127-131)

Every connection datum has one writer and one reader per step, every agent sees its complete connection lists in the
global order, and reductions keep their fixed order: a P-part run is bit-identical to the single-engine run.

torch.distributed is the plumbing here (NCCL all-to-all / all-reduce on the engine's stream; gloo in the CPU tests);
the engine itself (csrc/engine.cu) only packs and unpacks the exchange buffers.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import engine as _engine

SELLER_WORDS = 9  # ACC_SELLER_WORDS, include/acclimate_b200.h


def assign_regions(arrays, world):
    """Contiguous blocks of regions per part, balanced by the number of connections that end in the block
    (the purchasing side is the expensive one).  Returns region -> part."""
    R = int(arrays["meta"][1])
    if world > R:
        raise ValueError("more parts than regions")
    reg = np.asarray(arrays["agent_region"], dtype=np.int64)
    load = np.bincount(reg[np.asarray(arrays["edge_to"], dtype=np.int64)], minlength=R).astype(np.float64)
    load += np.bincount(reg[np.asarray(arrays["edge_from"], dtype=np.int64)], minlength=R) * 0.25
    cum = np.cumsum(load)
    owner = np.minimum((cum - load / 2) * world / cum[-1], world - 1).astype(np.int64)
    # every part needs at least one region: repair empty parts by walking the boundaries
    for p in range(world):
        if not np.any(owner == p):
            owner = np.minimum(np.arange(R) * world // R, world - 1)
            break
    return owner


def assign_regions_uniform(n_regions, world):
    """Equal contiguous blocks of regions (what assign_regions gives for a dense network of equally sized regions)."""
    return np.minimum(np.arange(n_regions) * world // n_regions, world - 1)


class Part:
    """What one rank needs: its sub-network, ownership mask and halo lists (local connection ids grouped by peer).
    `local=True`: `arrays` already holds only this part's connections (synth.make_network(part=...))."""

    def __init__(self, arrays, region_owner, rank, local=False):
        a = arrays
        world = int(np.max(region_owner)) + 1
        agent_part = np.asarray(region_owner, dtype=np.int8)[np.asarray(a["agent_region"], dtype=np.int64)]
        self.rank, self.world = rank, world
        self.agent_part = agent_part
        self.owned = (agent_part == rank).astype(np.uint8)
        ef = np.asarray(a["edge_from"], dtype=np.int64)
        et = np.asarray(a["edge_to"], dtype=np.int64)
        pf, pt = agent_part[ef], agent_part[et]
        if local:
            if not np.all((pf == rank) | (pt == rank)):
                raise ValueError("a local snapshot may only hold connections of this part")
            self.edge_gid = None
            self.global_edges = int(np.asarray(a["global_edges"]).reshape(-1)[0])
            lf, lt = pf, pt
            self.arrays = a
            meta = np.asarray(a["meta"], dtype=np.int64)
        else:
            keep = (pf == rank) | (pt == rank)
            self.edge_gid = np.nonzero(keep)[0]          # local id -> global connection id (order preserved)
            self.global_edges = int(ef.size)
            lf, lt = pf[keep], pt[keep]
            local_arrays = {k: v for k, v in a.items()}
            local_arrays["edge_from"], local_arrays["edge_to"] = ef[keep], et[keep]
            local_arrays["edge_zq"] = np.asarray(a["edge_zq"], dtype=np.int64)[keep]
            meta = np.array(a["meta"], dtype=np.int64).copy()
            meta[3] = int(keep.sum())
            local_arrays["meta"] = meta
            self.arrays = local_arrays
        ids = np.arange(meta[3], dtype=np.int64)
        send, recv, self.send_counts, self.recv_counts = [], [], [], []
        for p in range(world):
            s = ids[(lf == rank) & (lt == p)] if p != rank else ids[:0]
            r = ids[(lt == rank) & (lf == p)] if p != rank else ids[:0]
            send.append(s); recv.append(r)
            self.send_counts.append(int(s.size)); self.recv_counts.append(int(r.size))
        self.send_edges = np.concatenate(send) if send else ids[:0]
        self.recv_edges = np.concatenate(recv) if recv else ids[:0]
        # connections whose seller / buyer this part owns (their seller-side / buyer-side state is authoritative here)
        self.sells = lf == rank
        self.buys = lt == rank


class PartitionedEngine(_engine.Engine):
    """One part of a region-partitioned model.  Needs an initialised torch.distributed process group with one rank
    per part (NCCL for the CUDA library; the CPU tests drive the emulation library over gloo)."""

    def __init__(self, arrays, region_owner=None, device=0, lib_path=None, group=None, tensor_device=None, local=False, plane=None, peer_stores=True):
        """plane: "nccl" -- the engine exchanges its halos itself (acc_comm_init / acc_step_partitioned: NCCL from C++ on the
        engine's stream; default when the process group's backend is NCCL); "host" -- this class moves the exchange buffers
        between the four stages with torch.distributed (gloo in the CPU tests, where the emulation library has no NCCL)."""
        import torch
        import torch.distributed as dist

        self.torch, self.dist, self.group = torch, dist, group
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        if region_owner is None:
            region_owner = assign_regions(arrays, world)
        if int(np.max(region_owner)) + 1 != world:
            raise _engine.EngineError("region_owner does not name one part per rank")
        self.part = Part(arrays, np.asarray(region_owner), rank, local=local)
        self.lib = lib = _engine.load_library(lib_path)
        self._bind()
        net, par, self._keep = _engine.network_struct(self.part.arrays)
        meta = self.part.arrays["meta"]
        self.n_sectors, self.n_regions, self.n_agents, self.n_edges = (int(x) for x in meta[:4])
        self._owned = np.ascontiguousarray(self.part.owned)
        self.h = self.lib.acc_create_partition(C.byref(net), C.byref(par), int(device), self._owned.ctypes.data)
        if not self.h:
            raise _engine.EngineError(self.lib.acc_last_error().decode())
        if tensor_device is None:
            tensor_device = "cuda:%d" % device
        self.tensor_device = torch.device(tensor_device)
        if plane is None:
            plane = "nccl" if (world > 1 and dist.get_backend(group) == "nccl" and self.tensor_device.type == "cuda") else "host"
        self.plane = plane
        p = self.part
        ns, nr = int(p.send_edges.size), int(p.recv_edges.size)
        self._send_ids = np.ascontiguousarray(p.send_edges)
        self._recv_ids = np.ascontiguousarray(p.recv_edges)
        if plane == "nccl":
            # the engine keeps the exchange buffers and talks to its peers itself: nothing of the data plane is left here
            self._check(self.lib.acc_set_halo(self.h, self._send_ids.ctypes.data, ns, self._recv_ids.ctypes.data, nr, None, None, None, None, None))
            ident = torch.zeros(128, dtype=torch.uint8)
            if rank == 0:
                buf = (C.c_ubyte * 128)()
                lib.acc_comm_unique_id.argtypes = [C.c_void_p]
                self._check(lib.acc_comm_unique_id(buf))
                ident = torch.tensor(list(buf), dtype=torch.uint8)
            ident = ident.to(self.tensor_device)
            dist.broadcast(ident, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            raw = bytes(ident.cpu().tolist())
            part32 = np.ascontiguousarray(p.agent_part, dtype=np.int32)
            sc = np.ascontiguousarray(p.send_counts, dtype=np.int64)
            rc = np.ascontiguousarray(p.recv_counts, dtype=np.int64)
            lib.acc_comm_init.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
            lib.acc_step_partitioned.argtypes = [C.c_void_p]
            self._check(lib.acc_comm_init(self.h, raw, rank, world, part32.ctypes.data, sc.ctypes.data, rc.ctypes.data))
            self.peer_stores = False
            if peer_stores:
                lib.acc_comm_enable_peer_stores.argtypes = [C.c_void_p]
                ok = torch.tensor([1 if lib.acc_comm_enable_peer_stores(self.h) == 0 else 0], dtype=torch.int64, device=self.tensor_device)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
                self.peer_stores = int(ok[0]) == 1
                if not self.peer_stores:  # some part could not map its peers: everybody stays with the NCCL exchanges
                    lib.acc_comm_disable_peer_stores.argtypes = [C.c_void_p]
                    lib.acc_comm_disable_peer_stores(self.h)
        else:
            mk = lambda rows, cols: torch.zeros((rows, cols), dtype=torch.int64, device=self.tensor_device)
            self.flow_send, self.flow_recv = mk(ns, 2), mk(nr, 2)
            self.demand_send, self.demand_recv = mk(nr, 2), mk(ns, 2)
            self.seller_table = mk(self.n_agents, SELLER_WORDS)
            self._check(self.lib.acc_set_halo(self.h, self._send_ids.ctypes.data, ns, self._recv_ids.ctypes.data, nr, self.flow_send.data_ptr(),
                                              self.flow_recv.data_ptr(), self.demand_send.data_ptr(), self.demand_recv.data_ptr(),
                                              self.seller_table.data_ptr()))
        # one solver group size for all parts: the reduction order is part of the arithmetic
        lib.acc_solver_group.argtypes = [C.c_void_p]
        lib.acc_set_solver_group.argtypes = [C.c_void_p, C.c_int]
        grp = torch.tensor([self.lib.acc_solver_group(self.h)], dtype=torch.int64)
        if world > 1:
            g_dev = grp.to(self.tensor_device) if dist.get_backend(group) == "nccl" else grp
            dist.all_reduce(g_dev, op=dist.ReduceOp.MAX, group=group)
            grp = g_dev.cpu()
        self._check(self.lib.acc_set_solver_group(self.h, int(grp[0])))
        self._stream = None
        self.time_exchanges = False      # bench.py: CUDA events around every exchange
        self._exchange_events = []
        if self.tensor_device.type == "cuda":
            self._stream = torch.cuda.ExternalStream(self.stream(), device=self.tensor_device)
            torch.cuda.synchronize(self.tensor_device)  # the exchange buffers were zeroed on torch's stream, not the engine's

    def _bind(self):
        lib = self.lib
        lib.acc_create_partition.restype = C.c_void_p
        lib.acc_create_partition.argtypes = [C.POINTER(_engine._Network), C.POINTER(_engine._Params), C.c_int, C.c_void_p]
        lib.acc_set_halo.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t] + [C.c_void_p] * 5
        lib.acc_stage.argtypes = [C.c_void_p, C.c_int]

    def _stage(self, k):
        self._check(self.lib.acc_stage(self.h, k))

    def _staged(self):
        # gloo moves host memory only: with device tensors (two parts sharing one GPU in the single-GPU test tier) the
        # exchange is staged through the host.  The NCCL path never takes this branch.
        return self.tensor_device.type == "cuda" and self.dist.get_backend(self.group) == "gloo"

    def _exchange(self, recv, send, recv_counts, send_counts):
        if self.world == 1:
            return
        if self._staged():
            self.sync()
            h_send, h_recv = send.cpu(), self.torch.empty(recv.shape, dtype=recv.dtype)
            self.dist.all_to_all_single(h_recv, h_send, output_split_sizes=recv_counts, input_split_sizes=send_counts, group=self.group)
            recv.copy_(h_recv)
            return
        self.dist.all_to_all_single(recv, send, output_split_sizes=recv_counts, input_split_sizes=send_counts, group=self.group)

    def _gather_sellers(self):
        if self.world == 1:
            return
        if self._staged():
            self.sync()
            h = self.seller_table.cpu()
            self.dist.all_reduce(h, group=self.group)
            self.seller_table.copy_(h)
            return
        self.dist.all_reduce(self.seller_table, group=self.group)

    @property
    def world(self):
        return self.part.world

    def step(self):
        """One timestep: four engine stages with the three exchanges between them, all ordered on the engine's stream."""
        p = self.part
        if self.plane == "nccl":
            self._check(self.lib.acc_step_partitioned(self.h))
            return
        if self._stream is not None:
            with self.torch.cuda.stream(self._stream):
                self._step(p)
        else:
            self._step(p)

    def _timed(self, fn, *args):
        if not (self.time_exchanges and self._stream is not None):
            return fn(*args)
        a, b = self.torch.cuda.Event(enable_timing=True), self.torch.cuda.Event(enable_timing=True)
        a.record(self._stream)
        fn(*args)
        b.record(self._stream)
        self._exchange_events.append((a, b))

    def exchange_ms(self, reset=True):
        """Device time spent in the three exchanges since the last reset (needs time_exchanges = True and a sync before)."""
        ms = sum(a.elapsed_time(b) for a, b in self._exchange_events)
        if reset:
            self._exchange_events = []
        return ms

    def _step(self, p):
        self._stage(0)  # possible production, production + distribution, pack shipments
        self._timed(self._exchange, self.flow_recv, self.flow_send, p.recv_counts, p.send_counts)
        self._stage(1)  # unpack shipments, storages + deliveries, expectation, pack supply parameters
        self._timed(self._gather_sellers)
        self._stage(2)  # unpack supply parameters, purchase, pack demand requests
        self._timed(self._exchange, self.demand_recv, self.demand_send, p.send_counts, p.recv_counts)
        self._stage(3)  # unpack demand requests, tick

    def halo_bytes_per_step(self):
        """bytes this rank sends + receives per step: two halos of 16 B per cross connection and the sellers' records (the
        all-gather of owned rows in the NCCL plane, the full-table all-reduce in the host plane)"""
        p = self.part
        return 2 * 16 * (int(p.send_edges.size) + int(p.recv_edges.size)) + 8 * SELLER_WORDS * self.n_agents

    # ---- reads: only what this part owns is meaningful ----
    def owned_agents(self):
        return np.nonzero(self.part.owned)[0]

    def read_owned(self, name, storage_agent=None):
        """Raw state array `name` with, next to it, the mask of entries this part is authoritative for and (for
        per-connection arrays) the global connection ids.  Per-agent arrays are indexed by global agent id."""
        x = self.read_state(name)
        p = self.part
        if name.startswith("edge_"):
            mask = p.buys if name.startswith("edge_D") else p.sells
            return x, mask, p.edge_gid
        if x.size == self.n_agents:
            return x, p.owned.astype(bool), None
        return x, None, None


def state_hash(E, arrays, part=None, group=None):
    """Order-free 64-bit fingerprint of the demand requests and storage contents of a run, the same number whether the
    network runs on one engine or split over P parts: sum over connections of (2 gid + 1) D_q plus sum over storages of
    (2 (sector n_agents + agent) + 1 + 2^40) S_q, modulo 2^64.  A part contributes the connections it buys on and the
    storages it owns; the parts' sums are added over the process group (as 32-bit halves: no overflow in the collective)."""
    import numpy as np

    nA = int(np.asarray(arrays["meta"])[2])
    D = E.read_state("edge_D_q").astype(np.uint64)
    S = E.read_state("sto_S_q").astype(np.uint64)
    sa, ss = E.read_state("sto_agent").astype(np.uint64), E.read_state("sto_sector").astype(np.uint64)
    if part is not None:
        if part.edge_gid is None:
            return None
        mine = part.buys
        gid = part.edge_gid[mine].astype(np.uint64)
        D = D[mine]
        own = part.owned[sa.astype(np.int64)].astype(bool)
        S, sa, ss = S[own], sa[own], ss[own]
    else:
        gid = np.arange(D.size, dtype=np.uint64)
    h = np.sum((np.uint64(2) * gid + np.uint64(1)) * D, dtype=np.uint64)
    h += np.sum((np.uint64(2) * (ss * np.uint64(nA) + sa) + np.uint64(1) + np.uint64(1 << 40)) * S, dtype=np.uint64)
    h = int(h)
    if part is not None and part.world > 1:
        import torch
        import torch.distributed as dist

        t = torch.tensor([h & 0xFFFFFFFF, h >> 32], dtype=torch.int64)
        if dist.get_backend(group) == "nccl":
            t = t.cuda()
        dist.all_reduce(t, group=group)
        lo, hi = int(t[0]), int(t[1])
        h = (lo + (hi << 32)) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h
