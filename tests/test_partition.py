"""Region-partitioned run on the CPU tier: world_size 2 and 3 over gloo, the kernel logic from the
host-emulation build, every rank checked against the reference oracle's golden dump of the UNPARTITIONED run --
so a pass means P parts == 1 part == reference, bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

import parity


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, case, lib, owner, out):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        problems, halo = parity.run_case_partitioned(case, lib, tensor_device="cpu", region_owner=owner)
        out.put((rank, problems, halo))
    finally:
        dist.destroy_process_group()


def _run(world, case, lib, owner=None):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, lib, owner, out)) for r in range(world)]
    for p in procs:
        p.start()
    results = [out.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    return sorted(results)


@pytest.mark.parametrize("case,world", [("severe_shock30", 2), ("passage25", 2), ("flags20", 3), ("baskets20", 2), ("routes25", 2)])
def test_parts_equal_reference(case, world, emu_lib):
    results = _run(world, case, emu_lib)
    problems = [m for _, p, _ in results for m in p]
    assert not problems, "\n".join(problems)
    assert all(h[0] > 0 and h[1] > 0 for _, _, h in results), "the partition has no cross-part connections: nothing was exchanged"


def test_interleaved_ownership(emu_lib):
    """Regions dealt out round-robin instead of in contiguous blocks: ownership is arbitrary per region."""
    net, _, _ = parity.load_case("outage25")
    owner = np.arange(int(net["meta"][1])) % 2
    results = _run(2, "outage25", emu_lib, owner)
    problems = [m for _, p, _ in results for m in p]
    assert not problems, "\n".join(problems)


def test_halo_lists_pair_up():
    """The send list of part p towards q names the same connections, in the same order, as q's receive list from p."""
    from acclimate_b200 import partition

    net, _, _ = parity.load_case("severe_shock30")
    owner = partition.assign_regions(net, 3)
    parts = [partition.Part(net, owner, r) for r in range(3)]
    for p in parts:
        so = np.concatenate([[0], np.cumsum(p.send_counts)])
        for q in parts:
            ro = np.concatenate([[0], np.cumsum(q.recv_counts)])
            sent = p.edge_gid[p.send_edges[so[q.rank]:so[q.rank + 1]]]
            received = q.edge_gid[q.recv_edges[ro[p.rank]:ro[p.rank + 1]]]
            assert np.array_equal(sent, received)
    # every connection is computed exactly once on each side
    n = int(net["meta"][3])
    sellers = np.zeros(n, dtype=int)
    buyers = np.zeros(n, dtype=int)
    for p in parts:
        sellers[p.edge_gid[p.sells]] += 1
        buyers[p.edge_gid[p.buys]] += 1
    assert (sellers == 1).all() and (buyers == 1).all()


def _hash_worker(rank, world, port, lib, out):
    import torch.distributed as dist

    from acclimate_b200 import engine, partition

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        net, _, _ = parity.load_case("severe_shock30")
        E = partition.PartitionedEngine(net, lib_path=lib, tensor_device="cpu")
        sc = engine.Scenario(net)
        for t in range(8):
            sc.apply(E, t)
            E.step()
        out.put((rank, partition.state_hash(E, net, part=E.part)))
        E.close()
    finally:
        dist.destroy_process_group()


def test_state_hash_is_partition_invariant(emu_lib):
    """bench.py's state_hash: the same fingerprint from one engine and from two parts (what the driver compares over N)."""
    import torch.multiprocessing as mp

    from acclimate_b200 import engine, partition

    net, _, _ = parity.load_case("severe_shock30")
    E = engine.Engine(net, lib_path=emu_lib)
    sc = engine.Scenario(net)
    hashes = []
    for t in range(8):
        sc.apply(E, t)
        E.step()
        hashes.append(partition.state_hash(E, net))
    E.close()
    assert len(set(hashes)) > 4  # the fingerprint follows the state
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_hash_worker, args=(r, 2, port, emu_lib, out)) for r in range(2)]
    for p in procs:
        p.start()
    got = [out.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert {h for _, h in got} == {hashes[-1]}
