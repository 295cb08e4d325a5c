"""GPU tier of the region-partitioned run: the CUDA library's pack / unpack kernels and staged step, two parts.
With two or more GPUs each part has its own device and the exchanges run over NCCL; on a single-GPU box both parts
share cuda:0 and the exchange is staged through the host over gloo (the kernels under test are the same)."""
import pytest

import parity
from test_partition import _free_port


def _worker(rank, world, port, case, lib, backend, out):
    import os

    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dev = rank if backend == "nccl" else 0
    torch.cuda.set_device(dev)
    if backend == "nccl":
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", dev))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        problems, halo = parity.run_case_partitioned(case, lib, device=dev, tensor_device="cuda:%d" % dev)
        out.put((rank, problems, halo))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["severe_shock30", "passage25"])
def test_two_parts_on_gpu(case, cuda_lib):
    import torch
    import torch.multiprocessing as mp

    world = 2
    backend = "nccl" if torch.cuda.device_count() >= world else "gloo"
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, cuda_lib, backend, out)) for r in range(world)]
    for p in procs:
        p.start()
    results = [out.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    problems = [m for _, pr, _ in results for m in pr]
    assert not problems, "\n".join(problems)
