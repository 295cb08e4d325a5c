"""GPU tier of the region-partitioned run: the CUDA library's pack / unpack kernels and staged step, two parts.
test_two_parts_on_gpu: on a single-GPU box both parts share cuda:0 and the exchange is staged through the host over gloo
(the kernels under test are the same).  test_two_parts_over_nccl needs two GPUs and SKIPS LOUDLY without them: only that
one exercises the NCCL data path (bench.py --gpus 2 prints the state_hash that proves it on the driver's scaling run)."""
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
    backend = "gloo"  # both parts on cuda:0, exchange staged through the host; the NCCL path is test_two_parts_over_nccl
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


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["severe_shock30", "routes25"])
def test_two_parts_over_nccl(case, cuda_lib):
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("NCCL HALO PATH NOT TESTED: this box has %d GPU(s), the two-part NCCL run needs 2 "
                    "(gpurun --gpus 2 -- python -m pytest tests/test_partition_gpu.py -m gpu)" % torch.cuda.device_count())
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, cuda_lib, "nccl", out)) for r in range(world)]
    for p in procs:
        p.start()
    results = [out.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    problems = [m for _, pr, _ in results for m in pr]
    assert not problems, "\n".join(problems)
