"""-m gpu, needs >= 2 GPUs (skipped otherwise): the destination-range partition over NCCL with the
real CUDA kernels equals the single-GPU backbone (same checks as tests/test_partition_gloo.py)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.test_partition_gloo import _free_port, run_case

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, kind):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        run_case(kind, rank, world, f"cuda:{rank}", emulate=False)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind", ["pna", "gine", "tabgnn"])
def test_partition_nccl(kind):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    mp.spawn(_worker, args=(2, _free_port(), kind), nprocs=2, join=True)
