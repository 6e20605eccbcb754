"""CPU, world_size 2, gloo: the destination-range partition (b200mp/parallel.py) gives the same
activations, input gradients, parameter gradients and BatchNorm running statistics as the
single-process backbone.  Kernels are the plain-PyTorch emulation (tests/emulator.py): this checks
the HOST logic of the multi-GPU path (partition, padded all-gather / reduce-scatter, sync BN,
gradient all-reduce); the -m gpu variant runs the real kernels over NCCL."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

WORLD = 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run_case(kind, rank, world, device, emulate):
    import contextlib

    import b200mp
    from b200mp import parallel as PL
    from tests.util import make_graph, rel_err
    ctx = contextlib.nullcontext()
    if emulate:
        from tests.emulator import emulated_kernels
        ctx = emulated_kernels()
    n, e, H = 240, 2600, 16 if emulate else 64
    with ctx:
        ei = make_graph(n, e, 5).to(device)
        deg = torch.bincount(torch.bincount(ei[1].cpu(), minlength=n))
        torch.manual_seed(3)
        if kind == "pna":
            model = b200mp.PNAS(num_features=H, num_gnn_layers=2, n_hidden=H, edge_updates=True, edge_dim=H, deg=deg)
        else:
            model = b200mp.GINe(num_features=H, num_gnn_layers=2, n_hidden=H, edge_updates=True, edge_dim=H)
        model = model.to(device).train()
        sd0 = {k: v.clone() for k, v in model.state_dict().items()}
        g = torch.Generator().manual_seed(11)
        x = torch.randn(n, H, generator=g).to(device)
        ea = torch.randn(e, H, generator=g).to(device)
        wx, we = torch.randn(n, H, generator=g).to(device), torch.randn(e, H, generator=g).to(device)

        # single-process truth
        x1, e1 = x.clone().requires_grad_(True), ea.clone().requires_grad_(True)
        ox, oe = model.message_passing(x1, ei, e1)
        ((ox * wx).sum() + (oe * we).sum()).backward()
        ref_g = {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
        ref_sd = {k: v.clone() for k, v in model.state_dict().items()}
        model.zero_grad(set_to_none=True)
        model.load_state_dict(sd0)

        # partitioned
        part = PL.partition_graph(ei, n, rank, world)
        assert int(part.bounds[-1]) == n and part.st.num_edges > 0
        xl = part.shard_nodes(x).requires_grad_(True)
        el = part.shard_edges(ea).requires_grad_(True)
        px, pe = model.message_passing_partitioned(xl, el, part)
        ((px * part.shard_nodes(wx)).sum() + (pe * part.shard_edges(we)).sum()).backward()
        PL.all_reduce_gradients(model, part)

        assert rel_err(px, ox[part.lo:part.hi]) < 2e-5, ("x", rel_err(px, ox[part.lo:part.hi]))
        assert rel_err(pe, oe[part.edge_ids]) < 2e-5
        assert rel_err(xl.grad, x1.grad[part.lo:part.hi]) < 5e-4
        assert rel_err(el.grad, e1.grad[part.edge_ids]) < 5e-4
        gmax = max(float(v.abs().max()) for v in ref_g.values())
        for k, p in model.named_parameters():
            if k in ref_g:
                assert rel_err(p.grad, ref_g[k], 1e-2 * gmax) < 5e-4, (k, rel_err(p.grad, ref_g[k], 1e-2 * gmax))
        for k, v in model.state_dict().items():
            if "running" in k:
                assert rel_err(v, ref_sd[k]) < 1e-5, k


def _worker(rank, world, port, kind):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        run_case(kind, rank, world, "cpu", emulate=True)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind", ["pna", "gine"])
def test_partition_matches_single_process(kind):
    mp.spawn(_worker, args=(WORLD, _free_port(), kind), nprocs=WORLD, join=True)
