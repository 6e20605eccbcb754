#!/usr/bin/env python
"""bench.py -- message-passing edges/s (fwd+bwd) of the L-layer PNA (or GINE) backbone on a
synthetic transaction graph of a named shape, one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload hi_medium|hi_small|tiny]
                    [--model pna|gine] [--impl b200|reference]

Prints ONE JSON line (rank 0).  `value` = edges/s with inputs resident in HBM and the graph
structure already built; `e2e` = the same through the public module API from pinned HOST buffers
(H2D of x / edge_index / edge features, CSC build, fwd+bwd, D2H of the loss) every step;
`roofline` = the dominant kernel's algorithmic bytes / CUDA-event time against the measured HBM
peak; `cpu_baseline` = the CPU oracle (port of the reference path) on a bounded sample.
`--impl reference` times that CPU path alone (the reference has no GPU-independent build here:
torch_geometric is absent).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {   # name: (nodes, edges, seed)  -- SURVEY.md section 8(d)
    "hi_medium": (2_077_023, 31_898_238, 20240721),
    "hi_small": (515_080, 5_078_345, 20240719),
    "tiny": (8_989, 5_198 * 16, 20240718),
}
def _hidden_from_argv(default=64):
    """--hidden is read before the per-kernel byte tables below are built (they are functions of H); 64 = the headline width"""
    for i, a in enumerate(sys.argv):
        if a == "--hidden" and i + 1 < len(sys.argv):
            return int(sys.argv[i + 1])
        if a.startswith("--hidden="):
            return int(a.split("=", 1)[1])
    return default


H, LAYERS = _hidden_from_argv(), 2
K_COLS = 5                     # edge columns of the AML table (3 categorical + 1 numerical + 1 timestamp)
CPU_SAMPLE_EDGES = 1_000_000


def synth_graph(n, e, seed, device):
    """src uniform; dst = P[floor(n u^3)] (heavy-tailed in-degree, hubs); 10 % self loops;
    generation order (unsorted)."""
    g = torch.Generator(device=device).manual_seed(seed)
    src = torch.randint(0, n, (e,), generator=g, device=device)
    u = torch.rand(e, generator=g, device=device, dtype=torch.float64)
    P = torch.randperm(n, generator=g, device=device)
    dst = P[(u ** 3 * n).long().clamp_(max=n - 1)]
    loops = torch.rand(e, generator=g, device=device) < 0.1
    dst = torch.where(loops, src, dst)
    return torch.stack([src, dst])


def make_model(kind, deg_hist, device):
    import b200mp
    torch.manual_seed(1234)
    if kind == "pna":
        m = b200mp.PNAS(num_features=H, num_gnn_layers=LAYERS, n_hidden=H, edge_updates=True, edge_dim=H, deg=deg_hist)
    elif kind == "tabgnn":      # BASELINE.json config 5: PNA + FT-Transformer (column attention over the K + 1 tokens of every row)
        m = b200mp.TABGNN(channels=H, num_layers=LAYERS, deg=deg_hist, node_dim=H, nhidden=H, edge_dim=K_COLS * H, nhead=8, dropout=0.0)
    else:
        m = b200mp.GINe(num_features=H, num_gnn_layers=LAYERS, n_hidden=H, edge_updates=True, edge_dim=H)
    return m.to(device).train()


def algorithmic_bytes_per_edge(kind):
    """compulsory [E,H]-row touches per layer fwd+bwd, fp32."""
    s = 4
    return (22 if kind in ("pna", "tabgnn") else 19) * s * H + 64


def tab_bytes_per_row(S):
    """Column-attention layer (tab_conv + tab_norm) per token row, fwd + bwd, with the row-chunked recompute: forward reads the
    tokens and writes the result (2 S C s); backward reads the incoming gradient and the saved input and writes the input
    gradient (3 S C s)."""
    return 5 * S * H * 4


# algorithmic bytes per edge of each kernel (DESIGN.md "kernels"): [E,H] rows read/written or
# gathered/scattered once, plus the int32 index reads
KERNEL_BYTES = {
    "pna_message_fwd": 3 * 4 * H + 8,      # read e, gather B[src], write m (+ A[dst] cached) ; src,dst ids
    "gine_message_fwd": 3 * 4 * H + 4,
    "pna_message_fwd_reduce": 3 * 4 * H + 12,   # the same pass with the segmented reduction fused in (m written, never re-read)
    "gine_message_fwd_reduce": 3 * 4 * H + 8,
    "segment_reduce_pna": 1 * 4 * H,       # read m
    "segment_reduce_sum": 1 * 4 * H,
    "edge_update_fwd": 4 * 4 * H + 8,      # read e, gather P[src], write h, write e'
    "edge_update_bwd": 5 * 4 * H + 8,      # read g_e', h, e; write g_e; scatter g_P[src]
    "pna_message_bwd": 4 * 4 * H + 8,      # read m, e; write g_e; scatter g_B[src]
    "gine_message_bwd": 4 * 4 * H + 8,
    "permute_rows": 2 * 4 * H + 8,
}


# bf16 storage: the [E,H] rows are 2 H bytes, the gathered / scattered node rows stay fp32
KERNEL_BYTES_BF16 = {
    "pna_message_fwd": 2 * 2 * H + 4 * H + 8, "gine_message_fwd": 2 * 2 * H + 4 * H + 4,
    "segment_reduce_pna": 2 * H, "segment_reduce_sum": 2 * H,
    "edge_update_fwd": 3 * 2 * H + 4 * H + 8, "edge_update_bwd": 4 * 2 * H + 4 * H + 8,
    "pna_message_bwd": 3 * 2 * H + 4 * H + 8, "gine_message_bwd": 3 * 2 * H + 4 * H + 8,
}


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel, workload, world):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the committed ncu
    --set full capture of this workload (profiles/traffic.json, written by scripts/ncu_traffic.py);
    None when there is no capture for this kernel / workload / partitioning."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        if t.get("workload") != workload or world != 1:
            return None
        return t["kernels"][kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


def cpu_reference_steps(kind, steps, warmup, threads=None):
    """The CPU restatement of the reference path (oracle/models.py: reference model files'
    layer loop over restated PyG ops), fwd+bwd, on a bounded sample of the workload shape:
    CPU_SAMPLE_EDGES edges at the workload's edge/node ratio.  Returns (edges/s, cores, sample)."""
    import b200mp
    from oracle import models as OM
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    n_full, e_full, seed = WORKLOADS["hi_medium"]
    e = CPU_SAMPLE_EDGES // 4 if kind == "tabgnn" else CPU_SAMPLE_EDGES      # the attention rows cost ~4x the graph half on a CPU
    n = max(2, int(round(n_full * e / e_full)))
    ei = synth_graph(n, e, seed, "cpu")
    deg_hist = torch.bincount(torch.bincount(ei[1], minlength=n))
    torch.manual_seed(1234)
    if kind == "pna":
        mod = b200mp.PNAS(num_features=H, num_gnn_layers=LAYERS, n_hidden=H, edge_updates=True, edge_dim=H, deg=deg_hist)
    elif kind == "tabgnn":
        mod = b200mp.TABGNN(channels=H, num_layers=LAYERS, deg=deg_hist, node_dim=H, nhidden=H, edge_dim=K_COLS * H, nhead=8, dropout=0.0)
    else:
        mod = b200mp.GINe(num_features=H, num_gnn_layers=LAYERS, n_hidden=H, edge_updates=True, edge_dim=H)
    p = {k: v.clone() for k, v in mod.state_dict().items()}
    for k, v in p.items():
        v.requires_grad_(v.is_floating_point() and "running" not in k and "avg_deg" not in k and not k.endswith("eps"))
    g = torch.Generator().manual_seed(seed)
    tab = kind == "tabgnn"
    x = torch.randn(n, *((1, H) if tab else (H,)), generator=g).requires_grad_(True)
    ea = torch.randn(e, *((K_COLS, H) if tab else (H,)), generator=g).requires_grad_(True)
    gx, ge = torch.randn(n, H, generator=g), torch.randn(e, H, generator=g)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        if tab:
            ox, oe = OM.tabgnn(p, x, ei, ea, LAYERS, 8)
        else:
            ox, oe = OM.mp_stack(p, kind, x, ei, ea, LAYERS, True, False, True)
        torch.autograd.backward([ox, oe], [gx, ge])
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    sample = f"{e} edges / {n} nodes (hi_medium edge/node ratio, same generator), {warmup} warm-up + {steps} timed fwd+bwd steps"
    return e * len(times) / sum(times), cores, sample, sum(times) / len(times), n, e


def run_reference(args, rank, world):
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 3)), max(1, min(args.warmup, 1))
    eps, cores, sample, sec, sample_n, sample_e = cpu_reference_steps(args.model, steps, warmup)
    line = {
        "impl": "reference", "metric": "message-passing edges/sec (fwd+bwd)", "value": eps, "unit": "edges/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(workload_config(args, world), sampled=True, sample_edges=sample_e, sample_nodes=sample_n,
                       sample_note="the CPU arm times a bounded sample of the workload (same generator, same edge/node "
                                   "ratio); edges/s is a rate, so the two arms compare as rates, not as equal problem sizes"),
        "cpu_baseline": {"value": eps, "unit": "edges/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": eps, "unit": "edges/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "CPU oracle port of the reference path (reference model files' layer loop over restated PyG 2.5.3 ops); "
                "the reference itself cannot be installed: torch_geometric / pytorch-frame fork are absent (DESIGN.md)",
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    n, e, _ = WORKLOADS[args.workload]
    what = (f"TABGNN: {LAYERS} column-attention layers (FT-Transformer, nhead 8, {K_COLS} columns + CLS) over node and edge tokens, "
            f"node/edge embeddings, {LAYERS} PNA layers (H={H}, edge updates)") if args.model == "tabgnn" else \
        f"{args.model.upper()} backbone ({LAYERS} layers, H={H}, edge updates)"
    return {"workload": f"{what} fwd+bwd on synthetic "
                        f"{args.workload} shape: {n} accounts, {e} transactions"
                        + ("" if world == 1 else f", ONE graph partitioned by destination range over {world} GPUs"),
            "nodes": n, "edges": e, "layers": LAYERS, "hidden": H, "model": {"pna": "PNAS", "gine": "GINe", "tabgnn": "TABGNN"}[args.model],
            "parallelism": "single GPU" if world == 1 else f"dst-range partition x{world}: per layer NCCL all-gather of source-side node tables, all-reduce of BatchNorm sums; backward reduce-scatter + weight-gradient all-reduce",
            "l2": "inputs exceed L2 (every [E,H] tensor >> 126 MB)"}


def run_eth_lp(args, rank, world, local_rank):
    """BASELINE.json config 4: link prediction with neighbour sampling on a synthetic ETH-phishing-shaped graph
    (N = 2 973 489, E = 13 551 303, K = 5: 4 numerical + 1 timestamp), the loop of fused.py:98-125 (train_lp): per step B = 200
    seed edges -> lp_inputs on the device (2-hop in-neighbour sample, fan-out [100, 100], relabel, x64 attribute replication, 64
    negatives per positive) -> column encoders -> TABGNNFused(lp=True) -> LinkPredHead -> lp_loss -> backward.  Reports the
    sampler (batch construction) and the model (encoders + fwd + bwd) separately, in edges/s of the sampled batch graphs.
    N > 1: mini-batches are independent -> replicas (every rank runs its own batches; value = sum over ranks)."""
    import torch.distributed as dist

    import b200mp
    from b200mp.batch import BatchBuilder
    from b200mp.nn import LinearEncoder, StatType, StypeWiseFeatureEncoder, TensorFrame, TimestampEncoder, stype

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    n, e, seed = 2_973_489, 13_551_303, 20240720
    B, NEG, FAN = 200, 64, (100, 100)
    ei = synth_graph(n, e, seed, dev)
    g = torch.Generator(device=dev).manual_seed(seed + 1)
    num = torch.rand(e, 4, generator=g, device=dev)
    ts7 = torch.stack([torch.randint(2015, 2020, (e,), generator=g, device=dev), torch.randint(0, 12, (e,), generator=g, device=dev),
                       torch.randint(0, 28, (e,), generator=g, device=dev), torch.randint(0, 7, (e,), generator=g, device=dev),
                       torch.randint(0, 24, (e,), generator=g, device=dev), torch.randint(0, 60, (e,), generator=g, device=dev),
                       torch.randint(0, 60, (e,), generator=g, device=dev)], 1).unsqueeze(1)
    names = {stype.numerical: ["value", "gas", "gas_price", "nonce"], stype.timestamp: ["timestamp"]}
    stats = {nm: {StatType.MEAN: 0.5, StatType.STD: 0.2887} for nm in names[stype.numerical]}
    stats["timestamp"] = {StatType.YEAR_RANGE: [2015, 2019]}
    tf = TensorFrame({stype.numerical: num, stype.timestamp: ts7}, names)
    bb = BatchBuilder(ei, n, tf, FAN)
    deg_hist = torch.bincount(torch.bincount(ei[1], minlength=n)).cpu()
    torch.manual_seed(1234)
    enc = StypeWiseFeatureEncoder(H, stats, names, {stype.numerical: LinearEncoder(), stype.timestamp: TimestampEncoder()}).to(dev).train()
    model = b200mp.TABGNNFused(channels=H, num_layers=LAYERS, encoder=enc, deg=deg_hist, node_dim=1, nhidden=H, edge_dim=K_COLS * H,
                               nhead=8, dropout=0.0).to(dev).train()
    head = b200mp.LinkPredHead(n_classes=1, n_hidden=H, dropout=0.0).to(dev).train()
    params = [*enc.parameters(), *model.parameters(), *head.parameters()]
    opt = torch.optim.Adam(params, lr=5e-4)
    gs = torch.Generator(device=dev).manual_seed(seed + 7 + rank)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    stats_k = b200mp.functional.KernelStats(timing=False)

    def step(timers=None):
        ids = torch.randint(0, e, (B,), generator=gs, device=dev)
        seeds = torch.stack([ei[0, ids], ei[1, ids], ids], 1)
        t0, t1, t2 = ev(), ev(), ev()
        t0.record()
        node_feats, _, _, nei, nea, tei, tea = bb.lp_inputs(seeds, NEG, ego=False)
        t1.record()
        nea_t, _ = enc(nea)
        tea_t, _ = enc(tea)
        x_gnn, _, tgt = model(node_feats, nei, nea_t, tei, tea_t, True)
        pos, neg = head(x_gnn, tei[:, :B], tgt[:B], tei[:, B:], tgt[B:])
        loss = b200mp.SSLoss.lp_loss(pos, neg)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        t2.record()
        if timers is not None:
            timers.append((t0, t1, t2, int(nei.shape[1]) + int(tei.shape[1]), int(node_feats.shape[0])))
        return loss

    for _ in range(args.warmup):
        step()
    clocks = ClockSampler(local_rank)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks.start()
    b200mp.functional.set_kernel_stats(stats_k)
    timers = []
    a, b = ev(), ev()
    a.record()
    for _ in range(args.steps):
        loss = step(timers)
    lval = float(loss.item())                   # D2H of the step result
    b.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    b200mp.functional.set_kernel_stats(None)
    clk = clocks.stop()
    ms = a.elapsed_time(b)
    ms_s = sum(t0.elapsed_time(t1) for t0, t1, _, _, _ in timers)
    ms_m = sum(t1.elapsed_time(t2) for _, t1, t2, _, _ in timers)
    edges = sum(t[3] for t in timers)
    nodes = sum(t[4] for t in timers)
    tot = torch.tensor([ms, ms_s, ms_m, float(edges), float(nodes)], device=dev, dtype=torch.float64)
    if world > 1:
        mx = tot[:3].clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tot[3:].clone()
        dist.all_reduce(sm)
        tot = torch.cat([mx, sm])
    ms, ms_s, ms_m, edges, nodes = (float(v) for v in tot)
    if rank != 0:
        return
    line = {
        "metric": "message-passing edges/sec (fwd+bwd)", "value": edges / (ms_m * 1e-3), "unit": "edges/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"link prediction with neighbour sampling (fused.py train_lp loop) on a synthetic ETH-phishing-shaped graph: {n} "
                               f"accounts, {e} transactions, K = {K_COLS} columns; B = {B} seed edges / step, fan-out {list(FAN)}, {NEG} negatives per "
                               "positive; TABGNNFused(lp) + LinkPredHead, Adam", "nodes": n, "edges": e, "layers": LAYERS, "hidden": H,
                   "model": "TABGNNFused + LinkPredHead", "parallelism": "single GPU" if world == 1 else f"{world} replicas (independent mini-batches)",
                   "l2": "batch graphs are small (tens of thousands of edges): L2-resident by nature, as in the reference's loop"},
        "clocks": clk,
        "batch": {"edges_per_step": edges / args.steps / world, "nodes_per_step": nodes / args.steps / world},
        "sampler": {"value": edges / (ms_s * 1e-3), "unit": "edges/s", "ms_per_step": ms_s / args.steps,
                    "what": "BatchBuilder.lp_inputs: 2-hop CSC neighbour sample, seed drop, relabel, attribute row gathers, x64 replication, negative sampling"},
        "model": {"value": edges / (ms_m * 1e-3), "unit": "edges/s", "ms_per_step": ms_m / args.steps,
                  "what": "column encoders (neighbour + target rows), TABGNNFused fwd+bwd, LinkPredHead, lp_loss, Adam step"},
        "e2e": {"value": edges / (ms * 1e-3), "unit": "edges/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4,
                "includes": "sampler + model + the loss read on the host; the transaction table lives on the device, nothing is copied per step"},
        "gpu_launches": stats_k.launches, "loss": lval, "cpu_baseline": None,
        "note": "extra line (BASELINE.json config 4); the headline line is the default workload",
    }
    print(json.dumps(line), flush=True)


def run_b200(args, rank, world, local_rank):
    import torch.distributed as dist

    import b200mp
    from b200mp import functional as BF
    from b200mp import graph as BG

    from b200mp import parallel as PL

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    n, e, seed = WORKLOADS[args.workload]
    ei = synth_graph(n, e, seed, dev)            # every rank generates the same graph, keeps its shard
    part = None
    if world > 1:
        part = PL.partition_graph(ei, n, rank, world)
        deg = torch.bincount(ei[1], minlength=n)
        deg_hist = torch.bincount(deg)
        st = None
        n_loc, e_loc = part.n_local, part.st.num_edges
    else:
        st = b200mp.edge_structure(ei, n)
        _, deg_hist = b200mp.degree_histogram(st.fwd.ptr)
        n_loc, e_loc = n, e
    model = make_model(args.model, deg_hist.cpu(), dev)
    bf16 = args.dtype == "bf16"
    if bf16:
        model.set_edge_storage(torch.bfloat16)
    # inputs are a function of the GLOBAL row id (every rank draws the full tensor from the same seed and keeps its
    # shard), so the N = 1 / 2 / 4 / 8 runs compute the same function and their checksums are comparable
    g = torch.Generator(device=dev).manual_seed(seed)

    tab = args.model == "tabgnn"

    def draw(rows, shard, inner=(H,)):
        full = torch.randn(rows, *inner, generator=g, device=dev)
        if part is None:
            return full
        out = shard(full)
        del full
        return out

    # tabgnn: the inputs are the column TOKENS (stand-in for the encoder output): x [N, 1, C], edge_attr [E, K, C]
    x = draw(n, lambda t: part.shard_nodes(t), (1, H) if tab else (H,)).requires_grad_(True)
    ea = draw(e, lambda t: part.shard_edges(t), (K_COLS, H) if tab else (H,)).requires_grad_(True)
    gx = draw(n, lambda t: part.shard_nodes(t))
    ge = draw(e, lambda t: part.shard_edges(t))
    if bf16:
        ge = ge.to(torch.bfloat16)
        if not tab:                           # the embedded edge features arrive in storage precision (the cast belongs to edge_emb)
            ea = ea.detach().to(torch.bfloat16).requires_grad_(True)
    torch.cuda.empty_cache()

    def step(keep=False):
        if tab:
            ox, oe = model(x, ei, ea) if part is None else model.forward_partitioned(x, ea, part)
        elif part is None:
            ox, oe = model.message_passing(x, ei, ea)
        else:
            ox, oe = model.message_passing_partitioned(x, ea, part)
        torch.autograd.backward([ox, oe], [gx, ge])
        if part is not None:
            PL.all_reduce_gradients(model, part)
        res = None
        if keep:          # checksum of one step (outside the timed region): fp64 sums over ALL ranks' rows
            gw = torch.cat([p.grad.reshape(-1) for p in model.parameters() if p.grad is not None]).double()
            loc = torch.stack([ox.detach().double().sum(), ox.detach().double().abs().sum(), oe.detach().double().sum(),
                               oe.detach().double().abs().sum(), x.grad.double().abs().sum(), ea.grad.double().abs().sum()])
            if world > 1:
                dist.all_reduce(loc)
            res = {"sum_x": float(loc[0]), "abs_x": float(loc[1]), "sum_e": float(loc[2]), "abs_e": float(loc[3]),
                   "abs_grad_x": float(loc[4]), "abs_grad_e": float(loc[5]), "weight_grad_norm": float(gw.norm())}
        x.grad = None
        ea.grad = None
        model.zero_grad(set_to_none=True)
        return res

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    stats = BF.KernelStats(timing=True)
    clocks = ClockSampler(local_rank)
    barrier()
    clocks.start()
    BF.set_kernel_stats(stats)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    marks[0].record()
    for k in range(args.steps):
        step()
        marks[k + 1].record()
    barrier()
    BF.set_kernel_stats(None)
    clk = clocks.stop()
    ms = marks[0].elapsed_time(marks[-1])
    per_step = sorted(marks[k].elapsed_time(marks[k + 1]) for k in range(args.steps))
    ms_median = per_step[len(per_step) // 2] if len(per_step) % 2 else 0.5 * (per_step[len(per_step) // 2 - 1] + per_step[len(per_step) // 2])
    if world > 1:
        t = torch.tensor([ms, ms_median], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_median = float(t[0].item()), float(t[1].item())
    ms_step = ms / args.steps
    value = e / (ms_step * 1e-3)                 # N > 1: ONE graph partitioned over the ranks (strong scaling)
    checksum = step(keep=True)
    torch.cuda.synchronize()
    if os.environ.get("B200MP_TORCH_PROFILE") and rank == 0:     # diagnosis only (after the timed region): one more step under torch.profiler
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            step()
            torch.cuda.synchronize()
        with open(os.environ["B200MP_TORCH_PROFILE"], "w") as f:
            f.write(prof.key_averages().table(sort_by="cuda_time_total", row_limit=60, max_name_column_width=90))

    # whole model ("report whole-model numbers separately"): node_emb / edge_emb included, N = 1, f32 line only
    whole = None
    if part is None and not tab and not bf16 and not args.no_e2e:
        try:
            xw = x.detach().clone().requires_grad_(True)
            ew = ea.detach().clone().requires_grad_(True)

            def full_step():
                ox, oe = model(xw, ei, ew.view(e, 1, H))
                torch.autograd.backward([ox, oe], [gx, ge])
                xw.grad = None
                ew.grad = None
                model.zero_grad(set_to_none=True)

            full_step()
            w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            nw = max(1, min(args.steps, 3))
            w0.record()
            for _ in range(nw):
                full_step()
            w1.record()
            torch.cuda.synchronize()
            wms = w0.elapsed_time(w1) / nw
            whole = {"ms_per_step": wms, "value": e / (wms * 1e-3), "unit": "edges/s",
                     "includes": "node_emb and edge_emb (64 -> 64 linears over the N / E rows, fwd + bwd) around the timed layer loop"}
            del xw, ew
        except Exception as ex:
            whole = {"error": f"{type(ex).__name__}: {ex}"[:200]}

    # per-kernel accounting -> roofline of the dominant kernel
    summ = stats.summary()
    peak, peak_src = measured_peak_gbs()
    kern = {}
    KB = KERNEL_BYTES_BF16 if bf16 else KERNEL_BYTES
    for name, (calls, total_ms) in summ.items():
        per = total_ms / calls
        kern[name] = {"calls_per_step": calls / args.steps, "ms_per_call": round(per, 4),
                      "share_of_step": round(total_ms / ms, 4),
                      "achieved_gbs": round(KB[name] * e_loc / (per * 1e-3) / 1e9, 1) if name in KB else None}
    top = max(((k, v) for k, v in summ.items() if k in KB), key=lambda kv: kv[1][1])[0]
    achieved = kern[top]["achieved_gbs"]
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": ncu_traffic(top, args.workload, world),
                "traffic_source": "profiles/traffic.json: ncu --set full dram__bytes_read+write per launch, captured on the build named in that file (not this run)",
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": KB[top] * e_loc}
    step_bytes = LAYERS * (e_loc * algorithmic_bytes_per_edge(args.model) + n_loc * 12 * 4 * H)
    formula = "L*(E*(22*s*H+64) + N*12*s*H), SURVEY.md 8(d)"
    if bf16:      # SURVEY's bf16 figure: s = 2 for the 22 [E,H] touches (it also halves the node-table gathers, which stay fp32 here)
        step_bytes = LAYERS * (e_loc * (22 * 2 * H + 64) + n_loc * 12 * 4 * H)
        formula = "L*(E*(22*2*H+64) + N*12*4*H): SURVEY.md 8(d) with s = 2 for the edge terms"
    if tab:      # + L column-attention layers over the (K+1)-token edge rows and the 2-token node rows, + edge_emb / node_emb
        step_bytes += LAYERS * (e_loc * tab_bytes_per_row(K_COLS + 1) + n_loc * tab_bytes_per_row(2))
        step_bytes += 2 * 4 * (e_loc * ((K_COLS + 1) * H + H) + n_loc * (2 * H + H))
        formula += " + L*(E*5*(K+1)*C*s + N*5*2*C*s) [column attention, recomputing] + 2*s*(E*((K+1)*C+H) + N*(2*C+H)) [embeddings fwd+bwd]"
    step_roofline = {"algorithmic_bytes_per_step": step_bytes, "achieved_gbs": round(step_bytes / (ms_step * 1e-3) / 1e9, 1),
                     "frac": round(step_bytes / (ms_step * 1e-3) / 1e9 / peak, 4), "formula": formula}

    # end-to-end: pinned host buffers -> device -> CSC build -> fwd+bwd -> loss to host, every step
    e2e = None
    try:
        if args.no_e2e:
            raise RuntimeError("skipped by --no-e2e (profiling run)")
        del st
        BG.clear_cache()
        hx, hea, hei = x.detach().cpu().pin_memory(), ea.detach().cpu().pin_memory(), ei.cpu().pin_memory()
        del ea, ge, part, gx
        torch.cuda.empty_cache()
        e2e_steps = max(1, min(args.steps, 3))

        from b200mp import DeviceFeeder
        feeder = DeviceFeeder(dev)
        host = (hx, hea, hei) if world == 1 else (hx, hea)     # N > 1: the rank's shard; the partition is built once

        def e2e_step():
            # this step's inputs were staged during the previous step; start the NEXT step's copy now so that it
            # overlaps this step's kernels (one H2D of all inputs per step, on the copy stream)
            bufs = feeder.take()
            dx, dea, dei = bufs[0], bufs[1], (bufs[2] if world == 1 else None)
            feeder.stage(host)
            BG.clear_cache()
            dx, dea = dx.detach().requires_grad_(True), dea.detach().requires_grad_(True)
            if tab:
                ox, oe = model(dx, dei, dea) if world == 1 else model.forward_partitioned(dx, dea, e2e_part[0])
                p2 = e2e_part[0]
            elif world == 1:
                ox, oe = model.message_passing(dx, dei, dea)
            else:
                # the partition (which rank owns which destination range, local CSC slice) is a property of the graph:
                # built once, outside the timed steps; the per-step work is the shard's H2D + fwd + bwd
                p2 = e2e_part[0]
                ox, oe = model.message_passing_partitioned(dx, dea, p2)
            loss = ox.square().mean() + oe.square().mean()
            loss.backward()
            if world > 1:
                PL.all_reduce_gradients(model, p2)
            model.zero_grad(set_to_none=True)
            feeder.release()
            return float(loss.item())          # D2H read of the step's result

        e2e_part = [PL.partition_graph(ei, n, rank, world) if world > 1 else None]
        feeder.stage(host)
        e2e_step()
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        staged0 = feeder.bytes_staged
        a.record()
        for _ in range(e2e_steps):
            e2e_step()
        feeder.drain()                          # the last staged copy ends inside the timed region: K copies for K steps
        b.record()
        barrier()
        h2d = (feeder.bytes_staged - staged0) // e2e_steps
        ems = a.elapsed_time(b)
        if world > 1:
            t = torch.tensor([ems], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ems = float(t.item())
        e2e = {"value": e / (ems / e2e_steps * 1e-3), "unit": "edges/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": 4, "steps": e2e_steps, "ms_per_step": ems / e2e_steps,
               "includes": "H2D of x, edge_index, edge features from pinned host memory (one copy of all inputs per step, "
                           "double-buffered on a copy stream: step i+1's copy overlaps step i's kernels); CSC build; fwd+bwd; loss D2H"
                           if world == 1 else "H2D of this rank's shard of x and the edge features from pinned host memory "
                           "(double-buffered); fwd+bwd over the partition (built once, outside the timed steps); loss D2H"}
    except Exception as ex:          # reported, never silently dropped
        e2e = {"value": None, "unit": "edges/s", "error": f"{type(ex).__name__}: {ex}"[:300]}

    if rank != 0:
        return
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        eps, cores, sample, *_ = cpu_reference_steps(args.model, 2, 1)
        cpu = {"value": eps, "unit": "edges/s", "cores": cores, "kind": "port", "sample": sample}
        if cores > 4:          # BASELINE.md / benchmark.py:50: the reference's own benchmark pins 4 threads
            eps4, *_ = cpu_reference_steps(args.model, 1, 1, threads=4)
            cpu["value_4_threads"] = eps4
    line = {
        "metric": "message-passing edges/sec (fwd+bwd)", "value": value, "unit": "edges/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "ms_per_step_median": ms_median,
        "value_median": e / (ms_median * 1e-3), "higher_is_better": True,
        "scaling": "weak" if world == 1 else "strong", "checksum": checksum,
        "vs_baseline": None, "dtype": "bf16 storage of the [E,H] tensors, fp32 arithmetic (extra line)" if bf16 else "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "clocks": clk, "e2e": e2e, "gpu_launches": stats.launches, "roofline": roofline, "step_roofline": step_roofline,
        "kernels": kern, "cpu_baseline": cpu, "whole_model": whole,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS) + ["eth_lp"],
                    help="default: hi_medium (BASELINE.json's metric config); tabgnn on ONE GPU: hi_small -- hi_medium's [E, 6, 64] "
                         "token tensors (49 GB each) only fit partitioned over 8 GPUs, which is how config 5 is worded")
    ap.add_argument("--model", default="pna", choices=["pna", "gine", "tabgnn"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--dtype", default="f32", choices=["f32", "bf16"],
                    help="bf16: EXTRA line (BASELINE.json config 3) -- every [E,H] edge tensor stored as bf16, fp32 arithmetic; the f32 line is the headline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--hidden", type=int, default=64, choices=(64, 128),
                    help="hidden width H (extra lines; 64 is the headline configuration, 128 the reference's PNAS default: pna.py:50)")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the host-buffer arm")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.workload is None:
        args.workload = "hi_small" if (args.model == "tabgnn" and world < 8) else "hi_medium"
    if args.impl == "reference":
        if args.workload == "eth_lp":
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": "the eth_lp extra line has no CPU arm: the oracle's sampler restatement covers "
                                  "the take-all regime only (pyg-lib's sub-sampling RNG is not reproducible, SURVEY.md App. A.3)"}), flush=True)
            return
        run_reference(args, rank, world)
        return
    if args.warmup < 3:
        args.warmup = 3
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL's own log (whatever level the caller asked for with NCCL_DEBUG) goes to stderr: stdout carries the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        # torch itself prints "NCCL version ..." on STDOUT when NCCL_DEBUG is set (communicator creation): keep stdout for the JSON
        # line by pointing fd 1 at stderr while the communicator comes up (eager with device_id; the barrier covers a lazy one)
        sys.stdout.flush()
        keep_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(keep_fd, 1)
            os.close(keep_fd)
    try:
        if args.workload == "eth_lp":
            run_eth_lp(args, rank, world, local_rank)
        else:
            run_b200(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
