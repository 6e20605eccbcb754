"""Shared harness: run a b200mp drop-in on a golden fixture (tests/golden/*.npz, produced by the
reference's own model files) and compare outputs, input/parameter gradients and BatchNorm
running statistics.

Tolerance (BASELINE.json north_star): 1e-5 relative for fp32.  PNA's std gradient is
ill-conditioned near its 1e-5 variance clamp (d std / d m = (m - mean) / (deg std): a node whose
two messages nearly coincide in a channel amplifies one-ulp differences of m by ~1e2-1e3), so where
fp32 itself cannot hold 1e-5 the tolerance is 4x the fp32 CONDITIONING of the quantity, measured
two ways against the fp64 oracle: the fixture's own distance from fp64,
and the distance of the fp32 oracle re-run with every input and parameter perturbed by <= 1 ulp
(any re-association of the arithmetic, such as the hoisted node terms of PNAConv, is such a
perturbation).  The tolerance used is recorded per quantity and never exceeds 1e-3.
"""
from __future__ import annotations

import torch

import b200mp
from oracle import models as OM
from tests.util import load_golden, rel_err

FP32_TOL = 1e-5


def build_module(meta, rec):
    kind = meta["model"]
    H, K, Lyr = meta["H"], meta["K"], meta["L"]
    if kind == "GINe":
        m = b200mp.GINe(num_features=1, num_gnn_layers=Lyr, n_hidden=H, edge_updates=meta["edge_updates"],
                        edge_dim=K * H, reverse_mp=meta["reverse_mp"])
    elif kind == "PNAS":
        ei = rec["in"]["edge_index"]
        n = rec["in"]["x"].shape[0]
        deg = torch.bincount(torch.bincount(ei[1], minlength=n))
        m = b200mp.PNAS(num_features=H, num_gnn_layers=Lyr, n_hidden=H, edge_updates=meta["edge_updates"],
                        edge_dim=K * H, deg=deg, reverse_mp=meta["reverse_mp"])
    elif kind in ("CPNA", "CPNATAB"):
        ei = rec["in"]["edge_index"]
        n = rec["in"]["x"].shape[0]
        deg = torch.bincount(torch.bincount(ei[1], minlength=n))
        cls = b200mp.CPNA if kind == "CPNA" else b200mp.CPNATAB
        m = cls(num_features=H, num_gnn_layers=Lyr, n_hidden=H, edge_updates=meta["edge_updates"], edge_dim=K * H, deg=deg,
                reverse_mp=meta["reverse_mp"])
        for mod in m.modules():                      # as the fixture generator: the row attention's default dropout off
            if isinstance(mod, torch.nn.Dropout):
                mod.p = 0.0
            if isinstance(mod, torch.nn.MultiheadAttention):
                mod.dropout = 0.0
    elif kind in ("TABGNN", "TABGNNInterleaved", "TABGNNFused"):
        ei = rec["in"]["edge_index"]
        n = rec["in"]["x"].shape[0]
        deg = torch.bincount(torch.bincount(ei[1], minlength=n))
        C = meta["C"]
        kw = dict(channels=C, num_layers=Lyr, deg=deg, nhidden=H, edge_dim=K * C, nhead=meta["nhead"], dropout=0.0)
        if kind == "TABGNN":
            m = b200mp.TABGNN(node_dim=rec["in"]["x"].shape[1] * C, **kw)
        elif kind == "TABGNNInterleaved":
            m = b200mp.TABGNNInterleaved(node_dim=meta["node_dim"], **kw)
        else:
            m = b200mp.TABGNNFused(node_dim=meta["node_dim"], **kw)
    else:
        raise KeyError(kind)
    m.load_state_dict(rec["sd"], strict=True)
    return m


def oracle_fn(meta):
    if meta["model"] == "GINe":
        return lambda p, i: OM.gine(p, i["x"], i["edge_index"], i["edge_attr"], meta["L"], meta["edge_updates"], meta["reverse_mp"])
    if meta["model"] == "PNAS":
        return lambda p, i: OM.pnas(p, i["x"], i["edge_index"], i["edge_attr"], meta["L"], meta["edge_updates"], meta["reverse_mp"])
    if meta["model"] == "CPNA":
        return lambda p, i: OM.cpna(p, i["x"], i["edge_index"], i["edge_attr"], meta["L"], meta["edge_updates"], meta["reverse_mp"])
    if meta["model"] == "CPNATAB":
        return lambda p, i: OM.cpnatab(p, i["x"], i["edge_index"], i["edge_attr"], meta["L"], meta["edge_updates"], meta["reverse_mp"],
                                       nhead=meta["nhead"])
    if meta["model"] == "TABGNN":
        return lambda p, i: OM.tabgnn(p, i["x"], i["edge_index"], i["edge_attr"], meta["L"], meta["nhead"])
    if meta["model"] == "TABGNNInterleaved":
        return lambda p, i: OM.tabgnn_interleaved(p, i["x"], i["edge_index"], i["edge_attr"], meta["L"], meta["node_dim"], meta["nhead"])
    if meta["model"] == "TABGNNFused":
        return lambda p, i: OM.tabgnn_fused(p, i["x"], i["edge_index"], i["edge_attr"], i["target_edge_index"],
                                            i["target_edge_attr"], meta["L"], meta["node_dim"], lp=meta["lp"], nhead=meta["nhead"])
    raise KeyError(meta["model"])


def _jitter(t, gen):
    """Multiply by (1 + u * 2^-23), u ~ U(-1, 1): a one-ulp relative perturbation."""
    return t * (1 + (torch.rand(t.shape, generator=gen, dtype=torch.float64) * 2 - 1).to(t.dtype) * 2.0 ** -23)


def oracle_eval(rec, dtype, jitter_seed=None):
    """(outs, input grads, param grads) of the functional oracle in ``dtype`` on CPU.  With
    ``jitter_seed`` every float input / parameter is perturbed by <= 1 ulp (relative) first."""
    meta = rec["meta"]
    gen = torch.Generator().manual_seed(jitter_seed) if jitter_seed is not None else None
    jit = (lambda k, v: _jitter(v, gen) if (gen is not None and "running" not in k and "avg_deg" not in k) else v)
    p = {k: (jit(k, v.detach().clone()).to(dtype) if v.is_floating_point() else v.clone()) for k, v in rec["sd"].items()}
    for k, v in p.items():
        v.requires_grad_(v.is_floating_point() and "running" not in k and "avg_deg" not in k and not k.endswith("eps"))
    for k in list(p):
        if ".conv_back.nn." in k:
            p[k] = p[k.replace(".conv_back.nn.", ".conv_forw.nn.")]
    ins = {k: (jit(k, v.detach().clone()).to(dtype) if v.is_floating_point() else v.clone()) for k, v in rec["in"].items()}
    for k in rec["gin"]:
        ins[k].requires_grad_(True)
    outs = oracle_fn(meta)(p, ins)
    loss = sum((o * rec["w"][str(i)].to(dtype)).sum() for i, o in enumerate(outs))
    loss.backward()
    return outs, {k: ins[k].grad for k in rec["gin"]}, {k: v.grad for k, v in p.items() if v.grad is not None}


def run_dropin(rec, device):
    m = build_module(rec["meta"], rec).to(device)
    m.train()
    ins = {k: v.clone().to(device) for k, v in rec["in"].items()}
    for k in rec["gin"]:
        ins[k].requires_grad_(True)
    if rec["meta"]["model"] == "TABGNNFused":
        outs = m(ins["x"], ins["edge_index"], ins["edge_attr"], ins["target_edge_index"], ins["target_edge_attr"],
                 lp=rec["meta"]["lp"])
    else:
        outs = m(ins["x"], ins["edge_index"], ins["edge_attr"])
    loss = sum((o * rec["w"][str(i)].to(device)).sum() for i, o in enumerate(outs))
    loss.backward()
    gin = {k: ins[k].grad for k in rec["gin"]}
    gp = {k: p.grad for k, p in m.named_parameters(remove_duplicate=False) if p.grad is not None}
    return m, outs, gin, gp


def check_fixture(name, device):
    rec = load_golden(name)
    m, outs, gin, gp = run_dropin(rec, device)
    outs64, gin64, gp64 = oracle_eval(rec, torch.float64)
    jit = [oracle_eval(rec, torch.float32, jitter_seed=s) for s in (101, 102, 103)]
    worst = {}

    def yard(sel, truth, floor):
        return max(rel_err(sel(j), truth, floor) for j in jit)

    def check(got, want, truth, what, floor=1e-30, sel=None):
        cond = rel_err(want, truth, floor)
        if sel is not None:
            cond = max(cond, yard(sel, truth, floor))
        tol = max(FP32_TOL, 4 * cond)
        if what.endswith("bias") and what.startswith("grad."):
            tol = max(tol, 2 * FP32_TOL)      # biases ahead of a BatchNorm: an exactly-zero true gradient, judged against a floor
        err = rel_err(got, want, floor)
        worst[what] = (err, tol)
        assert tol < 1e-3, f"{name} {what}: fixture itself is {tol / 4:.2e} from fp64"
        assert err < tol, f"{name} {what}: rel err {err:.3e} (tol {tol:.1e})"

    for i, o in enumerate(outs):
        check(o, rec["out"][str(i)], outs64[i], f"out{i}", sel=lambda j, i=i: j[0][i])
    for k, g in rec["gin"].items():
        check(gin[k], g, gin64[k], f"grad_in.{k}", sel=lambda j, k=k: j[1][k])
    gmax = max(float(g.abs().max()) for g in rec["g"].values())
    for k, g in rec["g"].items():
        assert k in gp, f"{name}: no gradient for {k}"
        floor = (0.2 if k.endswith("bias") else 1e-3) * gmax
        check(gp[k], g, gp64.get(k, g.double()), f"grad.{k}", floor,
              sel=(lambda j, k=k: j[2][k]) if k in gp64 else None)
    sd = m.state_dict()
    for k, v in rec["sd_after"].items():
        assert rel_err(sd[k].to(v.dtype), v) < 2e-5, f"{name} running stat {k}"
    return worst
