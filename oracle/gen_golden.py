"""ORACLE (test infrastructure): generate tests/golden/*.npz by running the reference's OWN,
unmodified model files (loaded in place from /root/reference through oracle/ref_loader.py)
on seeded inputs.  Run in the build container only:

    python -m oracle.gen_golden

Each fixture holds: inputs, the initial state_dict (keys prefixed ``sd/``), forward outputs,
the loss weights, and gradients w.r.t. inputs (``gin/``) and parameters (``g/``) of
L = sum(out_x * w_x) + sum(out_e * w_e) [+ sum(out_t * w_t)], plus BatchNorm running stats
after the step (``sd_after/``).  The reference tests hold no golden vectors, so
these are the pins for oracle/models.py; PyG itself is restated (parity unpinned there).
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch

from . import ref_loader

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def make_graph(n, e, seed, hub=True):
    """Small multigraph with a hub, empty destinations, self-loops and duplicate edges."""
    g = torch.Generator().manual_seed(seed)
    src = torch.randint(0, n, (e,), generator=g)
    dst = torch.randint(0, n - n // 5, (e,), generator=g)          # last 20 % never a destination
    if hub:
        dst[: e // 6] = 3                                          # hub node
    src[e // 2: e // 2 + e // 10] = dst[e // 2: e // 2 + e // 10]  # self loops
    src[-5:], dst[-5:] = src[:5], dst[:5]                          # duplicate edges
    perm = torch.randperm(e, generator=g)
    return torch.stack([src[perm], dst[perm]])


def deg_hist(edge_index, n):
    d = torch.bincount(edge_index[1], minlength=n)
    return torch.bincount(d)


def _np(t):
    return t.detach().cpu().numpy()


def run_case(name, model, inputs, float_inputs, call, seed):
    """inputs: dict name->tensor; float_inputs: names needing grad; call(model, inputs)->tuple."""
    g = torch.Generator().manual_seed(seed + 1000)
    model.train()
    rec = {}
    for k, v in model.state_dict().items():
        rec["sd/" + k] = _np(v.clone())
    for k in float_inputs:
        inputs[k] = inputs[k].clone().requires_grad_(True)
    outs = call(model, inputs)
    loss = 0.0
    for i, o in enumerate(outs):
        w = torch.randn(o.shape, generator=g)
        rec[f"w/{i}"] = _np(w)
        rec[f"out/{i}"] = _np(o)
        loss = loss + (o * w).sum()
    loss.backward()
    rec["loss"] = np.asarray(float(loss))
    for k, v in inputs.items():
        rec["in/" + k] = _np(v)
        if k in float_inputs:
            rec["gin/" + k] = _np(v.grad)
    seen = set()
    for k, p in model.named_parameters(remove_duplicate=False):
        if p.grad is not None:
            rec["g/" + k] = _np(p.grad)
        seen.add(k)
    for k, v in model.state_dict().items():
        if "running_" in k or "num_batches" in k:
            rec["sd_after/" + k] = _np(v)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(f"{name}: {len(rec)} arrays, loss={float(loss):.6f}")


def main():
    R = ref_loader.load()
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(1)
    torch.use_deterministic_algorithms(True)
    meta = {}

    N, E, H, K = 83, 420, 32, 2
    ei = make_graph(N, E, 7)
    deg = deg_hist(ei, N)

    def rnd(shape, seed):
        return torch.randn(shape, generator=torch.Generator().manual_seed(seed))

    for rev in (False, True):
        tag = "_rev" if rev else ""
        torch.manual_seed(11)
        m = R.GINe(num_features=1, num_gnn_layers=2, n_hidden=H, edge_updates=True, edge_dim=K * H, reverse_mp=rev)
        run_case("gine_eu" + tag, m, {"x": rnd((N, 1), 1), "edge_index": ei, "edge_attr": rnd((E, K, H), 2)},
                 ["x", "edge_attr"], lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"]), 3)
        meta["gine_eu" + tag] = dict(model="GINe", L=2, H=H, K=K, edge_updates=True, reverse_mp=rev)
        torch.manual_seed(12)
        m = R.PNAS(num_features=H, num_gnn_layers=2, n_hidden=H, edge_updates=True, edge_dim=K * H, deg=deg, reverse_mp=rev)
        run_case("pnas_eu" + tag, m, {"x": rnd((N, 1, H), 4), "edge_index": ei, "edge_attr": rnd((E, K, H), 5)},
                 ["x", "edge_attr"], lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"]), 6)
        meta["pnas_eu" + tag] = dict(model="PNAS", L=2, H=H, K=K, edge_updates=True, reverse_mp=rev)

    torch.manual_seed(13)
    m = R.GINe(num_features=1, num_gnn_layers=3, n_hidden=H, edge_updates=False, edge_dim=K * H)
    run_case("gine_noeu", m, {"x": rnd((N, 1), 1), "edge_index": ei, "edge_attr": rnd((E, K, H), 2)},
             ["x", "edge_attr"], lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"]), 3)
    meta["gine_noeu"] = dict(model="GINe", L=3, H=H, K=K, edge_updates=False, reverse_mp=False)

    # one H=64 PNAS (the width of configs 1-5), forward+backward
    N2, E2, H2 = 61, 300, 64
    ei2 = make_graph(N2, E2, 9)
    torch.manual_seed(14)
    m = R.PNAS(num_features=H2, num_gnn_layers=1, n_hidden=H2, edge_updates=True, edge_dim=H2, deg=deg_hist(ei2, N2))
    run_case("pnas_h64", m, {"x": rnd((N2, 1, H2), 4), "edge_index": ei2, "edge_attr": rnd((E2, 1, H2), 5)},
             ["x", "edge_attr"], lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"]), 6)
    meta["pnas_h64"] = dict(model="PNAS", L=1, H=H2, K=1, edge_updates=True, reverse_mp=False)

    # tabular + graph models (dropout 0 so that the run is deterministic)
    C = 32
    torch.manual_seed(15)
    m = R.TABGNN(channels=C, num_layers=2, deg=deg, node_dim=1 * C, nhidden=H, edge_dim=K * C, dropout=0.0, nhead=8)
    run_case("tabgnn", m, {"x": rnd((N, 1, C), 4), "edge_index": ei, "edge_attr": rnd((E, K, C), 5)},
             ["x", "edge_attr"], lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"]), 6)
    meta["tabgnn"] = dict(model="TABGNN", L=2, H=H, C=C, K=K, nhead=8)

    torch.manual_seed(16)
    m = R.TABGNNInterleaved(channels=C, num_layers=2, deg=deg, node_dim=1, nhidden=C, edge_dim=K * C, dropout=0.0, nhead=8)
    run_case("interleaved", m, {"x": torch.ones(N, 1), "edge_index": ei, "edge_attr": rnd((E, K, C), 5)},
             ["edge_attr"], lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"]), 6)
    meta["interleaved"] = dict(model="TABGNNInterleaved", L=2, H=C, C=C, K=K, nhead=8, node_dim=1)

    B = 20
    C, Hf = 16, 16      # small widths keep the fuse-MLP (4*(C+2H))^2 weights out of the fixture size
    for lp in (False, True):
        torch.manual_seed(17)
        m = R.TABGNNFused(channels=C, num_layers=2, deg=deg, node_dim=1, nhidden=Hf, edge_dim=K * C, dropout=0.0, nhead=8)
        name = "fused_lp" if lp else "fused_mcm"
        ins = {"x": torch.ones(N, 1), "edge_index": ei[:, B:] if lp else ei,
               "edge_attr": rnd((E - B if lp else E, K, C), 5),
               "target_edge_index": ei[:, :B], "target_edge_attr": rnd((B, K, C), 8)}
        run_case(name, m, ins, ["edge_attr", "target_edge_attr"],
                 lambda mod, i, lp=lp: mod(i["x"], i["edge_index"], i["edge_attr"], i["target_edge_index"],
                                           i["target_edge_attr"], lp=lp), 6)
        meta[name] = dict(model="TABGNNFused", L=2, H=Hf, C=C, K=K, nhead=8, node_dim=1, lp=lp)

    # read-out heads (eval mode: dropout off)
    torch.manual_seed(18)
    head = R.ClassifierHead(n_classes=2, n_hidden=H, dropout=0.0)
    run_case("classifier_head", head, {"x": rnd((N, H), 1), "edge_index": ei[:, :B], "edge_attr": rnd((B, H), 2)},
             ["x", "edge_attr"], lambda mod, i: (mod(i["x"], i["edge_index"], i["edge_attr"]),), 3)
    torch.manual_seed(19)
    head = R.LinkPredHead(n_classes=1, n_hidden=H, dropout=0.0)
    run_case("linkpred_head", head, {"x": rnd((N, H), 1), "pos_ei": ei[:, :B], "pos_ea": rnd((B, H), 2),
                                     "neg_ei": ei[:, B:3 * B], "neg_ea": rnd((2 * B, H), 4)},
             ["x", "pos_ea", "neg_ea"],
             lambda mod, i: mod(i["x"], i["pos_ei"], i["pos_ea"], i["neg_ei"], i["neg_ea"]), 3)

    # the reference's only recorded real mini-batch (shape fixture for the C++ sampler,
    # src/primitives/negative_sampling/{edge_index,pos_edge_index}.json): structure only.
    base = os.path.join(ref_loader.REF_ROOT, "src", "primitives", "negative_sampling")
    with open(os.path.join(base, "edge_index.json")) as f:
        bei = np.asarray(json.load(f), dtype=np.int32)
    with open(os.path.join(base, "pos_edge_index.json")) as f:
        pei = np.asarray(json.load(f), dtype=np.int32)
    np.savez_compressed(os.path.join(OUT, "batch_fixture.npz"), edge_index=bei, pos_edge_index=pei)
    print("batch_fixture:", bei.shape, pei.shape)

    with open(os.path.join(OUT, "meta.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
