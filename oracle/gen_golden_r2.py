"""ORACLE (test infrastructure): round-2 fixtures, generated in the build container by running the reference's OWN files:

    python -m oracle.gen_golden_r2

  cpna.npz, cpnatab.npz   CPNA / CPNATAB (src/nn/gnn/pna.py:164-304) through oracle/ref_loader (PyG operators restated by the
                          shim, as for every model fixture; the row attention is torch's own TransformerEncoderLayer)
  ss_loss_metric.npz      SSLoss.lp_loss / mcm_loss (src/utils/loss.py) and SSMetric.mrr (src/utils/metric.py): those two files
                          are plain torch / numpy, so these vectors PIN the restatements in oracle/models.py to the reference
Existing fixtures and their meta entries are left untouched.
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch

from . import ref_loader
from .gen_golden import OUT, deg_hist, make_graph, run_case


def main():
    R = ref_loader.load()
    torch.set_num_threads(1)
    with open(os.path.join(OUT, "meta.json")) as f:
        meta = json.load(f)

    def rnd(shape, seed):
        return torch.randn(shape, generator=torch.Generator().manual_seed(seed))

    N, E, H, K, Lyr = 83, 420, 32, 2, 2
    ei = make_graph(N, E, 7)
    deg = deg_hist(ei, N)
    for name, cls, seed in (("cpna", R.CPNA, 31), ("cpnatab", R.CPNATAB, 35)):
        torch.manual_seed(seed)
        m = cls(num_features=H, num_gnn_layers=Lyr, n_hidden=H, edge_updates=True, edge_dim=K * H, deg=deg)
        for mod in m.modules():                       # deterministic fixture: the row attention's default dropout 0.1 off
            if isinstance(mod, torch.nn.Dropout):
                mod.p = 0.0
            if isinstance(mod, torch.nn.MultiheadAttention):
                mod.dropout = 0.0
        # the reference writes the updated columns into edge_attr IN PLACE (pna.py:230): hand it a non-leaf copy
        run_case(name, m, {"x": rnd((N, 1, H), 4), "edge_index": ei, "edge_attr": rnd((E, K, H), 5)}, ["x", "edge_attr"],
                 lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"] * 1.0), 6)
        meta[name] = dict(model=cls.__name__, L=Lyr, H=H, K=K, edge_updates=True, reverse_mp=False, nhead=8)

    # losses / metric from the reference's own loss.py / metric.py
    SSLoss, SSMetric = ref_loader.load_loss_metric()
    g = torch.Generator().manual_seed(77)
    B, NEG, NUM, CATS = 48, 64, 3, (5, 7, 4)
    pos = torch.sigmoid(torch.randn(B, 1, generator=g))
    neg = torch.sigmoid(torch.randn(B * NEG, 1, generator=g))
    neg[:NEG // 2] = pos[0]                                         # ties with the first positive
    cat_out = [torch.randn(B, c, generator=g) for c in CATS]
    num_out = torch.randn(B, NUM, generator=g)
    idx = torch.randint(0, NUM + len(CATS), (B,), generator=g)
    val = torch.where(idx >= NUM, torch.stack([torch.randint(0, CATS[max(int(i) - NUM, 0)], (1,), generator=g)[0] for i in idx]).float(),
                      torch.randn(B, generator=g))
    y = torch.stack([val, idx.float()], dim=1)
    L = SSLoss("cpu", num_numerical=NUM)
    total, (cl, tc, acc), (nl, tn) = L.mcm_loss(cat_out, num_out, y)
    mrr, hits = SSMetric.mrr(pos, neg, [1, 2, 5, 10], NEG)
    rec = {"pos": pos.numpy(), "neg": neg.numpy(), "num_out": num_out.numpy(), "y": y.numpy(),
           "lp_loss": np.asarray(float(SSLoss.lp_loss(pos, neg))), "mcm_total": np.asarray(float(total)),
           "mcm_cat_loss": np.asarray(float(cl)), "mcm_t_c": np.asarray(int(tc)), "mcm_acc": np.asarray(float(acc)),
           "mcm_num_loss": np.asarray(float(nl)), "mcm_t_n": np.asarray(int(tn)), "mrr": np.asarray(float(mrr)),
           "hits": np.asarray([hits[f"hits@{k}"] for k in (1, 2, 5, 10)]), "num_numerical": np.asarray(NUM), "num_neg": np.asarray(NEG)}
    for c, t in enumerate(cat_out):
        rec[f"cat_out_{c}"] = t.numpy()
    np.savez_compressed(os.path.join(OUT, "ss_loss_metric.npz"), **rec)
    print("ss_loss_metric: lp", rec["lp_loss"], "mcm", rec["mcm_total"], "mrr", rec["mrr"], "hits", rec["hits"])

    with open(os.path.join(OUT, "meta.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
