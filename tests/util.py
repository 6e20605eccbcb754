"""Shared helpers for the parity tests."""
import json
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    rec = {"sd": {}, "in": {}, "gin": {}, "g": {}, "w": {}, "out": {}, "sd_after": {}}
    for k in z.files:
        if "/" in k:
            grp, key = k.split("/", 1)
            rec[grp][key] = torch.from_numpy(z[k])
        else:
            rec[k] = z[k]
    with open(os.path.join(GOLDEN, "meta.json")) as f:
        rec["meta"] = json.load(f).get(name, {})
    return rec


def rel_err(a: torch.Tensor, b: torch.Tensor, floor: float = 1e-30) -> float:
    """max|a-b| / max(|b|): the 'relative' of the 1e-5 (fp32) / 2e-2 (bf16) tolerances in
    BASELINE.json's north_star, taken against the tensor's scale."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    if b.numel() == 0:
        return 0.0
    denom = max(float(b.abs().max()), floor)
    return float((a - b).abs().max()) / denom


def make_graph(n, e, seed, hub_frac=0.15, device="cpu"):
    """Multigraph with one hub, destinations that never appear (empty segments), self loops and
    duplicate edges, in random (unsorted) edge order."""
    g = torch.Generator().manual_seed(seed)
    src = torch.randint(0, n, (e,), generator=g)
    dst = torch.randint(0, max(n - n // 5, 1), (e,), generator=g)
    if e >= 8:
        dst[: int(e * hub_frac)] = min(3, n - 1)
        src[e // 2: e // 2 + e // 10] = dst[e // 2: e // 2 + e // 10]
        k = min(5, e // 4)
        src[-k:], dst[-k:] = src[:k].clone(), dst[:k].clone()
    perm = torch.randperm(e, generator=g)
    return torch.stack([src[perm], dst[perm]]).to(device)


def deg_hist(edge_index, n):
    d = torch.bincount(edge_index[1].cpu(), minlength=n)
    return torch.bincount(d)
