#!/usr/bin/env python
"""Micro-benchmarks of the tabular-side kernels through the public modules:
stype-wise column encoders and the column-attention layer over [rows, S = 6, C = 64] tokens, with torch's own
nn.TransformerEncoderLayer (library kernels, what the reference runs) timed beside it.

    python scripts/kbench_tab.py [--rows 4000000] [--reps 5]
"""
import argparse
import json
import os
import sys

import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200mp  # noqa: E402,F401
from b200mp import functional as BF  # noqa: E402
from tests.enc_util import make_encoder, make_frame  # noqa: E402


def timeit(fn, reps):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=4_000_000)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    R, S, C = args.rows, 6, 64
    dev = torch.device("cuda", 0)
    tf, stats, names = make_frame(R, seed=1)
    enc = make_encoder(C, stats, names).to(dev)
    tfd = tf.to(dev)
    cls = torch.randn(C, device=dev, requires_grad=True)
    tok_bytes = R * S * C * 4
    res = {}

    def rec(name, ms, nbytes):
        res[name] = {"ms": round(ms, 3), "GB/s": round(nbytes / ms / 1e6, 1)}
        print(f"{name:40s} {ms:8.3f} ms  {nbytes / ms / 1e6:8.1f} GB/s", flush=True)

    x, _ = enc(tfd, cls=cls)
    go = torch.randn_like(x)
    in_bytes = R * (2 * 4 + 2 * 8 + 7 * 8)
    rec("col_encode fwd (5 cols + CLS)", timeit(lambda: enc(tfd, cls=cls), args.reps), tok_bytes + in_bytes)
    rec("col_encode bwd (param grads)", timeit(lambda: x.backward(go, retain_graph=True), args.reps), tok_bytes + in_bytes)
    del x

    t = torch.randn(R, S, C, device=dev)
    q, k, v = (torch.randn(R, S, C, device=dev, requires_grad=True) for _ in range(3))
    rec("col_attention fwd (p=0)", timeit(lambda: BF._col_attention_fwd(q, k, v, 8, 0.0, 0), args.reps), 4 * tok_bytes)
    rec("col_attention fwd (p=0.5)", timeit(lambda: BF._col_attention_fwd(q, k, v, 8, 0.5, 7), args.reps), 4 * tok_bytes)
    rec("col_attention bwd (p=0.5)", timeit(lambda: BF._col_attention_bwd(q, k, v, go, 8, 0.5, 7), args.reps), 7 * tok_bytes)
    del q, k, v

    layer = nn.TransformerEncoderLayer(d_model=C, nhead=8, dim_feedforward=C, dropout=0.5, activation="relu", batch_first=True).to(dev)
    layer.train()
    tt = t.clone().requires_grad_(True)

    def ours():
        y = BF.transformer_encoder_layer(layer, tt)
        y.backward(go)

    from torch.nn.attention import SDPBackend, sdpa_kernel

    def lib():      # the fused SDP backends refuse batch > 65535 with dropout: math backend, as the reference's torch 2.0 ran it
        with sdpa_kernel(SDPBackend.MATH):
            y = layer(tt)
        y.backward(go)

    flops = R * (12 * S * C * C + 4 * S * S * C) * 3
    ms_o = timeit(ours, args.reps)
    ms_l = timeit(lib, args.reps)
    print(f"{'encoder layer fwd+bwd: b200mp kernels':40s} {ms_o:8.3f} ms  {flops / ms_o / 1e9:8.1f} TFLOP/s", flush=True)
    print(f"{'encoder layer fwd+bwd: torch library':40s} {ms_l:8.3f} ms  {flops / ms_l / 1e9:8.1f} TFLOP/s", flush=True)
    res["encoder_layer_fwd_bwd_b200mp_ms"] = round(ms_o, 3)
    res["encoder_layer_fwd_bwd_torch_ms"] = round(ms_l, 3)
    print(json.dumps({"rows": R, "S": S, "C": C, "results": res}))


if __name__ == "__main__":
    main()
