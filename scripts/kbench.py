#!/usr/bin/env python
"""Kernel micro-benchmarks through the C ABI: each per-edge kernel on the synthetic graph of a
named workload, in variants that switch the gathers / scatters off (null tables), to separate
the streaming + tensor-core pipeline from the random-access traffic.

    python scripts/kbench.py [--workload hi_small|hi_medium] [--reps 5]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import b200mp  # noqa: E402
from b200mp import _lib as L  # noqa: E402
from b200mp import functional as BF  # noqa: E402

H = 64


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
    ap.add_argument("--workload", default="hi_medium")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    n, e, seed = bench.WORKLOADS[args.workload]
    ei = bench.synth_graph(n, e, seed, dev)
    st = b200mp.edge_structure(ei, n)
    seg = st.fwd
    g = torch.Generator(device=dev).manual_seed(1)
    rn = lambda *s: torch.randn(*s, generator=g, device=dev)
    E = e
    ea, m, gout = rn(E, H), rn(E, H), rn(E, H)
    hh = torch.relu(rn(E, H))
    T1, T2 = rn(n, H), rn(n, H)
    W1, W2, b1, b2 = rn(H, H) / 8, rn(H, H) / 8, rn(H), rn(H)
    coef = rn(n, 7 * H)
    out1, out2 = torch.empty(E, H, device=dev), torch.empty(E, H, device=dev)
    gP, gQ = torch.zeros(n, H, device=dev), torch.zeros(n, H, device=dev)
    npart = L.n_part(E, H)
    gW1, gW2 = torch.empty(npart, H, H, device=dev), torch.empty(npart, H, H, device=dev)
    gb1, gb2 = torch.empty(npart, H, device=dev), torch.empty(npart, H, device=dev)
    X4, Z3, Wz = rn(n, 4 * H), rn(n, 3 * H), rn(3 * H, 4 * H) / 16
    lib, p, s = L.lib(), L.ptr, L.stream
    res = {}

    def fwd(mode, eid, t1, i1, t2, i2):
        L.check(lib.b200mp_edge_linear_fwd(mode, H, E, p(ea), p(eid), p(W1), p(b1), p(t1), p(i1), p(t2), p(i2),
                                           p(W2), p(b2), 1.0, 0.5, p(out1), p(out2 if mode == 2 else None), None, s()), "fwd")

    def updb(src, dst, gp, gq):
        L.check(lib.b200mp_edge_update_bwd(H, E, p(gout), p(hh), p(ea), p(src), p(dst), p(W1), p(W2), 1.0, 0.5,
                                           p(out1), p(gp), p(gq), p(gW1), p(gb1), p(gW2), p(gb2), npart, None, None, s()), "updb")

    def msgb(kind, node_in, gt1, eid=None):
        L.check(lib.b200mp_message_bwd(kind, H, E, p(m), p(ea), p(eid), p(node_in), p(seg.seg32), p(seg.nbr32), p(W1),
                                       p(out1), 0, p(gt1), p(gW1), p(gb1), npart, None, s()), "msgb")

    def segr():
        BF._segment_reduce(0, seg, m, H)

    cases = {
        "msg_fwd(A[dst]+B[src])": lambda: fwd(0, None, T1, seg.nbr32, T2, seg.seg32),
        "msg_fwd(no gathers)": lambda: fwd(0, None, None, None, None, None),
        "msg_fwd(B[src] only)": lambda: fwd(0, None, T1, seg.nbr32, None, None),
        "upd_fwd(P[src]+Q[dst])": lambda: fwd(2, None, T1, st.src32, T2, st.dst32),
        "upd_fwd(no gathers)": lambda: fwd(2, None, None, None, None, None),
        "msg_fwd(eid map)": lambda: fwd(0, seg.perm32(), T1, seg.nbr32, T2, seg.seg32),
        "upd_fwd(eid map)": lambda: fwd(2, seg.perm32(), T1, st.src32, T2, st.dst32),
        "upd_bwd(gP+gQ)": lambda: updb(st.src32, st.dst32, gP, gQ),
        "upd_bwd(gQ only)": lambda: updb(st.src32, st.dst32, None, gQ),
        "upd_bwd(gP only)": lambda: updb(st.src32, st.dst32, gP, None),
        "upd_bwd(no scatters)": lambda: updb(st.src32, st.dst32, None, None),
        "msg_bwd_pna(gB)": lambda: msgb(0, coef, gP),
        "msg_bwd_pna(no scatter)": lambda: msgb(0, coef, None),
        "msg_bwd_gine(gx)": lambda: msgb(1, T1, gP),
        "seg_reduce_pna": segr,
        "permute_rows": lambda: BF._permute(ea, seg.perm, False),
        "node_linear(256->192)": lambda: BF._node_linear(X4, Wz, None),
        "node_linear(192->256)": lambda: BF._node_linear(Z3, Wz.t(), None),
        "node_wgrad(256x192)": lambda: BF._node_wgrad(X4, Z3),
        "mm3 lib (256->192)": lambda: BF._mm3(BF._split_tf32(X4), BF._split_tf32(Wz), tb=True),
    }
    for name, fn in cases.items():
        if args.only and args.only not in name:
            continue
        res[name] = round(timeit(fn, args.reps), 3)
        print(f"{name:28s} {res[name]:8.3f} ms", flush=True)
    print(json.dumps({"workload": args.workload, "E": E, "N": n, "ms": res}))


if __name__ == "__main__":
    main()
