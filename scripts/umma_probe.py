"""Decode which shared-memory word the tcgen05 tensor core fetches for element (mn, k) of an
operand, per layout type / major-ness / LBO / SBO, with b200mp_umma_probe (one K=8 MMA).

The probed operand's image holds, in every 4-byte word, a code of its own word index (two passes:
index % 1024 + 1 and index // 1024 + 1, both exact in tf32); the other operand is one-hot over k.
Results (word index per (mn, k)) go to gpurun_out/umma_probe.npz and a summary is printed."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from b200mp import _lib as L


def desc(lbo, sbo, layout):
    return ((lbo >> 4) & 0x3FFF) << 16 | ((sbo >> 4) & 0x3FFF) << 32 | 1 << 46 | layout << 61


def idesc(M, N, a_mn, b_mn):
    return (1 << 4) | (2 << 7) | (2 << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def onehot_kmajor(rows):
    """[rows x 8] K-major interleaved image with element (r, k) = 1 if r == k (r < 8)."""
    img = np.zeros(8192, np.float32)
    plane = rows * 16 // 4
    for k in range(8):
        img[(k // 4) * plane + k * 4 + (k % 4)] = 1.0
    return img, desc(rows * 16, 128, 0)


def run(a_img, b_img, adesc, bdesc, idsc):
    a = torch.from_numpy(a_img).cuda()
    b = torch.from_numpy(b_img).cuda()
    D = torch.zeros(128, 64, device="cuda")
    L.check(L.lib().b200mp_umma_probe(L.ptr(a), a.numel() * 4, L.ptr(b), b.numel() * 4, C.c_uint64(adesc), C.c_uint64(bdesc),
                                      0, 0, 1, idsc, L.ptr(D), L.stream()), "probe")
    torch.cuda.synchronize()
    return D.cpu().numpy()


def codes():
    idx = np.arange(8192)
    return (idx % 1024 + 1).astype(np.float32), (idx // 1024 + 1).astype(np.float32)


def probe_b(layout, lbo, sbo, b_mn):
    """word index fetched for B element (n, k): D[r=k][n]."""
    a_img, adesc = onehot_kmajor(128)
    lo, hi = codes()
    i = idesc(128, 64, 0, b_mn)
    d1 = run(a_img, lo, adesc, desc(lbo, sbo, layout), i)
    d2 = run(a_img, hi, adesc, desc(lbo, sbo, layout), i)
    w = (d2[:8] - 1) * 1024 + (d1[:8] - 1)          # [k][n]
    return w.T, d1                                   # [n][k]


def probe_a(layout, lbo, sbo, a_mn, M):
    """word index fetched for A element (m, k): D[m][n=k]; returns raw lanes too."""
    b_img, bdesc = onehot_kmajor(64)
    lo, hi = codes()
    i = idesc(M, 64, a_mn, 0)
    d1 = run(lo, b_img, desc(lbo, sbo, layout), bdesc, i)
    d2 = run(hi, b_img, desc(lbo, sbo, layout), bdesc, i)
    return (d2[:, :8] - 1) * 1024 + (d1[:, :8] - 1), d1


def describe(w):
    """w[mn][k] word indices -> strides in bytes."""
    w = np.asarray(w)
    if not np.isfinite(w).all() or (w < 0).any():
        return "invalid (zeros / prefill)  sample=" + str(w[:2, :2].tolist())
    b = (w * 4).astype(np.int64)
    return (f"base {b[0, 0]}  mn+1 {b[1, 0] - b[0, 0]}  mn+4 {b[4, 0] - b[0, 0]}  mn+8 {b[8, 0] - b[0, 0]}  mn+32 {b[32, 0] - b[0, 0]}"
            f"  k+1 {b[0, 1] - b[0, 0]}  k+4 {b[0, 4] - b[0, 0]}  | row1: k+1 {b[1, 1] - b[1, 0]}")


if __name__ == "__main__" and len(sys.argv) > 1:
    kind, layout, lbo, sbo, mn, M = sys.argv[1], *map(int, sys.argv[2:7])
    if kind == "B":
        w, raw = probe_b(layout, lbo, sbo, mn)
        print(f"B major={'MN' if mn else 'K'} layout={layout} lbo={lbo} sbo={sbo}: {describe(w)}", flush=True)
    else:
        w, raw = probe_a(layout, lbo, sbo, mn, M)
        rows = np.where(raw[:, 0] > 0)[0]
        print(f"A M={M} major={'MN' if mn else 'K'} layout={layout} lbo={lbo} sbo={sbo}: lanes with data n={len(rows)} first {rows[:3].tolist()} "
              f"lanes16-18 {raw[16:19, 0].tolist()} lanes32-34 {raw[32:35, 0].tolist()} {describe(w[rows]) if len(rows) >= 33 else ''}", flush=True)
    np.save(f"gpurun_out/probe_{'_'.join(sys.argv[1:7])}.npy", w)
    sys.exit(0)

if __name__ == "__main__":
    import subprocess
    LBO, SBO = 4096, 512
    cfgs = []
    for layout in (0, 2, 1, 4, 6):
        for (l, s) in ((LBO, SBO), (SBO, LBO)):
            cfgs.append(("B", layout, l, s, 1, 128))
    for M in (128, 64):
        for layout in (0, 2):
            cfgs.append(("A", layout, LBO, SBO, 1, M))
            cfgs.append(("A", layout, SBO, LBO, 1, M))
    cfgs += [("B", 2, LBO, SBO, 0, 128), ("A", 0, LBO, SBO, 0, 64)]
    for c in cfgs:
        r = subprocess.run([sys.executable, __file__, *map(str, c)], capture_output=True, text=True, timeout=120)
        last = [l for l in (r.stdout + r.stderr).strip().splitlines() if l.strip()]
        print(last[0][:400] if r.returncode == 0 and last else f"{c}: FAILED {[l for l in last if 'rror' in l][:1]}", flush=True)
