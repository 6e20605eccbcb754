#!/usr/bin/env python
"""b200mp_row_linear at E rows, H = 128: time per launch of the plain product and of each epilogue form (CUDA events, 5 reps after 2)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200mp
from b200mp import functional as BF

E = int(sys.argv[1]) if len(sys.argv) > 1 else 5_078_345
N = int(sys.argv[2]) if len(sys.argv) > 2 else 515_080
H = int(sys.argv[3]) if len(sys.argv) > 3 else 128
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(E, H, device=dev, generator=g)
W = torch.randn(H, H, device=dev, generator=g) / H ** 0.5
b = torch.randn(H, device=dev, generator=g)
T1, T2 = torch.randn(N, H, device=dev, generator=g), torch.randn(N, H, device=dev, generator=g)
i1 = torch.randint(0, N, (E,), device=dev, generator=g).to(torch.int32)
i2 = torch.sort(torch.randint(0, N, (E,), device=dev, generator=g)).values.to(torch.int32)
R = torch.randn(E, H, device=dev, generator=g)
perm = torch.randperm(E, device=dev, generator=g).to(torch.int32)
out = torch.empty(E, H, device=dev)


def timeit(name, fn, bytes_):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        fn()
    c.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(c) / 5
    print(f"{name:34s} {ms:8.3f} ms  {bytes_ / ms / 1e6:8.1f} GB/s")


row = 4 * H * E
timeit("plain", lambda: BF._row_linear(x, W, b, out=out), 2 * row)
timeit("+T1[random] +T2[sorted]", lambda: BF._row_linear(x, W, b, T1=T1, idx1=i1, T2=T2, idx2=i2, out=out), 3 * row)
timeit("+T1[random] relu", lambda: BF._row_linear(x, W, b, T1=T1, idx1=i1, relu=True, out=out), 3 * row)
timeit("+R residual", lambda: BF._row_linear(x, W, b, R=R, alpha=0.5, beta=0.5, out=out), 3 * row)
timeit("+mask", lambda: BF._row_linear(x, W, None, mask=R, beta=0.5, out=out), 3 * row)
timeit("xrow=perm (random rows in)", lambda: BF._row_linear(x, W, b, xrow=perm, out=out), 2 * row)
timeit("orow=perm (random rows out)", lambda: BF._row_linear(x, W, b, out=out, orow=perm), 2 * row)
if H == 128:
    timeit("rows wgrad a^T b", lambda: BF._rows_wgrad(x, R), 2 * row)
Gm = torch.empty(E, H, device=dev)
coef = torch.randn(N, 7 * H, device=dev, generator=g)
import b200mp._lib as L
timeit("message_grad_rows pna +RED", lambda: L.check(L.lib().b200mp_message_grad_rows(0, H, E, L.ptr(x), L.ptr(coef), L.ptr(i2), L.ptr(i1), L.ptr(Gm), L.ptr(T1), L.stream()), "mgr"), 3 * row)
