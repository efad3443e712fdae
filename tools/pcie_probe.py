"""PCIe ceiling for the e2e leg of bench.py: pinned H2D 12 MB, D2H 25 MB, and both at once on two streams."""
import time
import torch
dev = torch.device("cuda", 0)
h_in = torch.empty(12_000_000, dtype=torch.uint8).pin_memory()
h_out = torch.empty(25_000_000, dtype=torch.uint8).pin_memory()
d_in = torch.empty_like(h_in, device=dev)
d_out = torch.empty_like(h_out, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def t(fn, n=20):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n


def h2d():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    h2d()
    d2h()


a, b, c = t(h2d), t(d2h), t(both)
print(f"H2D 12MB {a*1e3:.3f} ms {12e-3/a:.1f} GB/s | D2H 25MB {b*1e3:.3f} ms {25e-3/b:.1f} GB/s | both {c*1e3:.3f} ms -> {1e6/c/1e9:.2f} G queries/s ceiling")
