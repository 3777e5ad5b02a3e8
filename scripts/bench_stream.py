"""Weight-streaming GEMMs (deep UNet levels: few rows, huge K): sw_igemm vs cuBLAS with the weights
either cycling through a pool larger than L2 (HBM streaming) or held in L2 (pool of 1).  Graph-timed."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from chiaswarm_b200.ops import native as nat

dev = "cuda"
scratch = torch.zeros((64 << 20) + 16384, device=dev, dtype=torch.uint8)
REPS = 12


def graph_time(fns):
    for f in fns:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(REPS):
            fns[i % len(fns)]()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (5 * REPS) * 1e3


cfgs = [(0, 0, -1), (64, 1, 4), (64, 1, 8), (128, 1, 2), (128, 1, 4), (128, 1, 8), (256, 1, 2), (256, 1, 4), (256, 1, 8),
        (256, 2, 4), (256, 2, 8), (128, 2, 4), (64, 1, 0), (256, 1, 0)]
print("M,N,K pool | cublas | " + " ".join(f"{t}{'p' if pr == 2 else ''}/s{mk}" for t, pr, mk in cfgs))
for (m, n, k) in [(128, 1280, 11520), (512, 1280, 11520), (128, 1280, 23040), (512, 1280, 5120), (2048, 640, 5760)]:
    for pool in (1, 6):
        xs = [torch.randn(m, k, device=dev, dtype=torch.float16) for _ in range(pool)]
        ws = [torch.randn(n, k, device=dev, dtype=torch.float16) * k ** -0.5 for _ in range(pool)]
        outs = [torch.empty(m, n, device=dev, dtype=torch.float16) for _ in range(pool)]
        t_ref = graph_time([(lambda i=i: F.linear(xs[i], ws[i])) for i in range(pool)])
        row = []
        for t, pr, mk in cfgs:
            kw = dict(tile_n=t, cta_pair=pr)
            if mk != 0:
                kw.update(splitk_scratch=scratch, splitk_min_kb=max(mk, 0))
            try:
                row.append(graph_time([(lambda i=i: nat.igemm(xs[i], ws[i], outs[i], n=1, h=1, w=m, **kw)) for i in range(pool)]))
            except Exception:
                row.append(float("nan"))
        mb = n * k * 2 / 1e6
        print(f"{(m, n, k)} pool={pool} W={mb:.0f}MB | {t_ref:6.1f} | " + " ".join(f"{r:7.1f}" for r in row), flush=True)
