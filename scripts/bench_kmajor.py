"""Plain [cout, K] weights vs the k-block-major packing (ops/packing.pack_kmajor) on weight-streaming layers:
cold weights (a pool larger than L2 cycles through the graph), graph-timed, plus a numerics check."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from chiaswarm_b200.ops import native as nat
from chiaswarm_b200.ops.packing import pack_conv_weight, pack_kmajor

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


# (n, h, w, cin, cout, k)
shapes = [(2, 8, 8, 1280, 1280, 3), (2, 16, 16, 1280, 1280, 3), (2, 8, 8, 2560, 1280, 3), (2, 16, 16, 2560, 1280, 3),
          (2, 32, 32, 640, 640, 3), (2, 32, 32, 1280, 640, 3), (2, 64, 64, 320, 320, 3),
          (1, 1, 512, 1280, 1280, 1), (1, 1, 512, 5120, 1280, 1), (1, 1, 512, 1280, 3840, 1), (1, 1, 128, 1280, 1280, 1),
          (1, 1, 2048, 640, 640, 1), (1, 1, 8192, 320, 320, 1), (1, 1, 2048, 2560, 640, 1)]
print(f"{'shape':34s} {'W MB':>6s} {'torch':>7s} {'plain':>7s} {'kmajor':>7s}  GB/s(kmajor)  rel.err")
for (n, h, w, cin, cout, k) in shapes:
    wbytes = cout * cin * k * k * 2
    pool = max(2, int(200e6 // wbytes) + 1)
    pool = min(pool, 24)
    xs = [torch.randn(n, h, w, cin, device=dev, dtype=torch.float16) for _ in range(2)]
    w4 = [torch.randn(cout, cin, k, k, device=dev, dtype=torch.float16) * (cin * k * k) ** -0.5 for _ in range(pool)]
    wp = [pack_conv_weight(t) for t in w4]
    wk = [pack_kmajor(t, k * k, cin) for t in wp]
    bias = torch.randn(cout, device=dev, dtype=torch.float16)
    out = torch.empty(n * h * w, cout, device=dev, dtype=torch.float16)
    xn = [t.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last) for t in xs]
    t_ref = graph_time([(lambda i=i: F.conv2d(xn[i % 2], w4[i], bias, padding=k // 2)) for i in range(pool)])
    run = lambda ws, i: nat.igemm(xs[i % 2], ws[i], out, n=n, h=h, w=w, kh=k, kw=k, pad=k // 2, bias=bias,
                                  splitk_scratch=scratch)
    t_plain = graph_time([(lambda i=i: run(wp, i)) for i in range(pool)])
    t_km = graph_time([(lambda i=i: run(wk, i)) for i in range(pool)])
    run(wk, 0)
    ref = F.conv2d(xn[0].float(), w4[0].float(), bias.float(), padding=k // 2).permute(0, 2, 3, 1).reshape(-1, cout)
    err = ((out.float() - ref).abs().max() / ref.abs().max()).item()
    print(f"{str((n, h, w, cin, cout, k)):34s} {wbytes / 1e6:6.1f} {t_ref:7.1f} {t_plain:7.1f} {t_km:7.1f}  {wbytes / t_km / 1e3:8.0f}     {err:.1e}",
          flush=True)
