"""Device-time micro-benchmark of sw_igemm on the transformer / 1x1 GEMM shapes of the UNet steps.
Each configuration is captured ONCE into a CUDA graph of `REPS` launches cycling through a pool of
distinct operand sets (so weights do not sit in L2 by construction) and the graph replay is timed with
CUDA events: no host launch cost in the number (the plain loop of bench_igemm.py is host-bound below
~20 us).  cuBLAS (torch.nn.functional.linear) is timed the same way for reference.
  python scripts/bench_gemm_graph.py [filter]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from chiaswarm_b200.ops import native as nat

dev = "cuda"
REPS, POOL = 24, 6
scratch = torch.zeros((64 << 20) + 16384, device=dev, dtype=torch.uint8)


def graph_time(fns):
    """fns: list of callables (one per pool slot); returns us per launch."""
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


shapes = [  # (M, N, K, geglu)
    (128, 64, 64, 0), (128, 256, 64, 0),
    (8192, 320, 320, 0), (8192, 960, 320, 0), (8192, 2560, 320, 1), (8192, 320, 1280, 0),
    (2048, 640, 640, 0), (2048, 1920, 640, 0), (2048, 5120, 640, 1), (2048, 640, 2560, 0),
    (512, 1280, 1280, 0), (512, 3840, 1280, 0), (512, 10240, 1280, 1), (512, 1280, 5120, 0),
    (128, 1280, 1280, 0),
    (1024, 1280, 1280, 0), (1024, 3840, 1280, 0), (1024, 10240, 1280, 1), (1024, 1280, 5120, 0),
    (2048, 1280, 1280, 0), (2048, 1280, 5120, 0),
    (4096, 640, 640, 0), (4096, 640, 2560, 0), (4096, 5120, 640, 1),
]
cfgs = [  # (tile_n, cta_pair, splitk_min_kb (0 = no split-K), stages)
    (0, 0, -1, 0),      # what the engine uses today (auto + scratch)
    (64, 1, 0, 0), (128, 1, 0, 0), (256, 1, 0, 0), (128, 2, 0, 0), (256, 2, 0, 0),
    (128, 1, 4, 0), (256, 1, 4, 0), (128, 1, 8, 0), (256, 1, 8, 0), (256, 2, 8, 0),
]
flt = sys.argv[1] if len(sys.argv) > 1 else ""
if os.environ.get("GEGLU_SWEEP"):
    shapes = [sh for sh in shapes if sh[3]]
    cfgs = [(0, 0, -1, 0), (128, 1, 0, 2), (128, 1, 0, 3), (128, 1, 0, 4), (256, 1, 0, 2), (256, 1, 0, 3), (256, 2, 0, 2),
            (256, 2, 0, 3), (128, 2, 0, 2), (128, 2, 0, 3)]
hdr = " ".join(f"{('auto' if t == 0 else str(t)) + ('p' if pr == 2 else '') + ('' if mk <= 0 else '/s' + str(mk)) + ('' if not st else '@' + str(st)):>8s}" for t, pr, mk, st in cfgs)
print(f"{'M,N,K':22s} {'GF':>6s} {'cublas':>7s} {hdr}")
for (m, n, k, geglu) in shapes:
    if flt and flt not in f"{m},{n},{k}":
        continue
    xs = [torch.randn(m, k, device=dev, dtype=torch.float16) for _ in range(POOL)]
    ws = [torch.randn(n, k, device=dev, dtype=torch.float16) * k ** -0.5 for _ in range(POOL)]
    bs = [torch.randn(n, device=dev, dtype=torch.float16) for _ in range(POOL)]
    nout = n // 2 if geglu else n
    outs = [torch.empty(m, nout, device=dev, dtype=torch.float16) for _ in range(POOL)]
    gf = 2.0 * m * n * k / 1e9
    t_ref = graph_time([(lambda i=i: F.linear(xs[i], ws[i], bs[i])) for i in range(POOL)])
    row = []
    for t, pr, mk, st in cfgs:
        if geglu and t not in (0, 128, 256):
            row.append(float("nan"))
            continue
        if t > max(64, (n + 31) // 32 * 32):
            row.append(float("nan"))
            continue
        kw = dict(tile_n=t, cta_pair=pr, stages=st)
        if mk != 0:
            kw.update(splitk_scratch=scratch, splitk_min_kb=max(mk, 0))
        try:
            row.append(graph_time([(lambda i=i: nat.igemm(xs[i], ws[i], outs[i], n=1, h=1, w=m, bias=bs[i],
                                                          act=nat.ACT_GEGLU if geglu else nat.ACT_NONE, **kw))
                                   for i in range(POOL)]))
        except Exception as e:
            row.append(float("nan"))
    best = min(r for r in row if r == r)
    print(f"{str((m, n, k)) + ('g' if geglu else ''):22s} {gf:6.1f} {t_ref:7.1f} " + " ".join(f"{r:8.1f}" for r in row)
          + f"  best {gf / best * 1e3:5.0f} TF  cublas {gf / t_ref * 1e3:5.0f} TF", flush=True)
