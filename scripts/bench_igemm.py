"""Micro-benchmark of sw_igemm on the SD1.5 layer shapes (CFG batch 2, 512x512) against
cuDNN / cuBLAS through torch.  CUDA-event timing, 20 iterations after warm-up, inputs rotated
through a pool larger than L2 is NOT done here: these layers run back to back inside a UNet step,
so warm-L2 numbers are the relevant ones (weights stream from HBM either way)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from chiaswarm_b200.ops import native as nat
from chiaswarm_b200.ops.packing import pack_conv_weight

torch.backends.cudnn.benchmark = True
dev = "cuda"
scratch = torch.zeros((64 << 20) + 16384, device=dev, dtype=torch.uint8)


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


shapes = [
    # n, h, w, cin, cout, k
    (2, 64, 64, 320, 320, 3), (2, 64, 64, 640, 320, 3), (2, 64, 64, 960, 320, 3),
    (2, 32, 32, 640, 640, 3), (2, 32, 32, 1280, 640, 3), (2, 32, 32, 1920, 640, 3),
    (2, 16, 16, 1280, 1280, 3), (2, 16, 16, 2560, 1280, 3),
    (2, 8, 8, 1280, 1280, 3), (2, 8, 8, 2560, 1280, 3),
    (2, 64, 64, 320, 960, 1), (2, 64, 64, 320, 320, 1), (2, 64, 64, 1280, 320, 1),
    (2, 32, 32, 640, 1920, 1), (2, 32, 32, 2560, 640, 1), (2, 16, 16, 1280, 3840, 1), (2, 16, 16, 5120, 1280, 1),
    (1, 512, 512, 128, 128, 3), (1, 256, 256, 256, 256, 3), (1, 128, 128, 512, 512, 3),
]
tiles = [(0, 1), (0, 2)]
splitcfg = [(64, 1, 8), (64, 1, 16), (128, 1, 16), (0, 1, 16), (0, 1, 32), (0, 2, 16), (0, 2, 32), (256, 2, 16)]
only = sys.argv[1:]
print(f"{'shape':38s} {'GF':>6s} {'torch':>8s} " + " ".join(f"{(('bn'+str(t)) if t else 'auto')+('x2' if pr==2 else ''):>8s}" for t, pr in tiles) + "  | splitK: " + " ".join(f"{t}/{'x2' if pr==2 else 'x1'}/{mk:<3d}" for t, pr, mk in splitcfg))
for (n, h, w, cin, cout, k) in shapes:
    x = torch.randn(n, cin, h, w, device=dev, dtype=torch.float16)
    wt = torch.randn(cout, cin, k, k, device=dev, dtype=torch.float16) * 0.02
    b = torch.randn(cout, device=dev, dtype=torch.float16)
    xl = x.permute(0, 2, 3, 1).contiguous()
    wp = pack_conv_weight(wt)
    out = torch.empty(n, h, w, cout, device=dev, dtype=torch.float16)
    gflop = 2.0 * n * h * w * cin * cout * k * k / 1e9
    xc = x.contiguous(memory_format=torch.channels_last)
    t_torch = timeit(lambda: F.conv2d(xc, wt, b, padding=k // 2))
    res = []
    for t, pr in tiles:
        if t and t > max(32, (cout + 31) // 32 * 32):
            res.append(float("nan"))
            continue
        try:
            res.append(timeit(lambda: nat.igemm(xl, wp, out, n=n, h=h, w=w, kh=k, kw=k, pad=k // 2, bias=b, tile_n=t, cta_pair=pr)))
        except Exception as e:
            res.append(float("nan"))
    sp = []
    for t, pr, mk in splitcfg:
        if t and t > max(32, (cout + 31) // 32 * 32):
            sp.append(float("nan"))
            continue
        sp.append(timeit(lambda: nat.igemm(xl, wp, out, n=n, h=h, w=w, kh=k, kw=k, pad=k // 2, bias=b, tile_n=t,
                                           cta_pair=pr, splitk_scratch=scratch, splitk_min_kb=mk)))
    best = min(r for r in res + sp if r == r)
    print(f"{str((n,h,w,cin,cout,k)):38s} {gflop:6.1f} {t_torch:8.1f} " + " ".join(f"{r:8.1f}" for r in res)
          + "  | " + " ".join(f"{r:9.1f}" for r in sp) + f"   best {gflop/best*1e3:6.0f} TF vs torch {gflop/t_torch*1e3:6.0f} TF")
