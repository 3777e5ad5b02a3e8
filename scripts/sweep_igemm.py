"""Sweep (tile_n, split min k-blocks, stages) for the small-M / weight-streaming layers."""
import itertools, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chiaswarm_b200.ops import native as nat
from chiaswarm_b200.ops.packing import pack_conv_weight
scratch = torch.zeros((64 << 20) + 16384, device="cuda", dtype=torch.uint8)
junk = torch.empty(200 << 20, device="cuda", dtype=torch.uint8)

def timeit(fn, iters=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        junk.zero_()  # cold weights, as inside a UNet step
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2]

shapes = [(2, 8, 8, 1280, 1280, 3), (2, 8, 8, 2560, 1280, 3), (2, 16, 16, 1280, 1280, 3), (2, 16, 16, 2560, 1280, 3),
          (2, 32, 32, 640, 640, 3), (2, 32, 32, 1920, 640, 3), (2, 16, 16, 5120, 1280, 1), (2, 32, 32, 2560, 640, 1),
          (2, 16, 16, 1280, 1280, 1), (2, 8, 8, 1280, 1280, 1), (2, 64, 64, 320, 320, 3), (2, 64, 64, 320, 320, 1)]
for (n, h, w, cin, cout, k) in shapes:
    x = torch.randn(n, h, w, cin, device="cuda", dtype=torch.float16)
    wp = pack_conv_weight(torch.randn(cout, cin, k, k, device="cuda", dtype=torch.float16) * 0.02)
    b = torch.randn(cout, device="cuda", dtype=torch.float16)
    out = torch.empty(n, h, w, cout, device="cuda", dtype=torch.float16)
    res = []
    for bn, mk, st in itertools.product([64, 128, 256], [8, 16, 32, 1000], [0, 8]):
        if bn > cout:
            continue
        t = timeit(lambda: nat.igemm(x, wp, out, n=n, h=h, w=w, kh=k, kw=k, pad=k // 2, bias=b, tile_n=bn, cta_pair=1,
                                     splitk_scratch=scratch, splitk_min_kb=mk, stages=st))
        res.append((t, bn, mk, st))
    auto = timeit(lambda: nat.igemm(x, wp, out, n=n, h=h, w=w, kh=k, kw=k, pad=k // 2, bias=b, splitk_scratch=scratch))
    res.sort()
    print(f"{str((n,h,w,cin,cout,k)):32s} auto {auto:6.1f} | best " + "  ".join(f"{t:5.1f}(bn{bn},kb{mk},st{st})" for t, bn, mk, st in res[:5]), flush=True)
