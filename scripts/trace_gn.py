"""GroupNorm(+SiLU) on the UNet shapes: graph-timed sw_groupnorm vs an elementwise kernel over the same bytes
(sw_silu: the floor of one read + one write) vs torch, plus the per-CTA phase trace of k_gn_fused (probe build):
  SW_NATIVE_LIB=chiaswarm_b200/_lib/libswarm_sm100_trace.so python scripts/trace_gn.py"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.nn.functional as F

from chiaswarm_b200.ops import native as nat

dev = "cuda"
lib = nat.lib()
traced = "trace" in os.environ.get("SW_NATIVE_LIB", "")
if traced:
    lib.sw_trace_dump_gn.argtypes = [C.c_void_p, C.c_int]
    lib.sw_trace_dump_gn.restype = C.c_int


def graph_time(fn, reps=16):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps * 1e3)
    return best


shapes = [(2, 4096, 320, 0), (2, 4096, 320, 320), (2, 4096, 640, 320), (2, 1024, 640, 0), (2, 1024, 640, 640),
          (2, 1024, 1280, 640), (2, 256, 1280, 0), (2, 256, 1280, 1280), (2, 64, 1280, 0), (2, 64, 1280, 1280),
          (2, 16384, 320, 0), (2, 4096, 640, 0), (1, 16384, 512, 0), (1, 65536, 256, 0)]
for n, hw, c1, c2 in shapes:
    c = c1 + c2
    x1 = torch.randn(n, hw, c1, device=dev, dtype=torch.float16)
    x2 = torch.randn(n, hw, c2, device=dev, dtype=torch.float16) if c2 else None
    g = torch.randn(c, device=dev, dtype=torch.float16)
    b = torch.randn(c, device=dev, dtype=torch.float16)
    y = torch.empty(n, hw, c, device=dev, dtype=torch.float16)
    scratch = torch.zeros(nat.groupnorm_scratch_floats(n, hw, c, 32), device=dev, dtype=torch.float32)
    run = lambda: nat.groupnorm(x1, g, b, y, n=n, hw=hw, groups=32, eps=1e-5, silu=True, stats=scratch, x2=x2)
    t = graph_time(run)
    xe = torch.randn(n * hw * c, device=dev, dtype=torch.float16)
    ye = torch.empty_like(xe)
    t_el = graph_time(lambda: nat.silu(xe, ye)) if hasattr(nat, "silu") else float("nan")
    xt = torch.randn(n, c, hw, device=dev, dtype=torch.float16)
    t_torch = graph_time(lambda: F.silu(F.group_norm(xt, 32, g, b, 1e-5)))
    kc = lib.sw_groupnorm_kernel_count(n, hw, c, 32)
    line = f"({n},{hw},{c1}+{c2}) kernels {kc}: ours {t:5.1f} us  elementwise floor {t_el:5.1f} us  torch gn+silu {t_torch:5.1f} us  ({n*hw*c*4/1e6:5.1f} MB moved)"
    if traced and kc == 1:
        for _ in range(3):
            run()
        host = np.zeros(8192 * 16, dtype=np.uint64)
        assert lib.sw_trace_dump_gn(host.ctypes.data, 8192) == 0
        tr = host.reshape(8192, 16).astype(np.int64)
        gd = int(tr[0, 15])
        gx, gy, gz = gd & 0xFFFFF, (gd >> 20) & 0xFFFFF, gd >> 40
        nct = min(8192, gx * gy * gz)
        cfg = int(tr[0, 14])
        tr = tr[:nct]
        d = lambda a, c_: float(np.median(tr[:, a] - tr[:, c_]))
        span = (tr[:, 12].max() - tr[:, 11].min()) / 1e3
        line += (f"\n    grid {gx}x{gy}x{gz} threads {cfg & 0xFFFF} rows/cta {(cfg >> 16) & 0xFFFF} rows/pass {cfg >> 32} sms {len(set(tr[:,13].tolist()))} span {span:5.1f} us | clk: pdl {d(1,0):6.0f} "
                 f"load {d(2,1):6.0f} sync {d(3,2):6.0f} reduce {d(4,3):6.0f} cluster_sync {d(5,4):6.0f} apply {d(6,5):6.0f} total {d(6,0):7.0f} max {(tr[:,6]-tr[:,0]).max():7.0f}")
    print(line, flush=True)
