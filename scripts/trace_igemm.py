"""Per-CTA phase breakdown of sw_igemm on the transformer GEMM shapes (probe build, -DSW_TRACE).
  python -m chiaswarm_b200.build_native --trace
  SW_NATIVE_LIB=chiaswarm_b200/_lib/libswarm_sm100_trace.so python scripts/trace_igemm.py [filter]
Every CTA stamps clock64 at its phase boundaries (csrc/igemm.cu, TR(slot)); this script launches each
shape a few times back to back (PDL on, as inside a step) and prints medians over CTAs in SM clocks:
  setup   entry -> barriers/TMEM ready        pdl     -> predecessor drained
  tma0    -> producer issues its first load   lat0    first load issue -> first full barrier seen by the MMA thread
  kloop   first full barrier -> last commit   (and per k-block)      accw    last commit -> epilogue sees acc_full
  epi     epilogue body                       tail    -> final sync / TMEM dealloc
plus the launch's wall span from %globaltimer (min entry -> max exit) and the graph-timed cuBLAS number."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.nn.functional as F

from chiaswarm_b200.ops import native as nat

dev = "cuda"
scratch = torch.zeros((64 << 20) + 16384, device=dev, dtype=torch.uint8)
flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)

shapes = [  # (M, N, K, geglu, residual)
    (128, 256, 64, 0, 0),
    (8192, 320, 320, 0, 1), (8192, 960, 320, 0, 0), (8192, 2560, 320, 1, 0), (8192, 320, 1280, 0, 1),
    (2048, 640, 640, 0, 1), (2048, 1920, 640, 0, 0), (2048, 5120, 640, 1, 0), (2048, 640, 2560, 0, 1),
    (512, 1280, 1280, 0, 1), (512, 3840, 1280, 0, 0), (512, 10240, 1280, 1, 0), (512, 1280, 5120, 0, 1),
    (128, 1280, 1280, 0, 1),
]
# conv layers (weight streaming at the deep levels): (n, h, w, cin, cout, ksize) -- traced through the same path
conv_shapes = [(2, 8, 8, 1280, 1280, 3), (2, 16, 16, 1280, 1280, 3), (2, 8, 8, 2560, 1280, 3), (2, 32, 32, 640, 640, 3),
               (2, 64, 64, 320, 320, 3)]
flt = sys.argv[1] if len(sys.argv) > 1 else ""
lib = nat.lib()
lib.sw_trace_dump.argtypes = [C.c_void_p, C.c_int]
lib.sw_trace_dump.restype = C.c_int


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


def med(x):
    return float(np.median(x)) if len(x) else float("nan")


from chiaswarm_b200.ops.packing import pack_conv_weight

work = [("lin", sh) for sh in shapes] + [("conv", sh) for sh in conv_shapes]
for kind, sh in work:
    if kind == "lin":
        (m, n, k, geglu, res) = sh
        if flt and flt not in f"{m},{n},{k}":
            continue
        x = torch.randn(m, k, device=dev, dtype=torch.float16)
        w = torch.randn(n, k, device=dev, dtype=torch.float16) * k ** -0.5
        b = torch.randn(n, device=dev, dtype=torch.float16)
        nout = n // 2 if geglu else n
        out = torch.empty(m, nout, device=dev, dtype=torch.float16)
        r = torch.randn(m, nout, device=dev, dtype=torch.float16) if res else None
        act = nat.ACT_GEGLU if geglu else nat.ACT_NONE

        def run(**kw):
            return nat.igemm(x, w, out, n=1, h=1, w=m, bias=b, residual=r, act=act, splitk_scratch=scratch, **kw)

        t_cublas = graph_time(lambda: F.linear(x, w, b))
    else:
        (cn, ch, cw, cin, cout, ks) = sh
        if flt and flt not in f"conv{cn},{ch},{cw},{cin},{cout}":
            continue
        m, n, k, geglu, res = cn * ch * cw, cout, cin * ks * ks, 0, 0
        x = torch.randn(cn, ch, cw, cin, device=dev, dtype=torch.float16)
        w4 = torch.randn(cout, cin, ks, ks, device=dev, dtype=torch.float16) * k ** -0.5
        w = pack_conv_weight(w4)
        b = torch.randn(cout, device=dev, dtype=torch.float16)
        out = torch.empty(m, cout, device=dev, dtype=torch.float16)
        xn = x.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)

        def run(**kw):
            return nat.igemm(x, w, out, n=cn, h=ch, w=cw, kh=ks, kw=ks, pad=ks // 2, bias=b, splitk_scratch=scratch, **kw)

        t_cublas = graph_time(lambda: F.conv2d(xn, w4, b, padding=ks // 2))
    for label, kw in (("auto", {}), ("s6", dict(stages=6, cta_pair=1, splitk_min_kb=64, persist=1, tile_n=128))):
        if geglu and label != "auto":
            continue
        try:
            t_ours = graph_time(lambda: run(**kw))
        except Exception as e:
            print(f"({m},{n},{k}) {label}: {e}")
            continue
        for _ in range(3):
            run(**kw)
        flush.zero_()
        x.mul_(1)
        for _ in range(3):
            run(**kw)
        host = np.zeros(8192 * 16, dtype=np.uint64)
        rc = lib.sw_trace_dump(host.ctypes.data, 8192)
        assert rc == 0, lib.sw_last_error()
        tr = host.reshape(8192, 16).astype(np.int64)
        cfg = int(tr[0, 14])
        gd = int(tr[0, 15])
        gx, gy, gz = gd & 0xFFFFF, (gd >> 20) & 0xFFFFF, gd >> 40
        nct = min(8192, gx * gy * gz)
        tr = tr[:nct]
        bn, stages, splits = cfg & 0xFFFF, (cfg >> 16) & 0xFF, (cfg >> 24) & 0xFF
        pair, ck, tma_epi, pers = (cfg >> 32) & 1, (cfg >> 33) & 1, (cfg >> 34) & 1, (cfg >> 35) & 1
        kb = (k + 63) // 64
        d = lambda a, c: tr[:, a] - tr[:, c]
        span = (tr[:, 12].max() - tr[:, 11].min()) / 1e3
        kloop = d(6, 5)
        print(f"{kind}({m},{n},{k}){'g' if geglu else ''}{'+r' if res else ''} {label}: cublas {t_cublas:5.1f} us  ours {t_ours:5.1f} us  "
              f"span {span:5.1f} us | grid {gx}x{gy}x{gz} bn {bn} stages {stages} splits {splits} pair {pair} ck {ck} tma_epi {tma_epi} "
              f"persist {pers} sms {len(set(tr[:, 13].tolist()))}\n"
              f"    clk: setup {med(d(1, 0)):6.0f} pdl {med(d(2, 1)):6.0f} tma0 {med(d(3, 2)):6.0f} lat0 {med(d(5, 3)):6.0f} "
              f"kloop {med(kloop):7.0f} ({med(kloop) / max(1, kb // max(1, splits)):5.0f}/kb) prod_done-first {med(d(4, 3)):7.0f} "
              f"accw {med(d(8, 6)):6.0f} epi {med(d(9, 8)):6.0f} tail {med(d(10, 9)):6.0f} total {med(d(10, 0)):7.0f} "
              f"max total {d(10, 0).max():7.0f}", flush=True)
