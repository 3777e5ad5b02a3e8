"""Multi-GPU check + timing of the fused compute+collective kernels against the NCCL baseline and a
single-GPU oracle.  Run under torchrun (one process per GPU):
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/dist_check.py
Device-timed with CUDA events, max over ranks."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F

from chiaswarm_b200.ops import native as nat
from chiaswarm_b200.parallel import runtime
from chiaswarm_b200.parallel.collectives import FusedOps

ap = argparse.ArgumentParser()
ap.add_argument("--skew-ms", type=float, default=0.0,
                help="the last rank reaches every checked kernel this many ms late (device-side sleep on its stream)")
ap.add_argument("--no-timing", action="store_true")
args = ap.parse_args()

gang = runtime.init()
dev, W, R = gang.device, gang.world, gang.rank


def skewed(fn):
    """Run fn with the skewed arrival the flag protocol must tolerate (no-op without --skew-ms)."""
    if args.skew_ms > 0 and R == W - 1:
        torch.cuda._sleep(int(args.skew_ms * 1e-3 * 1.9e9))
    fn()
    if args.skew_ms > 0:  # and the other way round on the next call: the FIRST rank is late
        torch.cuda.synchronize()
        gang.barrier()
        if R == 0:
            torch.cuda._sleep(int(args.skew_ms * 1e-3 * 1.9e9))
        fn()
        torch.cuda.synchronize()
        gang.barrier()


arena = runtime.SymmetricArena(gang, 1 << 30)
ops = FusedOps(arena)
results = {}


def bcast(t):
    dist.broadcast(t, 0)
    return t


def timed(fn, iters=20):
    if args.no_timing:
        return 0.0
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    gang.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return gang.all_reduce_max(e0.elapsed_time(e1) / iters * 1e3)


def relerr(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6)).item()


def report(name, err_fused, err_nccl, t_fused, t_nccl, extra):
    err = gang.all_reduce_max(err_fused)
    errn = gang.all_reduce_max(err_nccl)
    results[name] = dict(rel_err_fused=err, rel_err_nccl=errn, us_fused=t_fused, us_nccl=t_nccl, **extra)
    if R == 0:
        print(f"{name:24s} err fused {err:.2e} nccl {errn:.2e} | fused {t_fused:8.1f} us  nccl {t_nccl:8.1f} us  {extra}", flush=True)


torch.manual_seed(0)
# ---------------------------------------------------------------- W11 all-gather -> GEMM (SDXL L1 FF-in shape)
m_local, k, n = 4096 // W * 2, 640, 5120
x_full = bcast(torch.randn(W * m_local, k, device=dev, dtype=torch.float16))
w_full = bcast(torch.randn(n, k, device=dev, dtype=torch.float16) * k ** -0.5)
b_full = bcast(torch.randn(n, device=dev, dtype=torch.float16))
w_cols = w_full[R * (n // W):(R + 1) * (n // W)].contiguous()  # column-parallel weight shard
b_cols = b_full[R * (n // W):(R + 1) * (n // W)].contiguous()
xs = arena.alloc(m_local, k)
xs.local.copy_(x_full[R * m_local:(R + 1) * m_local])
out_f = torch.empty(W * m_local, n // W, device=dev, dtype=torch.float16)
out_n = torch.empty_like(out_f)
arena.barrier()
skewed(lambda: ops.allgather_gemm_fused(xs, w_cols, out_f, bias=b_cols))
ops.allgather_gemm_nccl(xs.local, w_cols, out_n, bias=b_cols)
ref = x_full.float() @ w_cols.float().t() + b_cols.float()
tf1 = timed(lambda: ops.allgather_gemm_fused(xs, w_cols, out_f, bias=b_cols))
tn = timed(lambda: ops.allgather_gemm_nccl(xs.local, w_cols, out_n, bias=b_cols))
flops = 2.0 * W * m_local * k * (n // W)
link_bytes = (W - 1) * m_local * k * 2
err_v1 = relerr(out_f, ref)
# second generation: copy role inside the GEMM grid pushes the shard once, compute blocks read local memory.
# Inputs change on every call (the shard is rewritten, scaled by the call index) and the LAST result is checked, so a
# racing overwrite of a buffer a peer still reads cannot hide behind identical values.
out_f.zero_()
calls = [0]


def ag2():
    calls[0] += 1
    g = ops.gather_buffer(m_local, k)
    torch.mul(x_full[R * m_local:(R + 1) * m_local], float(calls[0] % 3 + 1), out=g.local[R * m_local:(R + 1) * m_local])
    ops.allgather_gemm(g, w_cols, out_f, bias=b_cols)


skewed(ag2)
for _ in range(5):  # back-to-back calls on changing inputs (parity buffers, no host sync in between)
    ag2()
torch.cuda.synchronize()
gang.barrier()
for pair_buf in ops._gather[(m_local, k)]:  # timing runs on fixed inputs: both parity buffers hold the shard
    pair_buf.local[R * m_local:(R + 1) * m_local].copy_(x_full[R * m_local:(R + 1) * m_local])
scratch_ag = torch.empty_like(out_f)
tf = timed(lambda: ops.allgather_gemm(ops.gather_buffer(m_local, k), w_cols, scratch_ag, bias=b_cols))
ref_last = (x_full.float() * float(calls[0] % 3 + 1)) @ w_cols.float().t() + b_cols.float()
# diagnostics: (a) the same kernel with every "peer" pointing at this rank's own memory and epoch 0 (no NVLink traffic,
# no waiting): what the two-role kernel costs by itself; (b) a bare push exchange of the same bytes
g_loc = ops.gather_buffer(m_local, k)
g_loc.local.copy_(x_full)
fl_loc = ops._flags_for("ag", W * (m_local // 128))
pf_loc = nat.make_peers(W, R, 0, [g_loc.peer_ptr(R)] * W, [fl_loc.peer_ptr(R)] * W)
t_ag_local = timed(lambda: nat.igemm_allgather(w_cols, out_n, pf_loc, m_local=m_local, k=k, bias=b_cols, tile_n=256))
xch = arena.alloc(W, m_local, k)
xfl = arena.alloc_flags(8)
xcalls = [0]


def push_only():
    xcalls[0] += 1
    pf = nat.make_peers(W, R, xcalls[0], None, xfl.peer_ptrs())
    nat.peer_exchange(pf, [(d, xs.local, xch.peer_ptr(d) + R * m_local * k * 2) for d in range(W) if d != R])


t_push = timed(push_only)
x_all = x_full.clone()
t_gemm = timed(lambda: nat.linear(x_all, w_cols, out=out_n, bias=b_cols))  # the math alone (no communication at all)
tf256 = timed(lambda: ops.allgather_gemm(ops.gather_buffer(m_local, k), w_cols, out_n, bias=b_cols, tile_n=256))
report("allgather_gemm", relerr(out_f, ref_last), relerr(out_n, ref) if args.no_timing else 0.0, tf, tn,
       dict(gflop=flops / 1e9, link_mb=link_bytes / 1e6, us_fused_gen1=tf1, rel_err_gen1=err_v1, us_local_gemm_only=t_gemm,
            us_fused_tile256=tf256, us_fused_tile256_local_only=t_ag_local, us_push_exchange_only=t_push))

# ---------------------------------------------------------------- W12 GEMM -> reduce-scatter (FF-out shape)
kk, n2 = 2560, 640
x2 = bcast(torch.randn(W * m_local, kk, device=dev, dtype=torch.float16))
w2 = bcast(torch.randn(n2, kk, device=dev, dtype=torch.float16) * kk ** -0.5)
b2 = bcast(torch.randn(n2, device=dev, dtype=torch.float16))
res_full = bcast(torch.randn(W * m_local, n2, device=dev, dtype=torch.float16))
ks = kk // W
x_ksh = x2[:, R * ks:(R + 1) * ks].contiguous()
w_ksh = w2[:, R * ks:(R + 1) * ks].contiguous()
res_loc = res_full[R * m_local:(R + 1) * m_local].contiguous()
staging = arena.alloc(W, m_local, n2)
o_f = torch.empty(m_local, n2, device=dev, dtype=torch.float16)
o_n = torch.empty_like(o_f)
arena.barrier()
skewed(lambda: ops.gemm_reducescatter_fused(x_ksh, w_ksh, staging, o_f, bias=b2, residual=res_loc))
ops.gemm_reducescatter_nccl(x_ksh, w_ksh, o_n, bias=b2, residual=res_loc)
ref2 = (x2.float() @ w2.float().t() + b2.float() + res_full.float())[R * m_local:(R + 1) * m_local]
tf1 = timed(lambda: ops.gemm_reducescatter_fused(x_ksh, w_ksh, staging, o_f, bias=b2, residual=res_loc))
tn = timed(lambda: ops.gemm_reducescatter_nccl(x_ksh, w_ksh, o_n, bias=b2, residual=res_loc))
err_v1 = relerr(o_f, ref2)
o_f.zero_()
x_var = x_ksh.clone()
rcalls = [0]


def rs2():
    rcalls[0] += 1
    torch.mul(x_ksh, float(rcalls[0] % 3 + 1), out=x_var)  # inputs differ call to call, the last result is checked
    ops.gemm_reducescatter(x_var, w_ksh, o_f, bias=b2, residual=res_loc)


skewed(rs2)
for _ in range(5):
    rs2()
torch.cuda.synchronize()
gang.barrier()
ref2_last = (x2.float() * float(rcalls[0] % 3 + 1) @ w2.float().t() + b2.float() + res_full.float())[R * m_local:(R + 1) * m_local]
st_loc = ops._stage[(m_local, n2)][0]
fl_rs = ops._flags_for("rs", W * (m_local // 128) * ((n2 + 127) // 128))
pf_rs = nat.make_peers(W, R, 0, [st_loc.peer_ptr(R)] * W, [fl_rs.peer_ptr(R)] * W)
scratch_out = torch.empty_like(o_f)
t_rs_local = timed(lambda: nat.igemm_reducescatter(x_ksh, w_ksh, pf_rs, scratch_out, m_local=m_local, ldc=n2, bias=b2,
                                                   residual=res_loc))
tf = timed(lambda: ops.gemm_reducescatter(x_ksh, w_ksh, scratch_out, bias=b2, residual=res_loc))
part = torch.empty(W * m_local, n2, device=dev, dtype=torch.float16)
t_gemm = timed(lambda: nat.linear(x_ksh, w_ksh, out=part))
report("gemm_reducescatter", relerr(o_f, ref2_last), relerr(o_n, ref2), tf, tn,
       dict(gflop=2.0 * W * m_local * ks * n2 / 1e9, link_mb=(W - 1) * m_local * n2 * 2 / 1e6, us_fused_gen1=tf1,
            rel_err_gen1=err_v1, us_local_gemm_only=t_gemm, us_fused_local_only=t_rs_local))

# ---------------------------------------------------------------- W9 attention with K/V gather (SDXL L1: S=4096, 10 heads, d=64)
b_, h_, d_, s_tot = 2, 10, 64, 4096
s_loc = s_tot // W
q_full = bcast(torch.randn(b_, s_tot, h_, d_, device=dev, dtype=torch.float16))
k_full = bcast(torch.randn(b_, s_tot, h_, d_, device=dev, dtype=torch.float16))
v_full = bcast(torch.randn(b_, s_tot, h_, d_, device=dev, dtype=torch.float16))
q_loc = q_full[:, R * s_loc:(R + 1) * s_loc].contiguous()
ksym, vsym = arena.alloc(b_, s_loc, h_, d_), arena.alloc(b_, s_loc, h_, d_)
ksym.local.copy_(k_full[:, R * s_loc:(R + 1) * s_loc])
vsym.local.copy_(v_full[:, R * s_loc:(R + 1) * s_loc])
a_f = torch.empty(b_, s_loc, h_ * d_, device=dev, dtype=torch.float16)
a_n = torch.empty_like(a_f)
arena.barrier()
skewed(lambda: ops.kv_gather_attention_fused(q_loc, ksym, vsym, a_f, batch=b_, heads=h_, head_dim=d_))
ops.kv_gather_attention_nccl(q_loc, ksym.local, vsym.local, a_n, batch=b_, heads=h_, head_dim=d_)
ref3 = F.scaled_dot_product_attention(q_loc.float().transpose(1, 2), k_full.float().transpose(1, 2),
                                      v_full.float().transpose(1, 2)).transpose(1, 2).reshape(b_, s_loc, h_ * d_)
tf = timed(lambda: ops.kv_gather_attention_fused(q_loc, ksym, vsym, a_f, batch=b_, heads=h_, head_dim=d_))
tn = timed(lambda: ops.kv_gather_attention_nccl(q_loc, ksym.local, vsym.local, a_n, batch=b_, heads=h_, head_dim=d_))
report("kv_gather_attention", relerr(a_f, ref3), relerr(a_n, ref3), tf, tn,
       dict(gflop=4.0 * b_ * h_ * s_loc * s_tot * d_ / 1e9, link_mb=(W - 1) * 2 * b_ * s_loc * h_ * d_ * 2 / 1e6))

# ---------------------------------------------------------------- W10 GroupNorm with stats all-reduce (SDXL L0: 128x128x320)
n_, hw_tot, c_, g_ = 2, 128 * 128, 320, 32
hw_loc = hw_tot // W
xg = bcast(torch.randn(n_, hw_tot, c_, device=dev, dtype=torch.float16) * 2 + 0.3)
gam = bcast(torch.randn(c_, device=dev, dtype=torch.float16))
bet = bcast(torch.randn(c_, device=dev, dtype=torch.float16))
xg_loc = xg[:, R * hw_loc:(R + 1) * hw_loc].contiguous()
g_f = torch.empty_like(xg_loc)
g_n = torch.empty_like(xg_loc)
arena.barrier()
skewed(lambda: ops.groupnorm_fused(xg_loc, gam, bet, g_f, n=n_, hw_local=hw_loc, groups=g_, eps=1e-5, silu=True))
ops.groupnorm_nccl(xg_loc, gam, bet, g_n, n=n_, hw_local=hw_loc, groups=g_, eps=1e-5, silu=True)
ref4 = F.silu(F.group_norm(xg.float().permute(0, 2, 1), g_, gam.float(), bet.float(), 1e-5)).permute(0, 2, 1)
ref4 = ref4[:, R * hw_loc:(R + 1) * hw_loc]
tf = timed(lambda: ops.groupnorm_fused(xg_loc, gam, bet, g_f, n=n_, hw_local=hw_loc, groups=g_, eps=1e-5, silu=True))
tn_eager = timed(lambda: ops.groupnorm_nccl(xg_loc, gam, bet, g_n, n=n_, hw_local=hw_loc, groups=g_, eps=1e-5, silu=True))
g_n.zero_()
ops.groupnorm_nccl_native(xg_loc, gam, bet, g_n, n=n_, hw_local=hw_loc, groups=g_, eps=1e-5, silu=True)
tn = timed(lambda: ops.groupnorm_nccl_native(xg_loc, gam, bet, g_n, n=n_, hw_local=hw_loc, groups=g_, eps=1e-5, silu=True))
report("groupnorm_allreduce", (g_f.float() - ref4).abs().max().item(), (g_n.float() - ref4).abs().max().item(), tf, tn,
       dict(mbytes=3 * xg_loc.numel() * 2 / 1e6, link_bytes=(W - 1) * n_ * g_ * 8, us_nccl_eager_torch=tn_eager,
            note="us_nccl = native partial-sum kernel + NCCL all-reduce + native apply kernel"))

gang.barrier()
if R == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/dist_check_w{W}{'_skew' if args.skew_ms else ''}.json", "w") as fh:
        json.dump(results, fh, indent=1)
    ok = all(v["rel_err_fused"] < 2e-2 for v in results.values())
    print("DIST_CHECK", "PASS" if ok else "FAIL", flush=True)
dist.destroy_process_group()
