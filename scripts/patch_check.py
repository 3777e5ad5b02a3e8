"""Patch-parallel engine against the single-GPU engine (same seed, same scheduler), plus latency.
Run under torchrun, one process per GPU:
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/patch_check.py \
      --arch sd15 --size 512 --steps 4 --guidance 1.0
guidance 1.0 -> one branch, every rank a strip (P = world); guidance > 1 -> two CFG branches on
disjoint halves of the gang (P = world / 2).  Device-timed (CUDA events), max over ranks."""
import argparse
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--arch", default="sd15")
ap.add_argument("--size", type=int, default=512)
ap.add_argument("--steps", type=int, default=4)
ap.add_argument("--guidance", type=float, default=1.0)
ap.add_argument("--runs", type=int, default=0, help="timed full-pipeline runs (0 = numerics only)")
ap.add_argument("--tol", type=float, default=3e-2)
ap.add_argument("--skip-single", action="store_true")
ap.add_argument("--nccl-baseline", action="store_true",
                help="also time one UNet forward of the NCCL-only patch-parallel path (parallel/patch.py, eager torch)")
args = ap.parse_args()

from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline
from chiaswarm_b200.parallel import runtime

gang = runtime.init()
dev = gang.device
pipe = StableDiffusionPipeline.from_architecture(args.arch)
pipe.backend = "sm100"
pipe.to(dev, torch.float16)
pipe.set_scheduler("dpmpp_2m", True)


def call(output_type="latent"):
    g = torch.Generator(device=dev).manual_seed(1234)
    return pipe(prompt="patch parallel probe", height=args.size, width=args.size, num_inference_steps=args.steps,
                guidance_scale=args.guidance, generator=g, output_type=output_type)


def timed(runs):
    ts = []
    for i in range(runs + 2):
        torch.cuda.synchronize()
        gang.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        call("device")
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(gang.all_reduce_max(e0.elapsed_time(e1)))
    return statistics.median(ts)


res = {"arch": args.arch, "size": args.size, "steps": args.steps, "guidance": args.guidance, "world": gang.world}
ref = None
if not args.skip_single:
    ref = call().latents.float()
    if args.runs:
        res["single_gpu_ms"] = timed(args.runs)
pipe.enable_patch_parallel(gang)
out = call().latents.float()
eng = pipe._denoiser()
sg = next(v for k, v in eng._graphs.items() if k[0] == "patch")
res.update(groups=sg.ctx.G, strips=sg.ctx.P, launches_per_step=sg.launches_per_step,
           arena_mb=round(sg.ctx.arena._cursor / 2 ** 20))
ok = bool(torch.isfinite(out).all())
if ref is not None:
    err = ((out - ref).abs().max() / ref.abs().max().clamp_min(1e-3)).item()
    res["rel_err_vs_single"] = gang.all_reduce_max(err)
    ok = ok and res["rel_err_vs_single"] < args.tol
if args.runs:
    res["patch_ms"] = timed(args.runs)
    # one denoising step of the captured plan, all ranks in lockstep
    torch.cuda.synchronize()
    gang.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        sg.replay()
    e1.record()
    torch.cuda.synchronize()
    res["patch_step_ms"] = gang.all_reduce_max(e0.elapsed_time(e1) / 10)
if args.nccl_baseline:
    from chiaswarm_b200.parallel.patch import PatchParallelUNet

    n_cond = 2 if args.guidance > 1.0 else 1
    lat = args.size // 8
    x = torch.randn(n_cond, 4, lat, lat, device=dev, dtype=torch.float16)
    ctx_dim = pipe.unet.config["cross_attention_dim"]
    ctx = torch.randn(n_cond, 77, ctx_dim, device=dev, dtype=torch.float16)
    added = None
    if pipe.unet.config.get("addition_embed_type") == "text_time":
        added = {"text_embeds": torch.randn(n_cond, 1280, device=dev, dtype=torch.float16),
                 "time_ids": torch.tensor([[args.size, args.size, 0, 0, args.size, args.size]] * n_cond, device=dev,
                                          dtype=torch.float32)}
    ppu = PatchParallelUNet(pipe.unet, gang)
    for _ in range(2):
        ppu(x, 500.0, ctx, added)
    torch.cuda.synchronize()
    gang.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ppu(x, 500.0, ctx, added)
    e1.record()
    torch.cuda.synchronize()
    res["nccl_patch_step_ms"] = gang.all_reduce_max(e0.elapsed_time(e1) / 5)
ok = gang.all_reduce_max(0.0 if ok else 1.0) == 0.0
if gang.rank == 0:
    print(json.dumps(res))
    print("PATCH_CHECK PASS" if ok else "PATCH_CHECK FAIL", flush=True)
torch.cuda.synchronize()
gang.barrier()
