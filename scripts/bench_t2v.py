"""Text-to-video clip latency (BASELINE.json config 5: zeroscope-shaped UNet3D, 576x320, 24 frames).
Device-timed (CUDA events, max over ranks); random-init weights of the named architecture, synthetic prompt.

  python scripts/bench_t2v.py --steps 25 --runs 3 [--impl torch]
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/bench_t2v.py --steps 25

Under torchrun the clip is frame-split over all ranks (engine/patch_engine.py::_FrameStepGraph: guidance branches on
disjoint halves of the gang, frames dealt out inside a half, VAE decode scattered by frame) and compared with the same
clip on one GPU: latents must agree (`rel_err_vs_single`), and both latencies are reported."""
import argparse
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--arch", default="zeroscope")
ap.add_argument("--width", type=int, default=576)
ap.add_argument("--height", type=int, default=320)
ap.add_argument("--frames", type=int, default=24)
ap.add_argument("--steps", type=int, default=25)
ap.add_argument("--runs", type=int, default=3)
ap.add_argument("--impl", default="ours", choices=["ours", "torch"])
ap.add_argument("--tol", type=float, default=3e-2)
ap.add_argument("--guidance", type=float, default=9.0, help="<= 1.0: one branch, so every rank takes a share of the frames")
ap.add_argument("--profile", action="store_true", help="print the per-op device-time breakdown of one step")
args = ap.parse_args()

from chiaswarm_b200.parallel import runtime
from chiaswarm_b200.video.pipeline_t2v import TextToVideoPipeline

gang = runtime.init()
dev = gang.device
pipe = TextToVideoPipeline.from_architecture(args.arch).to(dev, torch.float16)
pipe.backend = "sm100" if args.impl == "ours" else "torch"
pipe.set_scheduler("dpmpp_2m", True)


def call(g=None, output_type="np"):
    gen = torch.Generator(device=dev).manual_seed(0)
    return pipe(prompt="latency probe", height=args.height, width=args.width, num_frames=args.frames,
                num_inference_steps=args.steps, generator=gen, gang=g, output_type=output_type,
                guidance_scale=args.guidance)


def timed(g=None):
    times, out = [], None
    for i in range(args.runs + 1):
        torch.cuda.synchronize()
        gang.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = call(g)
        e1.record()
        torch.cuda.synchronize()
        if i >= 1:
            times.append(gang.all_reduce_max(e0.elapsed_time(e1)))
    return statistics.median(times), min(times), out


p50, best, out = timed()
res = {"arch": args.arch, "impl": args.impl, "width": args.width, "height": args.height, "frames": args.frames,
       "steps": args.steps, "guidance": args.guidance, "world": gang.world, "p50_ms": p50, "min_ms": best, "n_frames_out": len(out.frames)}
ok = True
if args.impl == "ours":
    eng = pipe._sm100()
    sg = next(iter(eng._graphs.values()))
    res["launches_per_step"] = sg.launches_per_step
    if args.profile and gang.rank == 0:
        from collections import defaultdict

        agg = defaultdict(float)
        for label, us in sg.rec.profile(3):
            agg[label.split(" ")[0]] += us
        res["step_us_by_op"] = {k: round(v) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])}
    if gang.world > 1:
        ref = call(None, "latent").latents.float()
        got = call(gang, "latent").latents.float()
        err = ((got - ref).abs().max() / ref.abs().max().clamp_min(1e-3)).item()
        res["rel_err_vs_single"] = gang.all_reduce_max(err if err == err else 1e9)
        ok = res["rel_err_vs_single"] < args.tol
        p50g, bestg, outg = timed(gang)
        psg = next(v for k, v in pipe._sm100(gang)._graphs.items() if k[0] == "patch")
        res.update(gang_p50_ms=p50g, gang_min_ms=bestg, speedup=p50 / p50g, gang_frames_out=len(outg.frames),
                   guidance_groups=psg.ctx.G, frame_shards=psg.ctx.P, gang_launches_per_step=psg.launches_per_step,
                   arena_mb=round((psg.ctx.arena._cursor - psg.ctx.arena._lo) / 2 ** 20))
        ok = ok and len(outg.frames) == args.frames
res["peak_mem_gb"] = torch.cuda.max_memory_allocated() / 2 ** 30
ok = gang.all_reduce_max(0.0 if ok else 1.0) == 0.0
if gang.rank == 0:
    print(json.dumps(res))
    print("T2V_CHECK PASS" if ok else "T2V_CHECK FAIL", flush=True)
torch.cuda.synchronize()
gang.barrier()
