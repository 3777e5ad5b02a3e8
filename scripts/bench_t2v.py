"""Text-to-video clip latency (BASELINE.json config 5: zeroscope-shaped UNet3D, 576x320, 24 frames).
Device-timed (CUDA events); random-init weights of the named architecture, synthetic prompt.
  python scripts/bench_t2v.py --steps 25 --runs 3 [--impl torch]"""
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
ap.add_argument("--profile", action="store_true", help="print the per-op device-time breakdown of one step")
args = ap.parse_args()

from chiaswarm_b200.video.pipeline_t2v import TextToVideoPipeline

dev = "cuda:0"
pipe = TextToVideoPipeline.from_architecture(args.arch).to(dev, torch.float16)
pipe.backend = "sm100" if args.impl == "ours" else "torch"
pipe.set_scheduler("dpmpp_2m", True)
gen = torch.Generator(device=dev).manual_seed(0)
kw = dict(prompt="latency probe", height=args.height, width=args.width, num_frames=args.frames,
          num_inference_steps=args.steps, generator=gen)
times = []
for i in range(args.runs + 1):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    out = pipe(**kw)
    e1.record()
    torch.cuda.synchronize()
    if i >= 1:
        times.append(e0.elapsed_time(e1))
res = {"arch": args.arch, "impl": args.impl, "width": args.width, "height": args.height, "frames": args.frames,
       "steps": args.steps, "p50_ms": statistics.median(times), "min_ms": min(times), "n_frames_out": len(out.frames),
       "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}
if args.impl == "ours":
    eng = pipe._sm100()
    sg = next(iter(eng._graphs.values()))
    res["launches_per_step"] = sg.launches_per_step
    if args.profile:
        from collections import defaultdict

        agg = defaultdict(float)
        for label, us in sg.rec.profile(3):
            agg[label.split(" ")[0]] += us
        res["step_us_by_op"] = {k: round(v) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])}
print(json.dumps(res))
