"""Single-image latency of a named architecture (BASELINE.json config 3: SDXL base 1024x1024;
config 4 shape: SD1.5 + ControlNet 768x768 batch 8).  Device-timed (CUDA events), p50 over runs.
  python scripts/bench_latency.py --arch sdxl --size 1024 --steps 30 --runs 5 [--impl torch]
Under torchrun with --patch the UNet runs patch-parallel through the NCCL baseline path."""
import argparse
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--arch", default="sdxl")
ap.add_argument("--size", type=int, default=1024)
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--runs", type=int, default=5)
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--impl", default="ours", choices=["ours", "torch"])
ap.add_argument("--controlnet", action="store_true")
args = ap.parse_args()

from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline, _init_module

dev = "cuda:0"
controlnet = None
if args.controlnet:
    from chiaswarm_b200.models import configs
    from chiaswarm_b200.models.controlnet import ControlNetModel

    controlnet = _init_module(ControlNetModel(**configs.get_controlnet("sd15")), 9)
pipe = StableDiffusionPipeline.from_architecture(args.arch, controlnet=controlnet,
                                                 family="sd_controlnet" if controlnet is not None else None)
pipe.backend = "sm100" if args.impl == "ours" else "torch"
pipe.to(dev, torch.float16)
pipe.set_scheduler("dpmpp_2m", True)
gen = torch.Generator(device=dev).manual_seed(0)
kw = dict(prompt="latency probe", height=args.size, width=args.size, num_inference_steps=args.steps,
          num_images_per_prompt=args.batch, generator=gen, output_type="device")
if controlnet is not None:
    from PIL import Image

    kw["image"] = Image.new("RGB", (args.size, args.size), (128, 128, 128))
times = []
for i in range(args.runs + 2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    pipe(**kw)
    e1.record()
    torch.cuda.synchronize()
    if i >= 2:
        times.append(e0.elapsed_time(e1))
print(json.dumps({"arch": args.arch, "impl": args.impl, "size": args.size, "steps": args.steps, "batch": args.batch,
                  "controlnet": bool(controlnet), "p50_ms": statistics.median(times), "min_ms": min(times),
                  "images_per_s": args.batch / (statistics.median(times) / 1e3)}))
