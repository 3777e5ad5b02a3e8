"""Per-stage device time of one pipeline call (SDAAS_TRACE stage timer): text encode / denoise / VAE decode.
  python scripts/stage_times.py --arch sdxl --size 1024 --steps 30 --guidance 7.5"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
os.environ["SDAAS_TRACE"] = "1"
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--arch", default="sdxl")
ap.add_argument("--size", type=int, default=1024)
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--guidance", type=float, default=7.5)
args = ap.parse_args()
from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

pipe = StableDiffusionPipeline.from_architecture(args.arch)
pipe.backend = "sm100"
pipe.to("cuda:0", torch.float16)
pipe.set_scheduler("dpmpp_2m", True)
for i in range(3):
    g = torch.Generator(device="cuda:0").manual_seed(0)
    out = pipe(prompt="stage probe", height=args.size, width=args.size, num_inference_steps=args.steps,
               guidance_scale=args.guidance, generator=g, output_type="device")
print(json.dumps({"arch": args.arch, "size": args.size, "steps": args.steps, "guidance": args.guidance,
                  "stages_ms": out.timings}))
