"""Warm per-kernel GPU time of the denoising loop via torch.profiler (CUPTI), plan executed eagerly."""
import os
import sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
os.environ["SDAAS_CUDA_GRAPH"] = "0"
import torch
from torch.profiler import ProfilerActivity, profile

from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

arch = sys.argv[1] if len(sys.argv) > 1 else "sd15"
size = int(sys.argv[2]) if len(sys.argv) > 2 else 512
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
pipe = StableDiffusionPipeline.from_architecture(arch)
pipe.backend = "sm100"
pipe.to("cuda:0", torch.float16)
pipe.set_scheduler("dpmpp_2m", True)
g = torch.Generator(device="cuda:0").manual_seed(0)
kw = dict(prompt="profile", height=size, width=size, generator=g, output_type="device")
pipe(num_inference_steps=steps, **kw)  # warm-up, builds plans
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    pipe(num_inference_steps=steps, **kw)
    torch.cuda.synchronize()
import json
import tempfile

trace = tempfile.mktemp(suffix=".json")
prof.export_chrome_trace(trace)
by_grid = defaultdict(lambda: [0, 0.0])
for ev in json.load(open(trace))["traceEvents"]:
    if ev.get("cat") != "kernel":
        continue
    nm = ev["name"].replace("void ", "").replace("(anonymous namespace)::", "").split("(")[0]
    if not (nm.startswith("k_igemm") or nm.startswith("k_flash") or nm.startswith("k_gn")):
        continue
    key = (nm, str(ev["args"].get("grid")))
    by_grid[key][0] += 1
    by_grid[key][1] += ev["dur"]
print("---- by grid (us total over the run)")
for key, (c, v) in sorted(by_grid.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{v:9.1f} us  n={c:5d} avg={v/c:7.1f} us  {key}")

agg = defaultdict(lambda: [0, 0.0])
for ev in prof.events():
    if ev.device_type.name != "CUDA":
        continue
    name = ev.name.replace("void ", "").replace("(anonymous namespace)::", "").replace("<unnamed>::", "").split("(")[0]
    agg[name][0] += 1
    agg[name][1] += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
total = sum(v for _, v in agg.values())
print(f"total kernel time {total/1e3:.2f} ms for {steps} steps + text encode + VAE decode")
for name, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"{v/1e3:9.3f} ms {100*v/total:5.1f}%  n={c:5d} avg={v/c:7.1f} us  {name[:80]}")
