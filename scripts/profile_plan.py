"""Per-op device time of the recorded UNet step plan (each op timed on its own, warm)."""
import os
import sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
import torch

from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

arch = sys.argv[1] if len(sys.argv) > 1 else "sd15"
size = int(sys.argv[2]) if len(sys.argv) > 2 else 512
pipe = StableDiffusionPipeline.from_architecture(arch)
pipe.backend = "sm100"
pipe.to("cuda:0", torch.float16)
pipe.set_scheduler("dpmpp_2m", True)
g = torch.Generator(device="cuda:0").manual_seed(0)
pipe(prompt="p", height=size, width=size, num_inference_steps=2, generator=g, output_type="latent")
eng = pipe._denoiser()
sg = next(iter(eng._graphs.values()))
res = sg.rec.profile()
total = sum(t for _, t in res)
print(f"plan: {len(res)} ops, sum of isolated op times {total/1e3:.2f} ms")
agg = defaultdict(lambda: [0, 0.0])
for label, t in res:
    agg[label][0] += 1
    agg[label][1] += t
for label, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print(f"{t:8.1f} us {100*t/total:5.1f}%  n={c:3d} avg={t/c:7.1f}  {label[:150]}")
