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
guidance = float(sys.argv[3]) if len(sys.argv) > 3 else 7.5
pipe = StableDiffusionPipeline.from_architecture(arch)
pipe.backend = "sm100"
pipe.to("cuda:0", torch.float16)
pipe.set_scheduler("dpmpp_2m", True)
g = torch.Generator(device="cuda:0").manual_seed(0)
pipe(prompt="p", height=size, width=size, num_inference_steps=2, generator=g, output_type="latent",
     guidance_scale=guidance)
eng = pipe._denoiser()
sg = next(iter(eng._graphs.values()))
res = sg.rec.profile()
total = sum(t for _, t in res)
print(f"plan: {len(res)} ops, sum of isolated op times {total/1e3:.2f} ms")


def flops(fn, args, kw):
    name = getattr(fn, "__name__", "")
    if name == "igemm":
        rows = kw["n"] * kw["h"] * kw["w"]
        return 2.0 * rows * args[1].numel()
    if name == "flash_attn":
        return 4.0 * kw["batch"] * kw["heads"] * kw["seq_q"] * kw["seq_kv"] * kw["head_dim"]
    return 0.0


agg = defaultdict(lambda: [0, 0.0, 0.0])
kinds = defaultdict(lambda: [0.0, 0.0])
for (label, t), (fn, args, kw) in zip(res, sg.rec.ops):
    f = flops(fn, args, kw)
    agg[label][0] += 1
    agg[label][1] += t
    agg[label][2] += f
    k = kinds[getattr(fn, "__name__", "op")]
    k[0] += t
    k[1] += f
for name, (t, f) in sorted(kinds.items(), key=lambda kv: -kv[1][0]):
    print(f"{t:9.1f} us {100*t/total:5.1f}%  {f/1e12:7.3f} TFLOP  {f/max(t,1e-9)/1e6:7.1f} TFLOP/s  {name}")
for label, (c, t, f) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:60]:
    print(f"{t:8.1f} us {100*t/total:5.1f}%  n={c:3d} avg={t/c:7.1f} {f/max(t,1e-9)/1e6:6.0f} TF/s  {label[:150]}")
