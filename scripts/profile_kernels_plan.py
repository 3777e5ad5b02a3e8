"""Per-op DEVICE time of one recorded UNet step: the plan runs once eagerly under CUPTI (torch.profiler,
SW_PDL=0 so kernels do not overlap), kernel records are matched to plan ops by launch order, then grouped
by op + shape with achieved TFLOP/s.  Caches are in their real per-step state (weights stream from HBM).
  SW_PDL=0 python scripts/profile_kernels_plan.py sdxl 1024 7.5
  SW_PDL=0 python scripts/profile_kernels_plan.py sd15 512 7.5 vae     (the VAE decode plan instead of the UNet step)"""
import os
import sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
os.environ.setdefault("SW_PDL", "0")
import torch
from torch.autograd import DeviceType
from torch.profiler import ProfilerActivity, profile

from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

arch = sys.argv[1] if len(sys.argv) > 1 else "sd15"
size = int(sys.argv[2]) if len(sys.argv) > 2 else 512
guidance = float(sys.argv[3]) if len(sys.argv) > 3 else 7.5
which = sys.argv[4] if len(sys.argv) > 4 else "unet"
pipe = StableDiffusionPipeline.from_architecture(arch)
pipe.backend = "sm100"
pipe.to("cuda:0", torch.float16)
pipe.set_scheduler("dpmpp_2m", True)
g = torch.Generator(device="cuda:0").manual_seed(0)
pipe(prompt="p", height=size, width=size, num_inference_steps=2, generator=g,
     output_type="device" if which == "vae" else "latent", guidance_scale=guidance)
eng = pipe._denoiser()
if which == "vae":
    rec = next(iter(eng._vae_graphs.values()))[2]
else:
    sg = next(iter(eng._graphs.values()))
    rec = sg.rec
rec.run()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    rec.run()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == DeviceType.CUDA and not e.name.startswith(("Memset", "Memcpy"))]
evs.sort(key=lambda e: e.time_range.start)
from chiaswarm_b200.ops import native as _nat


def nkernels(fn, args, kw):
    name = getattr(fn, "__name__", "")
    if name == "groupnorm":  # single-launch cluster kernel or statistics + apply, by shape
        c = args[0].shape[-1] + (kw["x2"].shape[-1] if kw.get("x2") is not None else 0)
        return _nat.lib().sw_groupnorm_kernel_count(kw["n"], kw["hw"], c, kw["groups"])
    return 2 if name == "groupnorm_dist" else 1


need = sum(nkernels(*op) for op in rec.ops)
print(f"plan: {len(rec.ops)} ops, {need} kernels expected, {len(evs)} kernel records")
if need != len(evs):
    names = defaultdict(int)
    for e in evs:
        names[e.name[:60]] += 1
    print("kernel/ops mismatch:", dict(names))
    sys.exit(1)


def flops(fn, args, kw):
    name = getattr(fn, "__name__", "")
    if name == "igemm":
        return 2.0 * kw["n"] * kw["h"] * kw["w"] * args[1].numel()
    if name == "flash_attn":
        return 4.0 * kw["batch"] * kw["heads"] * kw["seq_q"] * kw["seq_kv"] * kw["head_dim"]
    return 0.0


def label(fn, args, kw):
    shapes = [tuple(a.shape) for a in args if isinstance(a, torch.Tensor)][:3]
    extra = {k: kw[k] for k in ("n", "h", "w", "kh", "act", "heads", "seq_q", "seq_kv", "head_dim", "hw") if k in kw}
    return f"{getattr(fn, '__name__', 'op')} {shapes} {extra}"


agg = defaultdict(lambda: [0, 0.0, 0.0])
kinds = defaultdict(lambda: [0.0, 0.0])
i = 0
total = 0.0
for fn, args, kw in rec.ops:
    k = nkernels(fn, args, kw)
    t = sum(e.time_range.elapsed_us() for e in evs[i:i + k])
    i += k
    f = flops(fn, args, kw)
    a = agg[label(fn, args, kw)]
    a[0] += 1
    a[1] += t
    a[2] += f
    kk = kinds[getattr(fn, "__name__", "op")]
    kk[0] += t
    kk[1] += f
    total += t
print(f"sum of kernel device times {total / 1e3:.2f} ms")
for name, (t, f) in sorted(kinds.items(), key=lambda kv: -kv[1][0]):
    print(f"{t:9.1f} us {100 * t / total:5.1f}%  {f / 1e12:7.3f} TFLOP  {f / max(t, 1e-9) / 1e6:7.1f} TFLOP/s  {name}")
for lab, (c, t, f) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:70]:
    print(f"{t:8.1f} us {100 * t / total:5.1f}%  n={c:3d} avg={t / c:7.1f} {f / max(t, 1e-9) / 1e6:6.0f} TF/s  {lab[:170]}")
