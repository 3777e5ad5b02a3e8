"""Runs the recorded zeroscope step plan op by op (6 frames, one GPU) and reports the first op after which one of its
tensor arguments holds a non-finite value."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
os.environ["SDAAS_CUDA_GRAPH"] = "0"
import torch
from chiaswarm_b200.video.pipeline_t2v import TextToVideoPipeline

dev = "cuda:0"
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 6
pipe = TextToVideoPipeline.from_architecture("zeroscope").to(dev, torch.float16)
pipe.set_scheduler("dpmpp_2m", True)
pipe.backend = "sm100"
g = torch.Generator(device=dev).manual_seed(3)
pipe(prompt="six frames", height=320, width=576, num_frames=frames, num_inference_steps=1, guidance_scale=1.0, generator=g,
     output_type="latent")
eng = pipe._sm100()
sg = next(iter(eng._graphs.values()))
g = torch.Generator(device=dev).manual_seed(3)
# reload the job inputs, then walk the plan
from chiaswarm_b200.ops import schedulers
plan = schedulers.make_plan("dpmpp_2m", 1, pipe.scheduler_config, True)
x = torch.randn(frames, 4, 40, 72, generator=g, device=dev, dtype=torch.float16).float() * plan.init_noise_sigma
eng._load_job(sg, x, plan, None, None, None, None, 1.0, 0.0)
torch.cuda.synchronize()
assert torch.isfinite(sg.model_in.float()).all()


def label(fn, args, kw):
    shapes = [tuple(a.shape) for a in args if isinstance(a, torch.Tensor)][:4]
    extra = {k: v for k, v in kw.items() if not isinstance(v, torch.Tensor) and v is not None}
    return f"{getattr(fn, '__name__', fn)} {shapes} {extra}"


def finite(args, kw):
    return all(bool(torch.isfinite(a.float()).all()) for a in list(args) + list(kw.values())
               if isinstance(a, torch.Tensor) and a.is_floating_point())


ops = sg.rec.ops


def run(j, k):
    """ops [0, j) with a sync after each, ops [j, k) back to back (PDL may overlap them); True if all finite after."""
    eng._load_job(sg, x, plan, None, None, None, None, 1.0, 0.0)
    torch.cuda.synchronize()
    for fn, args, kw in ops[:j]:
        fn(*args, **kw)
        torch.cuda.synchronize()
    for fn, args, kw in ops[j:k]:
        fn(*args, **kw)
    torch.cuda.synchronize()
    return all(finite(a, kw) for _, a, kw in ops[max(0, j - 1):k])


n = len(ops)
_run1 = run


def run(j, k, tries=4):  # the race is intermittent: a window is bad if any try goes non-finite
    return all(_run1(j, k) for _ in range(tries))


first = run(0, n)
print("whole plan unsynced finite:", first, flush=True)
lo, hi = 0, n  # smallest k with run(0, k) non-finite
if not first:
    while hi - lo > 1:
        mid = (lo + hi) // 2
        if run(0, mid):
            lo = mid
        else:
            hi = mid
    k = hi
    lo2, hi2 = 0, k  # largest j with run(j, k) non-finite
    while hi2 - lo2 > 1:
        mid = (lo2 + hi2) // 2
        if run(mid, k):
            hi2 = mid
        else:
            lo2 = mid
    j = lo2
    print(f"smallest failing window: ops [{j}, {k})", flush=True)
    for t in range(max(0, j - 1), min(n, k + 1)):
        print("   ", t, label(*ops[t]), flush=True)

# which tensors are non-finite after the failing window?
for attempt in range(6):
    if not _run1(595, 613):
        for t in range(590, 613):
            fn, args, kw = ops[t]
            for ai, a in enumerate(list(args) + list(kw.values())):
                if isinstance(a, torch.Tensor) and a.is_floating_point() and not torch.isfinite(a.float()).all():
                    bad_rows = (~torch.isfinite(a.float())).reshape(-1, a.shape[-1]).any(dim=1).nonzero().flatten()
                    print(f"  op {t} arg {ai} shape {tuple(a.shape)} ptr {a.data_ptr():x}: {bad_rows.numel()} bad rows, first {bad_rows[:6].tolist()} last {bad_rows[-3:].tolist()}", flush=True)
        break
print("ptrs:", {t: [hex(a.data_ptr()) for a in ops[t][1] if isinstance(a, torch.Tensor)][:4] for t in (595, 596, 609, 610, 611, 612)})
