"""Builds `chiaswarm_b200/ops/igemm_tuning.json`: for every distinct implicit-GEMM shape in the step plans
of the supported pipelines, time the kernel's tile width / CTA-pair / split-K choices (CUDA graph of 8
launches of the real plan op, device-timed, best of 3 replays) and keep the fastest.
  python scripts/tune_igemm.py [--out gpurun_out/igemm_tuning.json]
Run on the target GPU; copy the result over chiaswarm_b200/ops/igemm_tuning.json."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SDAAS_RANDOM_INIT"] = "1"
os.environ["SW_IGEMM_TUNING"] = "0"  # measure against the built-in heuristics, not a previous table
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="gpurun_out/igemm_tuning.json")
ap.add_argument("--quick", action="store_true")
ap.add_argument("--only", default="", help="comma-separated plan names to (re)tune")
ap.add_argument("--strips", default="", help="also tune the 1/P-row strips of the single-image plans (patch-parallel "
                "ranks run exactly these shapes), e.g. 2,4")
ap.add_argument("--merge", default="", help="existing table to extend (its entries are kept unless re-measured)")
args = ap.parse_args()

from chiaswarm_b200.ops import native as nat
from chiaswarm_b200.ops import tuning

dev = "cuda:0"
REPS = 8


_flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)


def _replay_time(build):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REPS):
            build()
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / REPS * 1e3)
    return best


def graph_time(fn, acts=()):
    """us per launch in the state the op sees inside a step: weights cold (L2 flushed before every
    launch), activations warm (touched after the flush, as if the previous layer had just written them).
    The flush + touch cost is measured separately and subtracted."""
    fn()
    torch.cuda.synchronize()

    def prep():
        _flush.zero_()
        for a in acts:
            a.mul_(1)

    base = _replay_time(prep)
    return _replay_time(lambda: (prep(), fn())) - base


def plans():
    """(name, recorder) for every pipeline shape we serve by default."""
    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline, _init_module

    def sd(arch, size, guidance, batch=1, controlnet=False):
        cn = None
        if controlnet:
            from chiaswarm_b200.models import configs
            from chiaswarm_b200.models.controlnet import ControlNetModel

            cn = _init_module(ControlNetModel(**configs.get_controlnet("sd15")), 9)
        pipe = StableDiffusionPipeline.from_architecture(arch, controlnet=cn, family="sd_controlnet" if cn else None)
        pipe.backend = "sm100"
        pipe.to(dev, torch.float16)
        pipe.set_scheduler("dpmpp_2m", True)
        kw = dict(prompt="t", height=size, width=size, num_inference_steps=1, guidance_scale=guidance,
                  num_images_per_prompt=batch, generator=torch.Generator(device=dev).manual_seed(0), output_type="device")
        if cn is not None:
            from PIL import Image

            kw["image"] = Image.new("RGB", (size, size), (128, 128, 128))
        pipe(**kw)
        eng = pipe._denoiser()
        recs = [sg.rec for sg in eng._graphs.values()] + [e[2] for e in eng._vae_graphs.values()]
        return recs, pipe

    def t2v():
        from chiaswarm_b200.video.pipeline_t2v import TextToVideoPipeline

        pipe = TextToVideoPipeline.from_architecture("zeroscope").to(dev, torch.float16)
        pipe.backend = "sm100"
        pipe.set_scheduler("dpmpp_2m", True)
        pipe(prompt="t", height=320, width=576, num_frames=24, num_inference_steps=1,
             generator=torch.Generator(device=dev).manual_seed(0))
        eng = pipe._sm100()
        return [sg.rec for sg in eng._graphs.values()] + [e[2] for e in eng._vae_graphs.values()], pipe

    only = [x for x in args.only.split(",") if x]
    todo = [("sd15-512-cfg", lambda: sd("sd15", 512, 7.5))]
    if not args.quick:
        todo += [("sd15-512-nocfg", lambda: sd("sd15", 512, 1.0)),
                 ("sd15-768-cn-b8", lambda: sd("sd15", 768, 7.5, batch=8, controlnet=True)),
                 ("sdxl-1024-cfg", lambda: sd("sdxl", 1024, 7.5)),
                 ("sdxl-1024-nocfg", lambda: sd("sdxl", 1024, 1.0)),
                 ("sd21-768-cfg", lambda: sd("sd21", 768, 7.5)),
                 ("zeroscope-576x320x24", t2v)]
    for name, make in todo:
        if not only or name in only:
            yield name, make()


# (tile_n, cta_pair, splitk_min_kb): 64 k-blocks per split = effectively no split-K for the UNet shapes
# 4th field: 0 = kernel's default, 1 = one tile per CTA, 2 = persistent CTAs (double-buffered accumulators)
# 5th field: 2 = split K over a cluster of two CTAs (DSMEM exchange) when the launcher finds it eligible
CANDS = [(0, 0, 0, 0, 0), (64, 1, 64, 1, 0), (128, 1, 64, 1, 0), (256, 1, 64, 1, 0), (128, 2, 64, 1, 0), (256, 2, 64, 1, 0),
         (64, 1, 8, 1, 0), (128, 1, 8, 1, 0), (256, 1, 8, 1, 0), (128, 2, 8, 1, 0), (256, 2, 8, 1, 0), (64, 1, 16, 1, 0),
         (128, 1, 16, 1, 0), (256, 1, 16, 1, 0), (64, 1, 64, 2, 0), (128, 1, 64, 2, 0), (256, 1, 64, 2, 0),
         (64, 1, 64, 1, 2), (128, 1, 64, 1, 2), (256, 1, 64, 1, 2)]
configs, report = {}, []
kept = {}
if args.merge and os.path.isfile(args.merge):
    kept = {k: tuple(v) for k, v in json.load(open(args.merge)).get("configs", {}).items()}
def strip_of(a, kw, P):
    """The same op on the first 1/P of the rows (what a patch-parallel rank with P strips computes)."""
    if kw["n"] != 1:
        return None
    a1, wgt, out = a[0], a[1], a[2]
    if kw["h"] > 1:
        if kw["h"] % P:
            return None
        h, w = kw["h"] // P, kw["w"]
        rows = h * w
        k2 = dict(kw, h=h)
        a1s = a1.reshape(kw["h"], w, -1)[:h].reshape(1, h, w, -1)
        if kw.get("a2") is not None:
            k2["a2"] = kw["a2"].reshape(kw["h"], w, -1)[:h].reshape(1, h, w, -1)
    else:
        if kw["w"] % P or (kw["w"] // P) < 8:
            return None
        rows = kw["w"] // P
        k2 = dict(kw, w=rows)
        a1s = a1.reshape(kw["w"], -1)[:rows]
    if kw.get("residual") is not None:
        k2["residual"] = kw["residual"].reshape(-1, kw["residual"].shape[-1])[:rows]
    return (a1s, wgt, out.reshape(-1, out.shape[-1])[:rows]) + tuple(a[3:]), k2


strips = [int(x) for x in args.strips.split(",") if x]
for name, (recs, keep) in plans():
    seen = 0
    ops = []
    for rec in recs:
        for fn, a, kw in rec.ops:
            if getattr(fn, "__name__", "") != "igemm" or kw.get("tile_n"):
                continue
            ops.append((fn, a, kw))
            if "nocfg" in name:
                for P in strips:
                    so = strip_of(a, kw, P)
                    if so is not None:
                        ops.append((fn, so[0], so[1]))
    for rec in [None]:
        for fn, a, kw in ops:
            a1, wgt = a[0], a[1]
            c1 = a1.shape[-1]
            c2 = 0 if kw.get("a2") is None else kw["a2"].shape[-1]
            act = kw.get("act", 0)
            key = tuning.key_of(kw["n"], kw["h"], kw["w"], kw.get("kh", 1), kw.get("kw", 1), c1, c2, wgt.shape[0], act,
                                kw.get("residual") is not None)
            if key in configs:
                continue
            seen += 1
            times = []
            for t, pr, mk, ps, ck in CANDS:
                if act == nat.ACT_GEGLU and t not in (0, 128):
                    continue
                if t > max(64, (wgt.shape[0] + 31) // 32 * 32):
                    continue
                k2 = dict(kw, tile_n=t, cta_pair=pr, splitk_min_kb=mk, persist=ps, cluster_k=ck)
                try:
                    acts = [x for x in (a1, kw.get("a2"), kw.get("residual")) if x is not None]
                    times.append((graph_time(lambda: fn(*a, **k2), acts), (t, pr, mk, ps, ck)))
                except Exception:
                    pass
            times.sort()
            base = [x for x, c in times if c == (0, 0, 0, 0, 0)]
            configs[key] = times[0][1]
            report.append((name, key, times[0][1], times[0][0], base[0] if base else float("nan")))
    print(f"{name}: {seen} new shapes", flush=True)
    del recs, keep
    torch.cuda.empty_cache()
gain = sum(b - t for _, _, _, t, b in report if b == b)
print(f"{len(configs)} shapes; sum over distinct shapes: heuristic {sum(b for *_, b in report if b == b):.0f} us -> tuned "
      f"{sum(t for *_, t, _ in report):.0f} us (-{gain:.0f} us)")
for name, key, cfg, t, b in sorted(report, key=lambda r: r[3] - r[4])[:40]:
    print(f"  {name:18s} {key:44s} -> {cfg}  {b:7.1f} -> {t:7.1f} us")
os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
with open(args.out, "w") as f:
    kept.update(configs)
    json.dump({"device": torch.cuda.get_device_name(0), "configs": {k: list(v) for k, v in kept.items()}}, f, indent=0)
print("wrote", args.out)
