#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): images/sec/box, Stable Diffusion 1.5 txt2img 512x512 fp16,
50 DPM-Solver++ steps, data-parallel -- one resident pipeline per GPU, one process per GPU.

A "step" is one complete image job: text encode (2 prompts, CFG) -> 50 x (UNet + CFG + scheduler
update) -> VAE decode -> uint8 image.  Weights are random-init of the SD1.5 architecture and
prompts/latents are synthetic (there is no network for checkpoints); nothing is skipped or cached
between steps.

  python bench.py --gpus 1 --steps 4 --warmup 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port P bench.py --gpus N --steps K --warmup W

  --impl ours        native sm_100a engine (default)
  --impl torch       this repo's eager-PyTorch restatement of the same module graph (cuDNN /
                     cuBLAS / SDPA): the ops diffusers would issue; used for BASELINE.md rows
  --impl reference   the unmodified reference from baseline/_ref (unavailable: see DESIGN.md)
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "images/sec/box SD1.5 512x512 50-step txt2img"


def reference_unavailable():
    """Why the unmodified reference cannot be benchmarked here (one line), or None if it can."""
    why = ("reference is not installable: /root/reference has no setup.py/pyproject.toml (pip: not installable), and "
           "swarm/worker.py imports diffusers, accelerate and xformers, none of which are on the image or in /opt/wheelhouse")
    ref = os.path.join(ROOT, "baseline", "_ref", "swarm", "worker.py")
    if not os.path.isfile(ref):  # (this repo's own `swarm` alias package must not be mistaken for it)
        return why + "; baseline/_ref is empty"
    try:
        import importlib.util

        spec = importlib.util.spec_from_file_location("_reference_swarm_worker", ref)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return "reference imported but no driver for it exists in this bench (it needs diffusers pipelines and hub checkpoints)"
    except Exception as e:  # expected: ModuleNotFoundError: diffusers
        return f"{why} ({type(e).__name__}: {e})"


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        busy = sorted(sm)[len(sm) // 2:] if sm else []
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def torch_baseline(args, dev, prompts, gen, l2_flush, barrier, dist, world, images: int = 3):
    """img/s of the eager-PyTorch restatement (pipe.backend = "torch") on the same GPU, same clocks, same
    timed region as the headline (device-timed, L2 flushed, max over ranks).  The denominator of `vs_baseline`."""
    import torch

    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline

    pipe = StableDiffusionPipeline.from_architecture(args.arch)
    pipe.backend = "torch"
    pipe.to(dev, torch.float16)
    pipe.set_scheduler("dpmpp_2m", True)

    def one(i):
        return pipe(prompt=prompts[i % len(prompts)], negative_prompt="", height=args.size, width=args.size,
                    num_inference_steps=args.denoise_steps, guidance_scale=7.5, generator=gen, output_type="device")

    for i in range(3):
        one(i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(images):
        l2_flush.zero_()
        one(i)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    del pipe
    torch.cuda.empty_cache()
    return {"value": world * images / (float(t.item()) / 1e3), "unit": "images/s", "images": images,
            "what": "eager PyTorch restatement of the same graph in this process (cuDNN conv, cuBLAS GEMM, SDPA); "
                    "stands in for the reference, which cannot be installed here"}


def sdxl_patch_latency(args, dev, runs: int = 3):
    """One SDXL-base 1024x1024 image (30 DPM-Solver++ steps, CFG): latency on one GPU and patch-parallel over all
    N ranks of this job (engine/patch_engine.py), device-timed, max over ranks, median of `runs`."""
    import statistics

    import torch

    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline
    from chiaswarm_b200.parallel import runtime

    gang = runtime.init()
    pipe = StableDiffusionPipeline.from_architecture("sdxl")
    pipe.backend = "sm100"
    pipe.to(dev, torch.float16)
    pipe.set_scheduler("dpmpp_2m", True)
    steps = args.sdxl_steps

    def call(output_type):
        g = torch.Generator(device=dev).manual_seed(4321)
        return pipe(prompt="a single large image for the latency mode", negative_prompt="", height=1024, width=1024,
                    num_inference_steps=steps, guidance_scale=7.5, generator=g, output_type=output_type)

    def timed():
        ts = []
        for i in range(runs + 2):
            torch.cuda.synchronize()
            gang.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            call("device")
            e1.record()
            torch.cuda.synchronize()
            if i >= 2:
                ts.append(gang.all_reduce_max(e0.elapsed_time(e1)))
        return statistics.median(ts)

    ref = call("latent").latents.float()
    ms_1gpu = timed()          # every rank runs the same single-GPU job side by side; max over ranks
    pipe.enable_patch_parallel(gang)
    out = call("latent").latents.float()
    err = ((out - ref).abs().max() / ref.abs().max().clamp_min(1e-3)).item()
    err = gang.all_reduce_max(err if err == err else 1e9)
    ms = timed()
    eng = pipe._denoiser()
    sg = next(v for k, v in eng._graphs.items() if k[0] == "patch")
    torch.cuda.synchronize()
    gang.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        sg.replay()
    e1.record()
    torch.cuda.synchronize()
    step_ms = gang.all_reduce_max(e0.elapsed_time(e1) / 10)
    res = {"n": gang.world, "model": "sdxl-base architecture, 1024x1024, CFG 7.5, DPM-Solver++(2M) karras",
           "denoise_steps": steps, "ms": ms, "ms_1gpu": ms_1gpu, "speedup": ms_1gpu / ms, "ms_per_step": step_ms,
           "launches_per_step": sg.launches_per_step, "guidance_groups": sg.ctx.G, "strips": sg.ctx.P,
           "rel_err_vs_single": err,
           "timed_region": "text encode + denoise + VAE decode to uint8 (one image), CUDA events, max over ranks"}
    del pipe
    torch.cuda.empty_cache()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "torch", "reference"])
    ap.add_argument("--arch", default="sd15")
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--denoise-steps", type=int, default=50)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-torch-baseline", action="store_true",
                    help="skip the same-process eager-PyTorch comparator that fills vs_baseline")
    ap.add_argument("--no-sdxl-patch", action="store_true",
                    help="(N > 1) skip the SDXL 1024x1024 patch-parallel latency block")
    ap.add_argument("--sdxl-steps", type=int, default=30)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        why = reference_unavailable()
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": why or "unknown"}))
        return 0

    import torch

    os.environ["SDAAS_RANDOM_INIT"] = "1"
    os.environ["SDAAS_ROOT"] = os.environ.get("SDAAS_ROOT", "/tmp/sdaas_bench")
    if not torch.cuda.is_available():
        print(json.dumps({"metric": METRIC, "value": None, "error": "no CUDA device", "n_gpus": args.gpus}))
        return 1
    if args.steps < 1 or args.warmup < 3:
        args.warmup = max(args.warmup, 3)
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(dev))

    from chiaswarm_b200.diffusion.pipelines import StableDiffusionPipeline
    from chiaswarm_b200.ops import native

    backend = "sm100" if args.impl == "ours" else "torch"
    if backend == "sm100":
        native.lib()
    torch.backends.cudnn.benchmark = True  # as the reference sets it (swarm/worker.py:180)
    torch.backends.cuda.matmul.allow_tf32 = True
    pipe = StableDiffusionPipeline.from_architecture(args.arch)
    pipe.backend = backend
    pipe.to(dev, torch.float16)
    pipe.set_scheduler("dpmpp_2m", True)

    prompts = [f"a synthetic benchmark prompt number {i} with a few more words" for i in range(64)]
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    l2_flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def one_image(i, output_type="pil"):
        return pipe(prompt=prompts[i % len(prompts)], negative_prompt="", height=args.size, width=args.size,
                    num_inference_steps=args.denoise_steps, guidance_scale=7.5, generator=gen,
                    output_type=output_type)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-timed region: K images, CUDA events, L2 flushed between images
    for i in range(args.warmup):
        one_image(i, "device")
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        l2_flush.zero_()
        one_image(i, "device")
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * args.steps / (ms_max / 1e3)

    launches = None
    if backend == "sm100":
        eng = pipe._denoiser()
        launches = (eng.last_launches + getattr(eng, "last_vae_launches", 0)) * args.steps

    # ---------------- end-to-end through the public job API (host tokenise -> pinned H2D -> ... -> D2H image -> JPEG)
    e2e = None
    if not args.no_e2e:
        from chiaswarm_b200.generator import synchronous_do_work_function
        from chiaswarm_b200.gpu.device import Device
        from chiaswarm_b200.utils.weight_cache import pipeline_cache

        device = Device(local_rank)
        key_model = {"sd15": "runwayml/stable-diffusion-v1-5", "tiny-sd": "tiny-sd", "sdxl": "stabilityai/stable-diffusion-xl-base-1.0",
                     "sd21": "stabilityai/stable-diffusion-2-1"}.get(args.arch, args.arch)
        os.environ["SDAAS_KERNEL_BACKEND"] = backend

        def job(i):
            return {"id": f"bench-{rank}-{i}", "model_name": key_model, "prompt": prompts[i % len(prompts)],
                    "negative_prompt": "", "height": args.size, "width": args.size,
                    "num_inference_steps": args.denoise_steps, "guidance_scale": 7.5, "seed": 1000 + i,
                    "content_type": "image/jpeg", "outputs": ["primary"]}

        import contextlib
        import io

        sink = io.StringIO()
        with contextlib.redirect_stdout(sink):
            for i in range(max(1, min(args.warmup, 2))):
                r = synchronous_do_work_function(job(i), device)
                assert "error" not in r["pipeline_config"], r["pipeline_config"]
            barrier()
            t0 = time.perf_counter()
            for i in range(args.steps):
                r = synchronous_do_work_function(job(100 + i), device)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tok_bytes = 2 * 77 * 8  # cond + uncond token ids (int64), staged in pinned memory
        img_bytes = args.size * args.size * 3  # uint8 image read back per job
        e2e = {"value": world * args.steps / float(t.item()), "unit": "images/s", "h2d_bytes_per_step": tok_bytes,
               "d2h_bytes_per_step": img_bytes, "path": "generator.synchronous_do_work_function(job, Device) -> result dict (JPEG, base64, sha256)"}

        # the serve loop's shape (worker.py): the GPU lane hands each result to the result thread with its image encoding
        # still owed and starts the next job; K jobs are timed until the LAST result is fully encoded.  Extra key only:
        # `value` above stays the one-call-at-a-time number.  No collective inside the guarded part.
        loop_dt, loop_err = float("inf"), None
        barrier()
        try:
            from concurrent.futures import ThreadPoolExecutor

            from chiaswarm_b200.generator import materialize_result

            pool = ThreadPoolExecutor(max_workers=1)
            with contextlib.redirect_stdout(sink):
                t0 = time.perf_counter()
                futs = [pool.submit(materialize_result, synchronous_do_work_function(job(200 + i), device, defer_packaging=True))
                        for i in range(args.steps)]
                done = [f.result(timeout=300) for f in futs]
                torch.cuda.synchronize()
                loop_dt = time.perf_counter() - t0
            pool.shutdown(wait=False)
            if any("error" in r["pipeline_config"] or "blob" not in r["artifacts"]["primary"] for r in done):
                loop_dt, loop_err = float("inf"), "a job of the serving-loop run failed"
        except Exception as e:  # the headline line must still be printed
            loop_dt, loop_err = float("inf"), f"{type(e).__name__}: {e}"[:200]
        t2 = torch.tensor([loop_dt], device=dev, dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        if math.isfinite(float(t2.item())):
            e2e["serving_loop_value"] = world * args.steps / float(t2.item())
            e2e["serving_loop_path"] = "worker lane: do_work(defer_packaging) -> result thread: materialize_result (JPEG, thumbnail, base64, sha256)"
        elif loop_err:
            e2e["serving_loop_error"] = loop_err

    # ---------------- same-box comparator: this repo's eager-PyTorch restatement of the same module graph
    # (cuDNN / cuBLAS / SDPA -- the ops diffusers would issue).  NOT the reference (unavailable).
    vs_baseline = baseline = None
    if args.impl == "ours" and not args.no_torch_baseline:
        try:
            baseline = torch_baseline(args, dev, prompts, gen, l2_flush, barrier, dist, world)
        except Exception as e:  # the headline line must still be printed
            baseline = {"error": f"{type(e).__name__}: {e}"[:300]}
        vs_baseline = value / baseline["value"] if baseline and baseline.get("value") else None

    # ---------------- BASELINE.json's second half: single-image SDXL 1024x1024 latency, patch-parallel over the N ranks
    sdxl_patch = None
    if args.impl == "ours" and world > 1 and not args.no_sdxl_patch:
        del pipe
        torch.cuda.empty_cache()
        try:
            sdxl_patch = sdxl_patch_latency(args, dev)
        except Exception as e:  # the headline line must still be printed
            sdxl_patch = {"n": world, "error": f"{type(e).__name__}: {e}"[:300]}

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": vs_baseline, "dtype": "fp16", "data": "synthetic prompts + seeded latents; random-init weights of the named architecture",
            "impl": args.impl,
            "config": {"model": f"{args.arch} (Stable Diffusion 1.5 architecture)" if args.arch == "sd15" else args.arch,
                       "global_batch": world, "seq_len": (args.size // 8) ** 2, "parallelism": f"dp{world}",
                       "resolution": args.size, "denoise_steps": args.denoise_steps, "guidance": "cfg 7.5 (batch x2)",
                       "scheduler": "DPM-Solver++(2M) karras", "l2": "256 MiB buffer written between timed iterations",
                       "timed_region": "text encode + 50 x (UNet, CFG, scheduler) + VAE decode to uint8, per image"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
        }
        if baseline is not None:
            out["baseline_torch"] = baseline
        if sdxl_patch is not None:
            out["sdxl_patch"] = sdxl_patch
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
