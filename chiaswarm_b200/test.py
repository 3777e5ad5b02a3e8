"""Manual smoke run of one job of each workflow through the real executor (reference: swarm/test.py -- six literal job
dictionaries, only one of which is wired to `__main__`, no assertions, CUDA and network required).

    python -m chiaswarm_b200.test                      # every workflow, on cuda:0 when there is one, else on the CPU
    python -m chiaswarm_b200.test txt2img txt2vid      # some of them
    python -m chiaswarm_b200.test --tiny               # the toy architectures with random weights (no checkpoints needed)

Unlike the reference's script every job is run, each line says ok / the job's error, and the exit status is the number of
jobs that failed for a reason other than "the model is not on this box"."""
from __future__ import annotations

import argparse
import asyncio
import os
import sys


def jobs(tiny: bool):
    sd, p2p, t2v = ("tiny-sd", "tiny-pix2pix", "tiny-t2v") if tiny else (
        "stabilityai/stable-diffusion-2-1", "timbrooks/instruct-pix2pix", "cerspense/zeroscope_v2_576w")
    steps = 2 if tiny else 10
    side = 64 if tiny else 512
    return {
        "txt2img": {"model_name": sd, "prompt": "spoons", "num_inference_steps": steps, "height": side, "width": side},
        "txt2vid": {"model_name": t2v, "workflow": "txt2vid", "prompt": "dogs dancing", "num_inference_steps": steps,
                    "num_frames": 4 if tiny else 24, **({"height": 32, "width": 32} if tiny else {})},
        "txt2audio": {"model_name": "cvssp/audioldm", "workflow": "txt2audio", "num_inference_steps": steps,
                      "prompt": "Techno music with a strong, upbeat tempo and high melodic riffs"},
        "bark": {"model_name": "suno/bark", "workflow": "txt2audio", "prompt": "Hello, my name is Pepe. [laughs]"},
        "if": {"model_name": "DeepFloyd/IF-I-XL-v1.0", "prompt": "a green frog wearing blue sunglasses",
               "num_inference_steps": steps},
        "vid2vid": {"model_name": p2p, "workflow": "vid2vid", "prompt": "make it sunny", "num_inference_steps": steps,
                    "video_uri": os.environ.get("SDAAS_TEST_VIDEO_URI", "")},
    }


# the reference's literal job dictionaries, by their names there (swarm/test.py:7-59)
_full = jobs(False)
test_job, txt2audio_job, vid2vid_job = _full["txt2img"], _full["txt2audio"], _full["vid2vid"]
txt2vidJob, bark_job, if_job = _full["txt2vid"], _full["bark"], _full["if"]


async def run_test(job, device=None, name="job") -> str:
    """One job through the executor; returns the verdict line (the reference's `run_test(job)` prints it)."""
    from .generator import do_work
    from .gpu.device import Device

    if device is None:
        import torch

        device = Device(0) if torch.cuda.is_available() else Device("cpu")
    result = await do_work(dict(job, id=f"__test_{name}__", outputs=["primary"]), device)
    error = result["pipeline_config"].get("error")
    kind = result["artifacts"]["primary"]["content_type"]
    return f"error: {error}" if error else f"ok ({kind}, {len(result['artifacts']['primary']['blob']) * 3 // 4} bytes)"


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("names", nargs="*", help="workflows to run (default: all)")
    ap.add_argument("--tiny", action="store_true", help="toy architectures with random weights")
    args = ap.parse_args(argv)
    if args.tiny:
        os.environ.setdefault("SDAAS_RANDOM_INIT", "1")
    import torch

    from .gpu.device import Device

    device = Device(0) if torch.cuda.is_available() else Device("cpu")
    table = jobs(args.tiny)
    unknown = [n for n in args.names if n not in table]
    if unknown:
        ap.error(f"unknown workflow(s) {unknown}; choose from {sorted(table)}")
    failed = 0
    for name in args.names or list(table):
        if name == "vid2vid" and not table[name]["video_uri"]:
            print(f"{name:10} skipped (set SDAAS_TEST_VIDEO_URI to a video)")
            continue
        verdict = asyncio.run(run_test(table[name], device, name))
        print(f"{name:10} {verdict}")
        if verdict.startswith("error") and "not available locally" not in verdict and "local checkpoints" not in verdict:
            failed += 1
    return failed


if __name__ == "__main__":
    sys.exit(main())
