"""Gang serving on real GPUs: the GangScheduler (parallel/gang_worker.py) with one NCCL rank process per GPU runs
(a) one image patch-parallel over a leased gang, (b) a batch split over a gang, (c) two gangs at once when the box
has more GPUs than `--gang`, and compares every image with the same job on a single rank.

  python scripts/gang_check.py --world 2 --gang 2 [--arch sd15 --size 512 --steps 4]
"""
import argparse
import base64
import io
import json
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from PIL import Image

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--gang", type=int, default=2)
    ap.add_argument("--model", default="runwayml/stable-diffusion-v1-5")
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--tol", type=float, default=6.0, help="mean absolute 8-bit level difference allowed")
    ap.add_argument("--backend", default="sm100", help="torch = eager back end (CPU dry run of this script)")
    args = ap.parse_args()

    os.environ["SDAAS_RANDOM_INIT"] = "1"
    os.environ.setdefault("SDAAS_ROOT", "/tmp/sdaas_gang_check")
    env = {"SDAAS_RANDOM_INIT": "1", "SDAAS_ROOT": os.environ["SDAAS_ROOT"], "SDAAS_KERNEL_BACKEND": args.backend}

    from chiaswarm_b200.parallel.gang_worker import GangScheduler
    from chiaswarm_b200.settings import Settings


    def settings(mode, gang):
        s = Settings()
        s.parallel_mode, s.gang_size, s.patch_min_pixels = mode, gang, args.size * args.size
        return s


    def job(job_id, **kw):
        j = {"id": job_id, "model_name": args.model, "prompt": "gang check", "negative_prompt": "", "height": args.size,
             "width": args.size, "num_inference_steps": args.steps, "guidance_scale": 7.5, "seed": 77,
             "content_type": "image/png"}
        j.update(kw)
        return j


    def image(result):
        assert "error" not in result["pipeline_config"], result["pipeline_config"]
        return np.asarray(Image.open(io.BytesIO(base64.b64decode(result["artifacts"]["primary"]["blob"]))), dtype=np.float32)


    jobs = {"single": job("single"), "batch": job("batch", num_images_per_prompt=4, seed=5)}
    res = {"world": args.world, "gang": args.gang}

    sched = GangScheduler(args.world, settings("dp", 1), env=env, job_timeout_s=600)   # reference shape: one rank per job
    sched.start()
    try:
        ref = {k: image(sched.submit(j)) for k, j in jobs.items()}
    finally:
        sched.stop()

    sched = GangScheduler(args.world, settings("gang", args.gang), env=env, job_timeout_s=600)
    sched.start()
    ok = True
    try:
        for k, j in jobs.items():
            got = image(sched.submit(j))        # first call records + captures the patch plan
            t0 = time.time()
            got = image(sched.submit(j))
            res[f"{k}_wall_s"] = round(time.time() - t0, 3)
            diff = float(np.abs(got - ref[k]).mean())
            res[f"{k}_mean_abs_diff"] = round(diff, 3)
            ok = ok and got.shape == ref[k].shape and diff <= args.tol
        if args.world >= 2 * args.gang:  # two gangs side by side (disjoint aligned subsets, their own process groups)
            out = [None, None]

            def run(i):
                out[i] = image(sched.submit(job(f"pair{i}", seed=77)))

            ts = [threading.Thread(target=run, args=(i,)) for i in range(2)]
            [t.start() for t in ts]
            [t.join() for t in ts]
            d = max(float(np.abs(o - ref["single"]).mean()) for o in out)
            res["concurrent_gangs_mean_abs_diff"] = round(d, 3)
            ok = ok and d <= args.tol
        res["leases_by_size"] = sched.leases_by_size
        ok = ok and sched.leases_by_size.get(args.gang, 0) >= 4 and sched.restarts == 0
    finally:
        sched.stop()
    print(json.dumps(res))
    print("GANG_CHECK PASS" if ok else "GANG_CHECK FAIL", flush=True)


if __name__ == "__main__":  # rank processes are spawned: they re-import this module
    main()
