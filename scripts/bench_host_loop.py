"""Host side of the serve loop in isolation (no GPU): a local hive, the real `Worker` (poll, lanes, result worker, POST), and a
lane whose "GPU work" is a sleep of --gpu-ms that ends with a 512x512 image.  Reports jobs/s with the image encoding done on
the lane (what `synchronous_do_work_function` does for a direct caller) and with it owed to the result worker (what the serve
loop does, worker.py), i.e. how much of the packaging cost leaves the lane's critical path, and the host's own ceiling
(--gpu-ms 0).

  python scripts/bench_host_loop.py --jobs 40 --gpu-ms 243
"""
import argparse
import asyncio
import contextlib
import io
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
from aiohttp import web
from PIL import Image


def make_image(size):
    y, x = np.mgrid[0:size, 0:size]
    arr = np.stack([(np.sin(x / 40) + 1) * 127, (np.cos(y / 55) + 1) * 127, ((x + y) / 4) % 255], -1)
    arr = arr * 0.8 + np.random.default_rng(0).random((size, size, 3)) * 50      # texture: a realistic JPEG cost
    return Image.fromarray(arr.astype("uint8"))


class Hive:
    def __init__(self, jobs, per_poll):
        self.todo, self.per_poll, self.done = jobs, per_poll, 0
        self.first = self.last = None

    async def work(self, request):
        n = min(self.per_poll, self.todo)
        self.todo -= n
        base = self.todo
        return web.json_response({"jobs": [{"id": f"j{base + i}", "model_name": "stub"} for i in range(n)]})

    async def result(self, request):
        await request.read()
        now = time.perf_counter()
        self.first = self.first or now
        self.last = now
        self.done += 1
        return web.json_response({"ok": True})


async def run(args, deferred):
    from chiaswarm_b200.generator import materialize_result
    from chiaswarm_b200.gpu.device import Device
    from chiaswarm_b200.output_processor import OutputProcessor, deferring_packaging
    from chiaswarm_b200.settings import Settings
    from chiaswarm_b200.worker import Worker

    image = make_image(args.size)
    hive = Hive(args.jobs, args.per_poll or (args.lanes + args.prefetch))
    app = web.Application(client_max_size=1 << 28)
    app.router.add_get("/api/work", hive.work)
    app.router.add_post("/api/results", hive.result)
    runner = web.AppRunner(app, access_log=None)
    await runner.setup()
    site = web.TCPSite(runner, "127.0.0.1", 0)
    await site.start()
    port = site._server.sockets[0].getsockname()[1]

    def lane_job(job):
        time.sleep(args.gpu_ms / 1000.0)                      # the GPU's share of the job (the lane thread is blocked)
        with deferring_packaging(deferred):
            proc = OutputProcessor(["primary"], "image/jpeg")
            proc.add_outputs([image])
            artifacts = proc.get_results()
        return {"id": job["id"], "artifacts": artifacts, "nsfw": False, "worker_version": "bench", "pipeline_config": {}}

    class BenchWorker(Worker):
        async def execute(self, job, device):
            return await asyncio.get_running_loop().run_in_executor(None, lane_job, job)

    s = Settings()
    s.sdaas_uri, s.sdaas_token, s.result_spool = f"http://127.0.0.1:{port}", "t", ""
    s.prefetch_jobs, s.poll_found_s = args.prefetch, args.poll_found_s
    w = BenchWorker(settings=s, devices=[Device("cpu") for _ in range(args.lanes)], cpu_lane=False)

    async def no_wait(_seconds):
        await asyncio.sleep(0)

    async def stop_when_done():
        while hive.done < args.jobs:
            await asyncio.sleep(0.005)
        w.request_stop()

    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        if args.cadence:      # the real pauses between polls (poll_found_s after jobs, 11 s idle): ends when all jobs are back
            stopper = asyncio.create_task(stop_when_done())
            await w.run()
            await stopper
            dt = hive.last - t0
        else:                 # no pauses: the loop's own overheads only
            polls = -(-args.jobs // (args.per_poll or (args.lanes + args.prefetch))) + 1
            await w.run(max_polls=polls, sleep=no_wait)
            dt = time.perf_counter() - t0
    await runner.cleanup()
    assert hive.done == args.jobs, (hive.done, args.jobs)
    assert materialize_result is not None
    return args.jobs / dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--jobs", type=int, default=40)
    ap.add_argument("--gpu-ms", type=float, default=243.0, help="device time per job (SD1.5 512x512 on one B200: 243)")
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--lanes", type=int, default=1)
    ap.add_argument("--prefetch", type=int, default=0, help="settings.prefetch_jobs (0 = the reference: the poll holds a lane)")
    ap.add_argument("--per-poll", type=int, default=0, help="jobs the hive hands out per poll (default lanes + prefetch)")
    ap.add_argument("--poll-found-s", type=float, default=1.0, help="settings.poll_found_s")
    ap.add_argument("--cadence", action="store_true", help="keep the real pauses between polls")
    args = ap.parse_args()
    os.environ.setdefault("SDAAS_ROOT", "/tmp/bench_host_loop")
    inline = asyncio.run(run(args, deferred=False))
    deferred = asyncio.run(run(args, deferred=True))
    ideal = args.lanes * 1000.0 / args.gpu_ms if args.gpu_ms > 0 else float("inf")
    print(f"jobs {args.jobs}, lanes {args.lanes}, image {args.size}x{args.size} JPEG, device time {args.gpu_ms:g} ms per job, "
          f"prefetch {args.prefetch}, {args.per_poll or (args.lanes + args.prefetch)} jobs per poll, "
          f"{'poll pause ' + format(args.poll_found_s, 'g') + ' s' if args.cadence else 'no poll pauses'}")
    print(f"  encoding on the lane        : {inline:7.2f} jobs/s")
    print(f"  encoding on the result worker: {deferred:7.2f} jobs/s   ({deferred / inline:.3f}x)")
    print(f"  device-bound ceiling         : {ideal:7.2f} jobs/s")


if __name__ == "__main__":
    main()
