"""End-to-end worker loop against a fake hive: /api/work and
/api/results wire format, the 1 / 11 / 121 s cadence, 400 / 500 handling, bounded result retry."""
import asyncio
import json

import pytest
from aiohttp import web


class FakeHive:
    def __init__(self):
        self.jobs = []
        self.work_requests = []
        self.results = []
        self.work_status = 200
        self.fail_results = 0

    async def work(self, request):
        self.work_requests.append({"query": dict(request.query), "headers": dict(request.headers)})
        if self.work_status != 200:
            return web.json_response({"message": "slow worker"}, status=self.work_status)
        jobs, self.jobs = self.jobs, []
        return web.json_response({"jobs": jobs})

    async def result(self, request):
        if self.fail_results > 0:
            self.fail_results -= 1
            return web.Response(status=500, reason="boom")
        self.results.append(await request.json())
        return web.json_response({"ok": True})

    async def start(self):
        app = web.Application(client_max_size=64 * 1024 * 1024)
        app.router.add_get("/api/work", self.work)
        app.router.add_post("/api/results", self.result)
        self.runner = web.AppRunner(app)
        await self.runner.setup()
        site = web.TCPSite(self.runner, "127.0.0.1", 0)
        await site.start()
        port = site._server.sockets[0].getsockname()[1]
        return f"http://127.0.0.1:{port}"

    async def stop(self):
        await self.runner.cleanup()


@pytest.fixture(autouse=True)
def _env(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_RANDOM_INIT", "1")
    monkeypatch.setenv("SDAAS_KERNEL_BACKEND", "torch")


def _worker(uri, retries=3):
    from chiaswarm_b200.gpu.device import Device
    from chiaswarm_b200.settings import Settings
    from chiaswarm_b200.worker import Worker

    s = Settings()
    s.sdaas_uri, s.sdaas_token, s.worker_name, s.result_retries = uri, "secret", "testrig", retries
    return Worker(settings=s, devices=[Device("cpu")])


def test_poll_cadence_and_wire_format():
    from chiaswarm_b200 import __version__

    async def main():
        hive = FakeHive()
        uri = await hive.start()
        try:
            w = _worker(uri)
            await w.startup()
            assert await w.ask_for_work() == 11  # idle
            hive.jobs = [{"id": "a", "workflow": "stitch", "model_name": "x", "jobs": []}]
            assert await w.ask_for_work() == 1  # work found -> ask again right away
            assert w.cpu_queue.qsize() == 1  # CPU-only workflow goes to the CPU lane
            hive.work_status = 400
            assert await w.ask_for_work() == 121
            hive.work_status = 503
            assert await w.ask_for_work() == 121
            req = hive.work_requests[0]
            assert req["query"] == {"worker_version": __version__, "worker_name": "testrig"}
            assert req["headers"]["Authorization"] == "Bearer secret"
            assert req["headers"]["User-Agent"] == f"chiaSWARM.worker/{__version__}"
            assert w.available_gpus._value == 1  # the poll released its lane every time
        finally:
            await hive.stop()

    asyncio.run(main())


def test_unreachable_hive_backs_off():
    async def main():
        w = _worker("http://127.0.0.1:9")
        await w.startup()
        assert await w.ask_for_work() == 121

    asyncio.run(main())


def test_full_loop_generates_and_posts_results():
    async def main():
        hive = FakeHive()
        uri = await hive.start()
        hive.jobs = [
            {"id": "img", "model_name": "tiny-sd", "prompt": "p", "height": 64, "width": 64, "num_inference_steps": 2,
             "seed": 3},
            {"id": "bad", "model_name": "tiny-sd", "height": 5000, "width": 64},
        ]
        hive.fail_results = 1  # first POST gets a 500 and must be retried, not dropped
        try:
            w = _worker(uri)
            slept = []

            async def fast_sleep(s):
                slept.append(s)
                await asyncio.sleep(0)

            w.result_worker.__func__.__defaults__ = (fast_sleep,)
            await w.run(max_polls=2, sleep=fast_sleep)
            ids = sorted(r["id"] for r in hive.results)
            assert ids == ["bad", "img"]
            by_id = {r["id"]: r for r in hive.results}
            assert by_id["bad"]["fatal_error"] is True
            assert set(by_id["img"]["artifacts"]["primary"]) == {"blob", "content_type", "thumbnail", "sha256_hash"}
            assert by_id["img"]["pipeline_config"]["seed"] == 3
            assert slept[0] == 1 and 11 in slept
        finally:
            await hive.stop()

    asyncio.run(main())


def test_metrics_are_counted_and_served():
    """utils/metrics.py: the same loop as above, then GET /metrics of the worker's own endpoint."""
    import socket

    import aiohttp

    from chiaswarm_b200.utils import metrics as M

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]

    async def main():
        hive = FakeHive()
        uri = await hive.start()
        hive.jobs = [
            {"id": "img", "model_name": "tiny-sd", "prompt": "p", "height": 64, "width": 64, "num_inference_steps": 1, "seed": 3},
            {"id": "bad", "model_name": "tiny-sd", "height": 5000, "width": 64},
            {"id": "st", "workflow": "stitch", "model_name": "x", "jobs": []},
        ]
        hive.fail_results = 1
        try:
            w = _worker(uri)
            w.settings.metrics_port = port

            async def fast_sleep(s):
                await asyncio.sleep(0)

            w.result_worker.__func__.__defaults__ = (fast_sleep,)
            page = {}

            async def scrape_then_sleep(s):  # scrape while the worker is up (drain() closes the endpoint)
                if w.polls == 2 and not page:
                    await w.work_queue.join()
                    await w.cpu_queue.join()
                    await w.result_queue.join()
                    async with aiohttp.ClientSession() as session:
                        async with session.get(f"http://127.0.0.1:{port}/metrics") as r:
                            assert r.status == 200 and r.headers["Content-Type"].startswith("text/plain")
                            page["text"] = await r.text()
                        async with session.get(f"http://127.0.0.1:{port}/healthz") as r:
                            page["health"] = await r.json()
                await asyncio.sleep(0)

            await w.run(max_polls=2, sleep=scrape_then_sleep)
            m = w.metrics
            assert m.value("swarm_polls_total", {"outcome": "jobs"}) == 1 and m.value("swarm_polls_total", {"outcome": "idle"}) == 1
            assert m.value("swarm_jobs_total", {"workflow": "image", "outcome": "ok"}) == 1
            assert m.value("swarm_jobs_total", {"workflow": "image", "outcome": "fatal"}) == 1
            assert m.value("swarm_jobs_total", {"workflow": "stitch", "outcome": "ok"}) \
                + m.value("swarm_jobs_total", {"workflow": "stitch", "outcome": "error"}) \
                + m.value("swarm_jobs_total", {"workflow": "stitch", "outcome": "fatal"}) == 1
            assert m.value("swarm_job_seconds", {"workflow": "image"}) == 2
            assert m.value("swarm_results_posted_total", {"outcome": "ok"}) == 3
            assert m.value("swarm_results_posted_total", {"outcome": "rejected"}) == 1
            assert w._busy == 0 and w._metrics_server is None
            text = page["text"]
            assert page["health"] == {"status": "ok"}
            assert "# TYPE swarm_jobs_total counter" in text and 'swarm_jobs_total{outcome="ok",workflow="image"} 1' in text
            assert 'swarm_job_seconds_bucket{workflow="image",le="+Inf"} 2' in text and "swarm_job_seconds_sum" in text
            assert 'swarm_queue_depth{queue="gpu"} 0' in text and "swarm_lanes_busy 0" in text
            # every line is `name{labels} value` or a comment: the exposition format Prometheus parses
            import re

            for line in text.strip().splitlines():
                assert line.startswith("# ") or re.fullmatch(r'[a-z_]+(\{[^}]*\})? [-+0-9.eE]+(inf)?', line), line
        finally:
            await hive.stop()

    asyncio.run(main())
    # the registry on its own: labels are escaped, histograms are cumulative
    m = M.new_registry()
    m.inc("swarm_jobs_total", {"workflow": 'we"ird\n', "outcome": "ok"})
    for v in (0.05, 0.3, 400.0, 5000.0):
        m.observe("swarm_job_seconds", v)
    text = m.render()
    assert 'workflow="we\\"ird\\n"' in text
    assert 'swarm_job_seconds_bucket{le="0.1"} 1' in text and 'swarm_job_seconds_bucket{le="0.5"} 2' in text
    assert 'swarm_job_seconds_bucket{le="900"} 3' in text and 'swarm_job_seconds_bucket{le="+Inf"} 4' in text


def test_undelivered_results_are_spooled_and_resent(tmp_path):
    """The reference drops a result whose POST fails (swarm/worker.py:131-142); here it survives a hive outage and a
    worker restart."""
    async def main():
        hive = FakeHive()
        uri = await hive.start()
        try:
            async def fast_sleep(s):
                await asyncio.sleep(0)

            w = _worker(uri, retries=2)
            w.result_worker.__func__.__defaults__ = (fast_sleep,)
            hive.jobs = [{"id": "lost/../one", "workflow": "stitch", "model_name": "x", "jobs": []}]
            hive.fail_results = 99                                   # the hive takes no results at all
            await w.run(max_polls=1, sleep=fast_sleep)
            spool = tmp_path / "spool"
            files = list(spool.glob("*.json"))
            assert hive.results == [] and len(files) == 1 and "/" not in files[0].name[len(str(spool)):]
            assert json.loads(files[0].read_text())["id"] == "lost/../one"
            assert w.metrics.value("swarm_results_posted_total", {"outcome": "spooled"}) == 1

            hive.fail_results = 0                                    # the hive is back; a NEW worker process starts
            w2 = _worker(uri)
            w2.result_worker.__func__.__defaults__ = (fast_sleep,)
            await w2.run(max_polls=1, sleep=fast_sleep)
            for _ in range(50):
                if hive.results:
                    break
                await asyncio.sleep(0.05)
            assert [r["id"] for r in hive.results] == ["lost/../one"] and not list(spool.glob("*.json"))
            # spooling can be switched off: then the result is dropped as in the reference
            w3 = _worker(uri, retries=1)
            w3.settings.result_spool = ""
            assert w3.spool_result({"id": "x"}) is False and await w3.flush_spool() == 0
        finally:
            await hive.stop()

    asyncio.run(main())


def test_gpu_gauges_come_from_nvml_and_survive_its_absence():
    from chiaswarm_b200.utils import metrics as M

    class Mem:
        used = 5 << 30

    class Util:
        gpu = 87

    class FakeNvml:
        NVML_CLOCK_SM = 1
        calls = 0

        @staticmethod
        def nvmlInit():
            pass

        @staticmethod
        def nvmlDeviceGetCount():
            return 2

        @staticmethod
        def nvmlDeviceGetHandleByIndex(i):
            return f"h{i}"

        @staticmethod
        def nvmlDeviceGetMemoryInfo(h):
            return Mem

        @staticmethod
        def nvmlDeviceGetUtilizationRates(h):
            return Util

        @staticmethod
        def nvmlDeviceGetClockInfo(h, which):
            if h == "h1":
                raise RuntimeError("GPU is lost")             # one broken gauge must not take the page down
            return 1965

        @staticmethod
        def nvmlDeviceGetCurrentClocksThrottleReasons(h):
            return 0

    m = M.new_registry()
    assert M.register_gpu_gauges(m, FakeNvml) == 2
    text = m.render()
    assert 'swarm_gpu_memory_used_bytes{gpu="1"} 5.36871e+09' in text and 'swarm_gpu_utilization_ratio{gpu="0"} 0.87' in text
    assert 'swarm_gpu_sm_clock_mhz{gpu="0"} 1965' in text and 'swarm_gpu_sm_clock_mhz{gpu="1"}' not in text
    assert 'swarm_gpu_throttle_reasons{gpu="0"} 0' in text

    class NoDriver:
        @staticmethod
        def nvmlInit():
            raise OSError("libnvidia-ml.so.1: cannot open shared object file")

    assert M.register_gpu_gauges(M.new_registry(), NoDriver) == 0


def test_a_stop_request_finishes_the_jobs_in_flight():
    """SIGTERM (here: `request_stop`, and a real signal to this process) ends the polling at once -- even inside the 11 s
    idle back-off -- while the job already taken is finished and its result delivered."""
    import os
    import signal
    import time

    async def main(use_signal):
        hive = FakeHive()
        uri = await hive.start()
        hive.jobs = [{"id": "last", "model_name": "tiny-sd", "prompt": "p", "height": 64, "width": 64,
                      "num_inference_steps": 1, "seed": 1}]
        try:
            w = _worker(uri)

            async def stopper():
                while not hive.work_requests:
                    await asyncio.sleep(0.01)
                await asyncio.sleep(0.05)
                installed = signal.getsignal(signal.SIGTERM) not in (signal.SIG_DFL, signal.SIG_IGN, None)
                if use_signal and installed:            # (never send a signal nobody is listening for)
                    os.kill(os.getpid(), signal.SIGTERM)
                else:
                    w.request_stop()

            t0 = time.time()
            task = asyncio.create_task(stopper())
            await w.run()                                   # unbounded serve loop: only the stop request ends it
            await task
            assert time.time() - t0 < 60
            assert [r["id"] for r in hive.results] == ["last"] and len(hive.work_requests) <= 2
            assert w._tasks == []
        finally:
            await hive.stop()

    asyncio.run(main(False))
    asyncio.run(main(True))


def test_the_result_loop_outlives_a_result_it_cannot_handle(monkeypatch):
    async def main():
        hive = FakeHive()
        uri = await hive.start()
        try:
            w = _worker(uri)
            await w.startup()
            task = asyncio.create_task(w.result_worker())
            await w.result_queue.put(object())                       # not a result at all: materialising it blows up
            await w.result_queue.put({"id": "fine", "artifacts": {}, "pipeline_config": {}})
            await asyncio.wait_for(w.result_queue.join(), timeout=20)
            assert not task.done() and [r["id"] for r in hive.results] == ["fine"]
            task.cancel()
            await asyncio.gather(task, return_exceptions=True)
        finally:
            await hive.stop()

    asyncio.run(main())


def test_prefetch_polls_while_the_lane_is_busy():
    """Reference cadence (default): the poll holds a lane, so a busy single-lane worker does not poll.  With
    `settings.prefetch_jobs` the local queue is deeper than the lanes and the hive is asked while the lane works."""
    import threading

    async def main(prefetch):
        hive = FakeHive()
        uri = await hive.start()
        try:
            w = _worker(uri)
            w.settings.prefetch_jobs = prefetch
            gate = threading.Event()

            async def slow(job, device):
                await asyncio.get_running_loop().run_in_executor(None, gate.wait, 10)
                return {"id": job["id"], "artifacts": {}, "pipeline_config": {}}

            w.execute = slow
            await w.startup()
            assert w.work_queue.maxsize == 1 + prefetch
            lane = asyncio.create_task(w.device_worker(w.devices[0]))
            hive.jobs = [{"id": "a", "model_name": "m"}]
            assert await w.ask_for_work() == 1.0
            for _ in range(500):                                      # until the lane holds job "a" (and the only semaphore slot)
                if w._busy == 1:
                    break
                await asyncio.sleep(0.01)
            assert w._busy == 1
            hive.jobs = [{"id": "b", "model_name": "m"}]
            poll = asyncio.create_task(w.ask_for_work())
            done, _ = await asyncio.wait({poll}, timeout=(10 if prefetch else 0.5))
            polled_while_busy = bool(done)
            gate.set()
            await asyncio.wait_for(poll, 10)
            await asyncio.wait_for(w.work_queue.join(), 10)
            lane.cancel()
            await asyncio.gather(lane, return_exceptions=True)
            return polled_while_busy, len(hive.work_requests)
        finally:
            await hive.stop()

    assert asyncio.run(main(0)) == (False, 2)       # the second poll waited for the lane
    assert asyncio.run(main(2)) == (True, 2)        # the second poll went out while the lane was working
    # the pause after a productive poll is a setting (default: the reference's 1 s)
    from chiaswarm_b200.settings import Settings

    assert Settings().poll_found_s == 1.0 and Settings().prefetch_jobs == 0


def test_host_loop_benchmark_script_runs():
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "scripts", "bench_host_loop.py"), "--jobs", "6", "--gpu-ms", "5",
                          "--size", "64", "--prefetch", "2"], capture_output=True, text=True, timeout=240, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "encoding on the result worker" in out.stdout and "jobs/s" in out.stdout


def test_max_batch_merges_waiting_jobs_through_the_hive_loop(monkeypatch):
    """settings.max_batch + prefetch_jobs: jobs waiting in the local queue are taken together; those that differ only in
    prompt / seed are denoised in one pipeline call and every job still gets its own result, under its own id and seed."""
    from chiaswarm_b200.diffusion import pipelines

    calls = []
    real = pipelines.StableDiffusionPipeline.__call__

    def spy(self, *a, **k):
        calls.append(k.get("prompt"))
        return real(self, *a, **k)

    monkeypatch.setattr(pipelines.StableDiffusionPipeline, "__call__", spy)

    async def main():
        hive = FakeHive()
        uri = await hive.start()
        base = {"model_name": "tiny-sd", "height": 64, "width": 64, "num_inference_steps": 1}
        hive.jobs = [dict(base, id=f"j{i}", prompt=f"prompt {i}", seed=10 + i) for i in range(4)]
        try:
            w = _worker(uri)
            w.settings.prefetch_jobs, w.settings.max_batch = 4, 4

            async def fast_sleep(s):
                await asyncio.sleep(0)

            w.result_worker.__func__.__defaults__ = (fast_sleep,)
            await w.run(max_polls=2, sleep=fast_sleep)
            by_id = {r["id"]: r for r in hive.results}
            assert sorted(by_id) == ["j0", "j1", "j2", "j3"]
            assert [by_id[f"j{i}"]["pipeline_config"]["seed"] for i in range(4)] == [10, 11, 12, 13]
            assert all("blob" in r["artifacts"]["primary"] for r in hive.results)
            blobs = {r["artifacts"]["primary"]["sha256_hash"] for r in hive.results}
            assert len(blobs) == 4                                              # four different images
            merged = [c for c in calls if isinstance(c, list)]
            assert merged and sum(len(c) for c in merged) + sum(1 for c in calls if isinstance(c, str)) == 4
            assert len(calls) < 4                                               # fewer pipeline calls than jobs
            assert w.metrics.value("swarm_jobs_total", {"workflow": "image", "outcome": "ok"}) == 4
        finally:
            await hive.stop()

    asyncio.run(main())


def test_a_taken_metrics_port_does_not_stop_the_worker():
    import socket

    async def main():
        with socket.socket() as taken:
            taken.bind(("127.0.0.1", 0))
            taken.listen(1)
            w = _worker("http://127.0.0.1:9")
            w.settings.metrics_port = taken.getsockname()[1]
            await w.startup()                       # logs the failure, keeps going
            assert w._metrics_server is None and w.work_queue.maxsize == 1

    asyncio.run(main())
