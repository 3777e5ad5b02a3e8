"""Hive client and scheduler (reference: swarm/worker.py:34-195).

Same protocol and cadence: `GET {hive}/api/work` (10 s timeout; next poll after 1 s if jobs were
returned, 11 s if idle, 121 s after an error; 400 means the hive considers this worker
misbehaving), `POST {hive}/api/results` (60 s timeout), bearer token, `chiaSWARM.worker/<ver>`
user agent.  Same back-pressure: a queue and a semaphore both sized to the number of execution
lanes, and the poll itself holds a lane while it is in flight (Q11).

What differs from the reference:
  * settings are read when the worker starts, not at import;
  * a lane is a `Device` of this process (one GPU per job) plus one CPU lane for pure-CPU workflows
    (stitch) so they do not occupy a GPU (Q10); in gang mode (`settings.parallel_mode` != "dp",
    parallel/gang_worker.py) a lane instead leases 1 / 2 / 4 / 8 rank processes from the RankGroup per job;
  * results that fail to POST are retried a bounded number of times instead of dropped (Q14), then spooled to disk
    (`settings.result_spool`) and re-sent once the hive takes results again, also after a worker restart;
  * a lane does not encode its own results: `do_work(..., defer_packaging=True)` returns with the image encoding owed and
    the result worker finishes it (`generator.materialize_result`) on an executor thread while the lane is already
    denoising its next job (10-40 ms of JPEG / thumbnail / base64 / hash work per image off the GPU's critical path);
  * SIGTERM / SIGINT stop the polling, the jobs already taken are finished and their results delivered (or spooled)
    before the process leaves (`request_stop`); the reference dies with its jobs;
  * polls, jobs, job times, uploads and queue depths are counted (utils/metrics.py) and served as Prometheus text at
    `/metrics` when `settings.metrics_port` is set (the reference has no metrics at all);
  * `Worker` is an object (testable against a fake hive) -- the module-level functions
    `run_worker/startup/ask_for_work/device_worker/result_worker/submit_result` remain.
"""
from __future__ import annotations

import asyncio
import json
import logging
import time
from datetime import datetime
from typing import List, Optional

import aiohttp

from . import __version__
from .generator import do_work, materialize_result
from .gpu.device import Device
from .log_setup import setup_logging
from .settings import Settings, load_settings, resolve_path
from .utils import metrics as metrics_mod

POLL_FOUND_S = 1
POLL_IDLE_S = 11
POLL_ERROR_S = 121
WORK_TIMEOUT_S = 10
RESULT_TIMEOUT_S = 60
CPU_WORKFLOWS = ("stitch",)


class Worker:
    def __init__(self, settings: Optional[Settings] = None, devices: Optional[List[Device]] = None,
                 cpu_lane: bool = True):
        self.settings = settings or load_settings()
        self.hive_uri = f"{self.settings.sdaas_uri.rstrip('/')}/api"
        self.devices = devices
        self.cpu_lane = cpu_lane
        self.work_queue: asyncio.Queue
        self.cpu_queue: asyncio.Queue
        self.available_gpus: asyncio.Semaphore
        self.result_queue: asyncio.Queue = asyncio.Queue()
        self._tasks: list = []
        self.polls = 0
        self.metrics = metrics_mod.new_registry()
        self._busy = 0
        self._metrics_server = None

    # ------------------------------------------------------------------ lifecycle
    def _headers(self):
        return {
            "Content-type": "application/json",
            "Authorization": f"Bearer {self.settings.sdaas_token}",
            "user-agent": f"chiaSWARM.worker/{__version__}",
        }

    async def startup(self):
        import torch

        if self.devices is None:
            if not torch.cuda.is_available():
                raise Exception("CUDA not present. Quitting.")
            self.devices = [Device(i) for i in range(torch.cuda.device_count())]
        from packaging import version

        if version.parse(torch.__version__.split("+")[0]) < version.parse("2.0.0"):
            raise Exception(f"Pytorch must be 2.0 or greater (found {torch.__version__}). Quitting.")
        setup_logging(resolve_path(self.settings.log_filename), self.settings.log_level)
        logging.info(f"Version {__version__}")
        logging.debug(f"Torch version {torch.__version__}")
        torch.set_float32_matmul_precision("high")
        torch.backends.cudnn.benchmark = True
        torch.backends.cuda.matmul.allow_tf32 = True

        lanes = max(1, len(self.devices))
        print(f"Found {len(self.devices)} GPUs")
        if sum(1 for d in self.devices if not getattr(d, "is_cpu", True)) > 1:
            import os

            # several GPU lanes in ONE process: plan captures must not trip over the other lanes' allocations
            os.environ.setdefault("SW_GRAPH_CAPTURE_MODE", "thread_local")
        self.prefetch = max(0, int(getattr(self.settings, "prefetch_jobs", 0) or 0))
        self.work_queue = asyncio.Queue(maxsize=lanes + self.prefetch)
        self.cpu_queue = asyncio.Queue()
        self.available_gpus = asyncio.Semaphore(lanes)
        self.metrics.set("swarm_queue_depth", lambda: self.work_queue.qsize(), {"queue": "gpu"})
        self.metrics.set("swarm_queue_depth", lambda: self.cpu_queue.qsize(), {"queue": "cpu"})
        self.metrics.set("swarm_queue_depth", lambda: self.result_queue.qsize(), {"queue": "results"})
        self.metrics.set("swarm_lanes_busy", lambda: self._busy)
        from .utils.weight_cache import pipeline_cache

        for name, read in (("swarm_weight_cache_bytes", lambda: pipeline_cache().bytes_resident()),
                           ("swarm_weight_cache_entries", lambda: len(pipeline_cache())),
                           ("swarm_weight_cache_hits", lambda: pipeline_cache().hits),
                           ("swarm_weight_cache_misses", lambda: pipeline_cache().misses),
                           ("swarm_weight_cache_evictions", lambda: pipeline_cache().evictions)):
            self.metrics.set(name, read)
        port = int(getattr(self.settings, "metrics_port", 0) or 0)
        if port > 0 and self._metrics_server is None:
            metrics_mod.register_gpu_gauges(self.metrics)
            try:
                self._metrics_server = await metrics_mod.serve(self.metrics, port, getattr(self.settings, "metrics_host", "127.0.0.1"))
            except OSError as e:  # (port taken: serving jobs matters more than serving metrics)
                logging.error("metrics endpoint on port %d not started: %s", port, e)

    async def run(self, max_polls: Optional[int] = None, sleep=asyncio.sleep):
        """The serve loop. `max_polls` bounds it for tests."""
        await self.startup()
        logging.info("worker")
        for device in self.devices:
            self._tasks.append(asyncio.create_task(self.device_worker(device)))
            print(f"Started device {device.descriptor()}")
        if self.cpu_lane:
            self._tasks.append(asyncio.create_task(self.cpu_worker()))
        # one result worker keeps up with ~40 images/s of encoding (scripts/bench_host_loop.py); a process that serves many
        # lanes gets one per two lanes, at most four
        for _ in range(max(1, min(4, len(self.devices) // 2))):
            self._tasks.append(asyncio.create_task(self.result_worker()))
        if self._spool_dir() is not None and any(self._spool_dir().glob("*.json")):
            self._flush_task = asyncio.create_task(self.flush_spool())  # results a previous run could not deliver
        self._stop = asyncio.Event()
        self._install_signal_handlers()
        try:
            while (max_polls is None or self.polls < max_polls) and not self._stop.is_set():
                while self.work_queue.full() and not self._stop.is_set():
                    # (a local check, no hive traffic: with a prefetch queue it is looked at often enough that a fast lane
                    # does not drain the queue and idle before the next poll goes out)
                    await self._nap(0.05 if getattr(self, "prefetch", 0) else 1, sleep)
                if self._stop.is_set():
                    break
                delay = await self.ask_for_work()
                await self._nap(delay, sleep)
        finally:
            if max_polls is not None or self._stop.is_set():
                # bounded runs (tests) and a requested stop (SIGTERM / SIGINT: `docker stop`, a service manager) finish the
                # jobs already taken from the hive and deliver -- or spool -- their results before leaving
                await self.drain()

    def request_stop(self):
        """Stop asking the hive for work; `run()` returns once the jobs in flight are finished and their results sent."""
        stop = getattr(self, "_stop", None)
        if stop is not None and not stop.is_set():
            print("stop requested: finishing the jobs in flight")
            stop.set()

    def _on_signal(self):
        """SIGTERM / SIGINT: the first one drains (`request_stop`), a second one leaves at once."""
        if getattr(self, "_signalled", False):
            import os

            print("second signal: leaving without draining")
            os._exit(130)
        self._signalled = True
        self.request_stop()

    def _install_signal_handlers(self):
        import signal
        import threading

        if threading.current_thread() is not threading.main_thread():
            return
        loop = asyncio.get_running_loop()
        for sig in (signal.SIGTERM, signal.SIGINT):
            try:
                loop.add_signal_handler(sig, self._on_signal)
            except (NotImplementedError, RuntimeError, ValueError):  # platforms / loops without signal support
                pass

    async def _nap(self, seconds, sleep):
        """`sleep(seconds)` that a stop request cuts short (the idle / error back-off is up to 121 s)."""
        if sleep is not asyncio.sleep:      # tests inject their own clock
            await sleep(seconds)
            return
        try:
            await asyncio.wait_for(self._stop.wait(), timeout=seconds)
        except asyncio.TimeoutError:
            pass

    async def drain(self):
        await self.work_queue.join()
        await self.cpu_queue.join()
        await self.result_queue.join()
        flush = getattr(self, "_flush_task", None)
        if flush is not None:  # let a start-up re-send finish: cancelling it between POST and unlink would send twice
            await asyncio.gather(flush, return_exceptions=True)
            self._flush_task = None
        for t in self._tasks:
            t.cancel()
        await asyncio.gather(*self._tasks, return_exceptions=True)
        self._tasks.clear()
        if self._metrics_server is not None:
            await self._metrics_server.cleanup()
            self._metrics_server = None

    # ------------------------------------------------------------------ hive I/O
    async def ask_for_work(self) -> int:
        # reference behaviour (Q11): the poll holds a lane, so there is no polling while every lane is busy.  With
        # `settings.prefetch_jobs` the local queue is deeper than the lanes and polling goes on while they work
        holds_lane = getattr(self, "prefetch", 0) == 0
        if holds_lane:
            await self.available_gpus.acquire()
        self.polls += 1
        print(f"{datetime.now()}: Asking for work from the hive at {self.hive_uri}...")
        try:
            timeout = aiohttp.ClientTimeout(total=WORK_TIMEOUT_S)
            async with aiohttp.ClientSession(timeout=timeout) as session:
                async with session.get(
                    f"{self.hive_uri}/work",
                    params={"worker_version": __version__, "worker_name": self.settings.worker_name},
                    headers=self._headers(),
                ) as response:
                    if response.status == 200:
                        body = await response.json()
                        found = False
                        for job in body["jobs"]:
                            print(f"Got job {job['id']}")
                            found = True
                            if self.cpu_lane and job.get("workflow") in CPU_WORKFLOWS:
                                await self.cpu_queue.put(job)
                            else:
                                await self.work_queue.put(job)
                        self.metrics.inc("swarm_polls_total", {"outcome": "jobs" if found else "idle"})
                        return float(getattr(self.settings, "poll_found_s", POLL_FOUND_S)) if found else POLL_IDLE_S
                    if response.status == 400:
                        body = await response.json()
                        print(f"{self.hive_uri} says {body.pop('message', 'bad worker')}")
                    else:
                        print(f"{self.hive_uri} returned {response.status}")
                    response.raise_for_status()
        except Exception as e:
            logging.exception(e)
            print(e)
            self.metrics.inc("swarm_polls_total", {"outcome": "error"})
            return POLL_ERROR_S
        finally:
            if holds_lane:
                self.available_gpus.release()
        return POLL_IDLE_S

    async def submit_result(self, result) -> bool:
        print("Result complete")
        timeout = aiohttp.ClientTimeout(total=RESULT_TIMEOUT_S)
        async with aiohttp.ClientSession(timeout=timeout) as session:
            async with session.post(f"{self.hive_uri}/results", data=json.dumps(result),
                                    headers=self._headers()) as response:
                if response.status == 500:
                    print(f"The hive returned an error: {response.reason}")
                    self.metrics.inc("swarm_results_posted_total", {"outcome": "rejected"})
                    return False
                print(f"Result {await response.json()}")
                self.metrics.inc("swarm_results_posted_total", {"outcome": "ok"})
                return True

    # ------------------------------------------------------------------ lanes
    async def execute(self, job, device):
        """One job on one lane.  Here: the in-process Device (one GPU per job, as in the reference);
        parallel.gang_worker.GangWorker overrides it to lease a gang of rank processes instead."""
        return await do_work(job, device, defer_packaging=True)  # encoded by the result worker, off this lane

    async def _counted(self, job, run):
        """Runs one job and accounts for it (count by workflow / outcome, wall time, busy lanes)."""
        workflow = str(job.get("workflow") or "image")
        t0 = time.monotonic()
        self._busy += 1
        try:
            result = await run()
        except BaseException:
            self.metrics.inc("swarm_jobs_total", {"workflow": workflow, "outcome": "crashed"})
            raise
        finally:
            self._busy -= 1
            self.metrics.observe("swarm_job_seconds", time.monotonic() - t0, {"workflow": workflow})
        self.metrics.inc("swarm_jobs_total", {"workflow": workflow, "outcome": metrics_mod.job_outcome(result)})
        return result

    async def execute_batch(self, jobs, device):
        """Several queued jobs on one lane in one call (`settings.max_batch` > 1): compatible ones share a denoising run
        (generator.synchronous_do_work_batch).  Results in job order."""
        from .generator import synchronous_do_work_batch

        return await asyncio.get_running_loop().run_in_executor(None, synchronous_do_work_batch, jobs, device, True)

    async def device_worker(self, device: Device):
        max_batch = max(1, int(getattr(self.settings, "max_batch", 1) or 1))
        while True:
            jobs = [await self.work_queue.get()]
            while len(jobs) < max_batch:  # what is ALREADY waiting (settings.prefetch_jobs fills the queue): never waits for more
                try:
                    jobs.append(self.work_queue.get_nowait())
                except asyncio.QueueEmpty:
                    break
            acquired = False
            try:
                await self.available_gpus.acquire()
                acquired = True
                if len(jobs) == 1:
                    results = [await self._counted(jobs[0], lambda: self.execute(jobs[0], device))]
                else:
                    t0 = time.monotonic()
                    self._busy += 1
                    try:
                        results = await self.execute_batch(jobs, device)
                    finally:
                        self._busy -= 1
                    for job, result in zip(jobs, results):
                        workflow = str(job.get("workflow") or "image")
                        self.metrics.observe("swarm_job_seconds", time.monotonic() - t0, {"workflow": workflow})
                        self.metrics.inc("swarm_jobs_total", {"workflow": workflow, "outcome": metrics_mod.job_outcome(result)})
                    self.metrics.inc("swarm_batches_total", {"jobs": len(jobs)})
                for result in results:
                    await self.result_queue.put(result)
            except asyncio.CancelledError:
                raise
            except Exception as e:
                logging.exception(e)
                print(f"device_worker {e}")
            finally:
                if acquired:
                    self.available_gpus.release()
                for _ in jobs:
                    self.work_queue.task_done()

    async def cpu_worker(self):
        cpu = Device("cpu")
        while True:
            job = await self.cpu_queue.get()
            try:
                await self.result_queue.put(await self._counted(job, lambda: do_work(job, cpu, defer_packaging=True)))
            except asyncio.CancelledError:
                raise
            except Exception as e:
                logging.exception(e)
                print(f"cpu_worker {e}")
            finally:
                self.cpu_queue.task_done()

    async def result_worker(self, backoff=asyncio.sleep):
        while True:
            result = await self.result_queue.get()
            try:
                # image encoding (grid, JPEG, thumbnail, base64, hash) runs here, on an executor thread, while the lane
                # that produced the result is already denoising its next job
                result = await asyncio.get_running_loop().run_in_executor(None, materialize_result, result)
                attempts = max(1, int(getattr(self.settings, "result_retries", 1)))
                delivered = False
                for attempt in range(attempts):
                    try:
                        if await self.submit_result(result):
                            delivered = True
                            break
                    except asyncio.CancelledError:
                        raise
                    except Exception as e:
                        logging.exception(e)
                        print(f"result_worker {e}")
                        self.metrics.inc("swarm_results_posted_total", {"outcome": "failed"})
                    if attempt + 1 < attempts:
                        await backoff(min(2 ** attempt, 30))
                if delivered:
                    await self.flush_spool()  # the hive is taking results again: hand over what it missed
                else:
                    self.spool_result(result)
            except asyncio.CancelledError:
                raise
            except Exception as e:  # the result loop must outlive any one result (reference: swarm/worker.py:137-142)
                logging.exception(e)
                print(f"result_worker {e}")
            finally:
                self.result_queue.task_done()

    # ------------------------------------------------------------------ undelivered results
    SPOOL_LIMIT = 256  # files; beyond it the oldest are dropped (a result is a few hundred KB of base64)

    def _spool_dir(self):
        name = getattr(self.settings, "result_spool", "") or ""
        return resolve_path(f"{name}/.keep").parent if name else None

    def spool_result(self, result) -> bool:
        """A result the hive did not take (outage, 500s) is written under `settings.result_spool` instead of being
        dropped as in the reference (swarm/worker.py:131-142); `flush_spool` re-sends it."""
        folder = self._spool_dir()
        if folder is None:
            return False
        try:
            safe = "".join(ch if ch.isalnum() or ch in "-_" else "_" for ch in str(result.get("id", "job")))[:64]
            path = folder / f"{time.time_ns()}-{safe}.json"
            path.write_text(json.dumps(result))
            files = sorted(folder.glob("*.json"))
            for old in files[:-self.SPOOL_LIMIT]:
                old.unlink(missing_ok=True)
            self.metrics.inc("swarm_results_posted_total", {"outcome": "spooled"})
            return True
        except Exception as e:  # a full disk must not take the result loop down
            logging.exception(e)
            return False

    async def flush_spool(self) -> int:
        """Re-sends spooled results, oldest first; stops at the first one the hive does not take."""
        folder = self._spool_dir()
        if folder is None:
            return 0
        sent = 0
        for path in sorted(folder.glob("*.json")):
            try:
                result = json.loads(path.read_text())
            except Exception:
                path.unlink(missing_ok=True)  # unreadable: nothing to deliver
                continue
            try:
                if not await self.submit_result(result):
                    break
            except asyncio.CancelledError:
                raise
            except Exception as e:
                logging.exception(e)
                break
            path.unlink(missing_ok=True)
            sent += 1
        return sent


# ---------------------------------------------------------------- module-level surface
_worker: Optional[Worker] = None


def _get() -> Worker:
    global _worker
    if _worker is None:
        _worker = Worker()
    return _worker


async def startup():
    await _get().startup()


async def ask_for_work():
    return await _get().ask_for_work()


async def device_worker(device: Device):
    await _get().device_worker(device)


async def result_worker():
    await _get().result_worker()


async def submit_result(result):
    return await _get().submit_result(result)


async def run_worker():
    await _get().run()


def __getattr__(name):
    """The reference's module-level state (`swarm.worker.settings`, `.hive_uri`, `.result_queue`, read at import there,
    swarm/worker.py:30-33): served from the process-wide Worker on first use, so importing this module has no side effects."""
    if name in ("settings", "hive_uri", "result_queue"):
        return getattr(_get(), name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def main() -> int:
    """Console entry point (`chiaswarm-worker`, `python -m chiaswarm_b200.worker`, `python -m swarm.worker`).

    With gangs configured (`parallel_mode` gang / patch / auto / tp and `gang_size` > 1) the reference's start command must
    not quietly serve one job per GPU: it hands over to the launcher, which owns the rank processes."""
    if _wants_gangs(load_settings()):
        from .parallel import launcher

        print("gang mode is configured: starting the launcher (one hive client, one rank process per GPU)")
        return int(launcher.main([]) or 0)
    asyncio.run(run_worker())
    return 0


def _wants_gangs(settings) -> bool:
    from .parallel.gang_worker import GANG_MODES

    try:
        return str(settings.parallel_mode).lower() in GANG_MODES and int(settings.gang_size) > 1
    except (TypeError, ValueError):
        return False


if __name__ == "__main__":
    main()
