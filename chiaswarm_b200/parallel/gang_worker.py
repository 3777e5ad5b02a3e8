"""Gang serving: hive jobs that span 2 / 4 / 8 GPUs.

The reference gives every job exactly one GPU: `run_worker` starts one `device_worker` task per device and
each pops jobs from a shared queue (swarm/worker.py:41-43,113-128); its device pool (swarm/gpu/device_pool.py:7-34)
is dead code.  Here the pool is alive: `gpu.device_pool.RankGroup` leases aligned gangs of ranks, and this module is
what leases them.

  hive  <--HTTP-->  GangWorker (this process: hive client, no CUDA context)
                       |  GangScheduler.submit(job): size = plan_gang_size(job); gang = pool.lease(size)
                       |  job + gang ranks -> the job queue of every rank of the gang
                       v
        rank 0 .. rank N-1   one process per GPU in ONE torch.distributed world (NCCL; gloo on CPU), every
                             aligned sub-gang {0-1, 2-3, .., 0-3, 4-7, 0-7} created at start-up.  A rank runs
                             `synchronous_do_work_function(job, Device)` with `parallel.context.job_gang()` set; the
                             callbacks then go patch-parallel (one large image), split the batch (ControlNet batch 8),
                             scatter frames (vid2vid) or split the clip along the frame axis (txt2vid).  The gang's first rank
                             returns the result.

Failure handling: a job that overruns `job_timeout_s` (utils.tracing.Watchdog) or loses a
rank process is answered with a non-fatal error (the hive re-issues it), the rank is withdrawn
(`RankGroup.mark_dead`), and the rank world is restarted as a whole -- a torch.distributed world cannot lose a member.
A rank that fails ALONE (out of memory, a device fault) says so when it returns; its peers, left waiting in a collective,
then get `straggler_grace_s` (30 s) rather than the whole job timeout before the same restart happens.

`settings.parallel_mode`: "dp" = every job one rank (the reference's behaviour, N independent lanes);
"gang" / "patch" / "auto" = jobs that profit lease up to `settings.gang_size` ranks, the rest one; "tp" = the same, with the
self-attention of row-split plans head-parallel over the gang (fused all-gather GEMM + GEMM reduce-scatter; faster than the
K/V gather on SD1.5-sized levels, slower on SDXL's: DESIGN.md).
"""
from __future__ import annotations

import asyncio
import logging
import multiprocessing as mp
import os
import queue
import random
import socket
import threading
import time
from typing import Dict, List, Optional

from ..gpu.device_pool import VALID_GANG_SIZES, RankGroup
from ..settings import Settings, load_settings
from ..utils.tracing import Watchdog
from ..worker import Worker

GANG_MODES = ("gang", "patch", "auto", "tp")
LARGE_IMAGE_PIXELS = 1024 * 1024


# ------------------------------------------------------------------------------------------ policy
def _pow2_floor(n: int) -> int:
    p = 1
    while p * 2 <= n:
        p *= 2
    return p


def plan_gang_size(job: dict, settings: Settings, world: int) -> int:
    """How many ranks this job should own.  Pure function of the job and the settings (tested on CPU)."""
    mode = str(getattr(settings, "parallel_mode", "dp")).lower()
    cap = _pow2_floor(max(1, min(int(getattr(settings, "gang_size", 1)), world)))
    if mode not in GANG_MODES or cap == 1:
        return 1
    workflow = job.get("workflow")
    if workflow in ("stitch", "img2txt", "txt2audio"):
        return 1
    if workflow == "txt2vid":
        return cap          # the clip is split along the frame axis over the gang (engine/patch_engine.py)
    if workflow == "vid2vid":
        return cap          # frames are independent: scattered over the gang (video/pix2pix.py)
    if str(job.get("model_name", "")).startswith("DeepFloyd/"):
        return 1
    batch = int(job.get("num_images_per_prompt", 1) or 1)
    if batch > 1:           # batch split: largest power of two that divides the batch
        size = cap
        while size > 1 and batch % size:
            size //= 2
        return size
    # one image: worth a gang only when it is large (SDXL 1024x1024; smaller images are launch-bound)
    h, w = int(job.get("height", 0) or 0), int(job.get("width", 0) or 0)
    params = job.get("parameters") or {}
    if "start_image_uri" in job or "mask_image_uri" in job or params.get("controlnet"):
        return 1            # single img2img / inpaint / ControlNet images stay on one rank
    if h * w >= int(getattr(settings, "patch_min_pixels", LARGE_IMAGE_PIXELS)):
        guided = float(job.get("guidance_scale", 7.5)) > 1.0
        size = cap
        if guided and size % 2:
            return 1
        return size
    return 1


def aligned_subsets(world: int) -> List[tuple]:
    """Every gang the pool can lease, in the one order all ranks create their process groups."""
    out = []
    for size in VALID_GANG_SIZES:
        if 1 < size <= world:
            out.extend(tuple(range(b, b + size)) for b in range(0, world - size + 1, size))
    return out


# ------------------------------------------------------------------------------------------ rank process
def rank_main(rank: int, world: int, port: int, job_q, result_q, use_cuda: bool, env: Optional[dict] = None):
    """Body of one rank process: join the world, pre-create the sub-gangs, serve jobs until told to stop."""
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank if use_cuda else 0),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    os.environ.update(env or {})
    if not use_cuda:
        os.environ["CUDA_VISIBLE_DEVICES"] = ""
    import torch

    from ..generator import synchronous_do_work_function
    from ..gpu.device import Device
    from . import context, runtime

    if not use_cuda:
        torch.set_num_threads(max(1, (os.cpu_count() or 2) // max(world, 1)))
    try:  # each rank logs to its own rotating file next to the hive client's (log/generator.rank3.log)
        from ..log_setup import setup_logging
        from ..settings import load_settings, resolve_path

        s = load_settings()
        base, ext = os.path.splitext(s.log_filename)
        setup_logging(resolve_path(f"{base}.rank{rank}{ext or '.log'}"), s.log_level)
    except Exception:
        pass  # logging must never keep a rank from coming up
    try:
        gang_world = runtime.init(None if use_cuda else "gloo")
        gangs: Dict[tuple, object] = {}
        for subset in aligned_subsets(world):
            g = gang_world.subgang(subset)
            if rank in subset:
                gangs[subset] = g
        device = Device(rank) if use_cuda else Device("cpu")
    except Exception as e:  # the scheduler sees a rank that never became ready
        result_q.put(("failed", rank, f"{type(e).__name__}: {e}"))
        raise
    result_q.put(("ready", rank, None))
    while True:
        msg = job_q.get()
        if msg is None:
            break
        lease_id, ranks, job = msg
        ranks = tuple(ranks)
        leader = ranks[0] == rank
        try:
            with context.running_on(gangs.get(ranks)):
                result = synchronous_do_work_function(dict(job), device)
        except BaseException as e:  # synchronous_do_work_function already folds job errors into results
            result = {"id": job.get("id"), "rank_error": f"{type(e).__name__}: {e}"}
        # a follower hands back only whether it failed: a rank that errors out alone leaves its peers inside a collective
        error = result.get("rank_error") or (result.get("pipeline_config") or {}).get("error")
        result_q.put(("result" if leader else "done", lease_id, result if leader else ({"error": error} if error else None)))
    try:
        import torch.distributed as dist

        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


class RankWorld:
    """The N rank processes and their queues."""

    def __init__(self, world: int, use_cuda: bool, env: Optional[dict] = None, target=rank_main):
        self.world, self.use_cuda, self.env, self.target = world, use_cuda, dict(env or {}), target
        self.ctx = mp.get_context("spawn")
        self.procs: List = []
        self.job_qs: List = []
        self.result_qs: List = []

    @staticmethod
    def _free_port() -> int:
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            return s.getsockname()[1]

    def start(self, ready_timeout_s: float = 600.0):
        port = self._free_port()
        self.job_qs = [self.ctx.Queue() for _ in range(self.world)]
        self.result_qs = [self.ctx.Queue() for _ in range(self.world)]
        self.procs = []
        for r in range(self.world):
            p = self.ctx.Process(target=self.target, name=f"swarm-rank{r}", daemon=True,
                                 args=(r, self.world, port, self.job_qs[r], self.result_qs[r], self.use_cuda, self.env))
            p.start()
            self.procs.append(p)
        deadline = time.time() + ready_timeout_s
        for r in range(self.world):
            while True:
                try:
                    kind, _, why = self.result_qs[r].get(timeout=0.5)
                except queue.Empty:
                    if not self.procs[r].is_alive() or time.time() > deadline:
                        self.stop()
                        raise RuntimeError(f"rank {r} did not come up")
                    continue
                if kind == "ready":
                    break
                self.stop()
                raise RuntimeError(f"rank {r} failed to start: {why}")

    def alive(self, rank: int) -> bool:
        return rank < len(self.procs) and self.procs[rank].is_alive()

    def stop(self):
        for q in self.job_qs:
            try:
                q.put_nowait(None)
            except Exception:
                pass
        deadline = time.time() + 5
        for p in self.procs:
            p.join(timeout=max(0.0, deadline - time.time()))
        for p in self.procs:  # exact processes this object started, never a pattern
            if p.is_alive():
                p.terminate()
        for p in self.procs:
            p.join(timeout=5)
            if p.is_alive():
                p.kill()
        self.procs = []


# ------------------------------------------------------------------------------------------ scheduler
class GangScheduler:
    """Leases gangs from the RankGroup and runs jobs on them.  `submit` blocks; call it from worker threads."""

    def __init__(self, world: int, settings: Optional[Settings] = None, use_cuda: Optional[bool] = None,
                 env: Optional[dict] = None, job_timeout_s: Optional[float] = None, target=rank_main):
        self.settings = settings or load_settings()
        if use_cuda is None:
            import torch

            use_cuda = torch.cuda.is_available()
        self.world = world
        self.pool = RankGroup(world)
        env = dict(env or {})
        if str(getattr(self.settings, "parallel_mode", "dp")).lower() == "tp":
            # row-split plans run self-attention head-parallel through the fused all-gather GEMM / GEMM reduce-scatter
            # (engine/layers.py::TransformerBlock._self_attention_tp) instead of gathering K / V inside the attention kernel
            env.setdefault("SW_PATCH_ATTN", "tp")
        self.ranks = RankWorld(world, use_cuda, env, target)
        self.job_timeout_s = float(job_timeout_s or getattr(self.settings, "job_timeout_s", 900))
        # a job error is symmetric (every rank of the gang refuses the same arguments) and then all ranks return at once;
        # a rank that fails ALONE (out of memory, a device fault) leaves the others waiting in a collective: give them
        # this long before the world is declared wedged, instead of the whole job timeout
        self.straggler_grace_s = 30.0
        self._lock = threading.Lock()
        self.generation = 0
        self.restarts = 0
        self.leases_by_size: Dict[int, int] = {}
        self.metrics = None  # utils.metrics.Metrics of the hive client that owns this scheduler (GangWorker sets it)

    def start(self):
        self.ranks.start()

    def stop(self):
        self.ranks.stop()

    def _restart(self, seen_generation: int):
        """Bring up a fresh rank world (once per failure, whoever notices first)."""
        with self._lock:
            if self.generation != seen_generation:
                return
            self.generation += 1
            self.restarts += 1
            if self.metrics is not None:
                self.metrics.inc("swarm_rank_world_restarts_total")
            logging.error("gang scheduler: restarting the rank world (generation %d)", self.generation)
            self.ranks.stop()
            self.pool = RankGroup(self.world)
            self.ranks.start()

    @staticmethod
    def _error_result(job: dict, message: str) -> dict:
        from .. import __version__
        from ..output_processor import make_text_result

        return {"id": job.get("id"), "artifacts": {"primary": make_text_result(message)}, "nsfw": False,
                "worker_version": __version__, "pipeline_config": {"error": message}}

    def submit(self, job: dict) -> dict:
        job = dict(job)
        size = plan_gang_size(job, self.settings, self.world)
        if "seed" not in job or job["seed"] is None:
            job["seed"] = random.getrandbits(63)  # every rank of a gang must draw the same noise
        with self._lock:  # one consistent view: a restart swaps the pool, the processes and their queues together
            generation, pool, ranks = self.generation, self.pool, self.ranks
            job_qs, result_qs, procs = ranks.job_qs, ranks.result_qs, ranks.procs
        gang = pool.lease(size)
        with self._lock:
            self.leases_by_size[size] = self.leases_by_size.get(size, 0) + 1
        if self.metrics is not None:
            self.metrics.inc("swarm_gang_leases_total", {"size": size})
        failure = None
        result = None
        lone_error = None  # (rank, message, time) of the first rank that came back with an error
        try:
            for r in gang.ranks:
                job_qs[r].put((gang.lease_id, gang.ranks, job))
            pending = set(gang.ranks)
            try:
                with Watchdog(self.job_timeout_s, f"job {job.get('id')} on ranks {gang.ranks}") as dog:
                    while pending and failure is None:
                        dog.check()  # TimeoutError once the job has overrun
                        if lone_error is not None and time.time() - lone_error[2] > self.straggler_grace_s:
                            failure = (f"rank {lone_error[0]} failed ({lone_error[1]}) and ranks {sorted(pending)} did not "
                                       f"return within {self.straggler_grace_s:g} s")
                            break
                        # every pending rank is looked at without blocking and the loop naps 2 ms when none had news:
                        # a blocking wait per rank would add up to (ranks x its timeout) to the latency of every job
                        progressed = False
                        for r in list(pending):
                            try:
                                kind, lease_id, payload = result_qs[r].get_nowait()
                            except queue.Empty:
                                if generation != self.generation:
                                    failure = "the rank world was restarted under this job"
                                    break
                                if not procs[r].is_alive():
                                    pool.mark_dead(r)
                                    failure = f"rank {r} died while running the job"
                                    break
                                continue
                            progressed = True
                            if lease_id != gang.lease_id:
                                continue  # stale message of an abandoned lease
                            pending.discard(r)
                            if kind == "result":
                                result = payload
                                payload = {"error": payload.get("rank_error") or (payload.get("pipeline_config") or {}).get("error")}
                            if lone_error is None and payload and payload.get("error"):
                                lone_error = (r, str(payload["error"])[:200], time.time())
                        if not progressed and failure is None:
                            time.sleep(0.002)
            except TimeoutError as e:
                failure = str(e)
        finally:
            if failure is None:
                pool.release(gang)
        if failure is not None:
            self._restart(generation)
            return self._error_result(job, failure)  # non-fatal: the hive may hand the job out again
        if result is None or "rank_error" in result:
            return self._error_result(job, (result or {}).get("rank_error", "the gang returned no result"))
        return result


# ------------------------------------------------------------------------------------------ hive client
class _Lane:
    """One execution lane of the hive client (a rank that may be leased); mirrors Device's descriptor API."""

    def __init__(self, index: int):
        self.index = index

    def descriptor(self):
        return f"rank:{self.index}"


class GangWorker(Worker):
    """`Worker` whose lanes execute through the GangScheduler instead of an in-process Device."""

    def __init__(self, settings: Optional[Settings] = None, world: Optional[int] = None,
                 scheduler: Optional[GangScheduler] = None, cpu_lane: bool = True):
        settings = settings or load_settings()
        if scheduler is None:
            if world is None:
                import torch

                world = torch.cuda.device_count()
            scheduler = GangScheduler(world, settings)
        self.scheduler = scheduler
        super().__init__(settings=settings, devices=[_Lane(i) for i in range(scheduler.world)], cpu_lane=cpu_lane)
        scheduler.metrics = self.metrics

    async def startup(self):
        await super().startup()
        if not self.scheduler.ranks.procs:
            await asyncio.get_running_loop().run_in_executor(None, self.scheduler.start)

    async def execute(self, job, lane):
        try:
            return await asyncio.get_running_loop().run_in_executor(None, self.scheduler.submit, job)
        except asyncio.CancelledError:
            raise
        except Exception as e:  # a scheduler fault is answered like any other transient failure: the hive re-issues the job
            logging.exception(e)
            return GangScheduler._error_result(job, f"gang scheduler: {type(e).__name__}: {e}")

    async def execute_batch(self, jobs, lane):
        """Gang jobs are not merged (a lease is per job): `settings.max_batch` only lets one lane take several at once."""
        return [await self.execute(job, lane) for job in jobs]

    async def drain(self):
        await super().drain()
        self.scheduler.stop()


def serve(settings: Optional[Settings] = None, world: Optional[int] = None):
    asyncio.run(GangWorker(settings, world).run())
