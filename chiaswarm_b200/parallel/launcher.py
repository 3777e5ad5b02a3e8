"""Process-per-GPU launcher for the serving worker (the reference runs one
asyncio task per GPU inside a single process and funnels the blocking pipeline calls through the
default thread pool, swarm/worker.py:41-43, swarm/generator.py:12-14 -- the GIL then serialises
host work for 8 B200s).  Here the parent supervises one child *process* per GPU; each child is a
complete `Worker` bound to its device (resident weights, its own CUDA graphs) that polls the hive
on its own.  A child that dies is restarted after a back-off (its in-flight job is simply not
reported, so the hive re-issues it -- the non-fatal path of the reference's error taxonomy).

With `settings.parallel_mode` = gang / patch / auto and `gang_size` > 1 the launcher instead starts the gang
server (parallel/gang_worker.py): one hive client, one rank process per GPU in a single torch.distributed world,
gangs of 1 / 2 / 4 / 8 ranks leased per job.

  python -m chiaswarm_b200.parallel.launcher            # all visible GPUs
  python -m chiaswarm_b200.parallel.launcher --gpus 0,1,2,3
  SDAAS_PARALLEL_MODE=gang SDAAS_GANG_SIZE=8 python -m chiaswarm_b200.parallel.launcher
"""
from __future__ import annotations

import argparse
import asyncio
import multiprocessing as mp
import os
import signal
import sys
import time
from typing import List, Optional


def _child(gpu: int, worker_name_suffix: bool):
    os.environ["CUDA_DEVICE_ORDER"] = "PCI_BUS_ID"
    os.environ["SDAAS_LOCAL_GPU"] = str(gpu)
    import torch

    from ..gpu.device import Device
    from ..settings import load_settings
    from ..worker import Worker

    torch.cuda.set_device(gpu)
    settings = load_settings()
    if worker_name_suffix:
        settings.worker_name = f"{settings.worker_name}-gpu{gpu}"
    base, ext = os.path.splitext(settings.log_filename)
    settings.log_filename = f"{base}.gpu{gpu}{ext or '.log'}"  # one rotating log per process (the handler is not multi-process safe)
    if int(getattr(settings, "metrics_port", 0) or 0) > 0:
        settings.metrics_port = int(settings.metrics_port) + gpu  # one /metrics endpoint per worker process: base port + GPU index
    worker = Worker(settings=settings, devices=[Device(gpu)], cpu_lane=(gpu == 0))
    asyncio.run(worker.run())


class Supervisor:
    """Keeps one worker process per GPU alive."""

    def __init__(self, gpus: List[int], restart_backoff_s: float = 5.0, name_suffix: bool = True, target=_child):
        self.gpus = gpus
        self.backoff = restart_backoff_s
        self.name_suffix = name_suffix
        self.target = target
        self.ctx = mp.get_context("spawn")
        self.procs = {}
        self.restarts = {g: 0 for g in gpus}
        self._stop = False

    def _start(self, gpu: int):
        p = self.ctx.Process(target=self.target, args=(gpu, self.name_suffix), name=f"swarm-gpu{gpu}", daemon=False)
        p.start()
        self.procs[gpu] = p
        print(f"[launcher] gpu {gpu}: worker pid {p.pid}")

    def poll_once(self) -> int:
        """Restarts dead children; returns how many were restarted."""
        n = 0
        for gpu in self.gpus:
            p = self.procs.get(gpu)
            if p is None or not p.is_alive():
                if p is not None:
                    print(f"[launcher] gpu {gpu}: worker exited with {p.exitcode}; restarting")
                    self.restarts[gpu] += 1
                    n += 1
                self._start(gpu)
        return n

    def run(self, max_seconds: Optional[float] = None):
        signal.signal(signal.SIGTERM, lambda *_: self.stop())
        t0 = time.time()
        for gpu in self.gpus:
            self._start(gpu)
        try:
            while not self._stop and (max_seconds is None or time.time() - t0 < max_seconds):
                time.sleep(self.backoff)
                self.poll_once()
        except KeyboardInterrupt:
            pass
        finally:
            self.stop()

    def stop(self, grace_s: float = 60.0):
        """SIGTERM to every worker: each stops polling, finishes the job it holds and delivers the result
        (worker.py::request_stop); a worker still alive after `grace_s` is killed."""
        self._stop = True
        for p in self.procs.values():
            if p.is_alive():
                p.terminate()
        deadline = time.time() + grace_s
        for p in self.procs.values():
            p.join(timeout=max(0.0, deadline - time.time()))
        for p in self.procs.values():  # exactly the processes this supervisor started
            if p.is_alive():
                p.kill()
                p.join(timeout=5)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="", help="comma-separated GPU indices (default: all)")
    args = ap.parse_args(argv)
    if args.gpus:
        gpus = [int(g) for g in args.gpus.split(",")]
    else:
        import torch

        gpus = list(range(torch.cuda.device_count()))
    if not gpus:
        print("CUDA not present. Quitting.")
        return 1
    from ..settings import load_settings
    from .gang_worker import GANG_MODES, serve

    settings = load_settings()
    if str(settings.parallel_mode).lower() in GANG_MODES and int(settings.gang_size) > 1:
        # gang mode: ONE hive client in this process, one rank process per GPU in one torch.distributed world;
        # jobs lease 1 / 2 / 4 / 8 ranks from the RankGroup (parallel/gang_worker.py)
        if args.gpus:
            os.environ["CUDA_VISIBLE_DEVICES"] = args.gpus
        serve(settings, len(gpus))
        return 0
    Supervisor(gpus).run()
    return 0


if __name__ == "__main__":
    sys.exit(main())
