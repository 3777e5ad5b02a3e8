"""Worker metrics in the Prometheus text format (the reference has no metrics endpoint; its only
accounting value is `pipeline_config["cost"]` of vid2vid, swarm/video/pix2pix.py:85).

A process-local registry of counters, gauges and histograms -- no dependency: the worker serves `render()` itself at
`GET /metrics` over aiohttp (already the hive client's HTTP stack) when `settings.metrics_port` is set.  What is counted:

    swarm_polls_total{outcome}                 hive polls: jobs | idle | error
    swarm_jobs_total{workflow,outcome}         finished jobs: ok | error (hive re-issues) | fatal (hive must not)
    swarm_job_seconds{workflow}                histogram of wall time per job, lane acquisition included
    swarm_results_posted_total{outcome}        POST /api/results: ok | rejected | failed
    swarm_queue_depth{queue}                   gauge: jobs waiting for a lane
    swarm_lanes_busy                           gauge
    swarm_gang_leases_total{size}              gang mode: jobs by number of ranks leased
    swarm_rank_world_restarts_total            gang mode: restarts after a dead rank / wedged gang
    swarm_weight_cache_{bytes,entries,hits,misses,evictions}   resident models of this process (utils/weight_cache.py)
    swarm_gpu_{memory_used_bytes,utilization_ratio,sm_clock_mhz,throttle_reasons}{gpu}   gauges read from NVML per scrape

Gauges may be callables, read when the page is rendered."""
from __future__ import annotations

import threading
from typing import Callable, Dict, Iterable, Optional, Tuple, Union

BUCKETS = (0.1, 0.25, 0.5, 1.0, 2.5, 5.0, 10.0, 30.0, 60.0, 120.0, 300.0, 900.0)

_Labels = Tuple[Tuple[str, str], ...]


def _key(labels: Optional[dict]) -> _Labels:
    return tuple(sorted((str(k), str(v)) for k, v in (labels or {}).items()))


def _fmt(name: str, labels: _Labels, extra: Iterable[Tuple[str, str]] = ()) -> str:
    pairs = list(labels) + list(extra)
    if not pairs:
        return name
    esc = lambda v: v.replace("\\", "\\\\").replace('"', '\\"').replace("\n", "\\n")  # noqa: E731
    return name + "{" + ",".join(f'{k}="{esc(v)}"' for k, v in pairs) + "}"


class Metrics:
    def __init__(self):
        self._lock = threading.Lock()
        self._help: Dict[str, Tuple[str, str]] = {}
        self._counters: Dict[str, Dict[_Labels, float]] = {}
        self._gauges: Dict[str, Dict[_Labels, Union[float, Callable[[], float]]]] = {}
        self._hists: Dict[str, Dict[_Labels, list]] = {}

    def describe(self, name: str, kind: str, text: str):
        self._help[name] = (kind, text)

    def inc(self, name: str, labels: Optional[dict] = None, by: float = 1.0):
        with self._lock:
            series = self._counters.setdefault(name, {})
            series[_key(labels)] = series.get(_key(labels), 0.0) + by

    def set(self, name: str, value: Union[float, Callable[[], float]], labels: Optional[dict] = None):
        with self._lock:
            self._gauges.setdefault(name, {})[_key(labels)] = value

    def observe(self, name: str, value: float, labels: Optional[dict] = None):
        with self._lock:
            h = self._hists.setdefault(name, {}).setdefault(_key(labels), [[0] * len(BUCKETS), 0, 0.0])
            for i, edge in enumerate(BUCKETS):
                if value <= edge:
                    h[0][i] += 1
            h[1] += 1
            h[2] += float(value)

    # ---- reading
    def value(self, name: str, labels: Optional[dict] = None) -> float:
        """Counter / gauge value, or the observation count of a histogram (tests, health checks)."""
        with self._lock:
            k = _key(labels)
            if name in self._counters:
                return self._counters[name].get(k, 0.0)
            if name in self._gauges and k in self._gauges[name]:
                v = self._gauges[name][k]
                return float(v() if callable(v) else v)
            if name in self._hists and k in self._hists[name]:
                return float(self._hists[name][k][1])
            return 0.0

    def render(self) -> str:
        out = []
        with self._lock:
            counters = {n: dict(s) for n, s in self._counters.items()}
            gauges = {n: dict(s) for n, s in self._gauges.items()}
            hists = {n: {k: [list(v[0]), v[1], v[2]] for k, v in s.items()} for n, s in self._hists.items()}

        def head(name, kind):
            known = self._help.get(name)
            if known:
                out.append(f"# HELP {name} {known[1]}")
            out.append(f"# TYPE {name} {kind}")

        for name in sorted(counters):
            head(name, "counter")
            for labels, v in sorted(counters[name].items()):
                out.append(f"{_fmt(name, labels)} {v:g}")
        for name in sorted(gauges):
            head(name, "gauge")
            for labels, v in sorted(gauges[name].items()):
                try:
                    val = float(v() if callable(v) else v)
                except Exception:  # a gauge whose source is gone must not take the page down
                    continue
                out.append(f"{_fmt(name, labels)} {val:g}")
        for name in sorted(hists):
            head(name, "histogram")
            for labels, (buckets, count, total) in sorted(hists[name].items()):
                for edge, n in zip(BUCKETS, buckets):
                    out.append(f"{_fmt(name + '_bucket', labels, [('le', f'{edge:g}')])} {n}")
                out.append(f"{_fmt(name + '_bucket', labels, [('le', '+Inf')])} {count}")
                out.append(f"{_fmt(name + '_sum', labels)} {total:g}")
                out.append(f"{_fmt(name + '_count', labels)} {count}")
        return "\n".join(out) + "\n"


_HELP = {
    "swarm_polls_total": ("counter", "Hive polls by outcome (jobs, idle, error)."),
    "swarm_jobs_total": ("counter", "Finished jobs by workflow and outcome (ok, error, fatal)."),
    "swarm_job_seconds": ("histogram", "Wall time per job in seconds."),
    "swarm_results_posted_total": ("counter", "Result uploads by outcome (ok, rejected, failed)."),
    "swarm_batches_total": ("counter", "Lane calls that took several waiting jobs at once, by number of jobs."),
    "swarm_queue_depth": ("gauge", "Jobs waiting for an execution lane."),
    "swarm_lanes_busy": ("gauge", "Execution lanes running a job."),
    "swarm_gang_leases_total": ("counter", "Gang mode: jobs by number of ranks leased."),
    "swarm_rank_world_restarts_total": ("counter", "Gang mode: rank-world restarts after a dead rank or a wedged gang."),
}


def new_registry() -> Metrics:
    m = Metrics()
    for name, (kind, text) in _HELP.items():
        m.describe(name, kind, text)
    return m


def register_gpu_gauges(metrics: Metrics, nvml=None) -> int:
    """Per-GPU gauges read from NVML when the page is rendered: memory in use, utilisation, SM clock, throttle reasons
    (the same fields `bench.py` samples while timing).  Returns the number of GPUs registered; 0 when NVML is not usable
    (no driver: the page then simply has no GPU section)."""
    try:
        if nvml is None:
            import pynvml as nvml
        nvml.nvmlInit()
        count = int(nvml.nvmlDeviceGetCount())
    except Exception:
        return 0
    for name, text in (("swarm_gpu_memory_used_bytes", "Device memory in use."),
                       ("swarm_gpu_utilization_ratio", "Fraction of the last sample period a kernel was running."),
                       ("swarm_gpu_sm_clock_mhz", "Current SM clock."),
                       ("swarm_gpu_throttle_reasons", "NVML clocks-event reason bit mask (0 = none).")):
        metrics.describe(name, "gauge", text)
    for i in range(count):
        handle = nvml.nvmlDeviceGetHandleByIndex(i)
        labels = {"gpu": i}
        metrics.set("swarm_gpu_memory_used_bytes", lambda h=handle: nvml.nvmlDeviceGetMemoryInfo(h).used, labels)
        metrics.set("swarm_gpu_utilization_ratio", lambda h=handle: nvml.nvmlDeviceGetUtilizationRates(h).gpu / 100.0, labels)
        metrics.set("swarm_gpu_sm_clock_mhz", lambda h=handle: nvml.nvmlDeviceGetClockInfo(h, nvml.NVML_CLOCK_SM), labels)
        metrics.set("swarm_gpu_throttle_reasons",
                    lambda h=handle: nvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                    if hasattr(nvml, "nvmlDeviceGetCurrentClocksEventReasons")
                    else nvml.nvmlDeviceGetCurrentClocksThrottleReasons(h), labels)
    return count


def job_outcome(result: dict) -> str:
    if result.get("fatal_error"):
        return "fatal"
    return "error" if (result.get("pipeline_config") or {}).get("error") else "ok"


async def serve(metrics: Metrics, port: int, host: str = "127.0.0.1"):
    """Starts `GET /metrics` (and `/healthz`) on the running loop; returns the aiohttp runner (call `.cleanup()`)."""
    from aiohttp import web

    async def page(_request):
        return web.Response(text=metrics.render(), content_type="text/plain", charset="utf-8")

    async def health(_request):
        return web.json_response({"status": "ok"})

    app = web.Application()
    app.add_routes([web.get("/metrics", page), web.get("/healthz", health)])
    runner = web.AppRunner(app, access_log=None)
    await runner.setup()
    await web.TCPSite(runner, host, int(port)).start()
    return runner
