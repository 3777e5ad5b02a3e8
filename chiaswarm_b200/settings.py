"""Worker settings (reference: swarm/settings.py:7-73).

Same file location (`$SDAAS_ROOT` or `~/.sdaas`/settings.json), same keys, same environment
overrides (SDAAS_TOKEN / SDAAS_URI / SDAAS_WORKERNAME). New knobs for the B200 runtime are extra
keys with safe defaults."""
from __future__ import annotations

import json
import os
from pathlib import Path
from typing import Union

_DEFAULTS = {
    "log_level": "WARN",
    "log_filename": "log/generator.log",
    "sdaas_token": "",
    "sdaas_uri": "http://localhost:9511",
    "worker_name": "worker",
    # --- B200 runtime knobs (not in the reference)
    "parallel_mode": "dp",       # dp (one GPU per job) | gang / patch / auto (parallel/gang_worker.py leases gangs) | tp (+ head-parallel attention)
    "gang_size": 1,              # most GPUs one job may lease in gang mode (1, 2, 4, 8)
    "patch_min_pixels": 1024 * 1024,  # gang mode: single images at least this large run patch-parallel
    "job_timeout_s": 900,        # gang mode: a job that overruns is failed non-fatally and the rank world restarted
    "kernel_backend": "sm100",   # sm100 | torch (oracle / baseline path)
    "model_root": "",            # directory holding diffusers-layout checkpoints (HF cache also searched)
    "weight_cache_gb": 120,      # resident weight cache budget per GPU
    "result_retries": 3,         # bounded retry for POST /api/results (reference drops on failure)
    # the reference's cadence (poll holds a lane; 1 s pause after a poll that returned jobs) caps a lane at one poll's worth
    # of jobs per (1 s + their run time) -- fine for a 10 s job, idle time for a 0.25 s one.  Both knobs default to the
    # reference's behaviour; raise / lower them when the hive allows
    "prefetch_jobs": 0,          # > 0: keep up to this many extra jobs queued locally and poll while the lanes are busy
    "poll_found_s": 1.0,         # pause after a poll that returned jobs
    "max_batch": 1,              # > 1: a lane takes up to this many WAITING jobs at once; those that differ only in prompt /
                                 # seed share one denoising run (generator.synchronous_do_work_batch).  Needs prefetch_jobs
    "result_spool": "spool",     # results the hive did not take are kept here (under SDAAS_ROOT) and re-sent later; "" = drop
    "metrics_port": 0,           # > 0: serve Prometheus text at http://<metrics_host>:<port>/metrics (utils/metrics.py)
    "metrics_host": "127.0.0.1",
}

_ENV = {"sdaas_token": "SDAAS_TOKEN", "sdaas_uri": "SDAAS_URI", "worker_name": "SDAAS_WORKERNAME",
        "parallel_mode": "SDAAS_PARALLEL_MODE", "gang_size": "SDAAS_GANG_SIZE", "job_timeout_s": "SDAAS_JOB_TIMEOUT_S", "patch_min_pixels": "SDAAS_PATCH_MIN_PIXELS",
        "kernel_backend": "SDAAS_KERNEL_BACKEND", "model_root": "SDAAS_MODEL_ROOT",
        "prefetch_jobs": "SDAAS_PREFETCH_JOBS", "poll_found_s": "SDAAS_POLL_FOUND_S", "max_batch": "SDAAS_MAX_BATCH", "result_spool": "SDAAS_RESULT_SPOOL", "metrics_port": "SDAAS_METRICS_PORT", "metrics_host": "SDAAS_METRICS_HOST"}


class Settings:
    huggingface_token: Union[bool, str] = True
    log_level: str = _DEFAULTS["log_level"]
    log_filename: str = _DEFAULTS["log_filename"]
    sdaas_token: str = _DEFAULTS["sdaas_token"]
    sdaas_uri: str = _DEFAULTS["sdaas_uri"]
    worker_name: str = _DEFAULTS["worker_name"]
    parallel_mode: str = _DEFAULTS["parallel_mode"]
    gang_size: int = _DEFAULTS["gang_size"]
    job_timeout_s: int = _DEFAULTS["job_timeout_s"]
    patch_min_pixels: int = _DEFAULTS["patch_min_pixels"]
    kernel_backend: str = _DEFAULTS["kernel_backend"]
    model_root: str = _DEFAULTS["model_root"]
    weight_cache_gb: int = _DEFAULTS["weight_cache_gb"]
    result_retries: int = _DEFAULTS["result_retries"]
    prefetch_jobs: int = _DEFAULTS["prefetch_jobs"]
    poll_found_s: float = _DEFAULTS["poll_found_s"]
    max_batch: int = _DEFAULTS["max_batch"]
    result_spool: str = _DEFAULTS["result_spool"]
    metrics_port: int = _DEFAULTS["metrics_port"]
    metrics_host: str = _DEFAULTS["metrics_host"]


def get_settings_dir() -> Path:
    return Path(os.environ.get("SDAAS_ROOT") or "~/.sdaas/").expanduser()


def resolve_path(path) -> Path:
    """Path under the settings dir; creates the parent directory (reference behaviour)."""
    full = get_settings_dir().joinpath(path)
    full.parent.mkdir(parents=True, exist_ok=True)
    return full


def get_settings_full_path() -> Path:
    return resolve_path("settings.json")


def settings_exist() -> bool:
    return get_settings_dir().joinpath("settings.json").is_file()


def load_settings() -> Settings:
    settings = Settings()
    stored = {}
    try:
        with open(get_settings_full_path(), "r") as fh:
            stored = json.load(fh)
            if not isinstance(stored, dict):
                raise json.JSONDecodeError("settings must be an object", "", 0)
    except FileNotFoundError:
        print("no settings file")
    except json.JSONDecodeError:
        print("invalid settings file")
        stored = {}
    for key, default in _DEFAULTS.items():
        value = stored.get(key, default)
        env = _ENV.get(key)
        if env and os.getenv(env) is not None:
            value = type(default)(os.getenv(env))
        setattr(settings, key, value)
    return settings


def save_settings(settings: Settings) -> None:
    with open(get_settings_full_path(), "w") as fh:
        json.dump(dict(settings.__dict__), fh, indent=2)


def save_file(data, filename) -> None:
    with open(resolve_path(filename), "w") as fh:
        json.dump(data, fh, indent=2)
