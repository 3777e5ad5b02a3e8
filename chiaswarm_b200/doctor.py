"""`python -m chiaswarm_b200.doctor [--json] [--gpu-test]`: is this box ready to serve?

The reference's only pre-flight is `startup()` refusing to run without CUDA or with torch < 2 (swarm/worker.py:161-176).
A B200 worker has more that can be wrong before the first job: the native library not built for this tree, a GPU that is
not sm_100, no peer access between GPUs for gang mode, settings pointing at a hive that does not answer, none of the
hive's models on the box.  Every check is independent, never raises, and says what to do about a failure.

Exit status: 0 when every REQUIRED check passed (warnings allowed), 1 otherwise."""
from __future__ import annotations

import argparse
import json
import sys
from typing import Callable, List, Optional


class Check:
    def __init__(self, name: str, status: str, detail: str, fix: str = ""):
        self.name, self.status, self.detail, self.fix = name, status, detail, fix  # status: ok | warn | fail

    def as_dict(self):
        return {"check": self.name, "status": self.status, "detail": self.detail, **({"fix": self.fix} if self.fix else {})}


def _guard(name: str, fn: Callable[[], Check]) -> Check:
    try:
        return fn()
    except Exception as e:  # a check that crashes is itself a finding
        return Check(name, "fail", f"{type(e).__name__}: {e}")


def check_torch() -> Check:
    import torch
    from packaging import version

    v = torch.__version__
    if version.parse(v.split("+")[0]) < version.parse("2.0.0"):
        return Check("torch", "fail", f"torch {v}", "torch >= 2.0 with a CUDA >= 12.8 build is required")
    return Check("torch", "ok", f"torch {v}, CUDA build {torch.version.cuda}")


def check_gpus() -> Check:
    import torch

    if not torch.cuda.is_available():
        return Check("gpus", "fail", "no CUDA device visible", "the worker refuses to start without CUDA (as the reference)")
    found = []
    for i in range(torch.cuda.device_count()):
        p = torch.cuda.get_device_properties(i)
        found.append((i, p.name, f"sm_{p.major}{p.minor}", p.total_memory >> 30, p.multi_processor_count))
    wrong = [f for f in found if f[2] != "sm_100"]
    text = "; ".join(f"cuda:{i} {n} {a} {g} GiB {s} SMs" for i, n, a, g, s in found)
    if wrong:
        return Check("gpus", "fail", text, "the kernels are built for sm_100a (B200) only; other GPUs need "
                                          "SDAAS_KERNEL_BACKEND=torch (the eager oracle path)")
    return Check("gpus", "ok", text)


def check_native_library() -> Check:
    from .ops import native

    if not native.LIB_PATH.exists():
        return Check("native library", "fail", f"{native.LIB_PATH} is not built", "python __graft_entry__.py  (or pip install .)")
    lib = native.lib()
    missing = [n for n in ("sw_igemm", "sw_flash_attn", "sw_groupnorm_nhwc", "sw_sched_step", "sw_igemm_allgather",
                           "sw_igemm_reducescatter", "sw_flash_attn_gather") if not hasattr(lib, n)]
    if missing:
        return Check("native library", "fail", f"entry points missing: {missing}", "rebuild: the .so is older than the tree")
    return Check("native library", "ok", f"{native.LIB_PATH.name}, {native.LIB_PATH.stat().st_size >> 10} KiB")


def check_kernels_run() -> Check:
    """One tiny GEMM and one attention call against torch (only with --gpu-test: it touches the device)."""
    import torch

    from .ops import native

    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(256, 128, device="cuda", dtype=torch.float16, generator=g)
    w = torch.randn(192, 128, device="cuda", dtype=torch.float16, generator=g) * 0.1
    got = native.linear(x, w).float()
    want = x.float() @ w.float().t()
    err = float((got - want).abs().max() / want.abs().max())
    if not err < 5e-3:
        return Check("kernels", "fail", f"sw_igemm differs from torch by {err:.2e}", "rebuild the library for this GPU")
    return Check("kernels", "ok", f"sw_igemm vs torch: {err:.1e}")


def check_peer_access() -> Check:
    import torch

    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n < 2:
        return Check("peer access", "warn", f"{n} GPU(s): gang mode (SDAAS_PARALLEL_MODE=gang) has nothing to lease",
                     "data-parallel serving needs no peer access")
    blocked = [(a, b) for a in range(n) for b in range(n) if a != b and not torch.cuda.can_device_access_peer(a, b)]
    if blocked:
        return Check("peer access", "fail", f"no P2P between {blocked[:4]}", "gang mode needs NVLink / NVSwitch peer "
                                                                             "mappings; serve with SDAAS_PARALLEL_MODE=dp")
    return Check("peer access", "ok", f"all {n} GPUs can map each other's memory")


def check_settings_and_hive(timeout: float = 5.0) -> Check:
    from . import __version__
    from .settings import load_settings

    s = load_settings()
    if not s.sdaas_token:
        return Check("hive", "warn", f"{s.sdaas_uri}: no token configured", "python -m chiaswarm_b200.initialize")
    import requests

    try:
        r = requests.get(f"{s.sdaas_uri.rstrip('/')}/api/models", timeout=timeout,
                         headers={"user-agent": f"chiaSWARM.worker/{__version__}"})
        if r.status_code != 200:
            return Check("hive", "fail", f"{s.sdaas_uri} answered {r.status_code}", "check SDAAS_URI / the hive's status")
        return Check("hive", "ok", f"{s.sdaas_uri} answers; mode {s.parallel_mode}, gang size {s.gang_size}")
    except Exception as e:
        return Check("hive", "fail", f"{s.sdaas_uri} unreachable: {type(e).__name__}", "check SDAAS_URI and the network")


def check_models() -> Check:
    from .settings import get_settings_dir, load_settings
    from .utils.checkpoints import locate_checkpoint

    path = get_settings_dir() / "models.json"
    if not path.is_file():
        return Check("models", "warn", "no models.json (the hive's model list has not been fetched)",
                     "python -m chiaswarm_b200.initialize [--download]")
    data = json.loads(path.read_text())
    names = [m["model_name"] for m in data.get("language_models", []) + data.get("models", [])]
    root = getattr(load_settings(), "model_root", "")
    have = [n for n in names if locate_checkpoint(n, model_root=root) is not None]
    status = "ok" if have else "warn"
    return Check("models", status, f"{len(have)} of the hive's {len(names)} models are on this box",
                 "" if len(have) == len(names) else "python -m chiaswarm_b200.initialize --download")


def check_environment() -> Check:
    """Probe-only switches that must not be set on a serving worker."""
    import os

    if os.environ.get("SW_FUSED_DBG", "0") not in ("", "0"):
        return Check("environment", "fail", f"SW_FUSED_DBG={os.environ['SW_FUSED_DBG']} switches synchronisation steps of the "
                     "fused GEMM collectives off (a timing probe): results are wrong while it is set", "unset SW_FUSED_DBG")
    if os.environ.get("SDAAS_KERNEL_BACKEND", "sm100") == "torch":
        return Check("environment", "warn", "SDAAS_KERNEL_BACKEND=torch: jobs run on the eager oracle path, not on the sm_100a kernels")
    return Check("environment", "ok", "no probe switches set")


REQUIRED = ("torch", "gpus", "native library", "environment")


def run(gpu_test: bool = False, hive: bool = True) -> List[Check]:
    checks = [_guard("torch", check_torch), _guard("gpus", check_gpus), _guard("native library", check_native_library)]
    if gpu_test and checks[1].status == "ok" and checks[2].status == "ok":
        checks.append(_guard("kernels", check_kernels_run))
    checks.append(_guard("environment", check_environment))
    checks.append(_guard("peer access", check_peer_access))
    if hive:
        checks.append(_guard("hive", check_settings_and_hive))
    checks.append(_guard("models", check_models))
    return checks


def main(argv: Optional[list] = None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--json", action="store_true", help="machine-readable report")
    ap.add_argument("--gpu-test", action="store_true", help="also launch one kernel and compare it with torch")
    ap.add_argument("--no-hive", action="store_true", help="skip the network check")
    args = ap.parse_args(argv)
    checks = run(args.gpu_test, not args.no_hive)
    ok = all(c.status != "fail" for c in checks if c.name in REQUIRED or c.name == "kernels")
    if args.json:
        print(json.dumps({"ready": ok, "checks": [c.as_dict() for c in checks]}))
    else:
        for c in checks:
            print(f"[{c.status:4}] {c.name}: {c.detail}" + (f"\n       -> {c.fix}" if c.fix and c.status != "ok" else ""))
        print("ready to serve" if ok else "NOT ready: fix the failed checks above")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
