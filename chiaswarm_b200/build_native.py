"""Builds libswarm_sm100.so (all hand-written sm_100a kernels) in-tree with nvcc.

The library has a plain C ABI (include/swarm_b200.h) and no torch / pybind dependency, so it
cross-compiles on a GPU-less machine in seconds and travels to the GPU box with the tree.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
LIB_DIR = Path(__file__).resolve().parent / "_lib"
LIB_PATH = LIB_DIR / "libswarm_sm100.so"
STAMP = LIB_DIR / "libswarm_sm100.stamp"

NVCC_FLAGS = [
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "-O3",
    "-std=c++17",
    "-lineinfo",
    "--use_fast_math",
    "-Xcompiler",
    "-fPIC",
    "-Xptxas",
    "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _digest() -> str:
    h = hashlib.sha256()
    for f in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h"))
                    + [ROOT / "include" / "swarm_b200.h"]):
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every .cu under csrc/ into one shared library. Returns its path."""
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    digest = _digest()
    if not force and LIB_PATH.exists() and STAMP.exists() and STAMP.read_text().strip() == digest:
        return LIB_PATH
    nvcc = _nvcc()
    obj_dir = LIB_DIR / "obj"
    obj_dir.mkdir(exist_ok=True)
    procs = []
    objs = []
    for src in sources():
        obj = obj_dir / (src.stem + ".o")
        objs.append(obj)
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(ROOT / "include"), "-c", str(src), "-o", str(obj)]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        log.append(f"== {src.name}\n{out}")
        if p.returncode != 0:
            failed = True
    (LIB_DIR / "build.log").write_text("\n".join(log))
    if failed or verbose:
        sys.stderr.write("\n".join(log))
    if failed:
        raise RuntimeError("nvcc failed; see chiaswarm_b200/_lib/build.log")
    link = [nvcc, "-shared", "-o", str(LIB_PATH), *map(str, objs), "-lcudart"]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("link failed")
    STAMP.write_text(digest)
    return LIB_PATH


def build_trace() -> Path:
    """Probe build: the same sources with -DSW_TRACE (per-CTA phase stamps in the GEMM kernels, read back with
    sw_trace_dump) as `_lib/libswarm_sm100_trace.so`.  Load it with SW_NATIVE_LIB=<path> (scripts/trace_igemm.py)."""
    nvcc = _nvcc()
    out = LIB_DIR / "libswarm_sm100_trace.so"
    obj_dir = LIB_DIR / "obj_trace"
    obj_dir.mkdir(parents=True, exist_ok=True)
    procs = []
    for src in sources():
        obj = obj_dir / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-DSW_TRACE", "-I", str(ROOT / "include"), "-c", str(src), "-o", str(obj)]
        procs.append((obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for obj, p in procs:
        log, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(log)
            raise RuntimeError("nvcc failed (trace build)")
    r = subprocess.run([nvcc, "-shared", "-o", str(out), *[str(o) for o, _ in procs], "-lcudart"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("link failed (trace build)")
    return out


if __name__ == "__main__":
    if "--trace" in sys.argv:
        print(build_trace())
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
