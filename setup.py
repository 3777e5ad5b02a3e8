"""Build hook: compile libswarm_sm100.so (nvcc, sm_100a) before the Python files are collected, so a plain
`pip install .` yields a working wheel.  The kernels have no torch / pybind dependency (C ABI + ctypes), which is
why this is an ordinary build_py step and not a torch cpp_extension.  Set SWARM_SKIP_NATIVE=1 to package without
nvcc (CPU-only tooling hosts: the job API, schedulers and the fake-hive tests still work)."""
import os
import shutil
import sys
from pathlib import Path

from setuptools import setup
from setuptools.command.build_py import build_py

ROOT = Path(__file__).resolve().parent


class BuildWithNative(build_py):
    def run(self):
        if os.environ.get("SWARM_SKIP_NATIVE") != "1":
            sys.path.insert(0, str(ROOT))
            from chiaswarm_b200 import build_native

            lib = build_native.build()
            print(f"built {lib}")
        super().run()
        # ship the sources next to the package so an installed copy can rebuild itself on the target machine
        out = Path(self.build_lib) / "chiaswarm_b200"
        for sub in ("csrc", "include"):
            dst = out / sub
            if dst.exists():
                shutil.rmtree(dst)
            shutil.copytree(ROOT / sub, dst)


setup(cmdclass={"build_py": BuildWithNative})
