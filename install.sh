#!/usr/bin/env bash
# Installs the chiaswarm_b200 worker into ./venv on a Linux host with an NVIDIA driver >= 570 and the CUDA
# toolkit >= 12.8 (nvcc is needed once, to build the sm_100a kernel library).  Reference: install.sh:1-231.
set -o errexit -o nounset -o pipefail
cd "$(dirname "$0")"

die() { echo "error: $*" >&2; exit 1; }

[ "$(uname -s)" = "Linux" ] || die "Linux only (Windows: Install.ps1 under WSL2)"
[ "$(uname -m)" = "x86_64" ] || [ "$(uname -m)" = "aarch64" ] || die "a 64-bit OS is required"

# ---- python 3.10 .. 3.13
PY="${INSTALL_PYTHON_PATH:-}"
if [ -z "$PY" ]; then
  for v in 3.12 3.13 3.11 3.10; do
    if command -v "python$v" >/dev/null 2>&1; then PY="python$v"; break; fi
  done
fi
[ -n "$PY" ] || die "no supported python found (3.10 - 3.13); set INSTALL_PYTHON_PATH"
"$PY" -c 'import sys; assert (3, 10) <= sys.version_info[:2] <= (3, 13), sys.version' || die "unsupported python"
echo "python: $("$PY" --version)"

# ---- CUDA toolchain and GPU
NVCC="${NVCC:-$(command -v nvcc || true)}"
[ -n "$NVCC" ] || NVCC=/usr/local/cuda/bin/nvcc
[ -x "$NVCC" ] || die "nvcc not found (install the CUDA toolkit >= 12.8 or set NVCC=/path/to/nvcc)"
"$NVCC" --list-gpu-arch | grep -q compute_100 || die "$NVCC cannot target sm_100a (CUDA >= 12.8 required)"
if command -v nvidia-smi >/dev/null 2>&1; then
  nvidia-smi --query-gpu=name,memory.total --format=csv,noheader || true
else
  echo "warning: nvidia-smi not found; building anyway (the worker needs a B200 at run time)"
fi
command -v ffmpeg >/dev/null 2>&1 || echo "warning: ffmpeg not found; mp3 / webm outputs need it"

# ---- virtual environment
rm -rf venv
"$PY" -m venv venv
[ -e activate ] || ln -s venv/bin/activate activate
# shellcheck disable=SC1091
. ./activate
python -m pip install --upgrade pip wheel setuptools
python -m pip install -r requirements.txt
NVCC="$NVCC" python -m pip install --no-build-isolation -e .
python __graft_entry__.py   # builds libswarm_sm100.so and imports the package

cat <<MSG

chiaswarm_b200 is installed.
  . ./activate
  python -m swarm.initialize          # hive URI + token, model report (add --download to fetch missing models)
  python -m chiaswarm_b200.doctor     # pre-flight: GPU architecture, kernel library, peer access, hive, models
  python -m swarm.worker              # one process, one job lane per GPU
  python -m chiaswarm_b200.parallel.launcher   # one rank process per GPU (+ gang jobs: SDAAS_PARALLEL_MODE=patch)
MSG
