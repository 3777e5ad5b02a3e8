# chiaswarm_b200 worker image: CUDA 12.9 toolchain (nvcc builds the sm_100a kernel library at image build
# time), one rank process per GPU under the gang launcher.  Reference: Dockerfile:1-42.
FROM nvidia/cuda:12.9.1-devel-ubuntu24.04

ENV DEBIAN_FRONTEND=noninteractive TZ=Etc/UTC \
    PIP_DISABLE_PIP_VERSION_CHECK=1 PIP_ROOT_USER_ACTION=ignore PIP_BREAK_SYSTEM_PACKAGES=1
RUN apt-get update && apt-get install -y --no-install-recommends \
        python3 python3-pip python3-venv python3-dev git ffmpeg libglib2.0-0 tzdata && \
    ln -sf /usr/bin/python3 /usr/local/bin/python && rm -rf /var/lib/apt/lists/*

WORKDIR /sdaas
COPY requirements.txt /sdaas/requirements.txt
RUN python -m pip install -r requirements.txt
COPY ./ /sdaas
# compile every CUDA extension for sm_100a and import the package (same entry the CI uses)
RUN python __graft_entry__.py

# model files come from the host: bind-mount the HuggingFace cache
RUN mkdir -p /root/.cache/huggingface
ENV SDAAS_ROOT=/sdaas/ \
    SDAAS_TOKEN= \
    SDAAS_URI=https://chiaswarm.ai \
    SDAAS_WORKERNAME=b200_worker \
    SDAAS_PARALLEL_MODE=dp \
    SDAAS_GANG_SIZE=1
# one process per visible GPU; SDAAS_PARALLEL_MODE=patch + SDAAS_GANG_SIZE=2|4|8 leases gangs for single-image jobs
# Prometheus text at :9100+<gpu>/metrics when SDAAS_METRICS_PORT=9100 (and SDAAS_METRICS_HOST=0.0.0.0) is set: one endpoint per
# worker process in data-parallel mode, one (the hive client's) in gang mode
EXPOSE 9100-9107
# pre-flight as the container's health: GPUs are sm_100, the kernel library loads, peers can map each other
HEALTHCHECK --interval=5m --timeout=90s --start-period=2m CMD python -m chiaswarm_b200.doctor --no-hive --json || exit 1
CMD ["python", "-m", "chiaswarm_b200.parallel.launcher"]

# docker build -t chiaswarm-b200 .
# docker run --gpus all --ipc=host --mount type=bind,src=$HOME/.cache/huggingface,target=/root/.cache/huggingface \
#     --env SDAAS_TOKEN=<token> chiaswarm-b200
