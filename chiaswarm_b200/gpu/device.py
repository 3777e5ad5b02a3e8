"""Per-GPU execution context (reference: swarm/gpu/device.py:7-50).

Same contract: >= 8 GB VRAM, a non-blocking mutex that raises "busy" on re-entry, a generator
seeded from the job (`torch.seed()` when absent) created on the device, the callback invoked as
`func(identifier, model_name, **kwargs)`, and the seed echoed in the returned pipeline config.
Differences: `Device("cpu")` exists so the whole job path runs in CPU tests (the reference cannot
start without CUDA, swarm/worker.py:167-168), and the returned config is a plain dict copy rather
than a mutated frozen mapping (Q6)."""
from __future__ import annotations

import contextlib
import logging
from threading import Lock

import torch

MIN_VRAM_BYTES = 8_000_000_000


class Device:
    def __init__(self, device_id) -> None:
        self.is_cpu = device_id == "cpu"
        if not self.is_cpu:
            device_id = int(device_id)
            total = torch.cuda.mem_get_info(device_id)[1]
            if total < MIN_VRAM_BYTES:
                raise Exception(f"Not enough memory on device {device_id}. At least 8GB VRAM is required")
        self.device_id = device_id
        self.mutex = Lock()

    def descriptor(self):
        return f"{self.idenitifier()}:{self.name()}"

    def idenitifier(self):  # sic -- spelling kept: it is part of the reference's public surface
        return "cpu" if self.is_cpu else f"cuda:{self.device_id}"

    identifier = idenitifier

    def name(self):
        return "cpu" if self.is_cpu else torch.cuda.get_device_name(self.device_id)

    def __call__(self, func, **kwargs):
        if not self.mutex.acquire(False):
            logging.error(f"Device {self.device_id} is busy but got invoked.")
            raise Exception("busy")
        try:
            self.log_device()
            model_name = kwargs.pop("model_name")
            seed = kwargs.pop("seed", None)
            if seed is None:
                seed = torch.seed()
            kwargs["generator"] = torch.Generator(device=self.idenitifier()).manual_seed(seed)
            # the callback runs on an executor thread whose current CUDA device is 0: make this lane's GPU current for it
            # (the native kernels launch on the current device's stream; one process serving several GPUs needs this)
            with (contextlib.nullcontext() if self.is_cpu else torch.cuda.device(self.device_id)):
                artifacts, pipeline_config = func(self.idenitifier(), model_name, **kwargs)
            pipeline_config = dict(pipeline_config)
            pipeline_config["seed"] = seed
            return artifacts, pipeline_config
        finally:
            self.mutex.release()

    def log_device(self):
        logging.debug(f"Using device# {self.descriptor()}")
