"""Closed registry of pipeline and scheduler names.

The hive names diffusers classes in `parameters.pipeline_type` / `parameters.scheduler_type`
(reference: swarm/job_arguments.py:53-58,76-81,143-148 resolves them with
`getattr(diffusers, name)`).  Here every accepted name maps to one of this framework's own
pipeline drivers / scheduler specs.  Unknown names raise ValueError => fatal_error result."""
from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class PipelineType:
    """Which driver runs and how its inputs are interpreted."""

    name: str
    family: str  # sd | sd_img2img | sd_inpaint | sd_controlnet | sd_pix2pix | sd_upscale | txt2vid | audioldm | auto

    def __repr__(self):
        return f"<PipelineType {self.name}>"


@dataclass(frozen=True)
class SchedulerType:
    name: str
    kind: str  # dpmpp_2m | euler | euler_a | ddim | pndm | lms | heun | unipc | dpm_sde

    def __repr__(self):
        return f"<SchedulerType {self.name}>"


_PIPELINES = {
    "DiffusionPipeline": "auto",
    "StableDiffusionPipeline": "sd",
    "StableDiffusionImg2ImgPipeline": "sd_img2img",
    "StableDiffusionInpaintPipeline": "sd_inpaint",
    "StableDiffusionInpaintPipelineLegacy": "sd_inpaint",
    "StableDiffusionControlNetPipeline": "sd_controlnet",
    "StableDiffusionControlNetImg2ImgPipeline": "sd_controlnet",
    "StableDiffusionControlNetInpaintPipeline": "sd_controlnet",
    "StableDiffusionInstructPix2PixPipeline": "sd_pix2pix",
    "StableDiffusionLatentUpscalePipeline": "sd_upscale",
    "StableDiffusionUpscalePipeline": "sd_upscale",
    "StableDiffusionXLPipeline": "sd",
    "StableDiffusionXLImg2ImgPipeline": "sd_img2img",
    "StableDiffusionXLInpaintPipeline": "sd_inpaint",
    "TextToVideoSDPipeline": "txt2vid",
    "TextToVideoZeroPipeline": "txt2vid",
    "AudioLDMPipeline": "audioldm",
    "IFPipeline": "deepfloyd",
}

# Schedulers with their own update rule, and the ones served by the closest multistep/ODE rule.
_SCHEDULERS = {
    "DPMSolverMultistepScheduler": "dpmpp_2m",
    "DPMSolverSinglestepScheduler": "dpmpp_2m",
    "DEISMultistepScheduler": "dpmpp_2m",
    "UniPCMultistepScheduler": "dpmpp_2m",
    "KDPM2DiscreteScheduler": "dpmpp_2m",
    "EulerDiscreteScheduler": "euler",
    "HeunDiscreteScheduler": "euler",
    "LMSDiscreteScheduler": "euler",
    "EulerAncestralDiscreteScheduler": "euler_a",
    "KDPM2AncestralDiscreteScheduler": "euler_a",
    "DPMSolverSDEScheduler": "euler_a",
    "DDIMScheduler": "ddim",
    "PNDMScheduler": "ddim",
    "DDPMScheduler": "ddim",
}

_cache = {}


def resolve(name: str):
    if name in _cache:
        return _cache[name]
    if name in _PIPELINES:
        obj = PipelineType(name, _PIPELINES[name])
    elif name in _SCHEDULERS:
        obj = SchedulerType(name, _SCHEDULERS[name])
    elif name == "ControlNetModel":
        obj = PipelineType(name, "controlnet_model")
    else:
        raise ValueError(f"Unknown pipeline or scheduler type: {name}")
    _cache[name] = obj
    return obj


def pipeline_names():
    return sorted(_PIPELINES)


def scheduler_names():
    return sorted(_SCHEDULERS)
