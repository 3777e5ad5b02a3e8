"""AudioLDM text-to-audio callback (reference: swarm/audio/audioldm.py:12-36): CLAP text ->
UNet2DCondition on mel latents -> VAE -> HiFi-GAN, 16 kHz, defaults 20 steps (25 through the
router) / 10 s.  The denoising UNet + VAE run on this framework's own UNet/VAE drivers; CLAP and
the HiFi-GAN vocoder come from `transformers` (ClapTextModelWithProjection, SpeechT5HifiGan) and
need the checkpoint locally -- absent weights are reported as a fatal job error."""
from __future__ import annotations

import numpy as np
import torch

from ..output_processor import make_result
from ..registry import resolve
from .common import encode_audio

SAMPLE_RATE = 16000


def txt2audio_diffusion_callback(device_identifier, model_name, **kwargs):
    scheduler_type = kwargs.pop("scheduler_type", resolve("DPMSolverMultistepScheduler"))
    kwargs.pop("pipeline_type", None)
    kwargs["num_inference_steps"] = kwargs.pop("num_inference_steps", 20)
    kwargs["audio_length_in_s"] = kwargs.pop("audio_length_in_s", 10)
    content_type = kwargs.pop("content_type", "audio/mpeg")
    kwargs.pop("outputs", ["primary"])

    from .pipeline_audioldm import AudioLDMPipeline

    from ..utils.weight_cache import pipeline_cache

    # resident between jobs (the reference reloads the pipeline from disk for every clip)
    pipeline = pipeline_cache().get(
        ("audioldm", model_name, str(device_identifier)),
        lambda: AudioLDMPipeline.from_pretrained(model_name, torch_dtype=torch.float16, device=device_identifier))
    pipeline.set_scheduler(scheduler_type)
    audio = pipeline(**kwargs).audios[0]
    buffer, ctype = encode_audio(np.asarray(audio), SAMPLE_RATE, content_type)
    return {"primary": make_result(buffer, None, ctype)}, dict(pipeline.config)
