"""Bark text-to-speech callback (reference: swarm/audio/bark.py:11-38 -- `bark.preload_models`
+ `generate_audio`).  Bark is three GPT-style decoders plus EnCodec (off the
named hot path).  The `bark` package is not installable here; `transformers` ships the same model
as BarkModel, which is used when the checkpoint is cached locally."""
from __future__ import annotations

import numpy as np
import torch

from ..output_processor import make_result
from ..utils import checkpoints
from .common import encode_audio

SAMPLE_RATE = 24000


def generate_audio(prompt, device, model_name="suno/bark"):
    root = checkpoints.locate_checkpoint(model_name)
    if root is None:
        raise ValueError(f"Model {model_name} is not available locally")
    from transformers import AutoProcessor, BarkModel

    from ..utils.weight_cache import pipeline_cache

    dtype = torch.float16 if str(device).startswith("cuda") else torch.float32
    # resident between jobs (the reference's `preload_models()` re-reads them for every job)
    processor, model = pipeline_cache().get(
        ("bark", str(root), str(device)),
        lambda: (AutoProcessor.from_pretrained(str(root)), BarkModel.from_pretrained(str(root), torch_dtype=dtype).to(device).eval()))
    inputs = processor(prompt).to(device)
    with torch.no_grad():
        audio = model.generate(**inputs)
    return audio[0].float().cpu().numpy()


def bark_diffusion_callback(device_identifier, model_name, **kwargs):
    content_type = kwargs.pop("content_type", "audio/mpeg")
    kwargs.pop("outputs", ["primary"])
    audio = generate_audio(kwargs["prompt"], device_identifier, model_name)
    buffer, ctype = encode_audio(np.asarray(audio), SAMPLE_RATE, content_type)
    return {"primary": make_result(buffer, None, ctype)}, {}
