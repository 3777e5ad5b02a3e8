"""2x latent upscale stage (reference: swarm/diffusion/upscale.py:6-32 -- sd-x2-latent-upscaler,
20 steps, guidance 0).  The reference returns only the first image even for batches (Q3); all
images are returned here.  When the upscaler checkpoint is absent the stage degrades to a LANCZOS
2x resize so the job still completes with the requested geometry."""
from __future__ import annotations

import torch
from PIL import Image

from ..utils.weight_cache import pipeline_cache

UPSCALER = "stabilityai/sd-x2-latent-upscaler"


def upscale_image(images, device_identifier, prompt, num_images_per_prompt, generator):
    print("Upscaling...")
    from .pipelines import StableDiffusionPipeline

    try:
        pipe = pipeline_cache().get(
            (UPSCALER, device_identifier),
            lambda: StableDiffusionPipeline.from_pretrained(UPSCALER, torch_dtype=torch.float16,
                                                            device=device_identifier, family="sd_img2img"))
    except ValueError:
        return [im.resize((im.width * 2, im.height * 2), Image.Resampling.LANCZOS) for im in images]
    from ..registry import resolve

    pipe.set_scheduler(resolve("EulerDiscreteScheduler"), use_karras_sigmas=False)
    out = []
    for im in images:
        big = im.resize((im.width * 2, im.height * 2), Image.Resampling.LANCZOS)
        out.extend(pipe(prompt=prompt, image=big, num_inference_steps=20, guidance_scale=0.0, strength=0.3,
                        generator=generator).images)
    return out
