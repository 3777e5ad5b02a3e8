"""DeepFloyd-IF cascade callback (reference: swarm/diffusion/diffusion_func_if.py:14-92):
IF stage I (64 px, T5-XXL text) -> IF-II-L (256 px) -> stable-diffusion-x4-upscaler (noise level
100).  Reference bugs fixed (Q5): the negative prompt is the job's negative prompt (not the
prompt), nothing is written to the working directory, steps / guidance are honoured, and the NSFW
flag comes from the pipeline output.

The pixel-space IF UNets are `models/unet_if.py` (the IF block family, checkpoint-compatible names, strict
loading); they and T5-XXL (~20 GB) are only runnable from local checkpoints.  `SDAAS_IF_SIZES` ("64,256") sets the
stage resolutions (tests run the cascade at toy sizes)."""
from __future__ import annotations

from ..output_processor import OutputProcessor
from ..utils import checkpoints

STAGE_2 = "DeepFloyd/IF-II-L-v1.0"
STAGE_3 = "stabilityai/stable-diffusion-x4-upscaler"


def diffusion_if_callback(device_identifier, model_name, **kwargs):
    missing = [m for m in (model_name, STAGE_2, STAGE_3) if checkpoints.locate_checkpoint(m) is None]
    if missing:
        raise ValueError(f"DeepFloyd IF needs local checkpoints; missing: {', '.join(missing)}")
    import os

    from .pipeline_if import run_if_cascade

    sizes = tuple(int(v) for v in os.environ.get("SDAAS_IF_SIZES", "64,256").split(","))
    images, nsfw = run_if_cascade(device_identifier, model_name, STAGE_2, STAGE_3, sizes=sizes,
                                  prompt=kwargs.get("prompt", ""), negative_prompt=kwargs.get("negative_prompt"),
                                  num_inference_steps=kwargs.get("num_inference_steps"),
                                  guidance_scale=kwargs.get("guidance_scale", 7.0), generator=kwargs.get("generator"))
    pipeline_config = {"nsfw": True} if nsfw else {}
    output_processor = OutputProcessor(kwargs.pop("outputs", ["primary"]), kwargs.pop("content_type", "image/jpeg"))
    output_processor.add_outputs(images)
    return output_processor.get_results(), pipeline_config
