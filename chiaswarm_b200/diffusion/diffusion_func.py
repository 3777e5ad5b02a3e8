"""Stable-Diffusion job callback: txt2img / img2img / inpaint / ControlNet / instruct-pix2pix /
LoRA / textual inversion / 2x upscale (reference: swarm/diffusion/diffusion_func.py:14-124).

Same signature `cb(device_identifier, model_name, **kwargs) -> (artifacts, pipeline_config)` and
same kwarg handling.  What changes on B200:
  * pipelines stay resident in a per-device weight cache keyed by (model, revision, variant,
    controlnet, lora, textual inversion) instead of `from_pretrained` on every job;
  * the reference's low-memory switches (xformers / VAE slicing+tiling / sequential CPU offload,
    triggered by VRAM < 12-16 GB or num_images_per_prompt > 1, lines 76-94) are accepted and
    ignored: 180 GB of HBM3e keeps everything resident (Q4);
  * `pipeline_config` is a plain dict copy of the model index (Q6);
  * a job may own a gang of 2 / 4 / 8 GPUs (`parallel.context.job_gang()`), see the call below.
"""
from __future__ import annotations

import logging

import torch

from ..output_processor import BatchArtifacts, OutputProcessor
from ..parallel.context import job_gang
from ..registry import resolve
from ..utils.weight_cache import pipeline_cache
from .pipelines import build_pipeline
from .upscale import upscale_image


def diffusion_callback(device_identifier, model_name, **kwargs):
    scheduler_type = kwargs.pop("scheduler_type", resolve("DPMSolverMultistepScheduler"))
    pipeline_type = kwargs.pop("pipeline_type", resolve("DiffusionPipeline"))
    num_images_per_prompt = kwargs.get("num_images_per_prompt", 1)
    upscale = kwargs.pop("upscale", False)
    textual_inversion = kwargs.pop("textual_inversion", None)
    lora = kwargs.pop("lora", None)
    kwargs.pop("supports_xformers", True)  # no xformers here: attention is our own kernel
    cross_attention_scale = kwargs.pop("cross_attention_scale", 1.0)
    batch_seeds = kwargs.pop("batch_seeds", None)  # several hive jobs in one denoising run: one seed and one result per job

    output_processor = OutputProcessor(kwargs.pop("outputs", ["primary"]), kwargs.pop("content_type", "image/jpeg"))

    controlnet_name = kwargs.pop("controlnet_model_name", None)
    controlnet_revision = kwargs.pop("controlnet_revision", "main")
    if controlnet_name is not None and kwargs.pop("save_preprocessed_input", False):
        output_processor.add_other_outputs("preprocessed_input", [kwargs.get("image")])
    kwargs.pop("save_preprocessed_input", None)

    revision = kwargs.pop("revision", "main")
    variant = kwargs.pop("variant", None)

    def load():
        controlnet = None
        if controlnet_name is not None:
            from .loaders import load_controlnet

            controlnet = load_controlnet(controlnet_name, controlnet_revision, device_identifier)
        pipe = build_pipeline(pipeline_type, model_name, revision=revision, variant=variant,
                              torch_dtype=torch.float16, controlnet=controlnet, device=device_identifier)
        if textual_inversion is not None:
            try:
                from .loaders import load_textual_inversion

                load_textual_inversion(pipe, textual_inversion)
            except Exception as e:
                raise ValueError(
                    f"Textual inversion\n{textual_inversion}\nis incompatible with\n{model_name}\n{lora}\n\n{e}") from e
        if lora is not None:
            try:
                from .loaders import load_lora

                load_lora(pipe, lora, cross_attention_scale)
            except Exception as e:
                raise ValueError(f"Could not load lora \n{lora}\nIt might be incompatible with {model_name}\n{e}") from e
        return pipe

    key = (model_name, revision, variant, getattr(pipeline_type, "name", str(pipeline_type)), controlnet_name,
           controlnet_revision, lora, cross_attention_scale if lora else None, textual_inversion, device_identifier)
    pipeline = pipeline_cache().get(key, load)
    pipeline.set_scheduler(scheduler_type, use_karras_sigmas=True)

    # a job that leased a gang (parallel/gang_worker.py) runs SPMD on every rank of it: one large image goes
    # patch-parallel, a batch is split across the ranks; without a lease this is the reference's one-GPU call
    if batch_seeds is not None:
        kwargs["seeds"] = batch_seeds
    p = pipeline(gang=job_gang(), **kwargs)

    pipeline_config = dict(pipeline.config)
    # any flagged image flags the whole result
    flagged = p.nsfw_content_detected
    if flagged is not None and ((isinstance(flagged, bool) and flagged) or (isinstance(flagged, list) and any(flagged))):
        pipeline_config["nsfw"] = True

    note = getattr(kwargs.get("image"), "info", {}).get("controlnet_annotator")
    if note:  # the control image came from a classical stand-in, not from the learned annotator (controlnet/input_processor.py)
        pipeline_config["controlnet_annotator"] = note
    if batch_seeds is not None:
        flags = p.nsfw_content_detected if isinstance(p.nsfw_content_detected, list) else [False] * len(p.images)
        pipeline_config.pop("nsfw", None)
        per_job = []
        for image in p.images:
            one = OutputProcessor(output_processor.output_list, output_processor.main_content_type)
            one.add_outputs([image])
            per_job.append(one.get_results())
        return BatchArtifacts(per_job, flags), pipeline_config
    images = p.images
    if upscale:
        images = upscale_image(images, device_identifier, kwargs.get("prompt", ""), num_images_per_prompt,
                               kwargs.get("generator"), info=pipeline_config)
    output_processor.add_outputs(images)
    logging.debug("diffusion_callback done: %d image(s)", len(images))
    return output_processor.get_results(), pipeline_config
