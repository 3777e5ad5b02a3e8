"""DeepFloyd-IF cascade (reference: swarm/diffusion/diffusion_func_if.py:16-65).  Stage I and II are
pixel-space UNets conditioned on T5-XXL (`models/unet_if.py`, sampled by `if_unet.IFStage`), stage III is the
SD x4 upscaler (`pipelines.py`, mode sd_upscale).  The stage checkpoints (and T5-XXL, ~20 GB) must be present
locally; the callback has already verified that.  The T5 encoder and its tokenizer come from transformers."""
from __future__ import annotations

import torch

from ..utils import checkpoints


def run_if_cascade(device, stage1, stage2, stage3, prompt, negative_prompt=None, num_inference_steps=None,
                   guidance_scale=7.0, generator=None, sizes=(64, 256)):
    from transformers import AutoTokenizer, T5EncoderModel

    from ..utils.weight_cache import pipeline_cache

    # the reference loads and deletes every stage per job to fit a 24 GB card; on 180 GB the T5 encoder (~10 GB in fp16),
    # both IF stages and the x4 upscaler stay resident in the weight cache between jobs
    cache = pipeline_cache()
    root1 = checkpoints.locate_checkpoint(stage1)
    on_gpu = str(device).startswith("cuda")

    def load_text():
        enc = T5EncoderModel.from_pretrained(str(root1 / "text_encoder"), torch_dtype=torch.float16 if on_gpu else torch.float32)
        return AutoTokenizer.from_pretrained(str(root1 / "tokenizer")), enc.to(device).eval()

    tok, enc = cache.get(("if-t5", str(root1), str(device)), load_text)

    def embed(text):
        ids = tok(text, max_length=77, padding="max_length", truncation=True, return_tensors="pt").to(device)
        with torch.no_grad():
            return enc(input_ids=ids["input_ids"], attention_mask=ids["attention_mask"]).last_hidden_state

    cond, unc = embed(prompt), embed(negative_prompt or "")
    from .if_unet import IFStage

    def stage(name):
        return cache.get(("if-stage", name, str(device)), lambda: IFStage.from_checkpoint(checkpoints.locate_checkpoint(name), device))

    s1 = stage(stage1)
    img = s1.sample(cond, unc, size=sizes[0], steps=num_inference_steps or 100, guidance_scale=guidance_scale,
                    generator=generator)
    s2 = stage(stage2)
    img = s2.sample(cond, unc, size=sizes[1], steps=num_inference_steps or 50, guidance_scale=min(guidance_scale, 4.0),
                    generator=generator, low_res=img, noise_level=250)
    from .pipelines import StableDiffusionPipeline

    # stage III: the SD x4 upscaler proper (noise level 100 as in the reference, diffusers' defaults for the rest):
    # the 256 px tensor goes in as the noised conditioning channels, not through an img2img approximation
    up = cache.get(("if-x4", stage3, str(device)),
                   lambda: StableDiffusionPipeline.from_pretrained(stage3, torch_dtype=torch.float16, device=device, family="sd_upscale"))
    up.set_scheduler("ddim", False)
    out = up(prompt=prompt, negative_prompt=negative_prompt, image=img.to(torch.float16), noise_level=100,
             num_inference_steps=num_inference_steps or 75, guidance_scale=9.0, generator=generator)
    return out.images, any(out.nsfw_content_detected or [False])
