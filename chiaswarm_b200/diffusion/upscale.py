"""2x latent upscale stage (reference: swarm/diffusion/upscale.py:6-32 -- `StableDiffusionLatentUpscalePipeline` on
`stabilityai/sd-x2-latent-upscaler`, 20 steps, guidance 0, run on the images of the job that asked for `upscale`).

`LatentUpscalePipeline` is that pipeline on this framework's parts: CLIP text encoder (last layer BEFORE the final norm as
the cross-attention context, the pooled EOS state as part of the time condition), the checkpoint's own VAE, the K-diffusion
UNet of models/unet_k.py (strict key matching) and an Euler sampler over the checkpoint's schedule:

    low  = vae.encode(image) * scaling                       (or the latents handed in)
    cond = nearest_2x(low) / sqrt(noise_level^2 + 1)         noise_level = 0, as in diffusers
    tc   = [ones(64) ; zeros(64) ; pooled]                   the noise-level features of level 0 and the CLIP pooled state
    x    = randn * sigma_max
    per step:  F  = unet([x / sqrt(sigma^2 + 1) ; cond], t = log(sigma) / 4, text, tc)[:, :-1]     (variance channel dropped)
               x0 = x / (sigma^2 + 1) + sigma / sqrt(sigma^2 + 1) * F                              (Karras preconditioning)
               x  = euler_step(x, x0)
    image = vae.decode(x / scaling)

The reference returns only the first image even for batches; every image is returned here.  The
reference reloads the pipeline from disk for every job and runs it with sequential CPU offload; here it stays resident in
the per-device weight cache.  When no checkpoint of that name is on the box (and random init is not allowed) the stage
degrades to a LANCZOS 2x resize so the job still completes with the requested geometry, and says so in the log."""
from __future__ import annotations

import logging
import math
from typing import List

import torch
import torch.nn.functional as F
from PIL import Image

from ..models import configs as arch_configs
from ..models.clip_text import CLIPTextModel
from ..models.unet_k import KUNet
from ..models.vae import AutoencoderKL
from ..ops import schedulers
from ..utils import checkpoints
from ..utils.tokenizer import load_tokenizer
from ..utils.weight_cache import pipeline_cache

UPSCALER = "stabilityai/sd-x2-latent-upscaler"
K_SCHEDULER = dict(num_train_timesteps=1000, beta_start=0.0001, beta_end=0.02, beta_schedule="linear",
                   prediction_type="sample", steps_offset=0)


class LatentUpscalePipeline:
    def __init__(self, unet: KUNet, vae: AutoencoderKL, text_encoder: CLIPTextModel, tokenizer, scheduler_config=None,
                 config=None):
        self.unet, self.vae, self.text_encoder, self.tokenizer = unet, vae, text_encoder, tokenizer
        self.scheduler_config = {**K_SCHEDULER, **{k: v for k, v in (scheduler_config or {}).items() if k in K_SCHEDULER}}
        self.scheduler_config["prediction_type"] = "sample"  # the preconditioned UNet output below IS the denoised sample
        self.config = dict(config or {})
        self.device, self.dtype = torch.device("cpu"), torch.float32

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_pretrained(cls, model_name=UPSCALER, torch_dtype=torch.float16, device="cpu", random_init=None, seed=0,
                        model_root=""):
        from .pipelines import _init_module, _random_init_allowed

        root = checkpoints.locate_checkpoint(model_name, "main", None, model_root)
        if root is not None and (root / "model_index.json").is_file():
            pipe = cls._load(root)
        else:
            key = arch_configs.architecture_for(model_name)
            if random_init is None:
                random_init = _random_init_allowed()
            if key is None or not random_init:
                raise ValueError(f"Model {model_name} is not available locally")
            a = arch_configs.get_architecture(key)
            unet = _init_module(KUNet(**a["unet"]), seed + 1)
            with torch.no_grad():  # (the Fourier frequencies are a stored random vector, not a gain)
                unet.time_proj.weight.copy_(torch.randn(unet.time_proj.weight.shape,
                                                        generator=torch.Generator().manual_seed(seed + 4)))
            pipe = cls(unet, _init_module(AutoencoderKL(**a["vae"]), seed + 2),
                       _init_module(CLIPTextModel(**a["text_encoder"]), seed + 3),
                       load_tokenizer(None, a["text_encoder"]["vocab_size"]), a["scheduler"],
                       {"_class_name": a["_class_name"], "_random_init": True, "_name_or_path": model_name})
        return pipe.to(device, torch_dtype)

    @classmethod
    def _load(cls, root):
        def module(component, ctor, remap=None, strict=False):
            cfg = checkpoints.component_config(root, component)
            if cfg is None:
                raise ValueError(f"no {component}/config.json under {root}")
            m = ctor(**{k: v for k, v in cfg.items() if not k.startswith("_")})
            w = checkpoints.find_weights(root / component, "fp16") or checkpoints.find_weights(root / component)
            if w is None:
                raise ValueError(f"no weights for {component} under {root}")
            sd = checkpoints.load_state_dict(w)
            sd = remap(sd) if remap else sd
            if strict:
                try:
                    m.load_state_dict(sd, strict=True)
                except RuntimeError as e:
                    raise ValueError(f"{component} of {root} does not match the K-diffusion UNet this worker implements: "
                                     f"{str(e)[:400]}") from e
            else:
                checkpoints.load_into(m, sd, strict=False)
            return m

        unet = module("unet", KUNet, strict=True)
        vae = module("vae", AutoencoderKL, checkpoints.remap_vae_keys)
        te = module("text_encoder", CLIPTextModel)
        tok = load_tokenizer(root / "tokenizer", te.config["vocab_size"])
        return cls(unet, vae, te, tok, checkpoints.component_config(root, "scheduler"),
                   checkpoints.read_json(root / "model_index.json"))

    def to(self, device, dtype=None):
        self.device = torch.device(device)
        if dtype is not None:
            self.dtype = dtype if self.device.type == "cuda" else torch.float32
        for m in (self.unet, self.vae, self.text_encoder):
            m.to(self.device, self.dtype).eval()
        return self

    # ------------------------------------------------------------------ conditioning
    def _encode_text(self, texts: List[str]):
        ids = self.tokenizer(texts).to(self.device)
        n_layers = len(self.text_encoder.text_model.encoder.layers)
        hidden, pooled = self.text_encoder(ids, output_hidden_state=n_layers)  # hidden_states[-1]: before the final norm
        return hidden.to(self.dtype), pooled.to(self.dtype)

    def _low_res_latents(self, image, generator) -> torch.Tensor:
        """PIL image(s) -> VAE latents * scaling; a [b, 4, h, w] tensor is taken as latents already."""
        from .pipelines import pil_to_tensor

        if isinstance(image, torch.Tensor):
            if image.dim() == 3:
                image = image[None]
            if image.shape[1] != 3:
                return image.to(self.device, self.dtype)
            pixels = image.to(self.device, self.dtype)
        else:
            images = list(image) if isinstance(image, (list, tuple)) else [image]
            vae_scale = 2 ** (len(self.vae.config["block_out_channels"]) - 1)
            levels = len(self.unet.config["block_out_channels"])
            fit = max(vae_scale, vae_scale * 2 ** (levels - 1) // 2)  # the 2x latents must halve cleanly at every level
            w, h = images[0].size
            w, h = max(2 * fit, w - w % fit), max(2 * fit, h - h % fit)
            pixels = torch.cat([pil_to_tensor(im.convert("RGB").resize((w, h), Image.Resampling.LANCZOS)
                                              if im.size != (w, h) else im.convert("RGB"), self.device, self.dtype)
                                for im in images])
        return self.vae.encode(pixels, generator, True) * self.vae.config["scaling_factor"]

    # ------------------------------------------------------------------ main entry
    @torch.no_grad()
    def __call__(self, prompt="", image=None, num_inference_steps=20, guidance_scale=0.0, negative_prompt=None,
                 generator=None, output_type="pil"):
        from .pipelines import PipelineOutput, tensor_to_pil

        if image is None:
            raise ValueError("the latent upscaler needs an `image` (PIL image, list of images, or latents)")
        low = self._low_res_latents(image, generator)
        b = low.shape[0]
        div = 2 ** (len(self.unet.config["block_out_channels"]) - 1)
        if any((2 * d) % div or (2 * d) // div < 2 for d in low.shape[-2:]):
            raise ValueError(f"latents of {tuple(low.shape[-2:])} cannot be upscaled: twice each side must be a multiple of "
                             f"{div} and at least {2 * div} (the UNet halves it {len(self.unet.config['block_out_channels']) - 1} times)")
        prompts = [prompt] * b if isinstance(prompt, str) else list(prompt)
        if len(prompts) == 1 and b > 1:
            prompts = prompts * b
        if len(prompts) != b:
            raise ValueError(f"`prompt` has batch size {len(prompts)} and `image` has batch size {b}")
        cfg = float(guidance_scale) > 1.0
        text, pooled = self._encode_text(prompts)
        if cfg:
            neg = negative_prompt if negative_prompt is not None else ""
            negs = [neg] * b if isinstance(neg, str) else list(neg)
            utext, upooled = self._encode_text(negs)
            text, pooled = torch.cat([utext, text]), torch.cat([upooled, pooled])
        n_cond = 2 if cfg else 1

        noise_level = 0.0
        cond = F.interpolate(low.float(), scale_factor=2, mode="nearest") / math.sqrt(noise_level ** 2 + 1)
        cond = torch.cat([cond] * n_cond).to(self.dtype)
        width = self.unet.config.get("time_cond_proj_dim") or 0
        feat = width - pooled.shape[-1]
        if width and (feat < 0 or feat % 2):
            raise ValueError(f"time_cond_proj_dim {width} does not fit a pooled text state of width {pooled.shape[-1]}")
        tc = None
        if width:  # the noise-level embedding of level 0, as the pipeline the checkpoint was trained with builds it
            tc = torch.cat([torch.ones(pooled.shape[0], feat // 2, device=self.device, dtype=self.dtype),
                            torch.zeros(pooled.shape[0], feat // 2, device=self.device, dtype=self.dtype), pooled], dim=1)

        plan = schedulers.make_plan("euler", int(num_inference_steps), self.scheduler_config, False)
        shape = (b, self.unet.config["out_channels"] - 1, low.shape[-2] * 2, low.shape[-1] * 2)
        x = torch.randn(shape, generator=generator, device=self.device, dtype=self.dtype).float() * plan.init_noise_sigma
        for i in range(plan.num_steps):
            sigma = float(plan.sigmas[i])
            c_in = 1.0 / math.sqrt(sigma * sigma + 1.0)
            inp = torch.cat([torch.cat([x] * n_cond).to(self.dtype) * c_in, cond], dim=1)
            out = self.unet(inp, math.log(sigma) * 0.25, text, tc)[:, :-1].float()
            x0 = torch.cat([x] * n_cond) * (c_in * c_in) + (sigma * c_in) * out
            x0 = schedulers.guided(x0, n_cond, float(guidance_scale))
            x, _, _ = schedulers.apply_step_torch(plan, i, x, x0, None, None)
        if output_type == "latent":
            return PipelineOutput(images=x.to(self.dtype))
        pixels = self.vae.decode((x / self.vae.config["scaling_factor"]).to(self.dtype))
        return PipelineOutput(images=tensor_to_pil(pixels))


def upscale_image(images, device_identifier, prompt, num_images_per_prompt, generator, info: dict = None) -> List[Image.Image]:
    """`info` (optional dict) receives `{"upscaler": "sd-x2-latent-upscaler" | "lanczos"}`: which of the two ran goes into
    the job's `pipeline_config`, so a resize standing in for the network is visible on the wire, not only in the log."""
    print("Upscaling...")
    info = {} if info is None else info
    info["upscaler"] = "sd-x2-latent-upscaler"
    try:
        pipe = pipeline_cache().get((UPSCALER, device_identifier),
                                    lambda: LatentUpscalePipeline.from_pretrained(UPSCALER, torch.float16, device_identifier))
    except ValueError as e:
        if "not available locally" not in str(e):
            raise
        logging.warning("%s is not on this worker: 2x LANCZOS resize instead of the latent upscaler", UPSCALER)
        info["upscaler"] = "lanczos"
        return [im.resize((im.width * 2, im.height * 2), Image.Resampling.LANCZOS) for im in images]
    return list(pipe(prompt=prompt, image=list(images), num_inference_steps=20, guidance_scale=0, generator=generator).images)
