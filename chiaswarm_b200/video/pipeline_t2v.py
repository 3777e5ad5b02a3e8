"""Text-to-video driver (zeroscope / modelscope `TextToVideoSDPipeline` in the reference,
swarm/video/tx2vid.py:24-57; spec in SURVEY.md appendix A.6): latents (B, 4, F, H/8, W/8), the
3-D UNet per step with classifier-free guidance, frames decoded independently by the VAE, output a
list of uint8 HWC frames (`p.frames`).  The VAE decode of the frames runs on the sm_100a engine
when the pipeline is on a GPU; the 3-D UNet runs its module graph (spatial layers are the 2-D
ones, the temporal layers are (3,1,1) convs and attention over the frame axis)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import numpy as np
import torch

from ..diffusion.pipelines import (_init_module, _random_init_allowed, default_backend)
from ..models import configs as arch_configs
from ..models.clip_text import CLIPTextModel
from ..models.unet3d import UNet3DConditionModel
from ..models.vae import AutoencoderKL
from ..ops import schedulers
from ..utils import checkpoints
from ..utils.tokenizer import load_tokenizer


@dataclass
class VideoOutput:
    frames: List[np.ndarray]
    latents: Optional[torch.Tensor] = None


class TextToVideoPipeline:
    def __init__(self, unet, vae, text_encoder, tokenizer, scheduler_config, config=None):
        self.unet, self.vae, self.text_encoder, self.tokenizer = unet, vae, text_encoder, tokenizer
        self.scheduler_config = dict(scheduler_config)
        self.scheduler_kind, self.use_karras_sigmas = "dpmpp_2m", False
        self.config = dict(config or {})
        self.device, self.dtype = torch.device("cpu"), torch.float32
        self.backend = "torch"
        self._vae_engine = None
        self._engine = None
        self.controlnet = None

    @classmethod
    def from_pretrained(cls, model_name, revision="main", variant=None, torch_dtype=torch.float16, device="cpu",
                        seed=0):
        root = checkpoints.locate_checkpoint(model_name, revision, variant)
        if root is not None and (root / "model_index.json").is_file():
            def module(component, ctor, remap=None):
                cfg = checkpoints.component_config(root, component)
                m = ctor(**{k: v for k, v in cfg.items() if not k.startswith("_")})
                sd = checkpoints.load_state_dict(checkpoints.find_weights(root / component, variant))
                checkpoints.load_into(m, remap(sd) if remap else sd, strict=False)
                return m

            unet = module("unet", UNet3DConditionModel)
            vae = module("vae", AutoencoderKL, checkpoints.remap_vae_keys)
            te = module("text_encoder", CLIPTextModel)
            tok = load_tokenizer(root / "tokenizer", te.config["vocab_size"])
            sched = checkpoints.component_config(root, "scheduler") or schedulers.SD_SCHEDULER_CONFIG
            pipe = cls(unet, vae, te, tok, sched, checkpoints.read_json(root / "model_index.json"))
        else:
            key = arch_configs.architecture_for(model_name)
            if key not in ("zeroscope", "tiny-t2v"):
                key = "zeroscope" if _random_init_allowed() else None
            if key is None or not _random_init_allowed():
                raise ValueError(f"Model {model_name} is not available locally")
            pipe = cls.from_architecture(key, seed=seed)
            pipe.config["_name_or_path"] = model_name
        return pipe.to(device, torch_dtype)

    @classmethod
    def from_architecture(cls, key="zeroscope", seed=0, unet_overrides=None):
        a = arch_configs.get_architecture(key)
        ucfg = dict(a["unet"])
        ucfg.update(unet_overrides or {})
        unet = _init_module(UNet3DConditionModel(**ucfg), seed + 1)
        vae = _init_module(AutoencoderKL(**a["vae"]), seed + 2)
        te = _init_module(CLIPTextModel(**a["text_encoder"]), seed + 3)
        tok = load_tokenizer(None, a["text_encoder"]["vocab_size"])
        return cls(unet, vae, te, tok, a["scheduler"], {"_class_name": "TextToVideoSDPipeline", "_random_init": True})

    def to(self, device, dtype=None):
        self.device = torch.device(device)
        if dtype is not None:
            self.dtype = dtype if self.device.type == "cuda" else torch.float32
        for m in (self.unet, self.vae, self.text_encoder):
            m.to(self.device, self.dtype).eval()
        self.backend = default_backend(self.device)
        self._vae_engine = None
        self._engine = None
        return self

    def set_scheduler(self, scheduler_type, use_karras_sigmas=True):
        kind = getattr(scheduler_type, "kind", scheduler_type)
        if kind not in schedulers.KINDS:
            raise ValueError(f"Unsupported scheduler {scheduler_type}")
        self.scheduler_kind, self.use_karras_sigmas = kind, bool(use_karras_sigmas)

    def _decode(self, z):
        """z: (N, 4, h, w) -> uint8 (N, H, W, 3)."""
        if self.backend == "sm100" and self.device.type == "cuda":
            return self._sm100().decode(z)
        img = self.vae.decode(z.to(self.dtype)).float()
        return ((img / 2 + 0.5).clamp(0, 1) * 255).round().to(torch.uint8).permute(0, 2, 3, 1)

    def _sm100(self, gang=None):
        """Engine per gang identity (None = this GPU alone); a gang's engine frame-splits the clip
        (engine/patch_engine.py::_FrameStepGraph)."""
        if gang is not None and gang.world <= 1:
            gang = None
        if self._engine is None:
            self._engine = {}
        key = None if gang is None else (gang.global_ranks or ("world", gang.world))
        eng = self._engine.get(key)
        if eng is None:
            from ..engine.denoiser import SM100Denoiser

            if gang is None:
                eng = SM100Denoiser(self)
            else:
                from ..engine.patch_engine import PatchDenoiser

                eng = PatchDenoiser(self, gang)
            self._engine[key] = eng
        return eng

    def _denoise_sm100(self, x, embeds, plan, n_cond, guidance_scale, generator, gang=None):
        """Whole denoising loop on the engine: frames ride in the batch dimension of the captured plan."""
        nf = x.shape[2]
        frames = x[0].permute(1, 0, 2, 3).contiguous()                     # (F, 4, h, w)
        per_frame = embeds.repeat_interleave(nf, dim=0)                    # cond-major, one row per frame
        noise = None
        if plan.needs_noise:
            noise = torch.randn((plan.num_steps,) + tuple(frames.shape), generator=generator, device=self.device,
                                dtype=self.dtype)
        out = self._sm100(gang).run(frames, per_frame, plan, n_cond, float(guidance_scale), 0.0, None, None, 1.0, None,
                                    noise, None, num_frames=nf)
        return out.permute(1, 0, 2, 3).unsqueeze(0)

    @torch.no_grad()
    def __call__(self, prompt="", negative_prompt=None, height=None, width=None, num_frames=16,
                 num_inference_steps=50, guidance_scale=9.0, generator=None, output_type="np", gang=None, **ignored):
        """`gang` (parallel.runtime.Gang; every rank makes this call with the same seed): the clip is denoised
        frame-split over the gang and its frames are decoded by the VAE in per-rank shares, then all-gathered."""
        if int(num_frames) < 1 or int(num_inference_steps) < 1:
            raise ValueError(f"num_frames and num_inference_steps must be positive (got {num_frames}, {num_inference_steps})")
        vae_scale = 2 ** (len(self.vae.config["block_out_channels"]) - 1)
        sample = self.unet.config.get("sample_size", 32)
        height = int(height or sample * vae_scale)
        width = int(width or sample * vae_scale)
        if height % vae_scale or width % vae_scale:
            raise ValueError(f"`height` and `width` have to be divisible by {vae_scale}")
        fit = vae_scale * 2 ** (len(self.unet.config["block_out_channels"]) - 1)
        if height <= 0 or width <= 0 or height % fit or width % fit:
            # (a frame size the 3-D UNet cannot halve at every level would only fail later, inside the skip concatenation)
            raise ValueError(f"`height` and `width` have to be positive multiples of {fit} for this model (got {height} x {width})")
        do_cfg = guidance_scale > 1.0
        ids = self.tokenizer([prompt]).to(self.device)
        cond = self.text_encoder(ids)[0]
        if do_cfg:
            unc = self.text_encoder(self.tokenizer([negative_prompt or ""]).to(self.device))[0]
            embeds = torch.cat([unc, cond]).to(self.dtype)
        else:
            embeds = cond.to(self.dtype)
        plan = schedulers.make_plan(self.scheduler_kind, int(num_inference_steps), self.scheduler_config,
                                    self.use_karras_sigmas)
        shape = (1, self.unet.config["in_channels"], int(num_frames), height // vae_scale, width // vae_scale)
        x = torch.randn(shape, generator=generator, device=self.device, dtype=self.dtype).float() * plan.init_noise_sigma
        x0_prev, scale = None, plan.first_scale
        n_cond = 2 if do_cfg else 1
        engine = self.backend == "sm100" and self.device.type == "cuda"
        if engine:
            x = self._denoise_sm100(x, embeds, plan, n_cond, guidance_scale, generator, gang)
        for i in range(0 if engine else plan.num_steps):
            t = torch.tensor(float(plan.timesteps[i]), device=self.device)
            model_in = torch.cat([x] * n_cond).to(self.dtype) * scale
            out = self.unet(model_in, t, embeds)
            out = schedulers.guided(out.float(), n_cond, float(guidance_scale))
            noise = None
            if plan.needs_noise:
                noise = torch.randn(shape, generator=generator, device=self.device, dtype=self.dtype)
            x, x0_prev, scale = schedulers.apply_step_torch(plan, i, x, out, x0_prev, noise)
        if output_type == "latent":
            return VideoOutput(frames=[], latents=x)
        z = (x[0].permute(1, 0, 2, 3) / self.vae.scaling_factor).contiguous()  # (F, 4, h, w)
        frames = []
        if gang is not None and gang.world > 1:  # frames are independent for the VAE: every rank decodes its share
            mine = gang.shard(z.shape[0])
            part = self._decode(z[mine.start:mine.stop]) if len(mine) else \
                torch.zeros((0, height, width, 3), dtype=torch.uint8, device=self.device)
            return VideoOutput(frames=list(gang.all_gather_cat(part.contiguous()).cpu().numpy()), latents=x)
        for chunk in z.split(8 if not engine else 24):
            frames.extend(list(self._decode(chunk).cpu().numpy()))
        return VideoOutput(frames=frames, latents=x)
