"""AudioLDM driver (reference: `AudioLDMPipeline`, swarm/audio/audioldm.py:19-24).  CLAP text
embedding -> class-conditioned UNet2DCondition over mel-spectrogram latents (this framework's UNet:
`class_embed_type="simple_projection"`, `class_embeddings_concat`, no cross-attention width -- every attn2 is a second
self-attention, as `encoder_hidden_states=None` makes it in diffusers) -> VAE decode
to a mel spectrogram -> HiFi-GAN vocoder.  CLAP and the vocoder are transformers models
(ClapTextModelWithProjection, SpeechT5HifiGan) loaded from the local checkpoint; without a local
checkpoint the job fails as a fatal (non-resubmittable) error, exactly like any other missing model."""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from ..models.unet2d import UNet2DConditionModel
from ..models.vae import AutoencoderKL
from ..ops import schedulers
from ..utils import checkpoints


@dataclass
class AudioOutput:
    audios: np.ndarray


class AudioLDMPipeline:
    def __init__(self, root, device, dtype):
        from transformers import ClapTextModelWithProjection, RobertaTokenizerFast, SpeechT5HifiGan

        self.device, self.dtype = torch.device(device), dtype if str(device).startswith("cuda") else torch.float32
        self.config = checkpoints.read_json(root / "model_index.json")

        def module(component, ctor, remap=None):
            cfg = checkpoints.component_config(root, component)
            m = ctor(**{k: v for k, v in cfg.items() if not k.startswith("_")})
            sd = checkpoints.load_state_dict(checkpoints.find_weights(root / component))
            checkpoints.load_into(m, remap(sd) if remap else sd, strict=False)
            return m.to(self.device, self.dtype).eval()

        self.unet = module("unet", UNet2DConditionModel)
        self.vae = module("vae", AutoencoderKL, checkpoints.remap_vae_keys)
        self.tokenizer = RobertaTokenizerFast.from_pretrained(str(root / "tokenizer"))
        self.text_encoder = ClapTextModelWithProjection.from_pretrained(str(root / "text_encoder")).to(self.device, self.dtype)
        self.vocoder = SpeechT5HifiGan.from_pretrained(str(root / "vocoder")).to(self.device, self.dtype)
        if self.unet.class_embedding is None:
            raise ValueError("not an AudioLDM UNet: it has no class embedding for the CLAP text vector")
        self.scheduler_config = checkpoints.component_config(root, "scheduler") or schedulers.SD_SCHEDULER_CONFIG
        self.scheduler_kind = "dpmpp_2m"

    @classmethod
    def from_pretrained(cls, model_name, torch_dtype=torch.float16, device="cpu"):
        root = checkpoints.locate_checkpoint(model_name)
        if root is None or not (root / "model_index.json").is_file():
            raise ValueError(f"Model {model_name} is not available locally")
        return cls(root, device, torch_dtype)

    def set_scheduler(self, scheduler_type, use_karras_sigmas=False):
        self.scheduler_kind = getattr(scheduler_type, "kind", scheduler_type)

    @torch.no_grad()
    def __call__(self, prompt="", negative_prompt=None, num_inference_steps=10, audio_length_in_s=5.12,
                 guidance_scale=2.5, generator=None, **ignored):
        vocoder_up = int(np.prod(self.vocoder.config.upsample_rates)) / self.vocoder.config.sampling_rate
        vae_scale = 2 ** (len(self.vae.config["block_out_channels"]) - 1)
        height = int(audio_length_in_s / vocoder_up)
        height += (-height) % vae_scale
        n_mel = self.vocoder.config.model_in_dim

        def clap(text):
            tok = self.tokenizer([text], padding="max_length", truncation=True, return_tensors="pt").to(self.device)
            e = self.text_encoder(**tok).text_embeds
            return torch.nn.functional.normalize(e, dim=-1).to(self.dtype)

        do_cfg = guidance_scale > 1.0
        cond = clap(prompt)
        emb = torch.cat([clap(negative_prompt or ""), cond]) if do_cfg else cond
        plan = schedulers.make_plan(self.scheduler_kind, int(num_inference_steps), self.scheduler_config, False)
        shape = (1, self.unet.config["in_channels"], height // vae_scale, n_mel // vae_scale)
        x = torch.randn(shape, generator=generator, device=self.device, dtype=self.dtype).float() * plan.init_noise_sigma
        x0_prev, scale, n_cond = None, plan.first_scale, (2 if do_cfg else 1)
        for i in range(plan.num_steps):
            t = torch.tensor(float(plan.timesteps[i]), device=self.device)
            model_in = torch.cat([x] * n_cond).to(self.dtype) * scale
            # no text sequence: the CLAP vector is the class label (added or concatenated to the time embedding, as the
            # checkpoint's config says) and every attn2 is a second self-attention (encoder_hidden_states=None)
            out = self.unet(model_in, t, None, class_labels=emb)
            out = schedulers.guided(out.float(), n_cond, float(guidance_scale))
            x, x0_prev, scale = schedulers.apply_step_torch(plan, i, x, out, x0_prev, None)
        mel = self.vae.decode((x / self.vae.scaling_factor).to(self.dtype))
        wave = self.vocoder(mel.squeeze(1) if mel.dim() == 4 else mel).float().cpu().numpy()
        n = int(audio_length_in_s * self.vocoder.config.sampling_rate)
        return AudioOutput(audios=wave[:, :n])
