"""Stable Diffusion safety checker: CLIP ViT-L/14 image embedding against 17 concept and 3
"special care" embeddings; an image is flagged when any adjusted cosine score is positive.  The reference reads
the result as `p.nsfw_content_detected` and reports it on the wire (swarm/diffusion/diffusion_func.py:97-111);
diffusers also blacks the flagged images out, and so does `check()` here.

Checkpoint layout (`safety_checker/` of a diffusers repo): `vision_model.*`, `visual_projection.weight`,
`concept_embeds` [17, 768], `special_care_embeds` [3, 768], `concept_embeds_weights` [17],
`special_care_embeds_weights` [3]; `feature_extractor/preprocessor_config.json` gives the resize / crop /
normalisation of the CLIP image processor."""
from __future__ import annotations

from typing import List, Sequence

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from PIL import Image

from ..models.clip_vision import CLIPVisionModel

CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)


class ClipImageProcessor:
    """Resize the short side (bicubic), centre-crop, rescale to [0,1], normalise -- the CLIP preprocessing."""

    def __init__(self, size=224, crop_size=224, image_mean=CLIP_MEAN, image_std=CLIP_STD, **_ignored):
        def edge(v, default):
            if isinstance(v, dict):
                return int(v.get("shortest_edge") or v.get("height") or default)
            return int(v or default)

        self.size, self.crop = edge(size, 224), edge(crop_size, 224)
        self.mean = torch.tensor(list(image_mean), dtype=torch.float32).view(1, 3, 1, 1)
        self.std = torch.tensor(list(image_std), dtype=torch.float32).view(1, 3, 1, 1)

    def __call__(self, images: Sequence[Image.Image]) -> torch.Tensor:
        out = []
        for im in images:
            im = im.convert("RGB")
            w, h = im.size
            # short side to `size`, long side truncated (transformers' get_resize_output_image_size)
            w2, h2 = (self.size, int(self.size * h / w)) if w <= h else (int(self.size * w / h), self.size)
            im = im.resize((max(self.crop, w2), max(self.crop, h2)), Image.Resampling.BICUBIC)
            w, h = im.size
            l, t = (w - self.crop) // 2, (h - self.crop) // 2
            im = im.crop((l, t, l + self.crop, t + self.crop))
            out.append(torch.from_numpy(np.asarray(im, dtype=np.float32) / 255.0).permute(2, 0, 1))
        x = torch.stack(out)
        return (x - self.mean) / self.std


class StableDiffusionSafetyChecker(nn.Module):
    def __init__(self, vision_config=None, projection_dim=768, **_ignored):
        super().__init__()
        vc = dict(vision_config or {})
        vc = {k: v for k, v in vc.items() if k in ("hidden_size", "intermediate_size", "num_hidden_layers",
                                                   "num_attention_heads", "image_size", "patch_size", "hidden_act",
                                                   "layer_norm_eps")}
        self.vision_model = CLIPVisionModel(**vc)  # keys: vision_model.vision_model.* as in the checkpoint
        self.config = {"vision_config": self.vision_model.config, "projection_dim": projection_dim}
        self.visual_projection = nn.Linear(self.vision_model.config["hidden_size"], projection_dim, bias=False)
        self.concept_embeds = nn.Parameter(torch.ones(17, projection_dim), requires_grad=False)
        self.special_care_embeds = nn.Parameter(torch.ones(3, projection_dim), requires_grad=False)
        self.concept_embeds_weights = nn.Parameter(torch.ones(17), requires_grad=False)
        self.special_care_embeds_weights = nn.Parameter(torch.ones(3), requires_grad=False)

    @torch.no_grad()
    def forward(self, pixels: torch.Tensor) -> torch.Tensor:
        """pixels: CLIP-normalised [b, 3, 224, 224] -> bool [b] (True = flagged)."""
        emb = self.visual_projection(self.vision_model(pixels)).float()
        emb = F.normalize(emb, dim=-1)
        special = emb @ F.normalize(self.special_care_embeds.float(), dim=-1).t()
        concept = emb @ F.normalize(self.concept_embeds.float(), dim=-1).t()
        # images that trip a special-care concept are judged with a stricter threshold on the others
        special_hit = (special - self.special_care_embeds_weights.float() > 0).any(dim=1)
        adjustment = special_hit.float() * 0.01
        scores = concept - self.concept_embeds_weights.float() + adjustment[:, None]
        return (scores > 0).any(dim=1)


class SafetyChecker:
    """Feature extractor + checker bound to a device: `check(images)` -> (images with flagged ones blacked
    out, flags)."""

    def __init__(self, model: StableDiffusionSafetyChecker, processor: ClipImageProcessor, device="cpu",
                 dtype=torch.float32):
        self.model, self.processor = model, processor
        self.to(device, dtype)

    def to(self, device, dtype=None):
        self.device = torch.device(device)
        if dtype is not None:
            self.dtype = dtype if self.device.type == "cuda" else torch.float32
        self.model.to(self.device, self.dtype).eval()
        return self

    def __call__(self, images: List[Image.Image]) -> List[bool]:
        if not images:
            return []
        pixels = self.processor(images).to(self.device, self.dtype)
        return [bool(f) for f in self.model(pixels).cpu().tolist()]

    def check(self, images: List[Image.Image]):
        flags = self(images)
        cleaned = [Image.new("RGB", im.size, (0, 0, 0)) if f else im for im, f in zip(images, flags)]
        return cleaned, flags
