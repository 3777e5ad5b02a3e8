"""ControlNet / LoRA / textual-inversion loaders (the toolbox model loaders of the north star;
reference call sites swarm/diffusion/diffusion_func.py:29-34,48-68)."""
from __future__ import annotations

import os
import re
from pathlib import Path

import torch

from ..models import configs as arch_configs
from ..models.controlnet import ControlNetModel
from ..utils import checkpoints


def load_controlnet(name, revision="main", device="cpu", dtype=torch.float16, seed=7):
    root = checkpoints.locate_checkpoint(name, revision)
    if root is not None and (root / "config.json").is_file():
        cfg = checkpoints.read_json(root / "config.json")
        model = ControlNetModel(**{k: v for k, v in cfg.items() if not k.startswith("_")})
        weights = checkpoints.find_weights(root)
        if weights is None:
            raise ValueError(f"no ControlNet weights under {root}")
        checkpoints.load_into(model, checkpoints.load_state_dict(weights), strict=False)
    else:
        from .pipelines import _init_module, _random_init_allowed

        if not _random_init_allowed():
            raise ValueError(f"ControlNet {name} is not available locally")
        model = _init_module(ControlNetModel(**arch_configs.get_controlnet(name)), seed)
    dev = torch.device(device)
    return model.to(dev, dtype if dev.type == "cuda" else torch.float32).eval()


def _find_file(ref, names):
    root = checkpoints.locate_checkpoint(ref)
    if root is None:
        p = Path(ref).expanduser()
        return p if p.is_file() else None
    if root.is_file():
        return root
    for n in names:
        if (root / n).is_file():
            return root / n
    return None


_LORA_KEY = re.compile(r"(.*)\.processor\.(to_q|to_k|to_v|to_out)_lora\.(down|up)\.weight$")


def load_lora(pipe, lora, scale=1.0):
    """Merges attention-processor LoRA weights (`pytorch_lora_weights.*`, keys
    `...attn{1,2}.processor.to_{q,k,v,out}_lora.{down,up}.weight`) into the UNet projections:
    W += scale * up @ down.  Merging (instead of a runtime side path) keeps the fused QKV GEMM."""
    path = _find_file(lora, ("pytorch_lora_weights.safetensors", "pytorch_lora_weights.bin"))
    if path is None:
        raise ValueError(f"LoRA weights {lora} not found locally")
    sd = checkpoints.load_state_dict(path)
    pairs = {}
    for k, v in sd.items():
        m = _LORA_KEY.match(k.replace("unet.", "", 1) if k.startswith("unet.") else k)
        if not m:
            continue
        pairs.setdefault((m.group(1), m.group(2)), {})[m.group(3)] = v
    if not pairs:
        raise ValueError("no attention-processor LoRA keys found")
    modules = dict(pipe.unet.named_modules())
    merged = 0
    for (attn_path, proj), w in pairs.items():
        attn = modules.get(attn_path)
        if attn is None or "down" not in w or "up" not in w:
            raise ValueError(f"LoRA target {attn_path} does not exist in this UNet")
        lin = attn.to_out[0] if proj == "to_out" else getattr(attn, proj)
        delta = (w["up"].float() @ w["down"].float()) * float(scale)
        if delta.shape != lin.weight.shape:
            raise ValueError(f"LoRA shape {tuple(delta.shape)} does not match {attn_path}.{proj} {tuple(lin.weight.shape)}")
        lin.weight.data += delta.to(lin.weight.device, lin.weight.dtype)
        merged += 1
    pipe._engine = None  # packed weights are stale
    return merged


def load_textual_inversion(pipe, ref):
    """Adds learned token embeddings (`learned_embeds.bin|safetensors`: {token: vector}) by
    overwriting unused rows at the end of the embedding table and teaching the tokenizer the
    token when it supports it."""
    path = _find_file(ref, ("learned_embeds.safetensors", "learned_embeds.bin"))
    if path is None:
        raise ValueError(f"textual inversion {ref} not found locally")
    sd = checkpoints.load_state_dict(path)
    if "string_to_param" in sd:  # A1111 .pt
        sd = {k: v for k, v in zip(sd.get("string_to_token", {"*": 0}).keys(), sd["string_to_param"].values())}
    emb = pipe.text_encoder.text_model.embeddings.token_embedding
    dim = emb.weight.shape[1]
    slot = emb.weight.shape[0] - 3
    added = {}
    for token, vec in sd.items():
        vec = torch.as_tensor(vec).reshape(-1, vec.shape[-1])
        if vec.shape[-1] != dim:
            raise ValueError(f"embedding dim {vec.shape[-1]} does not match the text encoder ({dim})")
        emb.weight.data[slot] = vec[0].to(emb.weight.device, emb.weight.dtype)
        added[token] = slot
        slot -= 1
    pipe.textual_inversion_tokens = added
    return added
