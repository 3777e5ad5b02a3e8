"""ControlNet / LoRA / textual-inversion loaders (the toolbox model loaders of the north star;
reference call sites swarm/diffusion/diffusion_func.py:29-34,48-68)."""
from __future__ import annotations

import math
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
    if root.is_dir():  # a folder holding one weight file under another name (Kohya exports)
        files = [p for p in sorted(root.iterdir()) if p.suffix in (".safetensors", ".bin", ".pt")]
        if len(files) == 1:
            return files[0]
    return None


# ---- LoRA file formats.  Every format boils down to (component, target module, {down, up, alpha}); they differ in how the
# target is spelled.  `unet.load_attn_procs(lora)` of the reference (swarm/diffusion/diffusion_func.py:58-60) accepts all of
# these through the diffusers releases its unpinned installer may pull.
_DOWN = ("lora.down.weight", "lora_linear_layer.down.weight", "lora_A.weight", "lora_down.weight")
_UP = ("lora.up.weight", "lora_linear_layer.up.weight", "lora_B.weight", "lora_up.weight")
_PROCESSOR = re.compile(r"(.*)\.processor\.(to_q|to_k|to_v|to_out)_lora\.(down|up)\.weight$")
_COMPONENTS = (("unet.", "unet"), ("text_encoder_2.", "te2"), ("text_encoder.", "te"),
               ("lora_unet_", "unet"), ("lora_te2_", "te2"), ("lora_te1_", "te"), ("lora_te_", "te"))


def _parse_lora_key(key: str):
    """-> (component, target, flat, role) or None.  `flat`: the target is a Kohya name (dots written as underscores)."""
    comp, rest, flat = "unet", key, False
    for prefix, c in _COMPONENTS:
        if key.startswith(prefix):
            comp, rest, flat = c, key[len(prefix):], prefix.startswith("lora_")
            break
    m = _PROCESSOR.match(rest)
    if m:  # diffusers <= 0.16 attention processors
        return comp, f"{m.group(1)}.{'to_out.0' if m.group(2) == 'to_out' else m.group(2)}", False, m.group(3)
    if rest.endswith(".alpha"):
        return comp, rest[:-len(".alpha")], flat, "alpha"
    for role, suffixes in (("down", _DOWN), ("up", _UP)):
        for suffix in suffixes:
            if rest.endswith("." + suffix):
                return comp, rest[:-len(suffix) - 1], flat, role
    return None


def _lora_delta(w: dict, shape, scale: float) -> torch.Tensor:
    down, up = w["down"].float(), w["up"].float()
    rank = down.shape[0]
    if "alpha" in w:  # Kohya: the update is scaled by alpha / rank
        scale = scale * float(w["alpha"]) / rank
    delta = up.flatten(1) @ down.flatten(1)                     # linear, or conv with a 1x1 `up`
    if delta.numel() != math.prod(shape):
        raise ValueError(f"update of shape {tuple(up.shape)} x {tuple(down.shape)} does not fit a weight of {tuple(shape)}")
    return delta.reshape(shape) * scale


def load_lora(pipe, lora, scale=1.0):
    """Merges LoRA weights into the model: W += scale * (alpha / rank) * up @ down.  Merging (instead of a runtime side path)
    keeps the fused QKV GEMM and the packed weights of the engine.  Formats: attention-processor files of diffusers <= 0.16
    (`...attn1.processor.to_q_lora.down.weight`), the per-layer diffusers layout (`unet.<module>.lora.down.weight`,
    `text_encoder.<module>.lora_linear_layer.up.weight`), PEFT (`lora_A` / `lora_B`) and Kohya `.safetensors`
    (`lora_unet_<module with underscores>.lora_down.weight` + `.alpha`, `lora_te_...` for the text encoder) as long as the
    module names are the diffusers ones.  Targets may be attention projections, feed-forward layers, proj_in / proj_out and
    convolutions.  A key that names no module of this model is an error, not a silent skip."""
    path = _find_file(lora, ("pytorch_lora_weights.safetensors", "pytorch_lora_weights.bin"))
    if path is None:
        raise ValueError(f"LoRA weights {lora} not found locally")
    sd = checkpoints.load_state_dict(path)
    groups = {}
    for k, v in sd.items():
        parsed = _parse_lora_key(k)
        if parsed is not None:
            comp, target, flat, role = parsed
            groups.setdefault((comp, target, flat), {})[role] = v
    if not groups:
        raise ValueError("no LoRA keys found (expected attention-processor, diffusers, PEFT or Kohya names)")
    roots = {"unet": pipe.unet, "te": pipe.text_encoder, "te2": getattr(pipe, "text_encoder_2", None)}
    tables = {}
    merged = 0
    for (comp, target, flat), w in sorted(groups.items(), key=lambda kv: kv[0][:2]):
        root = roots.get(comp)
        if root is None:
            raise ValueError(f"LoRA targets {comp} but this pipeline has no such component")
        if comp not in tables:
            mods = {n: m for n, m in root.named_modules() if isinstance(m, (torch.nn.Linear, torch.nn.Conv2d))}
            tables[comp] = (mods, {n.replace(".", "_"): m for n, m in mods.items()})
        mod = tables[comp][1 if flat else 0].get(target)
        if mod is None:
            raise ValueError(f"LoRA target {target} does not exist in this {comp}")
        if "down" not in w or "up" not in w:
            raise ValueError(f"LoRA target {target} has only one of its two factors")
        try:
            delta = _lora_delta(w, mod.weight.shape, float(scale))
        except ValueError as e:
            raise ValueError(f"LoRA for {target}: {e}") from e
        mod.weight.data += delta.to(mod.weight.device, mod.weight.dtype)
        merged += 1
    pipe._engine = None  # packed weights are stale
    pipe.__dict__.pop("_text_engines", None)
    return merged


def _grow_embedding(text_encoder, rows: int):
    """`resize_token_embeddings`: the table gains rows for the placeholder ids; existing rows are kept."""
    emb_mod = text_encoder.text_model.embeddings
    old = emb_mod.token_embedding
    if rows <= old.weight.shape[0]:
        return old
    new = torch.nn.Embedding(rows, old.weight.shape[1], device=old.weight.device, dtype=old.weight.dtype)
    with torch.no_grad():
        new.weight.zero_()
        new.weight[: old.weight.shape[0]] = old.weight
    new.weight.requires_grad_(False)
    emb_mod.token_embedding = new
    if isinstance(getattr(text_encoder, "config", None), dict):
        text_encoder.config["vocab_size"] = rows
    return new


def _textual_inversion_entries(sd, default_token: str):
    """-> [(component, token, vectors [n, dim])].  Formats: diffusers `learned_embeds` ({token: [dim] or [n, dim]}),
    AUTOMATIC1111 `.pt` ({"string_to_param": {"*": [n, dim]}, "name": token}), SDXL pairs ({"clip_l": .., "clip_g": ..})."""
    if "string_to_param" in sd:
        vec = next(iter(sd["string_to_param"].values()))
        return [("te", str(sd.get("name") or default_token), torch.as_tensor(vec))]
    if "clip_l" in sd or "clip_g" in sd:
        out = []
        if "clip_l" in sd:
            out.append(("te", default_token, torch.as_tensor(sd["clip_l"])))
        if "clip_g" in sd:
            out.append(("te2", default_token, torch.as_tensor(sd["clip_g"])))
        return out
    return [("te", str(token), torch.as_tensor(vec)) for token, vec in sd.items() if torch.is_tensor(vec)]


def load_textual_inversion(pipe, ref):
    """`pipeline.load_textual_inversion(ref)` of the reference (swarm/diffusion/diffusion_func.py:48-50): the learned
    vectors become new rows of the token-embedding table, the placeholder becomes a token of the tokenizer (a placeholder
    that stands for n vectors is written as n tokens at prompt time), so a prompt that names the placeholder conditions on
    the learned vectors.  Returns {placeholder: first token id}."""
    path = _find_file(ref, ("learned_embeds.safetensors", "learned_embeds.bin"))
    if path is None:
        raise ValueError(f"textual inversion {ref} not found locally")
    sd = checkpoints.load_state_dict(path)  # (pickles are read with weights_only=True: tensors and plain containers only)
    entries = _textual_inversion_entries(sd, f"<{path.stem}>")
    if not entries:
        raise ValueError("no embeddings found in the textual-inversion file")
    parts = {"te": (pipe.text_encoder, pipe.tokenizer),
             "te2": (getattr(pipe, "text_encoder_2", None), getattr(pipe, "tokenizer_2", None))}
    added = {}
    for comp, token, vec in entries:
        encoder, tokenizer = parts[comp]
        if encoder is None:
            raise ValueError("the file carries embeddings for a second text encoder this pipeline does not have")
        vec = vec.reshape(-1, vec.shape[-1])
        dim = encoder.text_model.embeddings.token_embedding.weight.shape[1]
        if vec.shape[-1] != dim:
            raise ValueError(f"embedding dim {vec.shape[-1]} does not match the text encoder ({dim})")
        ids = tokenizer.register_placeholder(token, vec.shape[0])
        emb = _grow_embedding(encoder, max(max(ids) + 1, len(tokenizer)))
        with torch.no_grad():
            for row, v in zip(ids, vec):
                emb.weight[row] = v.to(emb.weight.device, emb.weight.dtype)
        added.setdefault(token, ids[0])
    pipe.textual_inversion_tokens = added
    pipe.__dict__.pop("_text_engines", None)  # (the recorded text-encoder plans keep no copy of the table, but start clean)
    return added
