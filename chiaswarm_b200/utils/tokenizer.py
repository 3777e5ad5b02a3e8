"""Prompt tokenisation. With a checkpoint's `tokenizer/` folder present the real CLIP BPE
tokenizer (transformers) is used; otherwise -- no network, random-init timing runs -- a
deterministic hashing tokenizer produces ids of the right shape (BOS, word ids, EOS, padding to
77), which is all a synthetic-prompt benchmark needs."""
from __future__ import annotations

import hashlib
import re
from pathlib import Path
from typing import List, Optional

import torch


class HashTokenizer:
    def __init__(self, vocab_size=49408, max_length=77):
        self.vocab_size = vocab_size
        self.model_max_length = max_length
        self.bos_token_id = vocab_size - 2
        self.eos_token_id = vocab_size - 1
        self.pad_token_id = vocab_size - 1

    def encode(self, text: str) -> List[int]:
        words = re.findall(r"[\w']+|[^\w\s]", text.lower())
        ids = [self.bos_token_id]
        for w in words[: self.model_max_length - 2]:
            h = int.from_bytes(hashlib.sha1(w.encode()).digest()[:4], "little")
            ids.append(h % (self.vocab_size - 2))
        ids.append(self.eos_token_id)
        ids += [self.pad_token_id] * (self.model_max_length - len(ids))
        return ids

    def __call__(self, prompts) -> torch.Tensor:
        if isinstance(prompts, str):
            prompts = [prompts]
        return torch.tensor([self.encode(p) for p in prompts], dtype=torch.long)


class ClipBpeTokenizer:
    def __init__(self, folder: Path):
        from transformers import CLIPTokenizer

        self.tok = CLIPTokenizer.from_pretrained(str(folder))
        self.model_max_length = min(self.tok.model_max_length, 77)
        self.eos_token_id = self.tok.eos_token_id

    def __call__(self, prompts) -> torch.Tensor:
        if isinstance(prompts, str):
            prompts = [prompts]
        out = self.tok(prompts, padding="max_length", max_length=self.model_max_length, truncation=True,
                       return_tensors="pt")
        return out.input_ids


def load_tokenizer(folder: Optional[Path], vocab_size: int, max_length: int = 77):
    if folder is not None and Path(folder).is_dir() and (Path(folder) / "vocab.json").is_file():
        try:
            return ClipBpeTokenizer(Path(folder))
        except Exception as e:  # pragma: no cover - depends on local files
            print(f"tokenizer at {folder} unusable ({e}); using hashing tokenizer")
    return HashTokenizer(vocab_size, max_length)
