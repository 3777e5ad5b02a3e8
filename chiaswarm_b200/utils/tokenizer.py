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


class _Placeholders:
    """Textual-inversion placeholders (diffusers `load_textual_inversion` + `maybe_convert_prompt`): a placeholder that
    stands for n learned vectors is written out as n tokens `tok tok_1 .. tok_{n-1}` before tokenisation."""

    def _init_placeholders(self):
        self.multi_vector = {}

    def register_placeholder(self, token: str, count: int) -> List[int]:
        names = [token] + [f"{token}_{i}" for i in range(1, count)]
        known = [n for n in names if self.has_token(n)]
        if known:
            raise ValueError(f"token {known[0]} is already in the tokenizer vocabulary: pick another placeholder")
        ids = self.add_tokens(names)
        if count > 1:
            self.multi_vector[token] = names
        return ids

    def expand(self, text: str) -> str:
        for token, names in self.multi_vector.items():
            text = re.sub(re.escape(token) + r"(?!_\d)", " ".join(names), text)
        return text


class HashTokenizer(_Placeholders):
    def __init__(self, vocab_size=49408, max_length=77):
        self.vocab_size = vocab_size
        self.model_max_length = max_length
        self.bos_token_id = vocab_size - 2
        self.eos_token_id = vocab_size - 1
        self.pad_token_id = vocab_size - 1
        self.added = {}
        self._init_placeholders()

    def __len__(self):
        return self.vocab_size + len(self.added)

    def has_token(self, token: str) -> bool:
        return token in self.added

    def add_tokens(self, tokens) -> List[int]:
        for t in tokens:
            self.added.setdefault(t, self.vocab_size + len(self.added))
        return [self.added[t] for t in tokens]

    def encode(self, text: str) -> List[int]:
        text = self.expand(text)
        if self.added:  # placeholders are matched literally (case kept), longest first, before the word split
            pattern = "|".join(re.escape(t) for t in sorted(self.added, key=len, reverse=True))
            pieces = re.split(f"({pattern})", text)
        else:
            pieces = [text]
        ids = [self.bos_token_id]
        for piece in pieces:
            if piece in self.added:
                ids.append(self.added[piece])
                continue
            for w in re.findall(r"[\w']+|[^\w\s]", piece.lower()):
                h = int.from_bytes(hashlib.sha1(w.encode()).digest()[:4], "little")
                ids.append(h % (self.vocab_size - 2))
        ids = ids[: self.model_max_length - 1]
        ids.append(self.eos_token_id)
        ids += [self.pad_token_id] * (self.model_max_length - len(ids))
        return ids

    def __call__(self, prompts) -> torch.Tensor:
        if isinstance(prompts, str):
            prompts = [prompts]
        return torch.tensor([self.encode(p) for p in prompts], dtype=torch.long)


class ClipBpeTokenizer(_Placeholders):
    def __init__(self, folder: Path):
        from transformers import CLIPTokenizer

        self.tok = CLIPTokenizer.from_pretrained(str(folder))
        self.model_max_length = min(self.tok.model_max_length, 77)
        self.eos_token_id = self.tok.eos_token_id
        self._init_placeholders()

    def __len__(self):
        return len(self.tok)

    def has_token(self, token: str) -> bool:
        return token in self.tok.get_vocab()

    def add_tokens(self, tokens) -> List[int]:
        self.tok.add_tokens(list(tokens))
        return [int(i) for i in self.tok.convert_tokens_to_ids(list(tokens))]

    def __call__(self, prompts) -> torch.Tensor:
        if isinstance(prompts, str):
            prompts = [prompts]
        prompts = [self.expand(p) for p in prompts]
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
