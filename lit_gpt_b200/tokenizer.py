"""Thin tokenizer wrapper with the reference's interface (lit_gpt/tokenizer.py): HF ``tokenizers`` or SentencePiece,
picked by which file the checkpoint directory holds.  Not on the hot path (benchmarks use synthetic token ids)."""
from __future__ import annotations

import json
from pathlib import Path
from typing import Optional

import torch


class Tokenizer:
    def __init__(self, checkpoint_dir: Path) -> None:
        self.use_bos = self.check_if_bos_token_used(checkpoint_dir)
        self.bos_id: Optional[int] = None
        self.eos_id: Optional[int] = None
        if (vocabulary_path := checkpoint_dir / "tokenizer.model").is_file():
            from sentencepiece import SentencePieceProcessor

            self.processor = SentencePieceProcessor(model_file=str(vocabulary_path))
            self.backend = "sentencepiece"
            self.bos_id = self.processor.bos_id()
            self.eos_id = self.processor.eos_id()
        elif (vocabulary_path := checkpoint_dir / "tokenizer.json").is_file():
            from tokenizers import Tokenizer as HFTokenizer

            self.processor = HFTokenizer.from_file(str(vocabulary_path))
            self.backend = "huggingface"
            if (special := checkpoint_dir / "tokenizer_config.json").is_file():
                with open(special) as fp:
                    cfg = json.load(fp)
                for attr, key in (("bos_id", "bos_token"), ("eos_id", "eos_token")):
                    tok = cfg.get(key)
                    if isinstance(tok, dict):
                        tok = tok.get("content")
                    if tok is not None:
                        setattr(self, attr, self.processor.token_to_id(tok))
            if (special := checkpoint_dir / "generation_config.json").is_file():
                with open(special) as fp:
                    cfg = json.load(fp)
                self.bos_id = cfg.get("bos_token_id", self.bos_id) if self.bos_id is None else self.bos_id
                self.eos_id = cfg.get("eos_token_id", self.eos_id) if self.eos_id is None else self.eos_id
        else:
            raise NotImplementedError(f"no tokenizer.model / tokenizer.json under {checkpoint_dir}")

    @property
    def vocab_size(self) -> int:
        if self.backend == "huggingface":
            return self.processor.get_vocab_size(with_added_tokens=False)
        return self.processor.vocab_size()

    @staticmethod
    def check_if_bos_token_used(checkpoint_dir: Path) -> bool:
        if not (path := checkpoint_dir / "tokenizer_config.json").is_file():
            return False
        with open(path) as fp:
            cfg = json.load(fp)
        # reference lit_gpt/tokenizer.py:78-87: either flag truthy -> BOS; else only a LlamaTokenizer WITHOUT an add_bos_token key
        if any(cfg.get(k, False) for k in ("add_bos_token", "add_prefix_space")):
            return True
        return cfg.get("add_bos_token") is None and cfg.get("tokenizer_class") == "LlamaTokenizer"

    def encode(self, string: str, device: Optional[torch.device] = None, bos: Optional[bool] = None, eos: bool = False,
               max_length: int = -1) -> torch.Tensor:
        tokens = self.processor.encode(string).ids if self.backend == "huggingface" else self.processor.encode(string)
        if bos or (bos is None and self.use_bos):
            if self.bos_id is None:
                raise NotImplementedError("This tokenizer does not have a defined a bos token")
            tokens = [self.bos_id] + tokens
        if eos:
            tokens = tokens + [self.eos_id]
        if max_length > 0:
            tokens = tokens[:max_length]
        return torch.tensor(tokens, dtype=torch.int, device=device)

    def decode(self, tensor: torch.Tensor) -> str:
        tokens = [tensor.item()] if tensor.ndim == 0 else tensor.tolist()
        return self.processor.decode(tokens)
