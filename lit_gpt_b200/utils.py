"""Host-side helpers the generate scripts use (subset of the reference's lit_gpt/utils.py on the hot path)."""
from __future__ import annotations

from pathlib import Path
from typing import Any, Dict

import torch
import torch.nn as nn

from .config import find_multiple  # noqa: F401  (re-export, as in the reference)


def check_valid_checkpoint_dir(checkpoint_dir: Path) -> None:
    """Same contract as lit_gpt/utils.py:93-124: the directory must hold the weights, config and a tokenizer."""
    files = {
        "lit_model.pth": (checkpoint_dir / "lit_model.pth").is_file(),
        "lit_config.json": (checkpoint_dir / "lit_config.json").is_file(),
        "tokenizer.json OR tokenizer.model": (checkpoint_dir / "tokenizer.json").is_file()
        or (checkpoint_dir / "tokenizer.model").is_file(),
        "tokenizer_config.json": (checkpoint_dir / "tokenizer_config.json").is_file(),
    }
    if checkpoint_dir.is_dir():
        if all(files.values()):
            return
        problem = f" is missing the files: {[f for f, exists in files.items() if not exists]!r}"
    else:
        problem = " is not a checkpoint directory"
    raise SystemExit(f"--checkpoint_dir {str(checkpoint_dir.absolute())!r}{problem}.")


def get_default_supported_precision(training: bool) -> str:
    """bf16 wherever the hardware has it (always on B200); lit_gpt/utils.py:334-347."""
    if torch.cuda.is_available() and not torch.cuda.is_bf16_supported():
        return "16-mixed" if training else "16-true"
    return "bf16-mixed" if training else "bf16-true"


def load_checkpoint(fabric: Any, model: nn.Module, checkpoint_path: Path, strict: bool = True) -> None:
    """lit_gpt/utils.py:350-356: plain ``weight``/``bias`` keys; QLinear quantizes on load when already on CUDA."""
    state_dict: Dict[str, Any] = torch.load(str(checkpoint_path), mmap=True, map_location="cpu", weights_only=True)
    state_dict = state_dict.get("model", state_dict)
    model.load_state_dict(state_dict, strict=strict)


def num_parameters(module: nn.Module, requires_grad: bool | None = None) -> int:
    total = 0
    for p in module.parameters():
        if requires_grad is None or p.requires_grad == requires_grad:
            qs = getattr(p, "quant_state", None)
            total += p.numel()
    for m in module.modules():  # packed 4-bit tensors hold two (nf4/fp4) or eight (int4) weights per element
        qs = getattr(m, "quant_state", None)
        if qs is not None:
            total += int(torch.Size(qs[1]).numel()) - m.weight.numel()
    return total
