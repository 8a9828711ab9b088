"""Minimal stand-in for the slice of ``lightning.Fabric`` the reference's generate scripts use
(generate/base.py:136-187, generate/tp.py:137-192): device selection, dtype/device init contexts, the precision plugin
hook-up, rank-zero printing, and one process per GPU over ``torch.distributed`` (NCCL on CUDA, gloo on CPU).
No Lightning dependency.
"""
from __future__ import annotations

import contextlib
import os
import sys
from typing import Any, Iterator, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from .quant import QuantPrecision

_PRECISION_TO_DTYPE = {"16-true": torch.float16, "bf16-true": torch.bfloat16, "32-true": torch.float32}


class Fabric:
    def __init__(self, devices: Any = 1, strategy: Optional[str] = None, precision: Optional[str] = None,
                 plugins: Optional[QuantPrecision] = None, accelerator: str = "auto") -> None:
        self.strategy = strategy
        self._precision = plugins
        self._dtype = plugins.dtype if plugins is not None else _PRECISION_TO_DTYPE.get(precision or "32-true", torch.float32)
        self.local_rank = int(os.environ.get("LOCAL_RANK", 0))
        self.global_rank = int(os.environ.get("RANK", 0))
        self.world_size = int(os.environ.get("WORLD_SIZE", 1)) if strategy == "ddp" else 1
        use_cuda = torch.cuda.is_available() and accelerator in ("auto", "cuda", "gpu")
        # B200_TP_SHARED_DEVICE=1: every rank drives cuda:0 and the control plane is gloo (two NCCL ranks cannot share a
        # device).  The peer-memory hand-off works unchanged between processes on one GPU (CUDA IPC, time-sliced contexts);
        # this is how the tensor-parallel decode path is tested on a one-GPU box (tests/test_gpu_tp_shared.py).
        self.shared_device = strategy == "ddp" and os.environ.get("B200_TP_SHARED_DEVICE", "0") == "1"
        index = self.local_rank if strategy == "ddp" and not self.shared_device else 0
        self.device = torch.device("cuda", index) if use_cuda else torch.device("cpu")
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)

    # ---- launching (generate/tp.py:137-138: "ddp" is used only for its launcher) -------------------------------
    def launch(self) -> None:
        if self.world_size > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            backend = "nccl" if self.device.type == "cuda" and not self.shared_device else "gloo"
            kw = {"device_id": self.device} if backend == "nccl" else {}
            dist.init_process_group(backend, rank=self.global_rank, world_size=self.world_size, **kw)

    def barrier(self) -> None:
        if self.world_size > 1 and dist.is_initialized():
            dist.barrier()

    def print(self, *args: Any, **kwargs: Any) -> None:
        if self.global_rank == 0:
            print(*args, **kwargs)

    # ---- init contexts ---------------------------------------------------------------------------------------
    @contextlib.contextmanager
    def init_tensor(self) -> Iterator[None]:
        prev = torch.get_default_dtype()
        torch.set_default_dtype(self._dtype)
        try:
            with torch.device(self.device):
                yield
        finally:
            torch.set_default_dtype(prev)

    @contextlib.contextmanager
    def init_module(self, empty_init: Optional[bool] = None) -> Iterator[None]:
        """Modules are created in the compute dtype; with a quantization plugin ``nn.Linear`` builds ``QLinear``.
        Tensors are allocated on ``fabric.device`` (``empty_init`` is accepted for signature parity: QLinear
        weights are always left uninitialised until a checkpoint or ``_init_weights`` fills them)."""
        with contextlib.ExitStack() as stack:
            if self._precision is not None:
                stack.enter_context(self._precision.module_init_context())
            else:
                prev = torch.get_default_dtype()
                torch.set_default_dtype(self._dtype)
                stack.callback(torch.set_default_dtype, prev)
            stack.enter_context(torch.device(self.device))
            yield

    # ---- module placement ------------------------------------------------------------------------------------
    def setup_module(self, module: nn.Module) -> nn.Module:
        if self._precision is not None:
            module = self._precision.convert_module(module)
        return self.to_device(module)

    setup = setup_module

    def to_device(self, obj: Any) -> Any:
        if isinstance(obj, nn.Module):
            if any(p.is_meta for p in obj.parameters()):
                # meta weights (empty_init): materialise uninitialised storage on the device; a checkpoint load follows
                obj.to_empty(device=self.device)
                for m in obj.modules():  # non-persistent buffers (rope tables, kv caches) must be rebuilt by the caller
                    if hasattr(m, "reset_parameters") and type(m).__name__ in ("GPT", "KVCache"):
                        with self.init_tensor():
                            m.reset_parameters()
                return obj
            return obj.to(self.device)
        if isinstance(obj, torch.Tensor):
            return obj.to(self.device)
        return obj


SEED_EPOCH = 0  # bumped by every seed_everything call: device-side sampler states reseed when it changes


def seed_everything(seed: int) -> int:
    import random

    global SEED_EPOCH
    SEED_EPOCH += 1

    import numpy as np

    random.seed(seed)
    np.random.seed(seed % (2**32))
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    return seed
