"""RMSNorm with the reference's constructor and numerics (lit_gpt/rmsnorm.py:6-28).

On CUDA bf16 the forward is the library's fused kernel (fp32 statistics, weight multiply in fp32, ONE cast — the
same rounding points as the reference); inside the fused decode step the norm does not run as a module at all: it is
folded into the prologue of the following dequant-GEMV.  Other dtypes / devices (the fp32 CPU plumbing config) use
the reference formulation in plain torch.
"""
from __future__ import annotations

import torch

from . import kernels as K


class RMSNorm(torch.nn.Module):
    def __init__(self, size: int, dim: int = -1, eps: float = 1e-5) -> None:
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(size))
        self.eps = eps
        self.dim = dim

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda and x.dtype == torch.bfloat16 and self.weight.dtype == torch.bfloat16 and self.dim in (-1, x.dim() - 1):
            if torch.compiler.is_compiling():  # traced: one opaque b200gpt::rmsnorm node (lit_gpt_b200/ops.py), no graph break
                from . import ops  # noqa: F401

                return torch.ops.b200gpt.rmsnorm(x, self.weight, self.eps)
            return K.rmsnorm(x, self.weight, self.eps)
        dtype = x.dtype
        x = x.float()
        norm_x = torch.mean(x * x, dim=self.dim, keepdim=True)
        x_normed = x * torch.rsqrt(norm_x + self.eps)
        return (self.weight * x_normed).to(dtype=dtype)

    def reset_parameters(self) -> None:
        torch.nn.init.ones_(self.weight)
