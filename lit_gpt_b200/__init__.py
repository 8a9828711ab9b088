"""b200gpt: B200-native (sm_100a) implementation of lit-gpt's quantized decode/prefill hot path.

Public surface mirrors the reference package (``from lit_gpt import GPT, Config``) plus the quantization plugin that
replaces ``lightning.fabric.plugins.BitsandbytesPrecision``.
"""
from lit_gpt_b200.config import Config
from lit_gpt_b200.model import GPT
from lit_gpt_b200.quant import QLinear, QuantPrecision
from lit_gpt_b200.rmsnorm import RMSNorm

__all__ = ["GPT", "Config", "QLinear", "QuantPrecision", "RMSNorm"]
__version__ = "0.1.0"
