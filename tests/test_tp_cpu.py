"""Tensor-parallel host logic (reference tests/test_generate_tp.py:14-103) + a world_size-2 gloo run on CPU."""
import os
import subprocess
import sys
from dataclasses import replace
from unittest.mock import Mock

import pytest
import torch

from lit_gpt_b200 import GPT
from lit_gpt_b200.generate.tp import tensor_parallel, tensor_parallel_linear

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_tensor_parallel_linear(golden):
    u = golden("units")
    fabric = Mock()
    fabric.world_size, fabric.global_rank = 4, 2

    def get_linear(bias=True):
        linear = torch.nn.Linear(8, 8, bias=bias)
        linear.weight.data = torch.arange(64, dtype=torch.float32).reshape(8, 8)
        if bias:
            linear.bias.data = torch.arange(8, dtype=torch.float32)
        return linear

    linear = get_linear()
    tensor_parallel_linear(fabric, linear, "colwise")
    torch.testing.assert_close(linear.weight, torch.arange(32, 48, dtype=torch.float32).reshape(2, 8))
    torch.testing.assert_close(linear.bias, torch.arange(4, 6, dtype=torch.float32))
    assert linear.out_features == 2
    assert torch.equal(linear.weight.detach(), torch.from_numpy(u["tp_col_w"]))  # what the reference itself produced

    linear = get_linear(bias=False)
    tensor_parallel_linear(fabric, linear, "rowwise")
    expected = torch.arange(4, 62, 8, dtype=torch.float32).reshape(8, 1)
    torch.testing.assert_close(linear.weight, torch.cat([expected, expected + 1], dim=1))
    assert linear.bias is None and linear.in_features == 2
    assert torch.equal(linear.weight.detach(), torch.from_numpy(u["tp_row_w"]))

    with pytest.raises(ValueError, match="not evenly divisible"):
        fabric.world_size = 3
        tensor_parallel_linear(fabric, get_linear(), "colwise")


def test_tensor_parallel_llama():
    fabric = Mock()
    fabric.world_size, fabric.global_rank = 8, 1
    with torch.device("meta"):
        model = GPT.from_name("Llama-2-70b-hf", n_layer=3)
    config = replace(model.config)
    model = tensor_parallel(fabric, model)
    for i in range(3):
        for mod in (model.transformer.h[i].attn, model.transformer.h[i].mlp):
            hooks = list(mod._forward_hooks.values())
            assert len(hooks) == 1 and hooks[0].func.__name__ == "all_reduce_output" and hooks[0].args == (8,)
    assert model.config.n_embd * 8 == config.n_embd
    assert model.config.n_head * 8 == config.n_head
    assert model.config.n_query_groups * 8 == config.n_query_groups == 8  # TP-8 flips 70B to MQA per rank
    blk = model.transformer.h[0]
    assert tuple(blk.attn.attn.weight.shape) == (1280, 8192) and tuple(blk.attn.proj.weight.shape) == (8192, 1024)
    assert tuple(blk.mlp.fc_1.weight.shape) == (3584, 8192) and tuple(blk.mlp.proj.weight.shape) == (8192, 3584)
    assert tuple(model.lm_head.weight.shape) == (32000, 8192)  # lm_head and wte are replicated


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from lit_gpt_b200 import GPT, Config
from lit_gpt_b200.fabric import Fabric
from lit_gpt_b200.generate.tp import tensor_parallel
import lit_gpt_b200.generate.base as gb
fabric = Fabric(devices="auto", strategy="ddp", precision="32-true", accelerator="cpu")
fabric.launch()
torch.manual_seed(0)
cfg = Config.from_name("Llama-2-7b-hf", n_layer=2, n_head=4, n_query_groups=2, n_embd=64, intermediate_size=96, padded_vocab_size=128, block_size=32)
full = GPT(cfg).eval()
full.apply(full._init_weights)
sd = {k: v.clone() for k, v in full.state_dict().items()}
x = torch.tensor([[5, 17, 3, 99, 42]])
full.max_seq_length = 16; full.set_kv_cache(1)
with torch.inference_mode():
    want = gb.generate(full, x[0], 14, temperature=0.0)
    want_logits = GPT(cfg).eval()
    want_logits.load_state_dict(sd)
    y_full = want_logits(x)
shard = GPT(Config.from_name("Llama-2-7b-hf", n_layer=2, n_head=4, n_query_groups=2, n_embd=64, intermediate_size=96, padded_vocab_size=128, block_size=32)).eval()
shard.load_state_dict(sd)
shard = tensor_parallel(fabric, shard)
shard.max_seq_length = 16; shard.set_kv_cache(1)
with torch.inference_mode():
    y = shard(x)
    got = gb.generate(shard, x[0], 14, temperature=0.0)
torch.testing.assert_close(y, y_full, rtol=1e-4, atol=1e-5)
assert got.tolist() == want.tolist(), (got, want)
dist.barrier()
if fabric.global_rank == 0:
    print("TP_OK", fabric.world_size)
dist.destroy_process_group()
"""


def test_tp_world_size_2_gloo(tmp_path):
    """Sharded model over 2 CPU ranks (gloo) == unsharded model: logits and a greedy rollout."""
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29611", str(script), ROOT]
    out = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "TP_OK 2" in out.stdout
