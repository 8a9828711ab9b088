"""Tensor-parallel decode parity on ONE GPU: WORLD processes share cuda:0 (B200_TP_SHARED_DEVICE=1, gloo control plane), so
the peer-memory hand-off inside the GEMV kernels (b200_tp_handoff), shard-then-quantize and the captured decode graph are
exercised on a one-GPU lease.  The checker is the oracle on identically sharded-then-quantized weights
(tests/tp_shared_worker.py).  Reference: generate/tp.py:28-92, tests/test_generate_tp.py:109-134."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _run(world, mode, n_head, n_groups, seq0=0, port=29641, inter=0, shard_head=False):
    env = dict(os.environ, B200_TP_SHARED_DEVICE="1", B200_TP_NCCL="0", OMP_NUM_THREADS="4", B200_TP_SHARD_LM_HEAD="1" if shard_head else "0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "tp_shared_worker.py"), ROOT, mode, str(n_head), str(n_groups), str(seq0), str(inter)]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert out.returncode == 0 and "TP_SHARED_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-6000:]
    print(out.stdout[-600:])


@pytest.mark.parametrize("mode", ["nf4", "int4"])
def test_tp2_handoff_against_oracle(mode):
    _run(2, mode, 8, 4)


def test_tp2_tag_wraparound():
    # both buffers' sequence counters start 20 hand-offs below the 16-bit tag wrap (tag = seq % 65535 + 1)
    _run(2, "nf4", 8, 4, seq0=65535 - 20, port=29642)


def test_tp4_per_rank_mqa_int4():
    # the Llama-2-70B pattern: n_query_groups == world, every rank is left with ONE kv group (MQA per rank)
    _run(4, "int4", 8, 4, port=29643)


def test_tp8_row_shard_not_block_aligned():
    # Llama-2-7B at world 8: mlp.proj is cut to 11008 / 8 = 1376 columns, not a multiple of the 64-element block -> the
    # folded tiled layout (QLinear.fold == 2) keeps that Linear on the streaming kernel and on the peer-memory hand-off
    _run(8, "nf4", 8, 8, port=29644, inter=11008)


@pytest.mark.parametrize("world,mode", [(2, "nf4"), (4, "int4")])
def test_tp_lm_head_sharded(world, mode):
    # SURVEY §8f-3 (beyond generate/tp.py:77-80): lm_head sharded by rows, the vocabulary slices all-gathered over a third
    # peer-memory buffer (b200_tp_gather); prefill through the all_gather hook; full logits on every rank, against the oracle
    _run(world, mode, 8, 4, port=29645 + world, shard_head=True)
