"""The streaming generator of chat/base.py:24-68 (second caller of next_token): the reference's own test vectors
(tests/test_chat.py:16-51) for the stop-sequence / hold-back logic, run against lit_gpt_b200.chat.base.generate."""
from itertools import repeat
from unittest.mock import MagicMock

import pytest
import torch


@pytest.mark.parametrize(
    ("generated", "stop_tokens", "expected"),
    [
        (repeat(1), (), [1] * 8),
        ([1, 2, 3, 0], ([0],), [1, 2, 3]),
        ([1, 2, 3, 0], ([9], [2, 4], [1, 2, 3, 0]), []),
        ([1, 2, 3, 0, 0], ([0, 0, 0], [0, 0]), [1, 2, 3]),
        ([3, 1, 2], ([1, 2], [3]), []),
        ([1, 2, 3, 0, 3, 2, 1, 0], ([4, 3, 2, 1], [2, 4]), [1, 2, 3, 0, 3, 2, 1, 0]),
    ],
)
def test_generate(monkeypatch, generated, stop_tokens, expected):
    import lit_gpt_b200.chat.base as chat
    import lit_gpt_b200.generate.base as generate

    input_idx = torch.tensor([5, 3])
    max_returned_tokens = len(input_idx) + 8
    model = MagicMock()
    model.config.block_size = 100
    model.max_seq_length = 100
    model.fused_decode_ready.return_value = False
    it = iter(generated)

    def fake_sample(*_, **__):
        return torch.tensor([next(it)])

    monkeypatch.setattr(generate, "sample", fake_sample)
    actual = list(chat.generate(model, input_idx, max_returned_tokens, stop_tokens=stop_tokens))
    assert len(actual) == len(expected)
    if actual:
        for t in actual:
            assert t.dtype == torch.long
        assert torch.cat(actual).tolist() == expected
