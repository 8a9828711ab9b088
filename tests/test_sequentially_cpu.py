"""generate/sequentially.py: the reference's own test vectors for the layer-to-device map and `replace_device`
(tests/test_generate_sequentially.py:18-83), and the one-device placement on the CPU (:86-139)."""
import itertools
from re import escape

import pytest
import torch


@pytest.mark.parametrize(
    ("n_layer", "devices", "expected"),
    [
        (6, 2, {0: 0, 1: 0, 2: 0, 3: 1, 4: 1, 5: 1}),
        (6, 3, {0: 0, 1: 0, 2: 1, 3: 1, 4: 2, 5: 2}),
        (6, 1, {0: 0, 1: 0, 2: 0, 3: 0, 4: 0, 5: 0}),
    ],
)
def test_layer_to_device(n_layer, devices, expected):
    from lit_gpt_b200.generate.sequentially import layer_to_device
    from lit_gpt_b200.model import GPT, Block

    with torch.device("meta"):
        model = GPT.from_name("pythia-14m", n_layer=n_layer)
    actual = layer_to_device(model, Block, chunk_size=n_layer // devices)
    assert actual == {f"transformer.h.{i}": v for i, v in expected.items()}


def path_to_device(model):
    return {k: str(v.device) for k, v in itertools.chain(model.named_parameters(), model.named_buffers())}


def test_replace_device():
    from lit_gpt_b200.generate.sequentially import replace_device

    class Submodule(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.register_buffer("foo", torch.tensor(1, device="cpu"))
            self.register_buffer("bar", torch.tensor(1, device="cpu"))

    class MyModel(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.modules = torch.nn.ModuleDict({
                "module1": torch.nn.Linear(1, 1, bias=True, device="meta"),
                "module2": torch.nn.Linear(1, 1, bias=False, device="cpu"),
            })
            self.submodule = Submodule()

    model = MyModel()
    model = replace_device(model, torch.device("cpu"), torch.device("meta"))
    assert path_to_device(model) == {
        "modules.module1.bias": "meta", "modules.module1.weight": "meta", "modules.module2.weight": "meta",
        "submodule.bar": "meta", "submodule.foo": "meta",
    }
    model = MyModel()
    model.submodule.bar = model.submodule.bar.to("meta")
    with pytest.raises(ValueError, match=escape("multiple devices: {'submodule.foo': device(type='cpu'), 'submodule.bar': device(type='meta')}")):
        replace_device(model, torch.device("cpu"), torch.device("meta"))


def test_unbalanced_partitioning_is_rejected():
    from lit_gpt_b200.generate.sequentially import sequential
    from lit_gpt_b200.model import GPT

    with torch.device("meta"):
        model = GPT.from_name("pythia-14m", n_layer=5)
    with pytest.raises(NotImplementedError, match="Only balanced partitioning is implemented: n_layer=5, devices 2"):
        sequential(model, torch.device("cpu"), 15, 2)


def test_model_1device_cpu():
    from lit_gpt_b200.generate.sequentially import sequential
    from lit_gpt_b200.model import GPT

    with torch.device("meta"):
        model = GPT.from_name("pythia-14m", n_layer=2)
    model = sequential(model, torch.device("cpu"), 15, 1)
    devs = path_to_device(model)
    assert set(devs.values()) == {"cpu"}
    for k in ("cos", "sin", "lm_head.weight", "transformer.wte.weight", "transformer.h.1.attn.kv_cache.k", "transformer.h.0.mlp.proj.bias"):
        assert k in devs
    assert model.max_seq_length == 15
