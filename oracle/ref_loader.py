"""Import the UNMODIFIED reference (``/root/reference``) in this container.

TEST INFRASTRUCTURE (see oracle/__init__.py).  ``/root/reference`` does not
exist on the GPU box, so nothing here may be used at GPU run time; it is used
only by ``tests/golden/make_golden.py`` (fixture generation), by the ``not gpu``
tests that cross-check the restatement when the reference is present, and by
``bench.py --impl reference`` when run inside this container.

The reference hard-fails at import without ``lightning`` (lit_gpt/__init__.py:9-16,
lit_gpt/utils.py:12-17).  None of the imported names is *executed* on the
unquantized model path, so a stub package providing the names is enough.
"""
from __future__ import annotations

import contextlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("B200GPT_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "lit_gpt", "model.py"))


def _mod(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []  # mark as package so submodule imports resolve
    sys.modules[name] = m
    return m


def _install_stubs() -> None:
    if "lightning" in sys.modules and not getattr(sys.modules["lightning"], "_b200_stub", False):
        return  # a real lightning is installed; use it

    class _Stub:  # placeholder classes: only their NAMES are needed
        def __init__(self, *a, **k):
            pass

        @staticmethod
        def is_available() -> bool:
            return False

    class Fabric(_Stub):
        pass

    def seed_everything(seed: int) -> int:
        import torch

        torch.manual_seed(seed)
        return seed

    def rank_zero_only(fn):
        return fn

    L = _mod("lightning", Fabric=Fabric, seed_everything=seed_everything, _b200_stub=True)
    fabric = _mod("lightning.fabric")
    L.fabric = fabric
    _mod("lightning.fabric.strategies", FSDPStrategy=type("FSDPStrategy", (_Stub,), {}))
    _mod("lightning.fabric.plugins", BitsandbytesPrecision=type("BitsandbytesPrecision", (_Stub,), {}))
    _mod("lightning.fabric.accelerators", MPSAccelerator=type("MPSAccelerator", (_Stub,), {}))
    _mod("lightning.fabric.utilities", rank_zero_only=rank_zero_only)
    _mod(
        "lightning.fabric.utilities.load",
        _lazy_load=lambda *a, **k: (_ for _ in ()).throw(NotImplementedError("stub")),
        _NotYetLoadedTensor=type("_NotYetLoadedTensor", (_Stub,), {}),
    )

    class RequirementCache:
        def __init__(self, *a, **k):
            pass

        def __bool__(self) -> bool:
            return True

        def __str__(self) -> str:
            return "stubbed"

    _mod("lightning_utilities")
    _mod("lightning_utilities.core")
    _mod("lightning_utilities.core.imports", RequirementCache=RequirementCache)


@contextlib.contextmanager
def _ref_on_path():
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        yield
    finally:
        with contextlib.suppress(ValueError):
            sys.path.remove(REFERENCE_ROOT)


_cached = None


def load_reference() -> types.SimpleNamespace:
    """Returns the reference's own GPT, Config, generate, sample, tensor_parallel*, build_rope_cache..."""
    global _cached
    if _cached is not None:
        return _cached
    if not reference_available():
        raise FileNotFoundError(f"reference not present at {REFERENCE_ROOT}")
    _install_stubs()
    # `generate` is also the name of OUR sub-package inside lit_gpt_b200 (never top-level), so no clash.
    with _ref_on_path():
        import lit_gpt  # noqa: F401  (reference package)
        import lit_gpt.model as ref_model
        import lit_gpt.rmsnorm as ref_rmsnorm
        from lit_gpt.config import Config as RefConfig
        import generate.base as ref_generate_base
        import generate.tp as ref_generate_tp
    _cached = types.SimpleNamespace(
        GPT=ref_model.GPT,
        Config=RefConfig,
        model=ref_model,
        RMSNorm=ref_rmsnorm.RMSNorm,
        generate=ref_generate_base.generate,
        sample=ref_generate_base.sample,
        next_token=ref_generate_base.next_token,
        generate_base=ref_generate_base,
        tensor_parallel=ref_generate_tp.tensor_parallel,
        tensor_parallel_linear=ref_generate_tp.tensor_parallel_linear,
        generate_tp=ref_generate_tp,
    )
    return _cached
