"""Shared fixtures.  `-m "not gpu"` tests run on the CPU container; `-m gpu` tests are the parity
tests proper and call the CUDA path through the C ABI (libq3a.so)."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import q3a_b200  # noqa: E402,F401  (registers the package under an importable name)


def _validated() -> dict:
    import json
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "validated_on_b200.json")) as f:
        return json.load(f)


def needs(feature: str):
    """skip marker for the tests of a round-2 path whose parity has not been confirmed on a B200 yet"""
    ok = bool(_validated().get(feature)) or bool(os.environ.get("Q3A_RUN_UNVALIDATED"))
    return pytest.mark.skipif(not ok, reason=f"'{feature}' is not validated on a B200 yet (tests/validated_on_b200.json); "
                                             "Q3A_RUN_UNVALIDATED=1 runs it anyway")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def lib():
    """libq3a.so, built in-tree if missing (nvcc cross-compiles without a GPU)."""
    from q3a_b200 import binding
    if not os.path.exists(binding.LIB_PATH):
        import importlib.util
        spec = importlib.util.spec_from_file_location("q3a_build", os.path.join(ROOT, "qwen3-asr.cpp_b200", "build.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build()
    return binding.load()


@pytest.fixture(scope="session")
def filters():
    from oracle import mel_oracle
    return mel_oracle.mel_filters_numpy()


@pytest.fixture(scope="session")
def tiny_models(tmp_path_factory):
    """Random-init tiny tower written as F32 / F16 / Q8_0 GGUF + the oracle's view of each file."""
    from q3a_b200 import synth
    from oracle import encoder_oracle as EO
    d = tmp_path_factory.mktemp("gguf")
    shape = synth.tiny_shape()
    w = synth.random_weights(shape, 7)
    out = {}
    for ftype in ("f32", "f16", "q8_0"):
        path = synth.write_gguf(str(d / f"tiny_{ftype}.gguf"), shape, w, ftype)
        stored = {k: v for k, v in synth.stored_weights(w, ftype).items() if "#" not in k}
        out[ftype] = (path, EO.Weights(stored, ftype))
    return shape, out


def compare(got, want):
    """(cosine, max-abs-error / max-abs-reference): the two figures the north star's tolerance names."""
    import numpy as np
    g = got.astype(np.float64).ravel()
    w = want.astype(np.float64).ravel()
    cos = float(g @ w / (np.linalg.norm(g) * np.linalg.norm(w) + 1e-300))
    rel = float(np.abs(g - w).max() / (np.abs(w).max() + 1e-300))
    return cos, rel
