"""Shared helpers for the parity tests (test infrastructure)."""
import os
import sys

import torch
from safetensors.torch import load_file

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")

from stylish_tts_b200 import fixture  # noqa: E402


def gold(name):
    return load_file(os.path.join(GOLD, name + ".safetensors"))


_CACHE = {}


def fixture_sds():
    if "sds" not in _CACHE:
        _CACHE["sds"] = fixture.fixture_nets(fixture.load_manifest(), seed=0)
    return _CACHE["sds"]


def gpu_nets():
    """The five module mirrors on cuda:0 with fixture weights (cached per process)."""
    if "nets" not in _CACHE:
        from stylish_tts_b200.config import ModelConfig
        from stylish_tts_b200.modules import build_model
        nets = build_model(ModelConfig())
        sds = fixture_sds()
        for k, m in nets.items():
            m.load_state_dict(sds[k])
            m.to("cuda:0").eval()
        _CACHE["nets"] = nets
    return _CACHE["nets"]


def synthesizer(precision="fp32"):
    key = "synth"
    if key not in _CACHE:
        from stylish_tts_b200.synth import Synthesizer
        _CACHE[key] = Synthesizer(gpu_nets(), device="cuda:0", precision=precision)
    _CACHE[key].set_precision(precision)
    return _CACHE[key]


def case_noise(n_frames, seed):
    return fixture.draw_noise(n_frames, torch.Generator().manual_seed(seed))


def maxabs(a, b):
    return float((a.double().cpu() - b.double().cpu()).abs().max())


def rel(a, b):
    return maxabs(a, b) / max(1e-12, float(b.double().abs().max()))
