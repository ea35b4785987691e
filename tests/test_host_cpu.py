"""CPU tests of the host layer: state_dict layout, no torch fallback in the product, and that the
C-ABI library loads and exports every symbol include/stts_b200.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest
import torch

from tests.util import ROOT, fixture_sds
from stylish_tts_b200 import fixture
from stylish_tts_b200.config import ModelConfig
from stylish_tts_b200.modules import build_model


@pytest.fixture(scope="module")
def cpu_nets():
    nets = build_model(ModelConfig())
    sds = fixture_sds()
    for k, m in nets.items():
        m.load_state_dict(sds[k])
        m.eval()
    return nets


def test_state_dict_layout_matches_reference_manifest():
    man = fixture.load_manifest()
    nets = build_model(ModelConfig())
    for name, m in nets.items():
        have = {k: list(v.shape) for k, v in m.state_dict().items()}
        want = {k: v["shape"] for k, v in man[name].items() if not k.startswith("posterior_encoder.")}
        assert have == want, name


def test_reference_checkpoint_with_posterior_encoder_loads():
    nets = build_model(ModelConfig())
    sd = dict(fixture_sds()["speech_predictor"])
    sd["posterior_encoder.pre.weight"] = torch.zeros(3)       # training-only keys are dropped
    nets["speech_predictor"].load_state_dict(sd)


def test_modules_hold_parameters_only(cpu_nets):
    """No torch statement of the arithmetic lives in the product package: on CPU every module forward
    must refuse (the CUDA library is the only implementation), and no module defines a torch fallback."""
    g = torch.Generator().manual_seed(5)
    t = fixture.synth_tokens(14, g)
    tk, ln = t[None], torch.tensor([14])
    for name in ("duration_predictor", "pe_text_encoder"):
        with pytest.raises(RuntimeError):
            cpu_nets[name](tk, ln)
    with pytest.raises(RuntimeError):
        cpu_nets["pe_text_style_encoder"](torch.zeros(1, 256, 14), ln)
    for m in cpu_nets.values():
        for sub in m.modules():
            assert not hasattr(sub, "forward_torch") and not hasattr(sub, "compute_cross"), type(sub)


def test_importing_the_package_does_not_touch_global_torch_flags():
    import subprocess
    import sys
    code = ("import torch; a=(torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32); "
            "import stylish_tts_b200.modules, stylish_tts_b200.synth; "
            "b=(torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32); assert a == b, (a, b)")
    subprocess.run([sys.executable, "-c", code], check=True, cwd=ROOT, timeout=300)


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "stts_b200.h")).read()
    declared = set(re.findall(r"\b(stts_[a-z0-9_]+)\s*\(", hdr))
    from stylish_tts_b200 import _lib
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    if not os.path.exists(_lib.LIB_PATH):
        from stylish_tts_b200.build import build
        build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    lib.stts_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.stts_version()


def test_no_cpu_fallback():
    from stylish_tts_b200.modules import Decoder
    d = Decoder(dim_in=128, style_dim=64, dim_out=512, hidden_dim=512, residual_dim=64)
    with pytest.raises(RuntimeError):
        d(torch.zeros(1, 128, 16), torch.zeros(1, 16), torch.zeros(1, 16), torch.zeros(1, 64))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "stylish_tts_b200")
    pat = re.compile(r"^\s*(from\s+oracle|import\s+oracle|from\s+\.+oracle)|oracle[/.]stylish_port|oracle/_ref", re.M)
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not pat.search(src), os.path.join(dp, f)
