"""CPU tests of the host layer: state_dict layout, token-rate mirrors vs the oracle, and that the
C-ABI library loads and exports every symbol include/stts_b200.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest
import torch

from tests.util import ROOT, fixture_sds
from oracle import stylish_port as port
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


def test_token_rate_mirrors_equal_oracle(cpu_nets):
    sds = fixture_sds()
    g = torch.Generator().manual_seed(5)
    t = fixture.synth_tokens(14, g)
    tk, ln = t[None], torch.tensor([14])
    with torch.no_grad():
        a = cpu_nets["duration_predictor"].forward_torch(tk, ln)
        b = port.duration_predictor(sds["duration_predictor"], tk, ln)
        assert float((a - b).abs().max()) <= 1e-6
        e1, _, _ = cpu_nets["pe_text_encoder"].forward_torch(tk, ln)
        e2, _, _ = port.text_encoder(sds["pe_text_encoder"], "", tk, ln)
        assert float((e1 - e2).abs().max()) <= 1e-6
        s1 = cpu_nets["pe_text_style_encoder"].forward_torch(e1, ln)
        s2 = port.text_style_encoder(sds["pe_text_style_encoder"], "", e2, ln)
        assert float((s1 - s2).abs().max()) <= 1e-6


def test_padded_batch_has_per_utterance_semantics(cpu_nets):
    sds = fixture_sds()
    g = torch.Generator().manual_seed(6)
    toks = [fixture.synth_tokens(n, g) for n in (7, 19, 12)]
    tk = torch.zeros(3, 19, dtype=torch.int64)
    for i, t in enumerate(toks):
        tk[i, : t.numel()] = t
    ln = torch.tensor([7, 19, 12])
    with torch.no_grad():
        a = cpu_nets["duration_predictor"].forward_torch(tk, ln)
        enc, _, _ = cpu_nets["pe_text_encoder"].forward_torch(tk, ln)
        sty = cpu_nets["pe_text_style_encoder"].forward_torch(enc, ln)
        for i, t in enumerate(toks):
            n = t.numel()
            b = port.duration_predictor(sds["duration_predictor"], t[None], torch.tensor([n]))
            assert float((a[i, :n] - b[0]).abs().max()) <= 2e-5
            e, _, _ = port.text_encoder(sds["pe_text_encoder"], "", t[None], torch.tensor([n]))
            s = port.text_style_encoder(sds["pe_text_style_encoder"], "", e, torch.tensor([n]))
            assert float((sty[i] - s[0]).abs().max()) <= 2e-5


def test_cross_attention_ragged_mode_equals_b1(cpu_nets):
    sds = fixture_sds()
    pep = cpu_nets["pitch_energy_predictor"]
    g = torch.Generator().manual_seed(7)
    Ts, Fs = (5, 16), (20, 50)      # T = 5 puts every token inside the +-5 band (all-masked rows)
    toks = [fixture.synth_tokens(n, g) for n in Ts]
    durs = [fixture.synth_durations(t, f, g) for t, f in zip(Ts, Fs)]
    tk = torch.zeros(2, 16, dtype=torch.int64)
    ft = torch.zeros(2, 50, dtype=torch.int64)
    for i in range(2):
        tk[i, : Ts[i]] = toks[i]
        ft[i, : Fs[i]] = port.duration_to_indices(durs[i].float())
    ln, fl = torch.tensor(Ts), torch.tensor(Fs)
    with torch.no_grad():
        enc, _, _ = cpu_nets["pe_text_encoder"].forward_torch(tk, ln)
        sty = cpu_nets["pe_text_style_encoder"].forward_torch(enc, ln)
        pros = pep.prosody_encoder.forward_torch(enc, sty, ln)
        x = pep.compute_cross(pros, ft, sty, ln, frame_lengths=fl)
        for i in range(2):
            taps = {}
            e, _, _ = port.text_encoder(sds["pe_text_encoder"], "", toks[i][None], ln[i:i + 1])
            s = port.text_style_encoder(sds["pe_text_style_encoder"], "", e, ln[i:i + 1])
            al = port.duration_to_alignment(durs[i].float())[None]
            port.pitch_energy_predictor(sds["pitch_energy_predictor"], e, ln[i:i + 1], al, s, taps)
            assert float((x[i, :, : Fs[i]] - taps["pe_cross"][0]).abs().max()) <= 2e-5


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
