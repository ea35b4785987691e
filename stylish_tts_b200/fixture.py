"""Deterministic fixture weights and synthetic inputs (shared by tests, bench and smoke).

The reference ships no checkpoint and plain random init leaves large parts of the path
degenerate (zero-initialised flow projections, GRN gamma/beta, unvoiced
pitch).  The fixture is therefore a *documented, seeded re-initialisation* (SURVEY section 8c):

  * ``tests/golden/manifest.json`` records, for every ``state_dict()`` key of the five hot
    modules, its shape and the mean / std of the reference's own default initialisation
    (written by ``oracle/make_golden.py`` from ``build_model``, models.py:27-103).
  * every tensor is drawn as ``mean + std_eff * N(0,1)`` from a CPU ``torch.Generator`` seeded
    with ``crc32(module/key) ^ seed``; ``std_eff = std`` unless the default init is constant
    (zeros / ones), in which case ``std_eff = 0.02`` so that nothing on the path is a no-op;
  * ``pitch_energy_predictor.F0_proj``: weight x400, bias = 100 so the predicted pitch crosses
    the 10 Hz / 20 Hz voicing thresholds of ``generate_pcph`` (generator.py:280-287).

Because it depends only on (manifest, seed) the same weights can be rebuilt on the GPU box,
where ``/root/reference`` does not exist.
"""
from __future__ import annotations

import json
import os
import zlib
from typing import Dict, List, Optional, Tuple

import torch

MODULES = (
    "duration_predictor",
    "pe_text_encoder",
    "pe_text_style_encoder",
    "pitch_energy_predictor",
    "speech_predictor",
)

_MANIFEST_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                              "tests", "golden", "manifest.json")


def load_manifest(path: Optional[str] = None) -> Dict[str, Dict[str, dict]]:
    with open(path or _MANIFEST_PATH) as f:
        return json.load(f)


def _gen(module: str, key: str, seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(f"{module}/{key}".encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFFFFFFFFFF)
    return g


def fixture_state_dict(module: str, manifest: Dict[str, Dict[str, dict]], seed: int = 0,
                       skip_prefixes: Tuple[str, ...] = ("posterior_encoder.",)) -> Dict[str, torch.Tensor]:
    """fp32 CPU tensors for every key of ``module`` (posterior encoder skipped: training only)."""
    out: Dict[str, torch.Tensor] = {}
    for key, info in manifest[module].items():
        if key.startswith(skip_prefixes):
            continue
        shape = tuple(info["shape"])
        mean, std = float(info["mean"]), float(info["std"])
        if not std > 1e-12:
            std = 0.02
        if module == "pitch_energy_predictor" and key == "F0_proj.weight":
            std *= 400.0
        if module == "pitch_energy_predictor" and key == "F0_proj.bias":
            mean, std = 100.0, 0.0
        t = torch.randn(shape, generator=_gen(module, key, seed), dtype=torch.float32)
        out[key] = (t * std + mean).contiguous()
    return out


def fixture_nets(manifest=None, seed: int = 0) -> Dict[str, Dict[str, torch.Tensor]]:
    manifest = manifest or load_manifest()
    return {m: fixture_state_dict(m, manifest, seed) for m in MODULES}


# ----------------------------------------------------------------------------- synthetic inputs
def synth_tokens(n_tokens: int, gen: torch.Generator, n_symbols: int = 178) -> torch.Tensor:
    """Uniform ids in [1, n_symbols) wrapped with pad id 0 at both ends (dataloader.py:178-179,
    test_onnx.py:50-53).  ``n_tokens`` counts the two pad tokens."""
    body = torch.randint(1, n_symbols, (max(n_tokens - 2, 0),), generator=gen, dtype=torch.int64)
    z = torch.zeros(1, dtype=torch.int64)
    return torch.cat([z, body, z])[:max(n_tokens, 1)]


def synth_durations(n_tokens: int, n_frames: int, gen: torch.Generator, max_dur: int = 46) -> torch.Tensor:
    """Integer durations >= 1, <= max_dur, summing exactly to ``n_frames``."""
    assert n_tokens <= n_frames <= n_tokens * max_dur, (n_tokens, n_frames)
    dur = torch.ones(n_tokens, dtype=torch.int64)
    extra = n_frames - n_tokens
    # multinomial allocation of the remaining frames, then clip + redistribute deterministically
    w = torch.rand(n_tokens, generator=gen) + 0.1
    add = torch.floor(w / w.sum() * extra).to(torch.int64)
    dur += add
    dur.clamp_(max=max_dur)
    i = 0
    while int(dur.sum()) < n_frames:
        if dur[i % n_tokens] < max_dur:
            dur[i % n_tokens] += 1
        i += 1
    return dur


def synth_utterance(seconds: float, gen: torch.Generator, tokens_per_sec: float = 24.0):
    """(tokens [T] int64, durations [T] int64, F) for an utterance of ``seconds`` (80 fps frames)."""
    n_frames = max(20, int(round(80.0 * seconds)))
    n_tok = max(4, int(round(tokens_per_sec * seconds)))
    n_tok = min(n_tok, 510, n_frames)
    n_tok = max(n_tok, -(-n_frames // 46))
    return synth_tokens(n_tok, gen), synth_durations(n_tok, n_frames, gen), n_frames


def synth_batch(seconds: List[float], seed: int):
    gen = torch.Generator(device="cpu")
    gen.manual_seed(1234 + seed)
    return [synth_utterance(s, gen) for s in seconds]


def mixed_lengths(batch: int, seed: int, lo: float = 0.25, hi: float = 10.0) -> List[float]:
    """BASELINE.json configs[1]/[3]: utterance lengths uniform in [lo, hi] seconds."""
    gen = torch.Generator(device="cpu")
    gen.manual_seed(4321 + seed)
    return (lo + (hi - lo) * torch.rand(batch, generator=gen)).tolist()


def draw_noise(n_frames: int, gen: torch.Generator) -> Dict[str, torch.Tensor]:
    """The three stochastic draws of one B = 1 reference forward, in call order
    (flow.py:314, generator.py:275-277, generator.py:306)."""
    f4 = 4 * n_frames
    return {
        "prior": torch.randn((1, 128, f4), generator=gen),
        "source": torch.randn((1, 1, f4 * 75), generator=gen),
        "phase": torch.rand((1, 1), generator=gen),
    }
