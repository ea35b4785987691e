"""Parity metrics shared by tests/ and bench.py's parity leg (TEST INFRASTRUCTURE ONLY)."""
from __future__ import annotations

import torch

_MEL = {}


def log_mel(audio: torch.Tensor) -> torch.Tensor:
    """log(1e-5 + mel) with the reference's training mel transform (train/train_context.py:151-157:
    torchaudio MelSpectrogram, n_mels 80, n_fft 2048, win 1200, hop 300, 24 kHz), on CPU in fp32."""
    import torchaudio
    if "t" not in _MEL:
        _MEL["t"] = torchaudio.transforms.MelSpectrogram(n_mels=80, n_fft=2048, win_length=1200, hop_length=300,
                                                         sample_rate=24000)
    return torch.log(1e-5 + _MEL["t"](audio.detach().reshape(1, -1).to("cpu", torch.float32)))


def mel_l1(a: torch.Tensor, b: torch.Tensor) -> float:
    """mean |log-mel(a) - log-mel(b)| of two waveforms (north_star: "bf16 mode ... with mel L1 reported")."""
    return float((log_mel(a) - log_mel(b)).abs().mean())


def maxabs(a: torch.Tensor, b: torch.Tensor) -> float:
    return float((a.detach().double().cpu().reshape(-1) - b.detach().double().cpu().reshape(-1)).abs().max())
