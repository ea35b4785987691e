"""Host-side mirrors of the two reference classes `north_star` names outside the default model.yml (SURVEY.md
section 8 f2): same constructor arguments, ``state_dict()`` keys and ``forward`` signatures; the arithmetic runs in
``csrc/variants.cu`` behind ``stts_source_sinegen`` / ``stts_generator_block_snake``.  No torch fallback.

  SineGenerator            /root/reference/src/stylish_tts/train/models/cfm/cfm_mel_decoder.py:16-101
  AdaptiveGeneratorBlock   /root/reference/src/stylish_tts/train/models/ada_norm.py:11-127
"""
from __future__ import annotations

import itertools
from typing import Optional, Sequence

import torch
from torch import nn
from torch.nn.utils.parametrizations import weight_norm

from ..engine import Engine

_ids = itertools.count()


_ENGINES = {}


def variant_engine(device: torch.device) -> Engine:
    """One context per device for these operators, separate from the synthesis contexts (loading a tensor marks a
    context's packed model as stale)."""
    if device.type != "cuda":
        raise RuntimeError("stylish_tts_b200 modules run on a B200 only: move the module to CUDA (no CPU fallback)")
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _ENGINES:
        _ENGINES[idx] = Engine(torch.device("cuda", idx))
    return _ENGINES[idx]


_engine_for = variant_engine


class SineGenerator(nn.Module):
    """cfm_mel_decoder.py:16-101.  ``forward(f0)`` takes the reference's [B, L, 1] per-sample f0 and returns
    [B, L, 1]; the random draws the reference makes inside forward (uniform initial phases :66, normal noise :97) come
    from torch's CUDA generator unless passed in (``rand_ini`` [B, dim], ``noise`` [B, L, dim]) -- parity tests pass
    them.  ``forward_ragged`` is the batched form for utterances of different lengths."""

    def __init__(self, samp_rate, harmonic_num=0, sine_amp=0.1, noise_std=0.003, voiced_threshold=0):
        super().__init__()
        self.sine_amp = sine_amp
        self.noise_std = noise_std
        self.harmonic_num = harmonic_num
        self.dim = harmonic_num + 1
        self.sampling_rate = samp_rate
        self.voiced_threshold = voiced_threshold
        self.merge = nn.Sequential(nn.Linear(self.dim, 1, bias=False), nn.Tanh())

    def forward_ragged(self, f0: Sequence[torch.Tensor], rand_ini: Optional[torch.Tensor] = None,
                       noise: Optional[Sequence[torch.Tensor]] = None, want_phase: bool = False):
        dev = self.merge[0].weight.device
        eng = _engine_for(dev)
        lens = [int(t.numel()) for t in f0]
        flat = torch.cat([t.detach().to(dev, torch.float32).reshape(-1) for t in f0])
        if rand_ini is None:
            rand_ini = torch.rand(len(lens), self.dim, device=dev)
        if noise is None:
            nz = torch.randn(sum(lens), self.dim, device=dev)
        else:
            nz = torch.cat([t.detach().to(dev, torch.float32).reshape(-1, self.dim) for t in noise])
        out, phase = eng.sinegen(flat, lens, rand_ini.detach().to(dev, torch.float32).contiguous(), nz.contiguous(),
                                 self.merge[0].weight, self.harmonic_num, self.sampling_rate, self.sine_amp,
                                 self.noise_std, self.voiced_threshold, want_phase=want_phase)
        outs = list(out.split(lens))
        return (outs, list(phase.split(lens))) if want_phase else outs

    def forward(self, f0: torch.Tensor, rand_ini: Optional[torch.Tensor] = None, noise: Optional[torch.Tensor] = None):
        if f0.dim() != 3 or f0.shape[2] != 1:
            raise ValueError("SineGenerator.forward expects f0 of shape [B, L, 1]")
        b, length, _ = f0.shape
        outs = self.forward_ragged([f0[i, :, 0] for i in range(b)], rand_ini,
                                   None if noise is None else [noise[i] for i in range(b)])
        return torch.stack(outs).view(b, length, 1)


class _AdaptiveInstance(nn.Module):
    def __init__(self, style_dim, num_features):
        super().__init__()
        self.fc = nn.Linear(style_dim, num_features * 2)


class AdaptiveGeneratorBlock(nn.Module):
    """ada_norm.py:11-127: three stages of AdaIN -> Snake1D -> conv(k, dilation d_i) -> AdaIN -> Snake1D -> conv(k) with
    a residual connection.  ``forward(x [B, C, L], s [B, style_dim])``; ``forward_ragged`` takes a list of [C, L_b]."""

    def __init__(self, channels, kernel_size=3, dilation=(1, 3, 5), style_dim=64):
        super().__init__()
        if style_dim != 64:
            raise NotImplementedError("the style Linear kernels are built for style_dim 64 (default model.yml)")
        if channels % 64 != 0 or kernel_size % 2 == 0 or kernel_size > 15:
            raise NotImplementedError("channels must be a multiple of 64 and kernel_size odd <= 15")
        self.channels, self.kernel_size, self.dilation = channels, kernel_size, tuple(int(d) for d in dilation)

        def conv(d):
            return weight_norm(nn.Conv1d(channels, channels, kernel_size, 1, dilation=d,
                                         padding=int((kernel_size * d - d) / 2)))

        self.convs1 = nn.ModuleList([conv(d) for d in self.dilation])
        self.convs2 = nn.ModuleList([conv(1) for _ in self.dilation])
        self.adain1 = nn.ModuleList([_AdaptiveInstance(style_dim, channels) for _ in range(3)])
        self.adain2 = nn.ModuleList([_AdaptiveInstance(style_dim, channels) for _ in range(3)])
        self.alpha1 = nn.ParameterList([nn.Parameter(torch.ones(1, channels, 1)) for _ in range(3)])
        self.alpha2 = nn.ParameterList([nn.Parameter(torch.ones(1, channels, 1)) for _ in range(3)])
        object.__setattr__(self, "_loaded", None)
        object.__setattr__(self, "_name", f"generator_block_{next(_ids)}")

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        object.__setattr__(self, "_loaded", None)
        return super().load_state_dict(state_dict, strict=strict, assign=assign)

    def _engine(self) -> Engine:
        dev = self.alpha1[0].device
        eng = _engine_for(dev)
        if self._loaded is not eng:
            eng.load_state(self._name, self.state_dict())
            object.__setattr__(self, "_loaded", eng)
        return eng

    def forward_ragged(self, x: Sequence[torch.Tensor], s: torch.Tensor):
        eng = self._engine()
        lens = [int(t.shape[1]) for t in x]
        flat = torch.cat([t.detach().to(eng.device, torch.float32).contiguous().view(-1) for t in x])
        out = eng.snake_block(self._name, self.channels, self.kernel_size, self.dilation, flat, lens,
                              s.detach().to(eng.device, torch.float32).contiguous())
        return [o.view(self.channels, n) for o, n in zip(out.split([self.channels * n for n in lens]), lens)]

    def forward(self, x: torch.Tensor, s: torch.Tensor):
        return torch.stack(self.forward_ragged([x[i] for i in range(x.shape[0])], s))


class STFT(nn.Module):
    """train/models/stft.py:6-195 -- STFT / iSTFT as conv1d / conv_transpose1d with a windowed DFT basis (the module the
    reference swaps in for its ONNX export).  Same constructor, the same five registered buffers (built the same way:
    float64 numpy, cast to float32), ``transform(waveform [B, T]) -> (magnitude, x, y)`` each [B, bins, frames] and
    ``inverse(magnitude, x, y, length=None) -> [B, 1, T]``; ``*_ragged`` take lists of utterances."""

    def __init__(self, filter_length=800, hop_length=200, win_length=800, window="hann", center=True,
                 pad_mode="replicate"):
        super().__init__()
        import numpy as np
        import torch.nn.functional as F
        if window != "hann" or not center:
            raise NotImplementedError("the kernels implement the hann window with center=True (the module's defaults)")
        self.filter_length, self.hop_length, self.win_length = filter_length, hop_length, win_length
        self.n_fft, self.center, self.pad_mode = filter_length, center, pad_mode
        self.freq_bins = self.n_fft // 2 + 1
        win = torch.hann_window(win_length, periodic=True, dtype=torch.float32)
        if win_length < self.n_fft:
            win = F.pad(win, (0, self.n_fft - win_length))
        elif win_length > self.n_fft:
            win = win[: self.n_fft]
        self.register_buffer("window", win)
        n, k = np.arange(self.n_fft), np.arange(self.freq_bins)
        ang = 2 * np.pi * np.outer(k, n) / self.n_fft
        w = win.numpy()
        self.register_buffer("weight_forward_real", torch.from_numpy(np.cos(ang) * w).float().unsqueeze(1))
        self.register_buffer("weight_forward_imag", torch.from_numpy(-np.sin(ang) * w).float().unsqueeze(1))
        ang_t = 2 * np.pi * np.outer(n, k) / self.n_fft
        inv_w = w * (1.0 / self.n_fft)
        self.register_buffer("weight_backward_real", torch.from_numpy(np.cos(ang_t).T * inv_w).float().unsqueeze(1))
        self.register_buffer("weight_backward_imag", torch.from_numpy(np.sin(ang_t).T * inv_w).float().unsqueeze(1))
        object.__setattr__(self, "_loaded", None)
        object.__setattr__(self, "_name", f"stft_conv_{next(_ids)}")

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        object.__setattr__(self, "_loaded", None)
        return super().load_state_dict(state_dict, strict=strict, assign=assign)

    def _engine(self) -> Engine:
        eng = _engine_for(self.window.device)
        if self._loaded is not eng:
            eng.load_state(self._name, {k: v for k, v in self.state_dict().items() if k.startswith("weight_")})
            object.__setattr__(self, "_loaded", eng)
        return eng

    def transform_ragged(self, waves: Sequence[torch.Tensor]):
        eng = self._engine()
        lens = [int(w.numel()) for w in waves]
        flat = torch.cat([w.detach().to(eng.device, torch.float32).reshape(-1) for w in waves])
        mag, x, y = eng.stft_conv_transform(self._name, self.n_fft, self.hop_length, self.win_length, flat, lens,
                                            self.pad_mode)
        sizes = [self.freq_bins * (n // self.hop_length + 1) for n in lens]
        shape = [(self.freq_bins, n // self.hop_length + 1) for n in lens]
        return tuple([p.view(sh) for p, sh in zip(t.split(sizes), shape)] for t in (mag, x, y))

    def transform(self, waveform: torch.Tensor):
        mag, x, y = self.transform_ragged([waveform[i] for i in range(waveform.shape[0])])
        return torch.stack(mag), torch.stack(x), torch.stack(y)

    def inverse_ragged(self, magnitude: Sequence[torch.Tensor], x: Sequence[torch.Tensor], y: Sequence[torch.Tensor],
                       length: Optional[Sequence[int]] = None):
        eng = self._engine()
        frames = [int(m.shape[1]) for m in magnitude]
        cat = lambda ts: torch.cat([t.detach().to(eng.device, torch.float32).contiguous().view(-1) for t in ts])
        ol = None
        if length is not None:
            ol = [min(int(n), (f - 1) * self.hop_length) for n, f in zip(length, frames)]
        wave, ol = eng.stft_conv_inverse(self._name, self.n_fft, self.hop_length, self.win_length, cat(magnitude), cat(x),
                                         cat(y), frames, ol)
        return list(wave.split(ol))

    def inverse(self, magnitude: torch.Tensor, x: torch.Tensor, y: torch.Tensor, length=None):
        b = magnitude.shape[0]
        outs = self.inverse_ragged([magnitude[i] for i in range(b)], [x[i] for i in range(b)], [y[i] for i in range(b)],
                                   None if length is None else [length] * b)
        return torch.stack(outs).unsqueeze(1)
