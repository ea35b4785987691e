"""CPU ORACLE for SURVEY.md section 8(f) rank 2: the source / activation variants `north_star` names that sit outside the
live inference path -- `SineGenerator` and the Snake `AdaptiveGeneratorBlock`.

TEST INFRASTRUCTURE ONLY (same rules as oracle/stylish_port.py): imported by tests/ and by the golden-vector script
oracle/make_golden_variants.py, never by the product package.  Functional restatements in plain torch ops on CPU, each
citing the reference lines it follows; pinned bit-exactly against the reference classes themselves (imported from
/root/reference in the build container) by oracle/make_golden_variants.py -> tests/golden/variants.safetensors.

Random draws are explicit inputs (the reference calls torch.rand / torch.randn_like inside forward); the golden script
replays the reference under a fixed torch seed and records the draws it made.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence

import torch
import torch.nn.functional as F


# ----------------------------------------------------------------------------------------------------------------
# SineGenerator                                   train/models/cfm/cfm_mel_decoder.py:16-101
# ----------------------------------------------------------------------------------------------------------------
def sinegen_phase(f0: torch.Tensor, rand_ini: torch.Tensor, samp_rate: float, harmonic_num: int) -> Dict[str, torch.Tensor]:
    """_f02sine (cfm_mel_decoder.py:58-83) with the harmonic expansion of forward (:87-91).

    f0 [B, L, 1] Hz per SAMPLE; rand_ini [B, dim] uniform draws (column 0 is overwritten with 0, :69).
    torch.cumsum on CPU accumulates fp32 inputs in fp64 and rounds every output to fp32 (ATen cumsum_cpu_kernel,
    acc_type<float, false> = double): the restatement just calls it; the CUDA kernel reproduces that contract."""
    B, L, _ = f0.shape
    dim = harmonic_num + 1
    f0_buf = torch.zeros(B, L, dim)
    f0_buf[:, :, 0] = f0[:, :, 0]
    for idx in range(harmonic_num):
        f0_buf[:, :, idx + 1] = f0_buf[:, :, 0] * (idx + 2)                       # :90-91
    rad = (f0_buf / samp_rate) % 1                                                # :64
    ini = rand_ini.clone()
    ini[:, 0] = 0                                                                 # :69
    rad[:, 0, :] = rad[:, 0, :] + ini                                             # :70
    c1 = torch.cumsum(rad, 1)
    over = c1 % 1                                                                 # :73
    wrap = (over[:, 1:, :] - over[:, :-1, :]) < 0                                 # :74
    shift = torch.zeros_like(rad)
    shift[:, 1:, :] = wrap * -1.0                                                 # :75-76
    c2 = torch.cumsum(rad + shift, dim=1)
    sines = torch.sin(c2 * 2 * math.pi)                                           # :78 (fp32 tensor x python scalars)
    return {"rad": rad, "c1": c1, "c2": c2, "sines": sines}


def sinegen(f0: torch.Tensor, rand_ini: torch.Tensor, noise: torch.Tensor, merge_weight: torch.Tensor, samp_rate: float,
            harmonic_num: int = 0, sine_amp: float = 0.1, noise_std: float = 0.003,
            voiced_threshold: float = 0.0) -> Dict[str, torch.Tensor]:
    """SineGenerator.forward (cfm_mel_decoder.py:85-101).  noise [B, L, dim] standard normal draws (the reference's
    torch.randn_like(sine_waves), :97); merge_weight [1, dim] = merge[0].weight (Linear without bias, then tanh)."""
    ph = sinegen_phase(f0, rand_ini, samp_rate, harmonic_num)
    sine_waves = ph["sines"] * sine_amp                                           # :93
    uv = torch.ones_like(f0) * (f0 > voiced_threshold)                            # :52-56
    noise_amp = uv * noise_std + (1 - uv) * sine_amp / 3                          # :96
    sine_waves = sine_waves * uv + noise_amp * noise                              # :97-99
    out = torch.tanh(F.linear(sine_waves, merge_weight))                          # :47-50, :101
    ph.update({"uv": uv, "sine_waves": sine_waves, "out": out})
    return ph


# ----------------------------------------------------------------------------------------------------------------
# AdaptiveGeneratorBlock (Snake1D)                train/models/ada_norm.py:11-127
# ----------------------------------------------------------------------------------------------------------------
def _weight(sd: Dict[str, torch.Tensor], p: str) -> torch.Tensor:
    """weight_norm parametrisation (dim 0): w = g * v / ||v||, norm over every dim but 0."""
    if p + "parametrizations.weight.original0" in sd:
        g, v = sd[p + "parametrizations.weight.original0"], sd[p + "parametrizations.weight.original1"]
        return torch._weight_norm(v, g, 0)
    if p + "weight_g" in sd:
        return torch._weight_norm(sd[p + "weight_v"], sd[p + "weight_g"], 0)
    return sd[p + "weight"]


def adaptive_instance(sd, p: str, x: torch.Tensor, s: torch.Tensor) -> torch.Tensor:
    """AdaptiveInstance.forward (ada_norm.py:136-140): InstanceNorm1d(affine=False, eps 1e-5), (1 + gamma) * n + beta."""
    h = F.linear(s, sd[p + "fc.weight"], sd[p + "fc.bias"])
    h = h.view(h.size(0), h.size(1), 1)
    gamma, beta = torch.chunk(h, chunks=2, dim=1)
    return (1 + gamma) * F.instance_norm(x, eps=1e-5) + beta


def snake_block(sd: Dict[str, torch.Tensor], x: torch.Tensor, s: torch.Tensor, kernel_size: int = 3,
                dilation: Sequence[int] = (1, 3, 5), trace: Optional[dict] = None) -> torch.Tensor:
    """AdaptiveGeneratorBlock.forward (ada_norm.py:109-120).  x [B, C, L], s [B, style_dim]."""
    def pad(k, d):
        return int((k * d - d) / 2)                                               # common.py:11-12

    for i in range(3):
        a1, a2 = sd[f"alpha1.{i}"], sd[f"alpha2.{i}"]
        xt = adaptive_instance(sd, f"adain1.{i}.", x, s)
        xt = xt + (1 / a1) * (torch.sin(a1 * xt) ** 2)                            # Snake1D, :114
        xt = F.conv1d(xt, _weight(sd, f"convs1.{i}."), sd[f"convs1.{i}.bias"], 1, pad(kernel_size, dilation[i]),
                      dilation[i])
        if trace is not None:
            trace[f"c1.{i}"] = xt
        xt = adaptive_instance(sd, f"adain2.{i}.", xt, s)
        xt = xt + (1 / a2) * (torch.sin(a2 * xt) ** 2)                            # :117
        xt = F.conv1d(xt, _weight(sd, f"convs2.{i}."), sd[f"convs2.{i}.bias"], 1, pad(kernel_size, 1), 1)
        x = xt + x                                                                # :119
        if trace is not None:
            trace[f"x.{i}"] = x
    return x


# ----------------------------------------------------------------------------------------------------------------
# STFT (conv1d / conv_transpose1d DFT, the ONNX-export stand-in)      train/models/stft.py:6-195
# ----------------------------------------------------------------------------------------------------------------
def stft_conv_weights(filter_length: int, win_length: int) -> Dict[str, torch.Tensor]:
    """The four registered buffers of STFT.__init__ (stft.py:38-96): periodic hann window zero-padded AT THE END to
    n_fft, forward basis window * (cos, -sin), backward basis window / n_fft * (cos, sin); built in float64 numpy and
    cast to float32 exactly like the reference."""
    import numpy as np
    n_fft = filter_length
    bins = n_fft // 2 + 1
    window = torch.hann_window(win_length, periodic=True, dtype=torch.float32)
    if win_length < n_fft:
        window = F.pad(window, (0, n_fft - win_length))                            # :43-46
    elif win_length > n_fft:
        window = window[:n_fft]                                                   # :47-48
    n, k = np.arange(n_fft), np.arange(bins)
    angle = 2 * np.pi * np.outer(k, n) / n_fft                                    # :55
    w = window.numpy()
    fwd_real = torch.from_numpy(np.cos(angle) * w).float().unsqueeze(1)           # :56-70
    fwd_imag = torch.from_numpy(-np.sin(angle) * w).float().unsqueeze(1)
    angle_t = 2 * np.pi * np.outer(n, k) / n_fft                                  # :79
    inv_w = w * (1.0 / n_fft)                                                     # :85
    bwd_real = torch.from_numpy(np.cos(angle_t).T * inv_w).float().unsqueeze(1)   # :80-96
    bwd_imag = torch.from_numpy(np.sin(angle_t).T * inv_w).float().unsqueeze(1)
    return {"window": window, "fwd_real": fwd_real, "fwd_imag": fwd_imag, "bwd_real": bwd_real, "bwd_imag": bwd_imag}


def stft_conv_transform(waveform: torch.Tensor, filter_length: int, hop_length: int, win_length: int,
                        pad_mode: str = "replicate"):
    """STFT.transform (stft.py:98-136): waveform [B, T] -> (magnitude, re / magnitude, im / magnitude), each
    [B, n_fft / 2 + 1, frames]; center padding n_fft / 2 in `pad_mode`; magnitude = sqrt(re^2 + im^2 + 1e-14)."""
    wts = stft_conv_weights(filter_length, win_length)
    pad_len = filter_length // 2
    x = F.pad(waveform.unsqueeze(1), (pad_len, pad_len), mode=pad_mode)            # :105-109 (pad on the 3-D view)
    real = F.conv1d(x, wts["fwd_real"], bias=None, stride=hop_length, padding=0)
    imag = F.conv1d(x, wts["fwd_imag"], bias=None, stride=hop_length, padding=0)
    mag = torch.sqrt(real ** 2 + imag ** 2 + 1e-14)                               # :127
    return mag, real / mag, imag / mag, real, imag


def stft_conv_inverse(magnitude: torch.Tensor, x: torch.Tensor, y: torch.Tensor, filter_length: int, hop_length: int,
                      win_length: int, length: Optional[int] = None) -> torch.Tensor:
    """STFT.inverse (stft.py:138-186): conv_transpose1d with the windowed / n_fft-scaled basis, real minus imaginary
    part, n_fft / 2 samples trimmed at both ends (no window-envelope normalisation, no DC / Nyquist weighting: the
    reference calls it an approximate reconstruction, :73-77)."""
    wts = stft_conv_weights(filter_length, win_length)
    real_rec = F.conv_transpose1d(magnitude * x, wts["bwd_real"], bias=None, stride=hop_length)
    imag_rec = F.conv_transpose1d(magnitude * y, wts["bwd_imag"], bias=None, stride=hop_length)
    wave = real_rec - imag_rec                                                    # :170
    pad_len = filter_length // 2
    wave = wave[..., pad_len:-pad_len]                                            # :173-177
    if length is not None:
        wave = wave[..., :length]
    return wave
