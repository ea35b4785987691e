"""Seeded inputs / weights for the SineGenerator and Snake AdaptiveGeneratorBlock parity cases (test infrastructure;
shared by oracle/make_golden_variants.py, which replays them through the REAL reference classes, and by the tests, which
rebuild the same tensors on the GPU box where /root/reference does not exist).  CPU torch.Generator streams are
platform-independent, so only the reference OUTPUTS need to be committed."""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import torch

SAMP_RATE = 24000
HARMONICS = 8            # dim = 9
STYLE_DIM = 64


def f0_curve(n_samples: int, seed: int, lo: float = 80.0, hi: float = 420.0, unvoiced_share: float = 0.3) -> torch.Tensor:
    """A per-sample f0 track [n_samples]: piecewise-smooth voiced stretches (frame-rate control points, linearly
    interpolated like the ringformer path's f0 upsampling) separated by unvoiced (0 Hz) gaps."""
    g = torch.Generator().manual_seed(seed)
    hop = 300
    n_frames = (n_samples + hop - 1) // hop + 1
    walk = torch.cumsum(torch.randn(n_frames, generator=g) * 0.04, 0)
    f = torch.exp(math.log(lo) + (math.log(hi) - math.log(lo)) * torch.sigmoid(walk + torch.randn(1, generator=g)))
    voiced = torch.ones(n_frames)
    pos = int(torch.randint(3, 12, (1,), generator=g))               # voiced and unvoiced stretches alternate
    while pos < n_frames:
        gap = int(torch.randint(2, max(3, int(20 * unvoiced_share / 0.3)), (1,), generator=g))
        voiced[pos: pos + gap] = 0
        pos += gap + int(torch.randint(4, 40, (1,), generator=g))
    f = f * voiced
    t = torch.arange(n_samples, dtype=torch.float32) / hop
    i0 = t.floor().long().clamp(max=n_frames - 2)
    w = t - i0
    out = f[i0] * (1 - w) + f[i0 + 1] * w
    out = out * ((voiced[i0] > 0) & (voiced[i0 + 1] > 0))           # no interpolation through a voicing boundary
    return out.contiguous()


def sinegen_case(lengths: Sequence[int], seed: int) -> Dict[str, object]:
    """Ragged batch: per-utterance f0 [L_b], uniform initial-phase draws [dim], normal draws [L_b, dim], and the
    merge weight [1, dim] (nn.Linear(dim, 1, bias=False))."""
    g = torch.Generator().manual_seed(seed)
    dim = HARMONICS + 1
    f0 = [f0_curve(n, seed * 131 + b) for b, n in enumerate(lengths)]
    rand_ini = [torch.rand(dim, generator=g) for _ in lengths]
    noise = [torch.randn(n, dim, generator=g) for n in lengths]
    merge = (torch.rand(1, dim, generator=g) * 2 - 1) / math.sqrt(dim)
    return {"f0": f0, "rand_ini": rand_ini, "noise": noise, "merge_weight": merge}


def snake_state_dict(channels: int, kernel_size: int, seed: int, legacy: bool = False) -> Dict[str, torch.Tensor]:
    """state_dict of AdaptiveGeneratorBlock(channels, kernel_size, dilation, STYLE_DIM) with the reference's key names
    (ada_norm.py:12-107).  Values are seeded: conv v ~ N(0, 0.01)-like magnitudes scaled so activations stay O(1),
    g near ||v||, alphas in [0.5, 2] (the reference initialises them to 1; a trained block does not keep them there)."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, torch.Tensor] = {}
    for grp in ("convs1", "convs2"):
        for i in range(3):
            v = torch.randn(channels, channels, kernel_size, generator=g) / math.sqrt(channels * kernel_size)
            norm = v.flatten(1).norm(dim=1).view(channels, 1, 1)
            gain = norm * (0.7 + 0.6 * torch.rand(channels, 1, 1, generator=g))
            if legacy:
                sd[f"{grp}.{i}.weight_g"], sd[f"{grp}.{i}.weight_v"] = gain, v
            else:
                sd[f"{grp}.{i}.parametrizations.weight.original0"] = gain
                sd[f"{grp}.{i}.parametrizations.weight.original1"] = v
            sd[f"{grp}.{i}.bias"] = torch.randn(channels, generator=g) * 0.05
    for grp in ("adain1", "adain2"):
        for i in range(3):
            sd[f"{grp}.{i}.fc.weight"] = torch.randn(2 * channels, STYLE_DIM, generator=g) / math.sqrt(STYLE_DIM) * 0.5
            sd[f"{grp}.{i}.fc.bias"] = torch.randn(2 * channels, generator=g) * 0.1
    for grp in ("alpha1", "alpha2"):
        for i in range(3):
            sd[f"{grp}.{i}"] = 0.5 + 1.5 * torch.rand(1, channels, 1, generator=g)
    return sd


def snake_case(channels: int, lengths: Sequence[int], seed: int) -> Tuple[List[torch.Tensor], torch.Tensor]:
    """x_b [C, L_b] per utterance and styles [B, STYLE_DIM]."""
    g = torch.Generator().manual_seed(seed + 7)
    x = [torch.randn(channels, n, generator=g) for n in lengths]
    s = torch.randn(len(lengths), STYLE_DIM, generator=g)
    return x, s


SINEGEN_GOLDEN = {"lengths": (9000, 4321), "seed": 3}
SNAKE_GOLDEN = [  # (channels, kernel, dilation, lengths, seed)
    (128, 3, (1, 3, 5), (150, 37), 11),
    (128, 7, (1, 3, 5), (150, 37), 12),
    (128, 11, (1, 3, 5), (150, 64), 13),
]


def wave_case(lengths: Sequence[int], seed: int) -> List[torch.Tensor]:
    """Speech-like test waveforms: a few decaying harmonics of a drifting f0 plus noise, |x| < 1."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for n in lengths:
        t = torch.arange(n, dtype=torch.float64) / SAMP_RATE
        f0 = 110.0 + 90.0 * float(torch.rand(1, generator=g)) + 20.0 * torch.sin(2 * math.pi * 1.7 * t)
        ph = 2 * math.pi * torch.cumsum(f0, 0) / SAMP_RATE
        x = sum((0.5 / h) * torch.sin(h * ph + float(torch.rand(1, generator=g))) for h in range(1, 9))
        x = 0.4 * x + 0.02 * torch.randn(n, generator=g, dtype=torch.float64)
        out.append(x.float().contiguous())
    return out


STFT_GOLDEN = [  # (filter_length, hop_length, win_length, lengths, seed); the module's defaults and the export's sizes
    (800, 200, 800, (3000, 1234), 41),
    (2048, 300, 1200, (6000, 2500), 42),
]
