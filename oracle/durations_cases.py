"""Seeded inputs for the duration-rule parity tests (TEST INFRASTRUCTURE ONLY): shared by
oracle/make_golden_durations.py (which runs the real reference on them) and tests/."""
from __future__ import annotations

import numpy as np
import torch

TABLE = [1, 2, 3, 4, 5, 6, 7, 9, 12, 15, 18, 22, 27, 32, 38, 46]  # train/utils.py:391-393


def random_logits(n: int = 1 << 20, seed: int = 77) -> torch.Tensor:
    """n logit rows: the first half peaked (sigma 3: argmax path and small soft durations), the second half
    flat (sigma 0.9: soft path with large durations)."""
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn((n, 16), generator=g) * 3.0
    logits[n // 2:] *= 0.3
    return logits


def tie_distance(logits: torch.Tensor):
    """(|frac(soft) - 0.5| with soft evaluated in fp64, bool: the row takes the soft path (table[argmax] >= 7))."""
    x = logits.double()
    p = torch.softmax(x, dim=-1)
    tab = torch.tensor(TABLE, dtype=torch.float64)
    soft = (p * tab).sum(-1)
    arg = tab[torch.argmax(logits, dim=-1)]
    return ((soft - torch.floor(soft)) - 0.5).abs(), arg >= 7


def adversarial_logits() -> torch.Tensor:
    """Rows whose fp32 result is fixed by IEEE arithmetic alone (at most two non-zero probabilities, so neither
    the summation order nor reciprocal-vs-division matters, and exp is evaluated at 0 and at -k * 2^-24 where any
    sane expf is exact):
      * exact ties: two equal maxima on classes (i, j) with table >= 7 -> soft = (d_i + d_j) / 2 = x.5 exactly
        -> round-half-to-even;
      * the same with one maximum lowered by k fp32 ulps of 0 (k = 1 .. 4096): x.5 -+ epsilon, epsilon from far
        below half an ulp of the result (rounds back onto the tie) to well above it;
      * one-hot saturated rows for every class, all-equal rows (first-maximum argmax), a row per (argmax < 7) class
        with a heavy tail that would give a large soft value (the rule must ignore it)."""
    rows = []
    neg = -1.0e4                                     # exp underflows to exactly 0
    ks = [0, 1, 2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64, 96, 128, 256, 512, 1024, 2048, 4096]
    for i in range(6, 16):
        for j in range(i + 1, 16):
            for k in ks:
                for lower in (i, j):
                    r = np.full(16, neg, dtype=np.float32)
                    r[i] = r[j] = 0.0
                    r[lower] = np.float32(-k * 2.0 ** -24)
                    rows.append(r)
    for c in range(16):
        r = np.zeros(16, dtype=np.float32)
        r[c] = 20.0
        rows.append(r)
        r = np.full(16, neg, dtype=np.float32)
        r[c] = 0.0
        rows.append(r)
    rows.append(np.zeros(16, dtype=np.float32))              # all equal: argmax = class 0 -> duration 1
    rows.append(np.full(16, 3.25, dtype=np.float32))
    for c in range(6):                                        # argmax < 7 wins whatever the soft value says
        r = np.full(16, 1.0, dtype=np.float32)
        r[c] = 1.0 + 2.0 ** -20
        r[15] = 1.0
        rows.append(r)
    return torch.from_numpy(np.stack(rows))
