"""CPU oracle for the Stylish-TTS synthesis forward path (TEST INFRASTRUCTURE ONLY).

This file is the *checker*, never the product: only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s CPU-baseline legs may import it.  The shipped path lives in
``stylish_tts_b200/`` and must fail loudly when its CUDA library is missing.

It is a functional restatement (plain functions over a ``state_dict``-style mapping of
fp32 tensors, torch CPU ops for the floating-point arithmetic) of the reference modules
listed below.  All citations are ``/root/reference/src/stylish_tts/train/...`` file:line.

Parity pin: the reference has no tests / golden vectors of its own, so
the pin is "outputs of the reference itself run here": ``oracle/make_golden.py`` imports the
reference, runs it on seeded fixture weights and inputs, checks that this port reproduces it
(max-abs printed, required <= 2e-6 on every recorded tensor) and writes ``tests/golden/*``.

Every tensor constructor takes its device from an input, so the same functions also serve as the
"stock PyTorch kernels on the B200" baseline of bench.py (state dicts and inputs moved to cuda);
on CPU tensors the arithmetic is unchanged.

Conventions: ``sd`` is a mapping key -> tensor using the reference ``state_dict()`` names,
``p`` is a key prefix ("decoder.", "flow.flows.14.", ...).  Tensors use the reference layout
([B, C, time]) unless stated otherwise.  Stochastic draws are *inputs* (``noise`` dict).
"""
from __future__ import annotations

import math
from typing import Dict, Mapping, Optional, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
SD = Mapping[str, Tensor]

CLASS_TO_DUR = [1, 2, 3, 4, 5, 6, 7, 9, 12, 15, 18, 22, 27, 32, 38, 46]  # utils.py:391-393
SAMPLE_RATE = 24000
N_FFT = 2048
WIN = 1200
HOP = 75  # generator hop = hop_length // 4, speech_predictor.py:77-83


# ----------------------------------------------------------------------------- weights
def wn_weight(sd: SD, p: str) -> Tensor:
    """Effective weight of a weight-normed layer, both dialects.

    parametrizations.weight_norm -> ``parametrizations.weight.original0`` (g) / ``original1`` (v)
    (ada_norm.py:158-163, decoder.py:35-45); legacy nn.utils.weight_norm -> ``weight_g`` /
    ``weight_v`` (flow.py:40,52,60).  Norm over every dim except 0.
    """
    if p + "parametrizations.weight.original0" in sd:
        g = sd[p + "parametrizations.weight.original0"]
        v = sd[p + "parametrizations.weight.original1"]
    elif p + "weight_g" in sd:
        g = sd[p + "weight_g"]
        v = sd[p + "weight_v"]
    else:
        return sd[p + "weight"]
    return torch._weight_norm(v, g, 0)


def _bias(sd: SD, p: str) -> Optional[Tensor]:
    return sd.get(p + "bias")


def seq_mask(lengths: Tensor, max_len: int) -> Tensor:
    """utils.py:52-56."""
    x = torch.arange(max_len, dtype=lengths.dtype, device=lengths.device)
    return x.unsqueeze(0) < lengths.unsqueeze(1)


# ----------------------------------------------------------------------------- token-rate blocks
def channel_layer_norm(sd: SD, p: str, x: Tensor, eps: float = 1e-4) -> Tensor:
    """text_encoder.py:15-33 -- LayerNorm over dim 1 with gamma/beta."""
    mean = torch.mean(x, 1, keepdim=True)
    var = torch.mean((x - mean) ** 2, 1, keepdim=True)
    x = (x - mean) * torch.rsqrt(var + eps)
    return x * sd[p + "gamma"].view(1, -1, 1) + sd[p + "beta"].view(1, -1, 1)


def ada_layer_norm(sd: SD, p: str, x: Tensor, s: Tensor, eps: float = 1e-5) -> Tensor:
    """ada_norm.py:185-201 -- x is [B, L, C]; style s is [B, S]."""
    h = F.linear(s, sd[p + "fc.weight"], sd[p + "fc.bias"])
    h = h.view(h.size(0), h.size(1), 1)
    gamma, beta = torch.chunk(h, chunks=2, dim=1)
    gamma, beta = gamma.transpose(1, 2), beta.transpose(1, 2)
    c = x.shape[-1]
    x = F.layer_norm(x, (c,), eps=eps)
    return (1 + gamma) * x + beta


def rope(x: Tensor, d: int, base: int = 10_000) -> Tensor:
    """text_encoder.py:89-168 -- rotate the first ``d`` features of every head; x is [b,h,t,c]."""
    x = torch.permute(x, (2, 0, 1, 3))
    seq_len = x.shape[0]
    theta = 1.0 / (base ** (torch.arange(0, d, 2, device=x.device).float() / d))
    seq_idx = torch.arange(seq_len, device=x.device).float()
    idx_theta = torch.einsum("n,d->nd", seq_idx, theta)
    idx_theta2 = torch.cat([idx_theta, idx_theta], dim=1)
    cos, sin = idx_theta2.cos()[:, None, None, :], idx_theta2.sin()[:, None, None, :]
    x_rope, x_pass = x[..., :d], x[..., d:]
    d2 = d // 2
    neg_half = torch.cat([-x_rope[:, :, :, d2:], x_rope[:, :, :, :d2]], dim=-1)
    x_rope = (x_rope * cos[:seq_len]) + (neg_half * sin[:seq_len])
    return torch.permute(torch.cat((x_rope, x_pass), dim=-1), (1, 2, 0, 3))


def mha(sd: SD, p: str, x: Tensor, c: Tensor, n_heads: int, mask: Optional[Tensor]) -> Tensor:
    """text_encoder.py:171-296 (use_sdpa=True branch).  ``mask`` truthy = attend, else -1e4."""
    q = F.conv1d(x, sd[p + "conv_q.weight"], sd[p + "conv_q.bias"])
    k = F.conv1d(c, sd[p + "conv_k.weight"], sd[p + "conv_k.bias"])
    v = F.conv1d(c, sd[p + "conv_v.weight"], sd[p + "conv_v.bias"])
    b, d, t_t = q.shape
    kc = d // n_heads

    def heads(z):
        z = torch.stack(z.chunk(chunks=n_heads, dim=1))
        return z.permute(1, 0, 3, 2)

    q, k, v = heads(q), heads(k), heads(v)
    rd = int(kc * 0.5)
    q, k = rope(q, rd), rope(k, rd)
    add = None
    if mask is not None:
        add = torch.zeros_like(mask, dtype=q.dtype)
        add.masked_fill_(~(mask.to(bool)), -1e4)
    o = F.scaled_dot_product_attention(q, k, v, attn_mask=add, dropout_p=0.0, is_causal=False)
    o = o.transpose(2, 3).contiguous().view(b, d, t_t)
    return F.conv1d(o, sd[p + "conv_o.weight"], sd[p + "conv_o.bias"])


def ffn(sd: SD, p: str, x: Tensor, x_mask: Tensor, k: int) -> Tensor:
    """text_encoder.py:305-329."""
    x = F.conv1d(x * x_mask, sd[p + "conv_1.weight"], sd[p + "conv_1.bias"], padding=k // 2)
    x = torch.relu(x)
    x = F.conv1d(x * x_mask, sd[p + "conv_2.weight"], sd[p + "conv_2.bias"], padding=k // 2)
    return x * x_mask


def text_encoder(sd: SD, p: str, tokens: Tensor, lengths: Tensor, n_layers: int = 8,
                 n_heads: int = 8, kernel: int = 3) -> Tuple[Tensor, Tensor, Tensor]:
    """text_encoder.py:396-462 (TextEncoder.forward) with ConvReluNorm 36-86, Encoder 332-393."""
    emb = sd[p + "emb.weight"]
    n_ch = emb.shape[1]
    x = F.embedding(tokens, emb) * math.sqrt(n_ch)
    x = torch.transpose(x, 1, -1)
    x_mask = torch.unsqueeze(seq_mask(lengths, x.size(2)), 1).to(x.dtype)
    # prenet: 3 x (conv k5, LN, ReLU) + zero-init 1x1 residual
    x_org = x
    i = 0
    while p + f"prenet.conv_layers.{i}.weight" in sd:
        w = sd[p + f"prenet.conv_layers.{i}.weight"]
        x = F.conv1d(x * x_mask, w, sd[p + f"prenet.conv_layers.{i}.bias"], padding=w.shape[2] // 2)
        x = channel_layer_norm(sd, p + f"prenet.norm_layers.{i}.", x)
        x = torch.relu(x)
        i += 1
    x = x_org + F.conv1d(x, sd[p + "prenet.proj.weight"], sd[p + "prenet.proj.bias"])
    x = x * x_mask
    # encoder
    attn_mask = x_mask.unsqueeze(2) * x_mask.unsqueeze(-1)
    for l in range(n_layers):
        x = x * x_mask
        y = mha(sd, p + f"encoder.attn_layers.{l}.", x, x, n_heads, attn_mask)
        x = channel_layer_norm(sd, p + f"encoder.norm_layers_1.{l}.", x + y)
        y = ffn(sd, p + f"encoder.ffn_layers.{l}.", x, x_mask, kernel)
        x = channel_layer_norm(sd, p + f"encoder.norm_layers_2.{l}.", x + y)
    x = x * x_mask
    mu = F.conv1d(x, sd[p + "proj_m.weight"], sd[p + "proj_m.bias"]) * x_mask
    return mu, x, x_mask


def grn(gamma: Tensor, beta: Tensor, x: Tensor) -> Tensor:
    """conv_next.py:4-15 / generator.py:488-499 -- x is [B, L, C]; L2 norm over time."""
    gx = torch.norm(x, p=2, dim=1, keepdim=True)
    nx = gx / (gx.mean(dim=-1, keepdim=True) + 1e-6)
    return gamma * (x * nx) + beta + x


def text_style_encoder(sd: SD, p: str, x: Tensor, lengths: Tensor) -> Tensor:
    """text_style_encoder.py:6-26 + BasicConvNeXtBlock conv_next.py:18-51."""
    x = F.conv1d(x, sd[p + "conv_in.weight"], sd[p + "conv_in.bias"], padding=3)
    i = 0
    while p + f"blocks.{i}.dwconv.weight" in sd:
        q = p + f"blocks.{i}."
        w = sd[q + "dwconv.weight"]
        r = x
        y = F.conv1d(x, w, sd[q + "dwconv.bias"], padding=w.shape[2] // 2, groups=w.shape[0])
        y = y.transpose(1, 2)
        y = F.layer_norm(y, (y.shape[-1],), sd[q + "norm.weight"], sd[q + "norm.bias"], eps=1e-6)
        y = F.linear(y, sd[q + "pwconv1.weight"], sd[q + "pwconv1.bias"])
        y = F.gelu(y)
        y = grn(sd[q + "grn.gamma"], sd[q + "grn.beta"], y)
        y = F.linear(y, sd[q + "pwconv2.weight"], sd[q + "pwconv2.bias"])
        x = r + y.transpose(1, 2)
        i += 1
    mask = seq_mask(lengths, x.shape[2]).unsqueeze(1).to(x.dtype)
    return (x * mask).sum(dim=2) / lengths.unsqueeze(1)


def prosody_encoder(sd: SD, p: str, x: Tensor, style: Tensor, lengths: Tensor,
                    n_layers: int, n_heads: int = 2) -> Tensor:
    """prosody_encoder.py:63-81.  Returns [B, T, d_model + sty]."""
    x_mask = torch.unsqueeze(seq_mask(lengths, x.size(2)), 1).to(x.dtype)
    attn_mask = x_mask.unsqueeze(2) * x_mask.unsqueeze(-1)
    s = style
    sty = style.unsqueeze(2).expand(x.shape[0], -1, x.shape[2])
    x = torch.cat([x, sty], dim=1)
    for i in range(n_layers):
        x = x * x_mask
        y = mha(sd, p + f"attn_layers.{i}.", x, x, n_heads, attn_mask)
        x = ada_layer_norm(sd, p + f"norm_layers_1.{i}.", (x + y).transpose(1, 2), s).transpose(1, 2)
        y = ffn(sd, p + f"ffn_layers.{i}.", x, x_mask, 1)
        x = ada_layer_norm(sd, p + f"norm_layers_2.{i}.", (x + y).transpose(1, 2), s).transpose(1, 2)
        x = F.conv1d(x, sd[p + f"proj_layers.{i}.weight"], sd[p + f"proj_layers.{i}.bias"])
        x = torch.cat([x, sty], dim=1)
    x = x * x_mask
    return x.transpose(-1, -2)


def duration_predictor(sd: SD, tokens: Tensor, lengths: Tensor, n_layer: int = 4) -> Tensor:
    """duration_predictor.py:30-36 -> logits [B, T, 16]."""
    enc, _, _ = text_encoder(sd, "text_encoder.", tokens, lengths)
    style = text_style_encoder(sd, "style_encoder.", enc, lengths)
    pros = prosody_encoder(sd, "prosody_encoder.", enc, style, lengths, n_layer)
    return F.linear(pros, sd["duration_proj.linear_layer.weight"], sd["duration_proj.linear_layer.bias"])


# ----------------------------------------------------------------------------- length regulation
def prediction_to_duration(pred: Tensor) -> Tensor:
    """utils.py:468-474 -- [T,16] logits -> integer-valued fp32 durations [T]."""
    table = torch.tensor(CLASS_TO_DUR, dtype=pred.dtype, device=pred.device)
    soft = (torch.softmax(pred, dim=-1) * table).sum(dim=-1).round().clamp(min=1)
    arg = table[torch.argmax(pred, dim=-1).long().clamp(min=0, max=16)]
    return (arg * (arg < 7)) + (soft * (arg >= 7))


def duration_to_indices(duration: Tensor) -> Tensor:
    """utils.py:481-484 -- frame -> token index (the non-zero row of every alignment column)."""
    return torch.repeat_interleave(torch.arange(duration.shape[0], device=duration.device), duration.to(torch.int))


def duration_to_alignment(duration: Tensor) -> Tensor:
    """utils.py:476-489 -- dense one-hot [T, sum(dur)]."""
    idx = duration_to_indices(duration)
    out = torch.zeros((duration.shape[0], idx.shape[0]), device=duration.device, dtype=duration.dtype)
    out[idx, torch.arange(idx.shape[0], device=duration.device)] = 1
    return out


# ----------------------------------------------------------------------------- AdaIN conv blocks
def adain(sd: SD, p: str, x: Tensor, s: Tensor) -> Tensor:
    """ada_norm.py:129-139 -- (1+gamma) * InstanceNorm1d(x) + beta; eps 1e-5, biased variance."""
    h = F.linear(s, sd[p + "fc.weight"], sd[p + "fc.bias"])
    h = h.view(h.size(0), h.size(1), 1)
    gamma, beta = torch.chunk(h, chunks=2, dim=1)
    return (1 + gamma) * F.instance_norm(x, eps=1e-5) + beta


def adaptive_decoder_block(sd: SD, p: str, x: Tensor, s: Tensor) -> Tensor:
    """ada_norm.py:142-182."""
    y = adain(sd, p + "norm1.", x, s)
    y = F.leaky_relu(y, 0.2)
    y = F.conv1d(y, wn_weight(sd, p + "conv1."), sd[p + "conv1.bias"], padding=1)
    y = adain(sd, p + "norm2.", y, s)
    y = F.leaky_relu(y, 0.2)
    y = F.conv1d(y, wn_weight(sd, p + "conv2."), sd[p + "conv2.bias"], padding=1)
    if p + "conv1x1.parametrizations.weight.original0" in sd:
        x = F.conv1d(x, wn_weight(sd, p + "conv1x1."))
    return (y + x) / math.sqrt(2)


def decoder(sd: SD, p: str, asr: Tensor, f0_curve: Tensor, n_curve: Tensor, s: Tensor) -> Tensor:
    """decoder.py:47-60 -> x [B, 512, L]."""
    f0 = F.conv1d(f0_curve.unsqueeze(1), wn_weight(sd, p + "F0_conv."), sd[p + "F0_conv.bias"], padding=1)
    n = F.conv1d(n_curve.unsqueeze(1), wn_weight(sd, p + "N_conv."), sd[p + "N_conv.bias"], padding=1)
    x = torch.cat([asr, f0, n], dim=1)
    x = adaptive_decoder_block(sd, p + "encode.", x, s)
    asr_res = F.conv1d(asr, wn_weight(sd, p + "asr_res.0."), sd[p + "asr_res.0.bias"])
    i = 0
    while p + f"decode.{i}.conv1.bias" in sd:
        x = torch.cat([x, asr_res, f0, n], dim=1)
        x = adaptive_decoder_block(sd, p + f"decode.{i}.", x, s)
        i += 1
    return x


# ----------------------------------------------------------------------------- pitch / energy
def band_mask(alignment: Tensor, text_mask: Tensor, window: int) -> Tensor:
    """pitch_energy_predictor.py:194-212.  True = "not allowed" per its docstring -- but the MHA
    (text_encoder.py:255-262) treats truthy as *allowed*: the inverted polarity is reproduced."""
    b, t, f = alignment.shape
    tau = alignment.argmax(dim=1)
    t_idx = torch.arange(t, device=alignment.device).view(1, 1, t)
    tau_exp = tau.unsqueeze(-1)
    band = (t_idx >= (tau_exp - window)) & (t_idx <= (tau_exp + window))
    full = (~band) | text_mask.unsqueeze(1).expand(b, f, t)
    return full.unsqueeze(1)


def pitch_energy_predictor(sd: SD, enc: Tensor, lengths: Tensor, alignment: Tensor,
                           style: Tensor, taps: Optional[Dict[str, Tensor]] = None):
    """pitch_energy_predictor.py:83-121 -> (F0 [B,F], N [B,F])."""
    t = enc.shape[2]
    mask = torch.gt(torch.arange(t, device=enc.device).unsqueeze(0).expand(lengths.shape[0], -1).type_as(lengths) + 1,
                    lengths.unsqueeze(1))  # utils.py:59-67
    pros = prosody_encoder(sd, "prosody_encoder.", enc, style, lengths, 3)
    base = torch.matmul(pros.transpose(1, 2), alignment)
    query = ada_layer_norm(sd, "query_norm.", base.transpose(1, 2), style).transpose(1, 2)
    key = ada_layer_norm(sd, "key_norm.", pros, style).transpose(1, 2)
    am = band_mask(alignment, mask, 5)
    att = mha(sd, "cross_attention.", query, key, 8, am)
    w0 = wn_weight(sd, "cross_post.0.")
    att = F.conv1d(att, w0, sd["cross_post.0.bias"], padding=2, groups=w0.shape[0])
    att = F.silu(att)
    att = F.conv1d(att, wn_weight(sd, "cross_post.2."), sd["cross_post.2.bias"])
    x = (base + att) / math.sqrt(2.0)
    if taps is not None:
        taps["pe_prosody"] = pros
        taps["pe_cross"] = x
    f0 = x
    for i in range(3):
        f0 = adaptive_decoder_block(sd, f"F0.{i}.", f0, style)
    f0 = F.conv1d(f0, sd["F0_proj.weight"], sd["F0_proj.bias"])
    n = x
    for i in range(3):
        n = adaptive_decoder_block(sd, f"N.{i}.", n, style)
    n = F.conv1d(n, sd["N_proj.weight"], sd["N_proj.bias"])
    return f0.squeeze(1), n.squeeze(1)


# ----------------------------------------------------------------------------- flow
def prior_encoder(sd: SD, p: str, x: Tensor, noise: Tensor):
    """flow.py:311-315 -- z = mean + noise * exp(logstd) with ``noise`` = the randn_like draw."""
    mean = F.linear(x.mT, sd[p + "proj_mean.weight"], sd[p + "proj_mean.bias"]).mT
    logstd = F.linear(x.mT, sd[p + "proj_logstd.weight"], sd[p + "proj_logstd.bias"]).mT
    return mean + noise * torch.exp(logstd), mean, logstd


def wavenet(sd: SD, p: str, x: Tensor, g: Tensor, n_layers: int = 4) -> Tensor:
    """flow.py:63-88 (WN.forward, x_mask == 1)."""
    hidden = x.shape[1]
    out = torch.zeros_like(x)
    g = F.linear(g.mT, wn_weight(sd, p + "cond_layer."), sd[p + "cond_layer.bias"]).mT
    for i in range(n_layers):
        w = wn_weight(sd, p + f"in_layers.{i}.")
        x_in = F.conv1d(x, w, sd[p + f"in_layers.{i}.bias"], padding=(w.shape[2] - 1) // 2)
        a = x_in + g[:, i * 2 * hidden:(i + 1) * 2 * hidden, :]
        acts = torch.tanh(a[:, :hidden, :]) * torch.sigmoid(a[:, hidden:, :])
        rs = F.linear(acts.mT, wn_weight(sd, p + f"res_skip_layers.{i}."), sd[p + f"res_skip_layers.{i}.bias"]).mT
        if i < n_layers - 1:
            x = x + rs[:, :hidden, :]
            out = out + rs[:, hidden:, :]
        else:
            out = out + rs
    return out


def flow_reverse(sd: SD, p: str, z: Tensor, cond: Tensor, n_flows: int = 8) -> Tensor:
    """flow.py:132-151 (reverse=True), 196-218, 221-226.  Only ``z`` is tracked: at inference the
    transformed mean/logstd are never consumed (speech_predictor.py:111)."""
    half = z.shape[1] // 2
    z0, z1 = torch.split(z, [half, half], 1)
    for i in reversed(range(n_flows)):
        z0, z1 = z1, z0  # Flip (flows[2i+1]) runs first in the reversed order
        q = p + f"flows.{2 * i}."
        h = F.linear(z0.mT, sd[q + "pre.weight"], sd[q + "pre.bias"]).mT
        h = wavenet(sd, q + "enc.", h, cond)
        m = F.linear(h.mT, sd[q + "proj_mean.weight"], sd[q + "proj_mean.bias"]).mT
        logs = F.linear(h.mT, sd[q + "proj_logstd.weight"], sd[q + "proj_logstd.bias"]).mT
        z1 = (z1 - m) * torch.exp(-logs)
    return torch.cat([z0, z1], 1)


# ----------------------------------------------------------------------------- source + STFT
def generate_pcph(f0: Tensor, voiced: Tensor, noise: Tensor, init_phase: Tensor,
                  hop: int = HOP, sr: int = SAMPLE_RATE, power_factor: float = 0.1) -> Tensor:
    """generator.py:247-315.  f0/voiced [B,1,L]; ``noise`` = the (already 0.01-scaled NOT applied)
    raw randn draw [B,1,L*hop]; ``init_phase`` = the torch.rand((1,1)) draw."""
    noise = 0.01 * noise
    if torch.all(voiced.round() <= 0.1):
        return noise
    vuv = voiced.round().bool()
    min_f0 = torch.min(f0[f0 > 20]).item()  # raises on empty, like the reference
    max_h = min(16, int((sr / 2) / min_f0))
    n_h = torch.ones_like(f0, dtype=torch.float)
    n_h[vuv] = sr / 2.0 / f0[vuv]
    idx = torch.arange(1, max_h + 1, device=f0.device).reshape(1, -1, 1)
    hmask = torch.repeat_interleave((f0 * idx) <= (sr / 2.0), hop, dim=2)
    amp = torch.repeat_interleave(vuv * power_factor * torch.sqrt(2.0 / n_h), hop, dim=2)
    f0r = torch.repeat_interleave(f0, hop, dim=2)
    rad = f0r.to(torch.float64) / sr
    rad[..., 0] += init_phase
    rad = torch.cumsum(rad, dim=2)
    harm = torch.sin(2.0 * torch.pi * rad * idx).to(torch.float32)
    harm = amp * torch.sum(hmask * harm, dim=1, keepdim=True)
    return harm + noise


def stft_mag_phase(x: Tensor) -> Tuple[Tensor, Tensor]:
    """generator.py:32-44 + 407-410: |X| and atan2 phase of the source, last frame dropped."""
    window = torch.hann_window(WIN, device=x.device).to(x.dtype)
    spec = torch.stft(x, N_FFT, HOP, WIN, window=window, return_complex=True)
    mag = torch.abs(spec) + 1e-9
    re = torch.real(spec) / mag
    im = torch.imag(spec) / mag
    return torch.abs(spec)[:, :, :-1], torch.atan2(im, re)[:, :, :-1]


def istft_tanh(logamp: Tensor, phase: Tensor) -> Tensor:
    """generator.py:425-433 + 46-56 -- logamp/phase already replicate-padded to L+1 frames."""
    window = torch.hann_window(WIN, device=logamp.device).to(logamp.dtype)
    spec = torch.exp(logamp)
    audio = torch.istft(spec * (torch.cos(phase) + torch.sin(phase) * 1j), N_FFT, HOP, WIN, window=window)
    return torch.tanh(audio.unsqueeze(-2))


def convnext_block(sd: SD, p: str, x: Tensor, style: Tensor) -> Tensor:
    """generator.py:441-485."""
    w = sd[p + "dwconv.weight"]
    y = F.conv1d(x, w, sd[p + "dwconv.bias"], padding=(w.shape[2] - 1) // 2, groups=w.shape[0])
    y = y.transpose(1, 2)
    y = ada_layer_norm(sd, p + "norm.", y, style, eps=1e-6)
    y = F.linear(y, sd[p + "pwconv1.weight"], sd[p + "pwconv1.bias"])
    y = F.silu(y)
    y = grn(sd[p + "grn.gamma"], sd[p + "grn.beta"], y)
    y = F.linear(y, sd[p + "pwconv2.weight"], sd[p + "pwconv2.bias"])
    return x + y.transpose(1, 2)


def exact_frame0_phase(har_phase: Tensor) -> Tensor:
    """Frame 0 of torch.stft(center=True, reflect) is centred on sample 0: the windowed frame is
    even about its centre, so X_k is exactly real and the reference's atan2 there is FFT rounding
    noise (0 or +-pi with a random sign).  This returns the mathematically exact phase for that
    frame (atan2(+0, Re): 0 for Re > 0, +pi for Re < 0) and leaves every other frame untouched.
    It is a *documented variant* of the oracle used where an implementation-independent target
    is needed (DESIGN.md "ill-conditioned phase"); the default oracle stays the raw reference."""
    out = har_phase.clone()
    f0 = out[:, :, 0]
    out[:, :, 0] = torch.where(f0.abs() > (math.pi / 2), torch.full_like(f0, math.pi), torch.zeros_like(f0))
    return out


def generator(sd: SD, p: str, mel: Tensor, style: Tensor, pitch: Tensor,
              noise: Tensor, init_phase: Tensor, taps: Optional[Dict[str, Tensor]] = None,
              exact_frame0: bool = False, stft_override: Optional[Tuple[Tensor, Tensor]] = None):
    """generator.py:402-438 -> (audio [B,1,75L], logamp [B,1025,L+1], phase [B,1025,L+1]).
    The ``energy`` argument of the reference is unused."""
    pit = pitch.unsqueeze(1)
    # the source is always evaluated in the reference's own arithmetic (fp32 with an fp64 phase); a float64 run of
    # this port (tests: "the reference against itself in higher precision") differs from it downstream only
    prior = generate_pcph(pit.float(), (pit > 10.0).float(), noise.float(), init_phase.float()).squeeze(1).to(mel.dtype)
    har_spec, har_phase = stft_mag_phase(prior)
    if stft_override is not None:
        har_spec, har_phase = stft_override
    if exact_frame0:
        har_phase = exact_frame0_phase(har_phase)
    la_prior = F.conv1d(har_spec, sd[p + "amp_prior_conv.weight"], sd[p + "amp_prior_conv.bias"], padding=3)
    ph_prior = F.conv1d(har_phase, sd[p + "phase_prior_conv.weight"], sd[p + "phase_prior_conv.bias"], padding=3)
    x = F.conv1d(torch.cat([mel, la_prior, ph_prior], dim=1), sd[p + "projector.weight"], sd[p + "projector.bias"])
    if taps is not None:
        taps.update(source=prior, har_spec=har_spec, har_phase=har_phase, gen_proj=x)
    for i in range(4):
        x = convnext_block(sd, p + f"convnext.{i}.", x, style)
    if taps is not None:
        taps["gen_convnext"] = x
    xt = x.transpose(1, 2)
    la = ada_layer_norm(sd, p + "amp_final_layer_norm.", xt, style).transpose(1, 2)
    la = F.conv1d(torch.cat([la, la_prior], dim=1), sd[p + "amp_output_conv.weight"],
                  sd[p + "amp_output_conv.bias"], padding=3)
    ph = ada_layer_norm(sd, p + "phase_final_layer_norm.", xt, style).transpose(1, 2)
    ph = F.conv1d(torch.cat([ph, ph_prior], dim=1), sd[p + "phase_output_conv.weight"],
                  sd[p + "phase_output_conv.bias"], padding=3)
    la = F.pad(la, pad=(0, 1), mode="replicate")
    ph = F.pad(ph, pad=(0, 1), mode="replicate")
    return istft_tanh(la, ph), la, ph


# ----------------------------------------------------------------------------- top level
def upsample4(x: Tensor) -> Tensor:
    """speech_predictor.py:76,89-90 -- nn.Upsample(scale_factor=4, mode="linear") on [B, F]."""
    return F.interpolate(x.unsqueeze(1), scale_factor=4, mode="linear").squeeze(1)


def speech_predictor(sd: SD, tokens: Tensor, lengths: Tensor, alignment: Tensor, pitch: Tensor,
                     energy: Tensor, noise: Dict[str, Tensor],
                     taps: Optional[Dict[str, Tensor]] = None, exact_frame0: bool = False):
    """speech_predictor.py:85-129, inference branch (audio_gt is None).

    ``noise`` holds the three reference draws in call order: "prior" [B,128,4F] (flow.py:314),
    "source" [B,1,300F] (generator.py:275-277), "phase" [1,1] (generator.py:306).
    Returns (audio [B,1,300F], logamp [B,1025,4F+1], phase [B,1025,4F+1]).
    """
    enc, _, _ = text_encoder(sd, "text_encoder.", tokens, lengths)
    style = text_style_encoder(sd, "style_encoder.", enc, lengths)
    al4 = alignment.repeat_interleave(4, dim=2)
    # fp32 upsample whatever the run's dtype: the source integrates pitch4, one ulp of it moves the waveform by 8e-3
    pitch4, energy4 = upsample4(pitch.float()).to(pitch.dtype), upsample4(energy.float()).to(energy.dtype)
    x = decoder(sd, "decoder.", enc @ al4, pitch4, energy4, style)
    cond = style.unsqueeze(-1)
    z, _, _ = prior_encoder(sd, "prior_encoder.", x, noise["prior"])
    z2 = flow_reverse(sd, "flow.", z, cond)
    mel = F.linear(z2.mT, sd["post_flow.weight"], sd["post_flow.bias"]).mT
    if taps is not None:
        taps.update(sp_text=enc, sp_style=style, dec_out=x, z_prior=z, z_flow=z2, mel=mel,
                    pitch4=pitch4, energy4=energy4)
    return generator(sd, "generator.", mel, style, pitch4, noise["source"], noise["phase"], taps,
                     exact_frame0=exact_frame0)


def synthesize(nets: Mapping[str, SD], tokens: Tensor, lengths: Tensor, noise: Dict[str, Tensor],
               duration: Optional[Tensor] = None, taps: Optional[Dict[str, Tensor]] = None,
               exact_frame0: bool = False):
    """The B = 1 inference recipe of stage_type.py:483-523 / export_model.py:35-45.

    ``nets`` maps the five module names to their state dicts.  If ``duration`` [T] is given it
    replaces the predicted durations (synthetic-duration timing runs); the
    duration network still runs.  Returns dict(audio, logamp, phase, duration, pitch, energy).
    """
    assert tokens.shape[0] == 1, "the reference has no batched mixed-length path (SURVEY 0.6)"
    logits = duration_predictor(nets["duration_predictor"], tokens, lengths)
    dur_pred = prediction_to_duration(logits[0])[: int(lengths[0])]
    dur = dur_pred if duration is None else duration
    alignment = duration_to_alignment(dur).unsqueeze(0)
    pe_enc, _, _ = text_encoder(nets["pe_text_encoder"], "", tokens, lengths)
    pe_style = text_style_encoder(nets["pe_text_style_encoder"], "", pe_enc, lengths)
    pitch, energy = pitch_energy_predictor(nets["pitch_energy_predictor"], pe_enc, lengths, alignment, pe_style, taps)
    audio, la, ph = speech_predictor(nets["speech_predictor"], tokens, lengths, alignment, pitch, energy, noise, taps,
                                     exact_frame0=exact_frame0)
    if taps is not None:
        taps.update(dur_logits=logits, pe_text=pe_enc, pe_style=pe_style)
    return dict(audio=audio, logamp=la, phase=ph, duration=dur_pred, pitch=pitch, energy=energy)
