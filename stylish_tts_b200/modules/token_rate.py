"""Token-rate (<= 510 phonemes) modules of the synthesis path, host-side mirror.

Same constructor arguments, ``forward`` signatures and ``state_dict()`` key layout as the
reference modules they replace:
  TextEncoder        train/models/text_encoder.py:396-462
  TextStyleEncoder   train/models/text_style_encoder.py:6-26
  ProsodyEncoder     train/models/prosody_encoder.py:10-81
  DurationPredictor  train/models/duration_predictor.py:8-36
They carry ~3 % of the path's FLOPs and, in this round, execute on the device
through torch's tensor ops (SURVEY 7.2 step 6 / 8f rank 1 keep the fully native token-rate
stack as the next row); the frame-rate mass of the path is in modules/frame_rate.py.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

# The reference path is fp32; PyTorch's default lets cuDNN convolutions run in TF32, which alone
# breaks the 1e-4 parity bar.  The token-rate modules therefore pin fp32.
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def length_mask(lengths: torch.Tensor, max_len: int) -> torch.Tensor:
    """[B, max_len] True inside the utterance (utils.py:52-56)."""
    pos = torch.arange(max_len, dtype=lengths.dtype, device=lengths.device)
    return pos[None, :] < lengths[:, None]


class ChannelNorm(nn.Module):
    """LayerNorm over the channel axis of [B, C, T] (text_encoder.py:15-33), eps 1e-4."""

    def __init__(self, channels: int, eps: float = 1e-4):
        super().__init__()
        self.eps = eps
        self.gamma = nn.Parameter(torch.ones(channels))
        self.beta = nn.Parameter(torch.zeros(channels))

    def forward(self, x):
        mu = x.mean(dim=1, keepdim=True)
        var = ((x - mu) ** 2).mean(dim=1, keepdim=True)
        y = (x - mu) * torch.rsqrt(var + self.eps)
        return y * self.gamma[None, :, None] + self.beta[None, :, None]


class StyleLayerNorm(nn.Module):
    """AdaptiveLayerNorm (ada_norm.py:185-201): LayerNorm over the last axis, style-conditioned
    affine (1 + gamma) * x + beta from one Linear(style)."""

    def __init__(self, style_dim: int, channels: int, eps: float = 1e-5):
        super().__init__()
        self.channels, self.eps = channels, eps
        self.fc = nn.Linear(style_dim, 2 * channels)

    def forward(self, x, s):
        gb = self.fc(s)
        gamma, beta = gb[:, None, : self.channels], gb[:, None, self.channels:]
        return (1 + gamma) * F.layer_norm(x, (self.channels,), eps=self.eps) + beta


def _partial_rope(x: torch.Tensor, d: int) -> torch.Tensor:
    """Rotary embedding on the first ``d`` features of each head; x is [B, H, T, c]
    (text_encoder.py:89-168: theta_i = 10000^(-2i/d), half-split rotation)."""
    t = x.shape[2]
    inv = 1.0 / (10_000 ** (torch.arange(0, d, 2, device=x.device).float() / d))
    ang = torch.arange(t, device=x.device).float()[:, None] * inv[None, :]
    ang = torch.cat([ang, ang], dim=1)
    cos, sin = ang.cos()[None, None], ang.sin()[None, None]
    xr, xp = x[..., :d], x[..., d:]
    h = d // 2
    rot = torch.cat([-xr[..., h:], xr[..., :h]], dim=-1)
    return torch.cat([xr * cos + rot * sin, xp], dim=-1)


class MultiHeadAttention(nn.Module):
    """text_encoder.py:171-296 (SDPA branch): 1x1-conv projections, RoPE on half of every head,
    additive mask of -1e4 wherever ``attn_mask`` is falsy."""

    def __init__(self, channels: int, out_channels: int, n_heads: int, p_dropout: float = 0.0, **_):
        super().__init__()
        assert channels % n_heads == 0
        self.n_heads = n_heads
        self.k_channels = channels // n_heads
        self.conv_q = nn.Conv1d(channels, channels, 1)
        self.conv_k = nn.Conv1d(channels, channels, 1)
        self.conv_v = nn.Conv1d(channels, channels, 1)
        self.conv_o = nn.Conv1d(channels, out_channels, 1)

    def forward(self, x, c, attn_mask=None, attn_bias=None):
        b, d, tq = x.shape
        split = lambda z: z.view(b, self.n_heads, self.k_channels, z.shape[2]).transpose(2, 3)
        q, k, v = split(self.conv_q(x)), split(self.conv_k(c)), split(self.conv_v(c))
        rd = int(self.k_channels * 0.5)
        q, k = _partial_rope(q, rd), _partial_rope(k, rd)
        bias = attn_bias  # ready-made additive bias (ragged batches), else built from the mask
        if bias is None and attn_mask is not None:
            bias = torch.zeros(attn_mask.shape, dtype=q.dtype, device=q.device)
            bias.masked_fill_(~attn_mask.to(torch.bool), -1e4)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=bias)
        return self.conv_o(o.transpose(2, 3).reshape(b, d, tq))


class FFN(nn.Module):
    """text_encoder.py:305-329."""

    def __init__(self, in_channels, out_channels, filter_channels, kernel_size, p_dropout=0.0):
        super().__init__()
        self.conv_1 = nn.Conv1d(in_channels, filter_channels, kernel_size, padding=kernel_size // 2)
        self.conv_2 = nn.Conv1d(filter_channels, out_channels, kernel_size, padding=kernel_size // 2)

    def forward(self, x, x_mask):
        return self.conv_2(torch.relu(self.conv_1(x * x_mask)) * x_mask) * x_mask


class _Prenet(nn.Module):
    """ConvReluNorm (text_encoder.py:36-86): n x (conv k, ChannelNorm, ReLU) + 1x1 residual."""

    def __init__(self, channels: int, kernel_size: int, n_layers: int):
        super().__init__()
        self.conv_layers = nn.ModuleList(
            [nn.Conv1d(channels, channels, kernel_size, padding=kernel_size // 2) for _ in range(n_layers)])
        self.norm_layers = nn.ModuleList([ChannelNorm(channels) for _ in range(n_layers)])
        self.proj = nn.Conv1d(channels, channels, 1)

    def forward(self, x, x_mask):
        y = x
        for conv, norm in zip(self.conv_layers, self.norm_layers):
            y = torch.relu(norm(conv(y * x_mask)))
        return (x + self.proj(y)) * x_mask


class _Transformer(nn.Module):
    """Encoder (text_encoder.py:332-393): post-norm blocks of MHA + conv FFN."""

    def __init__(self, hidden, filter_channels, n_heads, n_layers, kernel_size):
        super().__init__()
        self.attn_layers = nn.ModuleList([MultiHeadAttention(hidden, hidden, n_heads) for _ in range(n_layers)])
        self.norm_layers_1 = nn.ModuleList([ChannelNorm(hidden) for _ in range(n_layers)])
        self.ffn_layers = nn.ModuleList(
            [FFN(hidden, hidden, filter_channels, kernel_size) for _ in range(n_layers)])
        self.norm_layers_2 = nn.ModuleList([ChannelNorm(hidden) for _ in range(n_layers)])

    def forward(self, x, x_mask):
        pair = x_mask.unsqueeze(2) * x_mask.unsqueeze(-1)
        for attn, n1, ff, n2 in zip(self.attn_layers, self.norm_layers_1, self.ffn_layers, self.norm_layers_2):
            x = x * x_mask
            x = n1(x + attn(x, x, pair))
            x = n2(x + ff(x, x_mask))
        return x * x_mask


class TokenPack:
    """Host-known utterance lengths of a padded [B, Tmax] batch plus the device index vectors that
    move between the padded layout and the concatenated [sum T] layout the C library takes.  Built
    once per batch shape (H2D copies happen here, never inside a forward: the forwards stay free of
    host synchronisation and are CUDA-graph capturable)."""

    def __init__(self, lengths, tmax: int, device):
        self.lengths = [int(v) for v in lengths]
        self.tmax = int(tmax)
        idx = [b * self.tmax + t for b, n in enumerate(self.lengths) for t in range(n)]
        self.flat_idx = torch.tensor(idx, dtype=torch.int64).to(device)
        self.n = len(idx)

    @staticmethod
    def of(host_lengths, x_lengths, tmax, device) -> "TokenPack":
        if isinstance(host_lengths, TokenPack):
            return host_lengths
        lengths = host_lengths if host_lengths is not None else x_lengths.tolist()
        return TokenPack(lengths, tmax, device)

    def gather(self, padded: torch.Tensor) -> torch.Tensor:
        """[B, Tmax, ...] -> [sum T, ...]"""
        return padded.reshape((-1,) + tuple(padded.shape[2:])).index_select(0, self.flat_idx)

    def scatter(self, rows: torch.Tensor, b: int) -> torch.Tensor:
        """[sum T, C] -> zero-padded [B, Tmax, C]"""
        out = torch.zeros((b * self.tmax, rows.shape[1]), dtype=rows.dtype, device=rows.device)
        out.index_copy_(0, self.flat_idx, rows)
        return out.view(b, self.tmax, rows.shape[1])


class TextEncoder(nn.Module):
    """text_encoder.py:396-462.  The arithmetic runs in libstts_b200 (stts_text_encoder: embedding,
    prenet, 8 x [RoPE attention + LN, conv FFN + LN], proj_m on the packed token layout); this module
    holds the parameters in the reference layout.  ``forward_torch`` is the op-by-op torch statement
    of the same computation, kept for the CPU host tests of the state_dict layout only."""

    def __init__(self, *, inter_dim, config):
        super().__init__()
        self.n_channels = config.hidden_dim
        self.inter_dim = inter_dim
        self.emb = nn.Embedding(config.tokens, self.n_channels)
        self.prenet = _Prenet(self.n_channels, kernel_size=5, n_layers=3)
        self.encoder = _Transformer(self.n_channels, config.filter_channels, config.heads, config.layers,
                                    config.kernel_size)
        self.proj_m = nn.Conv1d(self.n_channels, inter_dim, 1)
        object.__setattr__(self, "_native", None)

    # ---- native engine plumbing
    def attach_native(self, engine, which: str):
        """Use a context that already holds this encoder's weights under module name ``which``."""
        object.__setattr__(self, "_native", (engine, which, False))

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        object.__setattr__(self, "_native", None)
        return super().load_state_dict(state_dict, strict=strict, assign=assign)

    def _engine(self):
        if self._native is None:
            from ..engine import Engine
            dev = self.emb.weight.device
            if dev.type != "cuda":
                raise RuntimeError("stylish_tts_b200 modules run on a B200 only (no CPU fallback)")
            eng = Engine(dev)
            eng.load_state("pe_text_encoder", self.state_dict())
            eng.finalize()
            object.__setattr__(self, "_native", (eng, "pe_text_encoder", True))
        return self._native[0], self._native[1]

    def encode_packed(self, tokens_cat: torch.Tensor, lengths):
        """tokens_cat int [sum T]; lengths: host ints.  -> token-major (mu [sum T, inter], h [sum T, 128])."""
        eng, which = self._engine()
        eng.set_token_layout(lengths)
        return eng.text_encoder(which, tokens_cat, self.inter_dim)

    def forward(self, x, x_lengths, spks=None, host_lengths=None):
        # host_lengths: the same lengths as Python ints or a TokenPack, when the caller has them
        # (saves the D2H sync of x_lengths.tolist())
        b, tmax = x.shape
        pack = TokenPack.of(host_lengths, x_lengths, tmax, x.device)
        mu_t, h_t = self.encode_packed(pack.gather(x), pack.lengths)
        valid = torch.arange(tmax, device=x.device)[None, :] < x_lengths[:, None]
        return (pack.scatter(mu_t, b).transpose(1, 2), pack.scatter(h_t, b).transpose(1, 2),
                valid.unsqueeze(1).to(torch.float32))

    def forward_torch(self, x, x_lengths, spks=None):
        h = (self.emb(x) * math.sqrt(self.n_channels)).transpose(1, 2)
        x_mask = length_mask(x_lengths, h.shape[2]).unsqueeze(1).to(h.dtype)
        h = self.encoder(self.prenet(h, x_mask), x_mask)
        return self.proj_m(h) * x_mask, h, x_mask


class _GRN(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.gamma = nn.Parameter(torch.zeros(1, 1, dim))
        self.beta = nn.Parameter(torch.zeros(1, 1, dim))

    def forward(self, x):  # [B, T, C]; L2 over time (conv_next.py:12-15)
        g = torch.norm(x, p=2, dim=1, keepdim=True)
        return self.gamma * (x * (g / (g.mean(dim=-1, keepdim=True) + 1e-6))) + self.beta + x


class _ConvNeXt1d(nn.Module):
    """BasicConvNeXtBlock (conv_next.py:18-51).  ``m`` ([B,1,T] 0/1) gives padded batches the
    per-utterance semantics of the B = 1 oracle (zero padding for the depthwise conv, GRN norm
    over the valid positions only); with m == 1 it is the reference computation."""

    def __init__(self, dim, intermediate_dim, kernel=7):
        super().__init__()
        self.dwconv = nn.Conv1d(dim, dim, kernel, padding=kernel // 2, groups=dim)
        self.norm = nn.LayerNorm(dim, eps=1e-6)
        self.pwconv1 = nn.Linear(dim, intermediate_dim)
        self.grn = _GRN(intermediate_dim)
        self.pwconv2 = nn.Linear(intermediate_dim, dim)

    def forward(self, x, m):
        y = self.norm(self.dwconv(x * m).transpose(1, 2))
        y = F.gelu(self.pwconv1(y)) * m.transpose(1, 2)
        y = self.pwconv2(self.grn(y))
        return (x + y.transpose(1, 2)) * m


class TextStyleEncoder(nn.Module):
    """text_style_encoder.py:6-26.  The arithmetic runs in libstts_b200 (stts_text_style_encoder: conv
    k7, ConvNeXt blocks with GRN over the utterance, mean over time); ``forward_torch`` is the torch
    statement of the same computation for the CPU host tests of the state_dict layout."""

    def __init__(self, inter_dim, style_dim, config):
        super().__init__()
        self.conv_in = nn.Conv1d(inter_dim, style_dim, kernel_size=7, padding=3)
        self.blocks = nn.ModuleList([_ConvNeXt1d(style_dim, style_dim * 4) for _ in range(config.layers)])
        object.__setattr__(self, "_native", None)

    def attach_native(self, engine, which: str):
        """Use a context that already holds this encoder's weights under module name ``which``."""
        object.__setattr__(self, "_native", (engine, which))

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        object.__setattr__(self, "_native", None)
        return super().load_state_dict(state_dict, strict=strict, assign=assign)

    def _engine(self):
        if self._native is None:
            from ..engine import Engine
            dev = self.conv_in.weight.device
            if dev.type != "cuda":
                raise RuntimeError("stylish_tts_b200 modules run on a B200 only (no CPU fallback)")
            eng = Engine(dev)
            eng.load_state("pe_text_style_encoder", self.state_dict())
            eng.finalize()
            object.__setattr__(self, "_native", (eng, "pe_text_style_encoder"))
        return self._native

    def forward(self, x, lengths, host_lengths=None):
        eng, which = self._engine()
        b, _, tmax = x.shape
        pack = TokenPack.of(host_lengths, lengths, tmax, x.device)
        eng.set_token_layout(pack.lengths)
        return eng.text_style_encoder(which, pack.gather(x.transpose(1, 2)))

    def forward_torch(self, x, lengths):
        m = length_mask(lengths, x.shape[2]).unsqueeze(1).to(x.dtype)
        x = self.conv_in(x * m) * m
        for blk in self.blocks:
            x = blk(x, m)
        return (x * m).sum(dim=2) / lengths.unsqueeze(1)


class ProsodyEncoder(nn.Module):
    def __init__(self, sty_dim, d_model, nlayers, dropout=0.1, n_heads=2, kernel_size=1, **_):
        super().__init__()
        hid = d_model + sty_dim
        self.n_layers = nlayers
        self.attn_layers = nn.ModuleList([MultiHeadAttention(hid, hid, n_heads) for _ in range(nlayers)])
        self.norm_layers_1 = nn.ModuleList([StyleLayerNorm(sty_dim, hid) for _ in range(nlayers)])
        self.ffn_layers = nn.ModuleList([FFN(hid, hid, hid * 2, kernel_size) for _ in range(nlayers)])
        self.norm_layers_2 = nn.ModuleList([StyleLayerNorm(sty_dim, hid) for _ in range(nlayers)])
        self.proj_layers = nn.ModuleList([nn.Conv1d(hid, d_model, 1) for _ in range(nlayers)])

    def attach_native(self, engine, which: str):
        """Use a context that holds this encoder's weights as ``which``/prosody_encoder.*"""
        object.__setattr__(self, "_native", (engine, which))

    def _engine(self):
        if getattr(self, "_native", None) is None:
            from ..engine import Engine
            dev = self.proj_layers[0].weight.device
            if dev.type != "cuda":
                raise RuntimeError("stylish_tts_b200 modules run on a B200 only (no CPU fallback)")
            eng = Engine(dev)
            eng.load_state("pitch_energy_predictor", self.state_dict(), prefix="prosody_encoder.")
            eng.finalize()
            object.__setattr__(self, "_native", (eng, "pitch_energy_predictor"))
        return self._native

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        object.__setattr__(self, "_native", None)
        return super().load_state_dict(state_dict, strict=strict, assign=assign)

    @property
    def hidden(self) -> int:
        return self.proj_layers[0].weight.shape[1]

    def forward(self, x, style, x_lengths, host_lengths=None):
        """prosody_encoder.py:63-81 in libstts_b200 (stts_prosody_encoder).  x [B, d_model, T] -> [B, T, hid]."""
        eng, which = self._engine()
        b, _, tmax = x.shape
        pack = TokenPack.of(host_lengths, x_lengths, tmax, x.device)
        eng.set_token_layout(pack.lengths)
        out_t, _ = eng.prosody_encoder(which, pack.gather(x.transpose(1, 2)), style, self.hidden)
        return pack.scatter(out_t, b)

    def forward_torch(self, x, style, x_lengths):
        x_mask = length_mask(x_lengths, x.shape[2]).unsqueeze(1).to(x.dtype)
        pair = x_mask.unsqueeze(2) * x_mask.unsqueeze(-1)
        sty = style.unsqueeze(2).expand(-1, -1, x.shape[2])
        x = torch.cat([x, sty], dim=1)
        for i in range(self.n_layers):
            x = x * x_mask
            y = self.attn_layers[i](x, x, pair)
            x = self.norm_layers_1[i]((x + y).transpose(1, 2), style).transpose(1, 2)
            y = self.ffn_layers[i](x, x_mask)
            x = self.norm_layers_2[i]((x + y).transpose(1, 2), style).transpose(1, 2)
            x = torch.cat([self.proj_layers[i](x), sty], dim=1)
        return (x * x_mask).transpose(1, 2)


class _LinearNorm(nn.Module):
    def __init__(self, in_dim, out_dim):
        super().__init__()
        self.linear_layer = nn.Linear(in_dim, out_dim)

    def forward(self, x):
        return self.linear_layer(x)


class DurationPredictor(nn.Module):
    def __init__(self, style_dim, inter_dim, text_config, style_config, duration_config):
        super().__init__()
        self.text_encoder = TextEncoder(inter_dim=inter_dim, config=text_config)
        self.style_encoder = TextStyleEncoder(inter_dim, style_dim, config=style_config)
        self.prosody_encoder = ProsodyEncoder(sty_dim=style_dim, d_model=inter_dim, nlayers=duration_config.n_layer,
                                              dropout=duration_config.dropout)
        self.duration_proj = _LinearNorm(inter_dim + style_dim, duration_config.duration_classes)

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        object.__setattr__(self, "_eng", None)
        return super().load_state_dict(state_dict, strict=strict, assign=assign)

    def attach_engine(self, engine):
        """Use a context that already holds this module's weights as "duration_predictor"."""
        object.__setattr__(self, "_eng", engine)

    def _engine(self):
        if getattr(self, "_eng", None) is None:
            from ..engine import Engine
            dev = self.duration_proj.linear_layer.weight.device
            if dev.type != "cuda":
                raise RuntimeError("stylish_tts_b200 modules run on a B200 only (no CPU fallback)")
            eng = Engine(dev)
            eng.load_state("duration_predictor", self.state_dict())
            eng.finalize()
            object.__setattr__(self, "_eng", eng)
        return self._eng

    def forward(self, texts, text_lengths, host_lengths=None):
        """duration_predictor.py:30-36: text encoder, style encoder, prosody encoder + duration_proj
        all run in libstts_b200 on the packed token layout."""
        eng = self._engine()
        self.text_encoder.attach_native(eng, "duration_predictor")
        b, tmax = texts.shape
        pack = TokenPack.of(host_lengths, text_lengths, tmax, texts.device)
        mu_t, _ = self.text_encoder.encode_packed(pack.gather(texts), pack.lengths)
        style = eng.text_style_encoder("duration_predictor", None)     # reads the mu rows left by the text encoder
        _, logits_t = eng.prosody_encoder("duration_predictor", mu_t, style, self.prosody_encoder.hidden,
                                          want_logits=True)
        return pack.scatter(logits_t, b)

    def forward_torch(self, texts, text_lengths):
        """Op-by-op torch statement of the same computation (CPU host tests of the mirror only)."""
        enc, _, _ = self.text_encoder.forward_torch(texts, text_lengths)
        style = self.style_encoder.forward_torch(enc, text_lengths)
        return self.duration_proj(self.prosody_encoder.forward_torch(enc, style, text_lengths))
