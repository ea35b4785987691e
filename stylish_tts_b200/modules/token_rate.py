"""Token-rate (<= 510 phonemes) modules of the synthesis path, host-side mirror.

Same constructor arguments, ``forward`` signatures and ``state_dict()`` key layout as the
reference modules they replace:
  TextEncoder        train/models/text_encoder.py:396-462
  TextStyleEncoder   train/models/text_style_encoder.py:6-26
  ProsodyEncoder     train/models/prosody_encoder.py:10-81
  DurationPredictor  train/models/duration_predictor.py:8-36
The modules only *hold* parameters in the reference layout: every forward is a call into
libstts_b200 (stts_text_encoder / stts_text_style_encoder / stts_prosody_encoder on the packed token
layout, csrc/token_pipeline.cu).  There is no torch statement of the arithmetic in this package --
the CPU restatement the tests check against lives outside the package (under the oracle directory).
"""
from __future__ import annotations

import torch
from torch import nn


class ChannelNorm(nn.Module):
    """Parameters of the channel LayerNorm of [B, C, T] (text_encoder.py:15-33), eps 1e-4."""

    def __init__(self, channels: int, eps: float = 1e-4):
        super().__init__()
        self.eps = eps
        self.gamma = nn.Parameter(torch.ones(channels))
        self.beta = nn.Parameter(torch.zeros(channels))


class StyleLayerNorm(nn.Module):
    """Parameters of AdaptiveLayerNorm (ada_norm.py:185-201): one Linear(style) -> (gamma, beta)."""

    def __init__(self, style_dim: int, channels: int, eps: float = 1e-5):
        super().__init__()
        self.channels, self.eps = channels, eps
        self.fc = nn.Linear(style_dim, 2 * channels)


class MultiHeadAttention(nn.Module):
    """Parameters of text_encoder.py:171-296: 1x1-conv q / k / v / o projections."""

    def __init__(self, channels: int, out_channels: int, n_heads: int, p_dropout: float = 0.0, **_):
        super().__init__()
        assert channels % n_heads == 0
        self.n_heads = n_heads
        self.k_channels = channels // n_heads
        self.conv_q = nn.Conv1d(channels, channels, 1)
        self.conv_k = nn.Conv1d(channels, channels, 1)
        self.conv_v = nn.Conv1d(channels, channels, 1)
        self.conv_o = nn.Conv1d(channels, out_channels, 1)


class FFN(nn.Module):
    """Parameters of text_encoder.py:305-329."""

    def __init__(self, in_channels, out_channels, filter_channels, kernel_size, p_dropout=0.0):
        super().__init__()
        self.conv_1 = nn.Conv1d(in_channels, filter_channels, kernel_size, padding=kernel_size // 2)
        self.conv_2 = nn.Conv1d(filter_channels, out_channels, kernel_size, padding=kernel_size // 2)


class _Prenet(nn.Module):
    """Parameters of ConvReluNorm (text_encoder.py:36-86): n x (conv k, ChannelNorm) + 1x1 residual."""

    def __init__(self, channels: int, kernel_size: int, n_layers: int):
        super().__init__()
        self.conv_layers = nn.ModuleList(
            [nn.Conv1d(channels, channels, kernel_size, padding=kernel_size // 2) for _ in range(n_layers)])
        self.norm_layers = nn.ModuleList([ChannelNorm(channels) for _ in range(n_layers)])
        self.proj = nn.Conv1d(channels, channels, 1)


class _Transformer(nn.Module):
    """Parameters of Encoder (text_encoder.py:332-393): post-norm blocks of MHA + conv FFN."""

    def __init__(self, hidden, filter_channels, n_heads, n_layers, kernel_size):
        super().__init__()
        self.attn_layers = nn.ModuleList([MultiHeadAttention(hidden, hidden, n_heads) for _ in range(n_layers)])
        self.norm_layers_1 = nn.ModuleList([ChannelNorm(hidden) for _ in range(n_layers)])
        self.ffn_layers = nn.ModuleList(
            [FFN(hidden, hidden, filter_channels, kernel_size) for _ in range(n_layers)])
        self.norm_layers_2 = nn.ModuleList([ChannelNorm(hidden) for _ in range(n_layers)])


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
    holds the parameters in the reference layout."""

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


class _GRN(nn.Module):
    """Parameters of GRN (conv_next.py:4-15)."""

    def __init__(self, dim):
        super().__init__()
        self.gamma = nn.Parameter(torch.zeros(1, 1, dim))
        self.beta = nn.Parameter(torch.zeros(1, 1, dim))


class _ConvNeXt1d(nn.Module):
    """Parameters of BasicConvNeXtBlock (conv_next.py:18-51)."""

    def __init__(self, dim, intermediate_dim, kernel=7):
        super().__init__()
        self.dwconv = nn.Conv1d(dim, dim, kernel, padding=kernel // 2, groups=dim)
        self.norm = nn.LayerNorm(dim, eps=1e-6)
        self.pwconv1 = nn.Linear(dim, intermediate_dim)
        self.grn = _GRN(intermediate_dim)
        self.pwconv2 = nn.Linear(intermediate_dim, dim)


class TextStyleEncoder(nn.Module):
    """text_style_encoder.py:6-26.  The arithmetic runs in libstts_b200 (stts_text_style_encoder: conv
    k7, ConvNeXt blocks with GRN over the utterance, mean over time)."""

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


class _LinearNorm(nn.Module):
    def __init__(self, in_dim, out_dim):
        super().__init__()
        self.linear_layer = nn.Linear(in_dim, out_dim)


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
