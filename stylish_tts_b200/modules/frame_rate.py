"""Frame-rate modules of the synthesis path: host-side mirror of the reference ``nn.Module``
interface (same constructor arguments, ``forward`` signatures, ``state_dict()`` keys -- SURVEY.md
section 8b) whose arithmetic runs in the sm_100a kernels of libstts_b200.so.

  AdaptiveDecoderBlock / AdaptiveInstance   train/models/ada_norm.py:129-182
  Decoder                                   train/models/decoder.py:6-60
  PriorEncoder, ResidualCouplingBlock, WN   train/models/flow.py:17-226, 296-315
  Generator (+ generate_pcph, TorchSTFT)    train/models/generator.py:23-56, 247-499
  PitchEnergyPredictor                      train/models/pitch_energy_predictor.py:11-121
  SpeechPredictor                           train/models/speech_predictor.py:15-129
  ExportModel                               train/models/export_model.py:5-45
  DurationProcessor, DecoderPrediction      train/utils.py:363-494

The modules only *hold* parameters in the reference layout; ``prepare()`` (implicit on first use)
hands the state dict to the C library, which folds weight-norm and prepacks.  A batch [B, C, L]
is B independent utterances (the reference has no masked mixed-length path); ragged
batches go through ``stylish_tts_b200.synth.Synthesizer``.  There is no PyTorch fallback.
"""
from __future__ import annotations

import warnings
from typing import Dict, Optional

import torch
from torch import nn
from torch.nn.utils.parametrizations import weight_norm

from ..config import check_supported
from ..engine import Engine
from .token_rate import (MultiHeadAttention, ProsodyEncoder, StyleLayerNorm, TextEncoder, TextStyleEncoder,
                         TokenPack)


def _legacy_weight_norm(m: nn.Module) -> nn.Module:
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return torch.nn.utils.weight_norm(m, name="weight")


class DecoderPrediction:
    """train/utils.py:363-382 (flow statistics are training-only and stay None)."""

    def __init__(self, *, audio, magnitude, phase, text_stats=None, text2mel_stats=None, mel_stats=None,
                 mel2text_stats=None):
        self.audio, self.magnitude, self.phase = audio, magnitude, phase
        self.text_stats, self.text2mel_stats = text_stats, text2mel_stats
        self.mel_stats, self.mel2text_stats = mel_stats, mel2text_stats


# ----------------------------------------------------------------------------- engine ownership
class _EngineOwner(nn.Module):
    """Mixin: lazily builds the C context from this module's own state dict.  A child module
    constructed inside a parent owner delegates to the parent (one context, one workspace)."""

    _module_name = "speech_predictor"
    _key_prefix = ""

    def _init_owner(self):
        object.__setattr__(self, "_eng", None)
        object.__setattr__(self, "_parent_owner", None)
        object.__setattr__(self, "noise_override", None)

    def _adopt(self, child: "_EngineOwner"):
        object.__setattr__(child, "_parent_owner", self)

    def attach_engine(self, engine: Engine):
        """Use a context whose weights were loaded by the caller (Synthesizer)."""
        object.__setattr__(self, "_eng", engine)

    def invalidate(self):
        object.__setattr__(self, "_eng", None)

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        sd = {k: v for k, v in state_dict.items() if not k.startswith("posterior_encoder.")}  # training-only
        self.invalidate()
        return super().load_state_dict(sd, strict=strict, assign=assign)

    def engine(self) -> Engine:
        if self._parent_owner is not None:
            return self._parent_owner.engine()
        if self._eng is None:
            self.prepare()
        return self._eng

    def prepare(self, precision: Optional[str] = None) -> Engine:
        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("stylish_tts_b200 modules run on a B200 only: move the module to CUDA (no CPU fallback)")
        eng = Engine(dev)
        eng.load_state(self._module_name, self.state_dict(), prefix=self._key_prefix)
        eng.finalize()
        if precision is not None:
            eng.set_precision(precision)
        object.__setattr__(self, "_eng", eng)
        return eng


def _dense(x: torch.Tensor) -> torch.Tensor:
    return x.detach().to(torch.float32).contiguous().view(-1)


# ----------------------------------------------------------------------------- parameter holders
class AdaptiveInstance(nn.Module):
    def __init__(self, style_dim, num_features):
        super().__init__()
        self.fc = nn.Linear(style_dim, num_features * 2)


class AdaptiveDecoderBlock(nn.Module):
    def __init__(self, dim_in, dim_out, style_dim=64, dropout_p=0.0, **_):
        super().__init__()
        self.learned_sc = dim_in != dim_out
        self.conv1 = weight_norm(nn.Conv1d(dim_in, dim_out, 3, 1, 1))
        self.conv2 = weight_norm(nn.Conv1d(dim_out, dim_out, 3, 1, 1))
        self.norm1 = AdaptiveInstance(style_dim, dim_in)
        self.norm2 = AdaptiveInstance(style_dim, dim_out)
        if self.learned_sc:
            self.conv1x1 = weight_norm(nn.Conv1d(dim_in, dim_out, 1, 1, 0, bias=False))


class Decoder(_EngineOwner):
    _key_prefix = "decoder."

    def __init__(self, *, dim_in, style_dim, dim_out, hidden_dim, residual_dim):
        super().__init__()
        self._init_owner()
        if (dim_in, style_dim, hidden_dim, residual_dim) != (128, 64, 512, 64):
            raise NotImplementedError("Decoder kernels are built for dim_in 128, style 64, hidden 512, residual 64")
        self.encode = AdaptiveDecoderBlock(dim_in + 2, hidden_dim, style_dim)
        self.decode = nn.ModuleList(
            [AdaptiveDecoderBlock(hidden_dim + 2 + residual_dim, hidden_dim, style_dim) for _ in range(4)])
        self.F0_conv = weight_norm(nn.Conv1d(1, 1, kernel_size=3, stride=1, groups=1, padding=1))
        self.N_conv = weight_norm(nn.Conv1d(1, 1, kernel_size=3, stride=1, groups=1, padding=1))
        self.asr_res = nn.Sequential(weight_norm(nn.Conv1d(dim_in, residual_dim, kernel_size=1)))

    def forward(self, asr, F0_curve, N, s):
        b, _, length = asr.shape
        eng = self.engine()
        eng.set_layout(None, [length] * b)
        out = eng.decoder(_dense(asr), _dense(F0_curve), _dense(N), _dense(s))
        return out.view(b, 512, length), F0_curve


class PriorEncoder(nn.Module):
    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.proj_mean = nn.Linear(in_channels, out_channels)
        self.proj_logstd = nn.Linear(in_channels, out_channels)


class WN(nn.Module):
    def __init__(self, hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=0, p_dropout=0):
        super().__init__()
        if dilation_rate != 1 or kernel_size != 5 or hidden_channels != 128 or n_layers != 4:
            raise NotImplementedError("WN kernels are built for hidden 128, kernel 5, dilation 1, 4 layers")
        self.in_layers = nn.ModuleList()
        self.res_skip_layers = nn.ModuleList()
        self.cond_layer = _legacy_weight_norm(nn.Linear(gin_channels, 2 * hidden_channels * n_layers))
        for i in range(n_layers):
            self.in_layers.append(_legacy_weight_norm(
                nn.Conv1d(hidden_channels, 2 * hidden_channels, kernel_size, padding=(kernel_size - 1) // 2)))
            rs = 2 * hidden_channels if i < n_layers - 1 else hidden_channels
            self.res_skip_layers.append(_legacy_weight_norm(nn.Linear(hidden_channels, rs)))


class ResidualCouplingLayer(nn.Module):
    def __init__(self, channels, hidden_channels, kernel_size, dilation_rate, n_layers, cond_channels=0):
        super().__init__()
        self.pre = nn.Linear(channels, hidden_channels)
        self.enc = WN(hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=cond_channels)
        self.proj_mean = nn.Linear(hidden_channels, channels)
        self.proj_logstd = nn.Linear(hidden_channels, channels)


class Flip(nn.Module):
    pass


class ResidualCouplingBlock(nn.Module):
    def __init__(self, channels, hidden_channels, kernel_size, dilation_rate, n_layers, n_flows=4, gin_channels=0,
                 use_transformer_flow=False):
        super().__init__()
        if use_transformer_flow or n_flows != 8 or channels != 128:
            raise NotImplementedError("flow kernels are built for 8 non-transformer coupling layers of 128 channels")
        self.flows = nn.ModuleList()
        for _ in range(n_flows):
            self.flows.append(ResidualCouplingLayer(channels // 2, hidden_channels, kernel_size, dilation_rate,
                                                    n_layers, cond_channels=gin_channels))
            self.flows.append(Flip())


class _GRNParams(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.gamma = nn.Parameter(torch.zeros(1, 1, dim))
        self.beta = nn.Parameter(torch.zeros(1, 1, dim))


class _AdaLNParams(nn.Module):
    def __init__(self, style_dim, channels):
        super().__init__()
        self.fc = nn.Linear(style_dim, channels * 2)


class ConvNeXtBlock(nn.Module):
    def __init__(self, dim, intermediate_dim, style_dim, kernel_size=7):
        super().__init__()
        self.dwconv = nn.Conv1d(dim, dim, kernel_size=kernel_size, padding=(kernel_size - 1) // 2, groups=dim)
        self.norm = _AdaLNParams(style_dim, dim)
        self.pwconv1 = nn.Linear(dim, intermediate_dim)
        self.grn = _GRNParams(intermediate_dim)
        self.pwconv2 = nn.Linear(intermediate_dim, dim)


class Generator(_EngineOwner):
    _key_prefix = "generator."

    def __init__(self, *, style_dim, n_fft, win_length, hop_length, config):
        super().__init__()
        self._init_owner()
        if (style_dim, n_fft, win_length, hop_length) != (64, 2048, 1200, 75):
            raise NotImplementedError("Generator kernels are built for style 64, n_fft 2048, win 1200, hop 75")
        hd = config.hidden_dim
        k = config.io_conv_kernel_size
        self.amp_output_conv = nn.Conv1d(hd + hd // 2, n_fft // 2 + 1, k, 1, padding=(k - 1) // 2)
        self.phase_output_conv = nn.Conv1d(hd + hd // 2, n_fft // 2 + 1, k, 1, padding=(k - 1) // 2)
        self.amp_final_layer_norm = _AdaLNParams(style_dim, hd)
        self.phase_final_layer_norm = _AdaLNParams(style_dim, hd)
        self.projector = nn.Conv1d(config.input_dim + hd, hd, 1, 1, 0)
        self.amp_prior_conv = nn.Conv1d(n_fft // 2 + 1, hd // 2, 7, 1, 3)
        self.phase_prior_conv = nn.Conv1d(n_fft // 2 + 1, hd // 2, 7, 1, 3)
        self.convnext = nn.ModuleList(
            [ConvNeXtBlock(hd, config.conv_intermediate_dim, style_dim, ks) for ks in (31, 15, 7, 3)])
        for m in self.modules():  # generator.py:397-400
            if isinstance(m, nn.Conv1d) and m.groups == 1:
                nn.init.trunc_normal_(m.weight, std=0.02)
                nn.init.constant_(m.bias, 0)

    def _draw(self, b, length, device):
        """Same torch draws, in the same order, as generate_pcph (generator.py:275-277, 306)."""
        if self.noise_override is not None:
            return self.noise_override["source"].to(device), self.noise_override["phase"].to(device)
        noise = torch.randn((b, 1, length * 75), device=device)
        return noise, torch.rand((1, 1), device=device)

    def forward(self, *, mel, style, pitch, energy):
        b, _, length = mel.shape
        eng = self.engine()
        eng.set_layout(None, [length] * b)
        noise, phase = self._draw(b, length, mel.device)
        init = phase.reshape(-1)[:1].expand(b).contiguous()  # one draw shared by the batch, as the reference
        audio, la, ph = eng.generator(_dense(mel), _dense(style), _dense(pitch), _dense(noise), _dense(init))
        eng.check_source_status()
        return DecoderPrediction(audio=audio.view(b, 1, 75 * length), magnitude=la.view(b, 1025, length + 1),
                                 phase=ph.view(b, 1025, length + 1))


# ----------------------------------------------------------------------------- pitch / energy
class PitchEnergyPredictor(_EngineOwner):
    _module_name = "pitch_energy_predictor"

    def __init__(self, style_dim, inter_dim, text_config, style_config, duration_config, pitch_energy_config):
        super().__init__()
        self._init_owner()
        if (style_dim, inter_dim) != (64, 256):
            raise NotImplementedError("PitchEnergyPredictor kernels are built for style 64, inter_dim 256")
        cc = inter_dim + style_dim
        self.prosody_encoder = ProsodyEncoder(sty_dim=style_dim, d_model=inter_dim, nlayers=3, dropout=0.2)
        self.query_norm = StyleLayerNorm(style_dim, cc)
        self.key_norm = StyleLayerNorm(style_dim, cc)
        self.cross_attention = MultiHeadAttention(channels=cc, out_channels=cc, n_heads=8)
        self.cross_window = 5
        self.cross_post = nn.Sequential(
            weight_norm(nn.Conv1d(cc, cc, kernel_size=5, padding=2, groups=cc)), nn.SiLU(),
            weight_norm(nn.Conv1d(cc, cc, kernel_size=1)))
        self.F0 = nn.ModuleList([AdaptiveDecoderBlock(cc, cc, style_dim) for _ in range(3)])
        self.N = nn.ModuleList([AdaptiveDecoderBlock(cc, cc, style_dim) for _ in range(3)])
        self.F0_proj = nn.Conv1d(cc, 1, 1, 1, 0)
        self.N_proj = nn.Conv1d(cc, 1, 1, 1, 0)

    def forward(self, text_encoding, text_lengths, alignment, style):
        """pitch_energy_predictor.py:104-119 in libstts_b200: prosody encoder, cross-attention block
        (stts_pe_cross) and the two towers."""
        b, _, t = text_encoding.shape
        f = alignment.shape[2]
        dev = text_encoding.device
        eng = self.engine()
        lengths = [int(v) for v in text_lengths.tolist()]
        eng.set_token_layout(lengths)
        eng.set_layout([f] * b, None)
        pack = TokenPack(lengths, t, dev)
        eng.prosody_encoder("pitch_energy_predictor", pack.gather(text_encoding.transpose(1, 2)), style,
                            self.prosody_encoder.hidden)
        idx = alignment.argmax(dim=1).to(torch.int32).contiguous().view(-1)   # one-hot columns -> token index
        cu_frames = torch.arange(b + 1, device=dev, dtype=torch.int32) * f
        eng.pe_cross(None, style, idx, cu_frames)
        pitch, energy = eng.pe_towers(None, _dense(style))
        return pitch.view(b, f), energy.view(b, f)


# ----------------------------------------------------------------------------- speech predictor
class SpeechPredictor(_EngineOwner):
    def __init__(self, model_config):
        super().__init__()
        self._init_owner()
        mc = model_config
        check_supported(mc)
        self.text_encoder = TextEncoder(inter_dim=mc.inter_dim, config=mc.text_encoder)
        self.style_encoder = TextStyleEncoder(mc.inter_dim, mc.style_dim, mc.style_encoder)
        self.decoder = Decoder(dim_in=mc.inter_dim, style_dim=mc.style_dim, dim_out=mc.generator.input_dim,
                               hidden_dim=mc.decoder.hidden_dim, residual_dim=mc.decoder.residual_dim)
        fh = mc.decoder.hidden_dim // 4
        self.prior_encoder = PriorEncoder(mc.decoder.hidden_dim, fh)
        self.flow = ResidualCouplingBlock(fh, fh, 5, 1, n_layers=4, n_flows=8, gin_channels=mc.style_dim,
                                          use_transformer_flow=False)
        self.post_flow = nn.Linear(fh, mc.decoder.hidden_dim)
        self.generator = Generator(style_dim=mc.style_dim, n_fft=mc.n_fft, win_length=mc.win_length,
                                   hop_length=mc.hop_length // 4, config=mc.generator)
        self._adopt(self.decoder)
        self._adopt(self.generator)

    def _draw(self, b, f4, device):
        """The three reference draws in call order: randn_like(mean) (flow.py:314), then the
        generator's randn / rand (generator.py:275-277, 306)."""
        if self.noise_override is not None:
            n = self.noise_override
            return n["prior"].to(device), n["source"].to(device), n["phase"].to(device)
        prior = torch.randn((b, 128, f4), device=device)
        source = torch.randn((b, 1, f4 * 75), device=device)
        return prior, source, torch.rand((1, 1), device=device)

    def forward(self, texts, text_lengths, alignment, pitch, energy, audio_gt=None):
        if audio_gt is not None:
            raise NotImplementedError("the posterior (training) branch is outside the synthesis forward path")
        self.text_encoder.attach_native(self.engine(), "speech_predictor")
        enc, _, _ = self.text_encoder(texts, text_lengths)
        self.style_encoder.attach_native(self.engine(), "speech_predictor")
        style = self.style_encoder(enc, text_lengths)
        b, _, t = enc.shape
        f = alignment.shape[2]
        dev = enc.device
        eng = self.engine()
        eng.set_layout([f] * b, [4 * f] * b)
        idx = alignment.argmax(dim=1).to(torch.int32).contiguous().view(-1)   # one-hot columns -> token index
        cu_tok = torch.arange(b + 1, device=dev, dtype=torch.int32) * t
        cu_frames = torch.arange(b + 1, device=dev, dtype=torch.int32) * f
        enc_tok = enc.transpose(1, 2).contiguous().view(b * t, 128)
        prior, source, phase = self._draw(b, 4 * f, dev)
        init = phase.reshape(-1)[:1].expand(b).contiguous()
        audio, la, ph = eng.speech_frames(_dense(enc_tok).view(b * t, 128), cu_tok, idx, cu_frames, _dense(pitch),
                                          _dense(energy), _dense(style), _dense(prior), _dense(source), _dense(init),
                                          want_spec=True)
        eng.check_source_status()
        return DecoderPrediction(audio=audio.view(b, 1, 300 * f), magnitude=la.view(b, 1025, 4 * f + 1),
                                 phase=ph.view(b, 1025, 4 * f + 1))


class ExportModel(nn.Module):
    """export_model.py:5-45: PE encoder -> PE style -> PE predictor -> speech predictor, B = 1."""

    def __init__(self, *, speech_predictor, duration_predictor, pitch_energy_predictor, pe_text_encoder,
                 pe_text_style_encoder, device, **kwargs):
        super().__init__()
        for m in (speech_predictor, duration_predictor, pitch_energy_predictor, pe_text_encoder,
                  pe_text_style_encoder):
            m.to(device).eval()
            for p in m.parameters():
                p.requires_grad = False
        self.speech_predictor = speech_predictor
        self.pitch_energy_predictor = pitch_energy_predictor
        self.pe_text_encoder = pe_text_encoder
        self.pe_text_style_encoder = pe_text_style_encoder

    @torch.no_grad()
    def forward(self, texts, text_lengths, alignment):
        pe_enc, _, _ = self.pe_text_encoder(texts, text_lengths)
        pe_style = self.pe_text_style_encoder(pe_enc, text_lengths)
        pitch, energy = self.pitch_energy_predictor(pe_enc, text_lengths, alignment, pe_style)
        pred = self.speech_predictor(texts, text_lengths, alignment, pitch, energy)
        return pred.audio.reshape(-1)


# ----------------------------------------------------------------------------- length regulation
_DEFAULT_ENGINES: Dict[int, Engine] = {}


def default_engine(device: torch.device) -> Engine:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _DEFAULT_ENGINES:
        _DEFAULT_ENGINES[idx] = Engine(torch.device("cuda", idx))
    return _DEFAULT_ENGINES[idx]


class DurationProcessor(nn.Module):
    """train/utils.py:385-494.  Same methods / return types (fp32 durations, dense one-hot
    alignment) for drop-in use; ``duration_to_indices`` is the form the kernels consume."""

    TABLE = [1, 2, 3, 4, 5, 6, 7, 9, 12, 15, 18, 22, 27, 32, 38, 46]

    def __init__(self, class_count, max_dur):
        super().__init__()
        if class_count != 16:
            raise NotImplementedError("duration kernel is built for 16 classes")
        self.class_count, self.max_dur = class_count, max_dur
        self.register_buffer("class_to_dur_table", torch.tensor(self.TABLE, dtype=torch.float32))

    def prediction_to_duration(self, pred, text_length=None):
        shape = pred.shape[:-1]
        dur = default_engine(pred.device).durations(pred.detach().to(torch.float32).reshape(-1, 16))
        return dur.to(torch.float32).view(shape)

    def duration_to_indices(self, duration):
        d = duration.to(torch.int32).contiguous().view(-1)
        eng = default_engine(d.device)
        cu_tok = torch.tensor([0, d.numel()], dtype=torch.int32, device=d.device)
        total = int(d.sum().item())  # output size is data dependent, as in the reference (utils.py:481-484)
        cu_frames, idx, status = eng.frame_index(d, cu_tok, total)
        return idx[:total].to(torch.int64)

    def duration_to_alignment(self, duration):
        idx = self.duration_to_indices(duration)
        out = torch.zeros((duration.shape[0], idx.shape[0]), device=duration.device)
        out[idx, torch.arange(idx.shape[0], device=duration.device)] = 1
        return out

    def forward(self, pred, text_length):
        return self.duration_to_alignment(self.prediction_to_duration(pred, text_length))


def build_model(model_config) -> Dict[str, nn.Module]:
    """The five hot modules with the reference's constructor arguments (models/models.py:32-57,83)."""
    mc = model_config
    check_supported(mc)
    from .token_rate import DurationPredictor
    return dict(
        duration_predictor=DurationPredictor(style_dim=mc.style_dim, inter_dim=mc.inter_dim,
                                             text_config=mc.text_encoder, style_config=mc.style_encoder,
                                             duration_config=mc.duration_predictor),
        pitch_energy_predictor=PitchEnergyPredictor(style_dim=mc.style_dim,
                                                    inter_dim=mc.pitch_energy_predictor.inter_dim,
                                                    text_config=mc.text_encoder, style_config=mc.style_encoder,
                                                    duration_config=mc.duration_predictor,
                                                    pitch_energy_config=mc.pitch_energy_predictor),
        pe_text_encoder=TextEncoder(inter_dim=mc.pitch_energy_predictor.inter_dim, config=mc.text_encoder),
        pe_text_style_encoder=TextStyleEncoder(mc.pitch_energy_predictor.inter_dim, mc.style_dim, mc.style_encoder),
        speech_predictor=SpeechPredictor(mc),
    )
