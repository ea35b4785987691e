"""Model hyper-parameters of the synthesis path (defaults = the reference's
train/config/model.yml).  Objects with the same attribute names -- e.g. the reference's pydantic
``ModelConfig`` (lib/config_loader.py:369-420) -- are accepted wherever a ModelConfig is."""
from __future__ import annotations

from dataclasses import dataclass, field


@dataclass
class TextEncoderConfig:
    tokens: int = 178
    hidden_dim: int = 128
    filter_channels: int = 512
    heads: int = 8
    layers: int = 8
    kernel_size: int = 3
    dropout: float = 0.2


@dataclass
class StyleEncoderConfig:
    layers: int = 2


@dataclass
class DecoderConfig:
    hidden_dim: int = 512
    residual_dim: int = 64


@dataclass
class GeneratorConfig:
    type: str = "freegan"
    input_dim: int = 512
    hidden_dim: int = 512
    conv_intermediate_dim: int = 1536
    io_conv_kernel_size: int = 7


@dataclass
class DurationPredictorConfig:
    n_layer: int = 4
    duration_classes: int = 16
    max_duration: int = 50
    dropout: float = 0.2
    last_dropout: float = 0.5


@dataclass
class PitchEnergyPredictorConfig:
    inter_dim: int = 256
    dropout: float = 0.2


@dataclass
class ModelConfig:
    sample_rate: int = 24000
    n_fft: int = 2048
    win_length: int = 1200
    hop_length: int = 300
    style_dim: int = 64
    inter_dim: int = 128
    text_encoder: TextEncoderConfig = field(default_factory=TextEncoderConfig)
    style_encoder: StyleEncoderConfig = field(default_factory=StyleEncoderConfig)
    decoder: DecoderConfig = field(default_factory=DecoderConfig)
    generator: GeneratorConfig = field(default_factory=GeneratorConfig)
    duration_predictor: DurationPredictorConfig = field(default_factory=DurationPredictorConfig)
    pitch_energy_predictor: PitchEnergyPredictorConfig = field(default_factory=PitchEnergyPredictorConfig)


def check_supported(mc) -> None:
    """The kernels are specialised for the default model.yml shapes; anything else must fail
    loudly rather than silently fall back."""
    want = dict(sample_rate=24000, n_fft=2048, win_length=1200, hop_length=300, style_dim=64, inter_dim=128)
    for k, v in want.items():
        if getattr(mc, k) != v:
            raise NotImplementedError(f"stylish_tts_b200 kernels are built for {k}={v}, got {getattr(mc, k)}")
    if mc.decoder.hidden_dim != 512 or mc.decoder.residual_dim != 64:
        raise NotImplementedError("decoder dims must be 512/64")
    g = mc.generator
    if getattr(g, "type", "freegan") != "freegan" or g.hidden_dim != 512 or g.input_dim != 512 or \
            g.conv_intermediate_dim != 1536 or g.io_conv_kernel_size != 7:
        raise NotImplementedError("generator must be the default 'freegan' 512/1536/k7 configuration")
    if mc.pitch_energy_predictor.inter_dim != 256:
        raise NotImplementedError("pitch_energy_predictor.inter_dim must be 256")
