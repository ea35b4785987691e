"""Host-side mirror of the reference nn.Module interface for the synthesis path."""
from .frame_rate import (AdaptiveDecoderBlock, Decoder, DecoderPrediction, DurationProcessor, ExportModel, Generator,
                         PitchEnergyPredictor, PriorEncoder, ResidualCouplingBlock, SpeechPredictor, build_model)
from .token_rate import DurationPredictor, ProsodyEncoder, TextEncoder, TextStyleEncoder
from .variants import STFT, AdaptiveGeneratorBlock, SineGenerator

__all__ = ["AdaptiveDecoderBlock", "Decoder", "DecoderPrediction", "DurationProcessor", "ExportModel", "Generator",
           "PitchEnergyPredictor", "PriorEncoder", "ResidualCouplingBlock", "SpeechPredictor", "build_model",
           "DurationPredictor", "ProsodyEncoder", "TextEncoder", "TextStyleEncoder", "AdaptiveGeneratorBlock",
           "SineGenerator", "STFT"]
