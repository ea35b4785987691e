"""stylish_tts_b200 -- B200-native (sm_100a) synthesis forward path of Stylish-TTS.

Public surface:
  stylish_tts_b200.modules   nn.Module mirrors of the reference interface (drop-in signatures)
  stylish_tts_b200.synth     Synthesizer: batched mixed-length synthesis on the packed layout
  stylish_tts_b200.engine    Engine: torch-facing wrapper of the C ABI (include/stts_b200.h)
The CUDA library is required; nothing here falls back to CPU or to PyTorch kernels for the
frame-rate path.
"""
__version__ = "0.1.0"
