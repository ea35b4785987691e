"""Is the localized 5e-3 difference on a 60 s utterance a 16-bit effect?  Compare the two fp32-grade modes
(fp32 CUDA cores vs tf32x3 tensor cores: different STFT rounding at the 1e-7 level) on the same inputs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from stylish_tts_b200 import fixture
from tests.util import synthesizer

for T, F, seed in ((510, 4800, 14), (120, 400, 17)):
    g = torch.Generator().manual_seed(seed)
    tokens, dur = fixture.synth_tokens(T, g), fixture.synth_durations(T, F, g)
    noise = [fixture.draw_noise(F, g)]
    syn = synthesizer("tf32x3")
    ref = syn.synthesize([tokens], durations=[dur], noise=noise)
    pe = [(ref["pitch"][0].clone(), ref["energy"][0].clone())]
    res = {}
    for mode in ("tf32x3", "fp32", "f16"):
        syn = synthesizer(mode)
        out = syn.synthesize([tokens], durations=[dur], noise=noise, pitch_energy=pe)
        res[mode] = out["audio"][0].clone()
    for a, b in (("fp32", "tf32x3"), ("f16", "fp32"), ("f16", "tf32x3")):
        e = (res[a] - res[b]).abs()
        print(f"T={T} {a} vs {b}: max {float(e.max()):.2e} at frame {int(e.argmax()) // 75}, frac>1e-3 {float((e > 1e-3).float().mean()):.2e}, "
              f"mean {float(e.mean()):.2e}")
