"""Where does the 16-bit mode differ from the tf32x3 mode on a 60 s utterance?  (diagnostic)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from stylish_tts_b200 import fixture
from tests.util import synthesizer

for T, F, seed in ((510, 4800, 14), (510, 4800, 15), (255, 2400, 16), (120, 400, 17)):
    g = torch.Generator().manual_seed(seed)
    tokens, dur = fixture.synth_tokens(T, g), fixture.synth_durations(T, F, g)
    noise = [fixture.draw_noise(F, g)]
    syn = synthesizer("tf32x3")
    ref = syn.synthesize([tokens], durations=[dur], noise=noise)
    pe = [(ref["pitch"][0].clone(), ref["energy"][0].clone())]
    ref = syn.synthesize([tokens], durations=[dur], noise=noise, pitch_energy=pe, want_spec=True)
    a_ref, la_ref, ph_ref = ref["audio"][0].clone(), ref["logamp"][0].clone(), ref["phase"][0].clone()
    syn = synthesizer("f16")
    out = syn.synthesize([tokens], durations=[dur], noise=noise, pitch_energy=pe, want_spec=True)
    e = (out["audio"][0] - a_ref).abs()
    i = int(e.argmax())
    q = torch.quantile(e[::7].float(), torch.tensor([0.5, 0.99, 0.9999], device=e.device))
    ela = (out["logamp"][0] - la_ref).abs()
    print(f"T={T} F={F}: audio max {float(e.max()):.2e} at sample {i} (frame {i // 75} of {4 * F}), "
          f"p50 {float(q[0]):.1e} p99 {float(q[1]):.1e} p99.99 {float(q[2]):.1e}, frac>1e-3 {float((e > 1e-3).float().mean()):.2e}, "
          f"|audio| max {float(a_ref.abs().max()):.3f}; logamp max err {float(ela.max()):.2e} at frame {int(ela.max(0).values.argmax())}, "
          f"pitch range {float(pe[0][0].min()):.0f}-{float(pe[0][0].max()):.0f}")
    w = e[max(0, i - 3000): i + 3000]
    print("   around the max: mean err", f"{float(w.mean()):.2e}", " whole-utterance mean", f"{float(e.mean()):.2e}")
