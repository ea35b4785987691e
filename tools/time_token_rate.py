"""Device time of each token-rate section of Synthesizer.synthesize on the bench batch (CUDA events)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import workload
from stylish_tts_b200 import fixture
from stylish_tts_b200.config import ModelConfig
from stylish_tts_b200.modules import build_model
from stylish_tts_b200.synth import Synthesizer

if len(sys.argv) > 1 and sys.argv[1] == "bench":
    torch.backends.cudnn.benchmark = True
sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
nets = build_model(ModelConfig())
for k, m in nets.items():
    m.load_state_dict(sds[k])
syn = Synthesizer(nets, device="cuda:0", precision="f16")
tokens, durs, frames = workload(0)
dev = "cuda:0"
B = len(tokens); T = [int(t.numel()) for t in tokens]; Tmax = max(T)
tok = torch.zeros((B, Tmax), dtype=torch.int64, device=dev)
for b, t in enumerate(tokens):
    tok[b, :T[b]] = t.to(dev)
lens = torch.tensor(T, device=dev)
nets = syn.nets

def timeit(name, fn, n=5):
    for _ in range(2):
        out = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        out = fn()
    e1.record(); torch.cuda.synchronize()
    print(f"{name:28s} {e0.elapsed_time(e1) / n:8.3f} ms")
    return out

with torch.no_grad():
    timeit("duration_predictor", lambda: nets["duration_predictor"](tok, lens))
    timeit("  dp.text_encoder", lambda: nets["duration_predictor"].text_encoder(tok, lens))
    pe_enc = timeit("pe_text_encoder", lambda: nets["pe_text_encoder"](tok, lens))[0]
    pe_style = timeit("pe_text_style_encoder", lambda: nets["pe_text_style_encoder"](pe_enc, lens))
    pep = nets["pitch_energy_predictor"]
    pros = timeit("pe prosody_encoder", lambda: pep.prosody_encoder(pe_enc, pe_style, lens))
    enc = timeit("sp.text_encoder", lambda: nets["speech_predictor"].text_encoder(tok, lens))[0]
    timeit("sp.style_encoder", lambda: nets["speech_predictor"].style_encoder(enc, lens))
    dd = [d.to(dev, torch.int32) for d in durs]
    timeit("full synthesize", lambda: syn.synthesize([t.to(dev) for t in tokens], durations=dd, frames=frames))
