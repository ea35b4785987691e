"""The other BASELINE.json configs on one B200 (parity-test cases, not bench.py lines): latency / throughput of
  C1  one 5 s utterance (120 phonemes)              -> p50 latency
  C3  one 60 s utterance (510 phonemes, max segment) -> latency
  C2  batch 32 mixed 0.25-10 s                      -> throughput (same as bench.py)
  C4s batch 256 mixed, one GPU's share at 8 GPUs = 32; here: the whole 256 on ONE GPU for reference
  C5  vocoder-only: decoder -> prior/flow -> generator on synthetic mel-side inputs, 64 x 10 s
Writes one JSON object per config to stdout.  usage: python tools/bench_configs.py [precision]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from stylish_tts_b200 import fixture  # noqa: E402
from stylish_tts_b200.config import ModelConfig  # noqa: E402
from stylish_tts_b200.modules import build_model  # noqa: E402
from stylish_tts_b200.synth import Synthesizer  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "f16"
dev = "cuda:0"
sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
nets = build_model(ModelConfig())
for k, m in nets.items():
    m.load_state_dict(sds[k])
syn = Synthesizer(nets, device=dev, precision=prec)


def run(name, items, reps, graph=True):
    tokens = [it[0].to(dev) for it in items]
    durs = [it[1].to(dev, torch.int32) for it in items]
    frames = [it[2] for it in items]
    audio_s = 300.0 * sum(frames) / 24000.0
    for _ in range(3):
        syn.synthesize(tokens, durations=durs, frames=frames, graph=graph)
    syn.check()
    torch.cuda.synchronize()
    lat = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        syn.synthesize(tokens, durations=durs, frames=frames, graph=graph)
        e1.record()
        torch.cuda.synchronize()
        lat.append((e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0)))
    dev_ms = sorted(x[0] for x in lat)
    wall_ms = sorted(x[1] for x in lat)
    p50, p50w = dev_ms[len(dev_ms) // 2], wall_ms[len(wall_ms) // 2]
    print(json.dumps({"config": name, "precision": prec, "launch": "graph" if graph else "eager",
                      "utterances": len(items), "audio_s": round(audio_s, 2),
                      "phonemes": sum(int(t.numel()) for t in tokens),
                      "p50_device_ms": round(p50, 3), "p95_device_ms": round(dev_ms[int(0.95 * (len(dev_ms) - 1))], 3),
                      "p50_wall_ms": round(p50w, 3), "audio_s_per_s": round(audio_s / (p50 / 1e3), 1),
                      "workspace_bytes": syn.engine.workspace_bytes()}), flush=True)


g = torch.Generator().manual_seed(13)
c1 = [(fixture.synth_tokens(120, g), fixture.synth_durations(120, 400, g), 400)]
run("C1: 1 x 5 s, 120 phonemes (latency)", c1, 30)
run("C1: 1 x 5 s, 120 phonemes (latency)", c1, 30, graph=False)
g = torch.Generator().manual_seed(14)
c3 = [(fixture.synth_tokens(510, g), fixture.synth_durations(510, 4800, g), 4800)]
run("C3: 1 x 60 s, 510 phonemes (latency)", c3, 10)
run("C2: batch 32 mixed 0.25-10 s", fixture.synth_batch(fixture.mixed_lengths(32, seed=0), seed=0), 10)
run("C4 on one GPU: batch 256 mixed 0.25-10 s", fixture.synth_batch(fixture.mixed_lengths(256, seed=0), seed=0), 5)

# ---- C5: vocoder-only (decoder -> prior/flow^-1 -> generator), 64 x 10 s, synthetic asr / pitch / energy / style
B, F4 = 64, 3200
eng = syn.engine
eng.set_layout(None, [F4] * B)
gen = torch.Generator(device=dev).manual_seed(5)
asr = torch.randn(B * 128 * F4, device=dev, generator=gen)
pitch = (120.0 * torch.exp2(torch.rand(B * F4, device=dev, generator=gen) - 0.5))
pitch = torch.where(torch.rand(B * F4, device=dev, generator=gen) < 0.3, torch.zeros_like(pitch), pitch)
energy = torch.rand(B * F4, device=dev, generator=gen)
style = torch.randn(B * 64, device=dev, generator=gen)
n_prior = torch.randn(B * 128 * F4, device=dev, generator=gen)
n_src = torch.randn(B * 75 * F4, device=dev, generator=gen)
phase = torch.rand(B, device=dev, generator=gen)


def c5():
    x = eng.decoder(asr, pitch, energy, style)
    mel, _, _ = eng.prior_flow(x, n_prior, style)
    return eng.generator(mel, style, pitch, n_src, phase, want_spec=False)


for _ in range(2):
    c5()
torch.cuda.synchronize()
ts = []
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    c5()
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ts.sort()
audio_s = B * F4 * 75 / 24000.0
print(json.dumps({"config": "C5: vocoder-only (decoder + flow + generator), 64 x 10 s", "precision": prec,
                  "launch": "eager, module-level C-ABI entries (incl. layout transposes)", "audio_s": audio_s,
                  "p50_device_ms": round(ts[len(ts) // 2], 3), "audio_s_per_s": round(audio_s / (ts[len(ts) // 2] / 1e3), 1),
                  "tflops_algorithmic": round(74.07e6 * B * F4 / (ts[len(ts) // 2] / 1e3) / 1e12, 1),
                  "workspace_bytes": eng.workspace_bytes()}), flush=True)
