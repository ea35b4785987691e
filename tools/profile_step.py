"""One synthesis step of the bench workload between cudaProfilerStart/Stop (for ncu
--profile-from-start off).  usage: python tools/profile_step.py [precision] [batch]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import workload  # noqa: E402
from stylish_tts_b200 import fixture  # noqa: E402
from stylish_tts_b200.config import ModelConfig  # noqa: E402
from stylish_tts_b200.modules import build_model  # noqa: E402
from stylish_tts_b200.synth import Synthesizer  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "f16"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 32
sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
nets = build_model(ModelConfig())
for k, m in nets.items():
    m.load_state_dict(sds[k])
syn = Synthesizer(nets, device="cuda:0", precision=prec)
tokens, durs, frames = workload(0, batch)
tok = [t.cuda() for t in tokens]
dd = [d.cuda().to(torch.int32) for d in durs]
for _ in range(2):
    syn.synthesize(tok, durations=dd, frames=frames)
torch.cuda.synchronize()
import time
torch.cuda.synchronize()
t0 = time.perf_counter()
syn.synthesize(tok, durations=dd, frames=frames)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host enqueue {1e3 * (t1 - t0):.2f} ms, until device idle {1e3 * (t2 - t0):.2f} ms")
layer_mode = bool(os.environ.get("STTS_PROFILE_LAYER"))   # the library brackets the named GEMM itself
if not layer_mode:
    torch.cuda.profiler.start()
syn.synthesize(tok, durations=dd, frames=frames)
torch.cuda.synchronize()
if not layer_mode:
    torch.cuda.profiler.stop()
print("done", sum(frames) * 300 / 24000.0, "audio-s")
