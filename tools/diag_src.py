"""Diagnostic (GPU box): source + STFT of one utterance against the oracle, bit level."""
import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from bench import workload
from oracle import stylish_port as port
from stylish_tts_b200 import fixture
from tests.util import fixture_sds, gpu_nets

i = int(sys.argv[1]) if len(sys.argv) > 1 else 12
tokens, durs, frames = workload(1)
g = torch.Generator().manual_seed(2025)
noise = [fixture.draw_noise(f, g) for f in frames]
taps = {}
with torch.no_grad():
    ref = port.synthesize(fixture_sds(), tokens[i][None], torch.tensor([tokens[i].numel()]), noise[i],
                          duration=durs[i].float(), exact_frame0=True, taps=taps)
eng = gpu_nets()["speech_predictor"].engine()
eng.set_precision("fp32")
L = 4 * frames[i]
eng.set_layout(None, [L])
dev = "cuda:0"
p4 = taps["pitch4"].reshape(-1).to(dev).contiguous()
src = eng.source_pcph(p4, noise[i]["source"].reshape(-1).to(dev), noise[i]["phase"].reshape(-1)[:1].to(dev)).cpu()
want = taps["source"].reshape(-1)
d = (src.double() - want.double()).abs()
print("source: max abs diff", float(d.max()), "n differing", int((src != want).sum()), "of", src.numel())
if float(d.max()) > 0:
    j = int(d.argmax()); print("  worst sample", j, "frame", j // 75, float(src[j]), float(want[j]), "pitch", float(p4[j // 75]))
    nz = torch.nonzero(src != want).flatten()
    print("  differing frames (first 10):", sorted(set((nz // 75).tolist()))[:10])
# exact DFT (fp64) of the fp32 products at one bin / frame from both sources
def dft(s, k, fr):
    x = torch.nn.functional.pad(s[None, None], (1024, 1024), mode="reflect")[0, 0]
    seg = x[75 * fr + 424: 75 * fr + 1624]
    w = torch.hann_window(1200)
    y = (seg * w).double()
    n = torch.arange(424, 1624, dtype=torch.float64)
    ang = -2 * math.pi * k * n / 2048
    return float((y * torch.cos(ang)).sum()), float((y * torch.sin(ang)).sum())
for k, fr in ((847, 1224),):
    print("bin", k, "frame", fr, "exact DFT of GPU source:", dft(src, k, fr), " of oracle source:", dft(want, k, fr))
    sp = torch.stft(want[None], 2048, 75, 1200, window=torch.hann_window(1200), return_complex=True)[0]
    print("   torch.stft fp32 of oracle source:", complex(sp[k, fr]))
mel = taps["mel"].to(dev).contiguous().view(-1)
eng.generator(mel, taps["sp_style"].to(dev).view(-1), p4, noise[i]["source"].reshape(-1).to(dev),
              noise[i]["phase"].reshape(-1)[:1].to(dev))
rows, starts = eng.layout_info(4)
hs = eng.debug_read("HS")[starts[0]: starts[0] + L, :1025].cpu()
hp = eng.debug_read("HP")[starts[0]: starts[0] + L, :1025].cpu()
print("ours |X|, phase at that bin:", float(hs[1224, 847]), float(hp[1224, 847]), " ref:", float(taps["har_spec"][0, 847, 1224]), float(taps["har_phase"][0, 847, 1224]))
