"""Diagnostic (GPU box): one utterance of the C2 batch, alone and inside the batch, stage by stage against the oracle."""
import json, math, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from bench import workload
from oracle import stylish_port as port
from stylish_tts_b200 import fixture
from tests.util import fixture_sds, synthesizer

mode = sys.argv[1] if len(sys.argv) > 1 else "tf32x3"
i = int(sys.argv[2]) if len(sys.argv) > 2 else 12
tokens, durs, frames = workload(1)
g = torch.Generator().manual_seed(2025)
noise = [fixture.draw_noise(f, g) for f in frames]
sds = fixture_sds()
taps = {}
with torch.no_grad():
    ref = port.synthesize(sds, tokens[i][None], torch.tensor([tokens[i].numel()]), noise[i], duration=durs[i].float(),
                          exact_frame0=True, taps=taps)
syn = synthesizer(mode)
eng = syn.engine
F4 = 4 * frames[i]

def report(tag, out_audio):
    e = (out_audio.double().cpu() - ref["audio"].double().reshape(-1)).abs()
    fr = e.reshape(-1, 75).max(dim=1).values
    bad = torch.nonzero(fr > 1e-4).flatten()
    print(tag, "max", float(e.max()), "mean", float(e.mean()), "frames>1e-4:", bad.numel(),
          "first/last", (int(bad[0]), int(bad[-1])) if bad.numel() else None)
    return bad

def stage_dump(tag):
    rows, starts = eng.layout_info(4)
    seg = 0 if tag == "alone" else i
    s0 = starts[seg]
    def rd(name, c):
        return eng.debug_read(name)[s0: s0 + F4, :c].t().cpu()
    for name, buf, c, key in (("mel", "CATP", 512, "mel"), ("har_spec", "HS", 1025, "har_spec"), ("gen_proj?XG_final", "XG", 512, "gen_convnext")):
        got, want = rd(buf, c), taps[key][0]
        d = (got.double() - want.double()).abs()
        print(f"  {tag} {name}: max {float(d.max()):.3e} rel {float(d.max())/float(want.abs().max()):.3e} worst frame {int(d.max(dim=0).values.argmax())}")
    got, want = rd("HP", 1025), taps["har_phase"][0]
    d = (got.double() - want.double()).abs()
    dw = torch.minimum(d, (d - 2 * math.pi).abs())
    fl = torch.nonzero(d[:, 1:] > 3.0)
    print(f"  {tag} har_phase: wrapped max {float(dw[:,1:].max()):.3e} flips(non frame0) {fl.shape[0]}", fl[:6].tolist())
    for k, fr in fl[:6].tolist():
        fr += 1
        print(f"     bin {k} frame {fr}: ours {float(got[k, fr]):.6f} ref {float(want[k, fr]):.6f} |X| ref {float(taps['har_spec'][0][k, fr]):.4e} ours {float(rd('HS',1025)[k, fr]):.4e}")
    la, ph = rd("LA", 1025), rd("PH", 1025)
    for name, got, want in (("logamp", la, ref["logamp"][0][:, :-1]), ("phase", ph, ref["phase"][0][:, :-1])):
        d = (got.double() - want.double()).abs()
        print(f"  {tag} {name}: max {float(d.max()):.3e} worst frame {int(d.max(dim=0).values.argmax())}")

pe = [(ref["pitch"], ref["energy"])]
if mode == "f16":
    os.environ.setdefault("X", "1")
one = syn.synthesize([tokens[i]], durations=[durs[i]], noise=[noise[i]], pitch_energy=pe, want_spec=True)
bad = report("alone   ", one["audio"][0])
stage_dump("alone")
first = syn.synthesize(tokens, durations=durs, noise=noise)
pes = [(first["pitch"][k].clone(), first["energy"][k].clone()) for k in range(len(tokens))]
pes[i] = pe[0]
full = syn.synthesize(tokens, durations=durs, noise=noise, pitch_energy=pes, want_spec=True)
bad = report("in batch", full["audio"][i])
stage_dump("batch")
print("alone vs batch", float((one["audio"][0] - full["audio"][i]).abs().max()))
