"""Reference-side numbers for BASELINE.md section 3 (C1, C3 as single utterances; C2 is bench.py's cpu_baseline /
torch_cuda_baseline): the oracle port (= the reference's torch ops, pinned bit-exact) on the box's host cores and on
the same B200 through stock cuDNN / cuBLAS / cuFFT.  Test / measurement infrastructure (imports oracle/).
usage: python tools/ref_configs.py > gpurun_out/ref_configs.jsonl"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from oracle import stylish_port as port  # noqa: E402
from stylish_tts_b200 import fixture  # noqa: E402


def case(n_tok, frames, seed):
    g = torch.Generator().manual_seed(seed)
    return fixture.synth_tokens(n_tok, g), fixture.synth_durations(n_tok, frames, g), fixture.draw_noise(frames, g)


def run(sds, tok, dur, noise, dev, reps):
    sd = {m: {k: v.to(dev) for k, v in d.items()} for m, d in sds.items()}
    nz = {k: v.to(dev) for k, v in noise.items()}
    ts = []
    with torch.no_grad():
        for i in range(reps + 1):
            if dev != "cpu":
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = port.synthesize(sd, tok[None].to(dev), torch.tensor([tok.numel()], device=dev), nz,
                                  duration=dur.float().to(dev))
            _ = out["audio"].float().cpu()
            ts.append(time.perf_counter() - t0)
    ts = sorted(ts[1:])
    return ts[len(ts) // 2]


def main():
    sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
    torch.set_num_threads(os.cpu_count() or 1)
    for name, (n_tok, frames, seed, reps_cpu) in {"C1": (120, 400, 13, 5), "C3": (510, 4800, 14, 2)}.items():
        tok, dur, noise = case(n_tok, frames, seed)
        audio_s = 300.0 * frames / 24000.0
        rec = {"config": name, "audio_s": audio_s, "cpu_threads": torch.get_num_threads()}
        t = run(sds, tok, dur, noise, "cpu", reps_cpu)
        rec["cpu_s"] = t
        rec["cpu_audio_s_per_s"] = audio_s / t
        if torch.cuda.is_available():
            for flag in (False, True):
                torch.backends.cuda.matmul.allow_tf32 = flag
                torch.backends.cudnn.allow_tf32 = flag
                t = run(sds, tok, dur, noise, "cuda:0", 8)
                rec["cuda_tf32_%s_ms" % ("on" if flag else "off")] = t * 1e3
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
