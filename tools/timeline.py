"""Kernel timeline of graph-replayed steps (CUPTI through torch.profiler -- there is no nsys in the image):
     python tools/timeline.py c1|c2 [precision] > gpurun_out/timeline.txt
Prints, for the LAST profiled replay, one line per kernel: start offset (us), duration (us), gap since the latest
end of any earlier kernel (us; > 0 = the GPU had nothing of this step running), stream, name.  A measurement aid only:
numbers under the profiler are never bench values."""
import json
import os
import re
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stylish_tts_b200 import fixture  # noqa: E402
from stylish_tts_b200.config import ModelConfig  # noqa: E402
from stylish_tts_b200.modules import build_model  # noqa: E402
from stylish_tts_b200.synth import Synthesizer  # noqa: E402


def short(name: str) -> str:
    name = re.sub(r"^void ", "", name).replace("(anonymous namespace)::", "")
    name = re.sub(r"\((?:int|bool|unsigned int|[\w:]+)\)(?=[-\w])", "", name)      # template-argument casts
    name = re.sub(r"\(.*$", "", name)
    return name.replace("stts::", "")[:80]


def pool_timeline(nets, dev, precision, lanes=3, steps=6):
    """Three batches in flight (the bench's throughput mode): every kernel of `steps` steady-state steps with its
    stream, so that the overlap of one lane's token-rate chain with the other lanes' frame-rate grids can be read."""
    import bench
    from stylish_tts_b200.synth import SynthesizerPool
    from torch.profiler import ProfilerActivity, profile
    pool = SynthesizerPool(nets, device=dev, precision=precision, lanes=lanes, use_graphs=True)
    toks, durs, frames = bench.workload(1)
    toks = [t.to(dev) for t in toks]
    durs = [d.to(dev, torch.int32) for d in durs]
    for _ in range(2 * lanes):
        pool.submit(toks, durations=durs, frames=frames)
    pool.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(steps):
            pool.submit(toks, durations=durs, frames=frames)
        pool.drain()
        torch.cuda.synchronize()
    path = os.path.join(tempfile.mkdtemp(), "trace.json")
    prof.export_chrome_trace(path)
    ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") == "kernel"]
    ev.sort(key=lambda e: e["ts"])
    t0 = ev[0]["ts"]
    print(f"# pool {precision}: {len(ev)} kernels over {steps} steps, span {ev[-1]['ts'] + ev[-1]['dur'] - t0:.1f} us")
    print("# start_us dur_us stream name")
    for e in ev:
        print(f"{e['ts'] - t0:10.1f} {e['dur']:8.1f} {e['args'].get('stream', -1):4d} {short(e['name'])}")


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c1"
    precision = sys.argv[2] if len(sys.argv) > 2 else "f16"
    dev = torch.device("cuda:0")
    sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
    nets = build_model(ModelConfig())
    for k, m in nets.items():
        m.load_state_dict(sds[k])
        m.eval()
    if which == "pool":
        return pool_timeline(nets, dev, precision)
    syn = Synthesizer(nets, device=dev, precision=precision, use_graphs=True)
    g = torch.Generator().manual_seed(13)
    if which == "c1":
        toks, durs, frames = [fixture.synth_tokens(120, g)], [fixture.synth_durations(120, 400, g).to(torch.int32)], [400]
    else:
        import bench
        toks, durs, frames = bench.workload(1)
    for _ in range(4):
        syn.synthesize(toks, durations=durs, frames=frames)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3):
            syn.synthesize(toks, durations=durs, frames=frames)
            torch.cuda.synchronize()
    path = os.path.join(tempfile.mkdtemp(), "trace.json")
    prof.export_chrome_trace(path)
    ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") == "kernel"]
    ev.sort(key=lambda e: e["ts"])
    # split into replays at gaps > 200 us
    runs, cur = [], [ev[0]]
    for a, b in zip(ev, ev[1:]):
        if b["ts"] - (a["ts"] + a["dur"]) > 200:
            runs.append(cur)
            cur = []
        cur.append(b)
    runs.append(cur)
    last = runs[-1]
    if len(runs) == 1 and len(last) % 3 == 0:          # short steps: replays closer than the gap threshold
        last = last[-(len(last) // 3):]
    t0 = last[0]["ts"]
    end = 0.0
    busy = 0.0
    print(f"# {which} {precision}: {len(last)} kernels, span {last[-1]['ts'] + last[-1]['dur'] - t0:.1f} us "
          f"({len(runs)} replays seen)")
    print("# start_us dur_us gap_us stream name")
    for e in last:
        s = e["ts"] - t0
        gap = s - end if end else 0.0
        print(f"{s:9.1f} {e['dur']:7.1f} {gap:7.1f} {e['args'].get('stream', -1):3d} {short(e['name'])}")
        if s + e["dur"] > end:
            busy += s + e["dur"] - max(end, s)
            end = s + e["dur"]
    print(f"# union busy {busy:.1f} us, idle {end - busy:.1f} us")


if __name__ == "__main__":
    main()
