#!/usr/bin/env python
"""Headline benchmark: audio-seconds synthesized per wall-second (1/RTF) of the Stylish-TTS
synthesis forward path on B200 (BASELINE.json metric).

Workload: ONE seeded batch of 32 x N mixed-length utterances (0.25-10 s), random-init fixture weights,
synthetic phonemes -- N = 1 is BASELINE configs[1] (batch 32 on one GPU), N = 8 is configs[3] (batch 256
sharded by utterance over 8 GPUs).  The batch is partitioned by LPT bin packing on frame counts
(stylish_tts_b200.shard), every rank synthesises its shard, and ONE gather of the waveforms to rank 0
(NCCL) closes the step -- the path's only collective.

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA path)
  python bench.py --impl reference ...                      the reference's CPU path (oracle port,
                                                            bounded sample) on the box's host cores
One JSON line on stdout from rank 0.  See DESIGN.md "Measurement" for the definitions.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "audio-sec synthesized per sec (1/RTF)"
UNIT = "audio_s/s"
BATCH = 32                                   # utterances per GPU
GFLOP_PER_AUDIO_S = 25.2                     # whole-path algorithmic FLOPs (flop counter on the reference, C1)


def workload(world: int = 1, batch_per_gpu: int = BATCH):
    """One seeded batch of `batch_per_gpu * world` utterances, lengths uniform in [0.25, 10] s
    (world = 1: BASELINE configs[1]; world = 8: configs[3]).  Identical on every rank."""
    from stylish_tts_b200 import fixture
    secs = fixture.mixed_lengths(batch_per_gpu * world, seed=0)
    items = fixture.synth_batch(secs, seed=0)
    return [it[0] for it in items], [it[1] for it in items], [it[2] for it in items]


def attention_flops(tokens, frames) -> float:
    """QK^T + PV of every attention on the path (not conv-as-GEMM launches, so not in the engine's
    per-layer profile): 3 text encoders x 8 layers x 128 ch, prosody encoders 4 x 192 + 3 x 320 ch,
    frame x token cross attention 320 ch."""
    fl = 0.0
    for t, f in zip(tokens, frames):
        T = int(t.numel())
        fl += 4.0 * T * T * (3 * 8 * 128 + 4 * 192 + 3 * 320) + 4.0 * f * T * 320
    return fl


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "bf16_burst": p["bf16_tflops"], "bf16_sustained": p["bf16_tflops_sustained"],
                "source": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def traffic_record():
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the step's largest kernel, from the
    committed `ncu --set full` capture (profiles/traffic.json names the capture it was read from)."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(path):
        return None
    return json.load(open(path))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", os.environ.get("STTS_CLOCK_MS", "20")],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm),
                "window": "nvidia-smi -lms 20 over the last untimed warm-up steps and the timed region"}


# ------------------------------------------------------------------------------------ baseline legs (oracle)
def _stratified(frames):
    order = sorted(range(len(frames)), key=lambda i: frames[i])
    pick = [order[len(order) // 2], order[len(order) // 4], order[(3 * len(order)) // 4], order[-1], order[0]]
    return pick + [i for i in range(len(frames)) if i not in pick]


def cpu_port_throughput(tokens, durs, frames, budget_s: float, seed: int = 0):
    """The reference's CPU path (oracle port = pinned restatement of the reference PyTorch forward,
    B = 1 loop exactly like stage_type.py:491-512) on a bounded sample of the same workload.  The three
    noise draws of every utterance are made outside its timed call (the reference draws inside: ~1 ms)."""
    from oracle import stylish_port as port
    from stylish_tts_b200 import fixture
    sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
    torch.set_num_threads(os.cpu_count() or 1)
    done_audio, t_total, used = 0.0, 0.0, []
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        i0 = min(range(len(frames)), key=lambda i: frames[i])       # warm-up: thread pools, allocator
        port.synthesize(sds, tokens[i0][None], torch.tensor([tokens[i0].numel()]), fixture.draw_noise(frames[i0], g),
                        duration=durs[i0].float())
        seq = _stratified(frames)
        k = 0
        while t_total < budget_s:
            i = seq[k % len(seq)]
            k += 1
            noise = fixture.draw_noise(frames[i], g)
            t0 = time.perf_counter()
            port.synthesize(sds, tokens[i][None], torch.tensor([tokens[i].numel()]), noise, duration=durs[i].float())
            t_total += time.perf_counter() - t0
            done_audio += 300.0 * frames[i] / 24000.0
            used.append(i)
    return done_audio / t_total, torch.get_num_threads(), \
        f"{len(used)} utterances of the batch ({done_audio:.1f} audio-s, B=1 loop), {t_total:.1f} s CPU wall"


def parity_against_oracle(syn, tokens, durs, frames, modes, n_utts: int = 6):
    """Checker leg: `n_utts` length-stratified utterances of the timed batch (the longest and the shortest
    included) through the CPU oracle and through the CUDA path on the SAME noise draws, the oracle's pitch /
    energy fed to the speech stage (the harmonic source integrates f0: waveform parity is defined for
    identical pitch input).  Targets: the reference with the exact frame-0 STFT phase
    ("x0": its raw frame-0 phase is FFT rounding noise) everywhere, and the raw reference beyond sample 7 500."""
    from oracle import stylish_port as port
    from oracle.metrics import maxabs, mel_l1
    from stylish_tts_b200 import fixture
    sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
    pick = _stratified(frames)[:n_utts]
    g = torch.Generator().manual_seed(4242)
    noise = {i: fixture.draw_noise(frames[i], g) for i in pick}
    ref = {}
    t0 = time.perf_counter()
    with torch.no_grad():
        for i in pick:
            a = (sds, tokens[i][None], torch.tensor([tokens[i].numel()]), noise[i])
            ref[i] = (port.synthesize(*a, duration=durs[i].float()),
                      port.synthesize(*a, duration=durs[i].float(), exact_frame0=True))
    cpu_s = time.perf_counter() - t0
    out = {}
    keep = syn.engine.precision
    for mode in modes:
        syn.set_precision(mode)
        o = syn.synthesize([tokens[i] for i in pick], durations=[durs[i] for i in pick],
                           noise=[noise[i] for i in pick], want_spec=True,
                           pitch_energy=[(ref[i][0]["pitch"], ref[i][0]["energy"]) for i in pick])
        syn.check()
        r = {"n_utts": len(pick), "seconds": [round(300.0 * frames[i] / 24000.0, 2) for i in pick]}
        err = torch.cat([(o["audio"][k].double().cpu() - ref[i][1]["audio"].double().reshape(-1)).abs()
                         for k, i in enumerate(pick)])
        # utterances beyond ~5 s contain isolated atan2 phase-wrap events of the source STFT (DESIGN.md section 4,
        # spot 3) that move ~7 frames of ANY implementation by 5e-3 .. 8e-3: max-abs is reported with the robust
        # statistics that bound everything outside those events
        r["mean_abs_x0"] = float(err.mean())
        r["q99_abs_x0"] = float(torch.quantile(err[::7].float(), 0.99))
        r["share_gt_2e-3"] = float((err > 2e-3).double().mean())
        r["max_abs_x0_le_5s"] = max([maxabs(o["audio"][k], ref[i][1]["audio"]) for k, i in enumerate(pick)
                                     if frames[i] <= 400] or [0.0])
        r["max_abs_x0"] = max(maxabs(o["audio"][k], ref[i][1]["audio"]) for k, i in enumerate(pick))
        r["max_abs_after_7500"] = max(maxabs(o["audio"][k][7500:], ref[i][0]["audio"].reshape(-1)[7500:])
                                      for k, i in enumerate(pick) if 300 * frames[i] > 7500)
        r["mel_l1"] = max(mel_l1(o["audio"][k], ref[i][1]["audio"]) for k, i in enumerate(pick))
        r["logamp_l1"] = max(float((o["logamp"][k].cpu() - ref[i][1]["logamp"][0]).abs().mean())
                             for k, i in enumerate(pick))
        r["pitch_max_abs_hz"] = max(maxabs(o["pitch"][k], ref[i][0]["pitch"]) for k, i in enumerate(pick))
        r["durations_bit_equal"] = all(
            torch.equal(o["duration_pred"][k].cpu().to(torch.int32), ref[i][0]["duration"].to(torch.int32))
            for k, i in enumerate(pick))
        out[mode] = r
    syn.set_precision(keep)
    out["target"] = ("reference (oracle port) with the exact real frame-0 STFT phase (audio_x0); raw reference beyond "
                     "sample 7500; oracle pitch/energy injected into the speech stage")
    out["oracle_cpu_s"] = round(cpu_s, 2)
    return out


def torch_cuda_baseline(tokens, durs, frames, dev, budget_s: float = 8.0):
    """BASELINE.md section 3 "the library kernels to beat on the same box": the same reference forward through
    stock PyTorch CUDA kernels (cuDNN / cuBLAS / cuFFT) on the B200 -- the oracle port's torch ops on cuda
    tensors, B = 1 loop (the reference's only mixed-length path), TF32 off (fp32 parity setting) and on
    (PyTorch's cuDNN default)."""
    from oracle import stylish_port as port
    from stylish_tts_b200 import fixture
    sds = {m: {k: v.to(dev) for k, v in sd.items()} for m, sd in
           fixture.fixture_nets(fixture.load_manifest(), seed=0).items()}
    g = torch.Generator().manual_seed(7)
    gc = torch.Generator().manual_seed(8)
    c1_tok, c1_dur, c1_f = fixture.synth_utterance(5.0, gc)
    keep = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    res = {}
    seq = _stratified(frames)

    def one(i_tok, i_dur, f):
        noise = {k: v.to(dev) for k, v in fixture.draw_noise(f, g).items()}
        o = port.synthesize(sds, i_tok[None].to(dev), torch.tensor([i_tok.numel()], device=dev), noise,
                            duration=i_dur.float().to(dev))
        return o["audio"]

    try:
        with torch.no_grad():
            for name, tf32 in (("tf32_off", False), ("tf32_on", True)):
                torch.backends.cudnn.allow_tf32 = tf32
                torch.backends.cuda.matmul.allow_tf32 = tf32
                for i in seq[:3]:                                   # warm-up: cuDNN heuristics, cuFFT plans
                    one(tokens[i], durs[i], frames[i])
                torch.cuda.synchronize(dev)
                t0 = time.perf_counter()
                audio_s, k = 0.0, 0
                while time.perf_counter() - t0 < budget_s:
                    i = seq[k % len(seq)]
                    k += 1
                    one(tokens[i], durs[i], frames[i]).sum().item()   # B = 1 loop with the result read back
                    audio_s += 300.0 * frames[i] / 24000.0
                dt = time.perf_counter() - t0
                lat = []
                for _ in range(12):
                    torch.cuda.synchronize(dev)
                    t1 = time.perf_counter()
                    one(c1_tok, c1_dur, c1_f).cpu()
                    lat.append((time.perf_counter() - t1) * 1e3)
                lat.sort()
                res[name] = {"value": audio_s / dt, "unit": UNIT, "utterances": k,
                             "c1_latency_p50_ms": lat[len(lat) // 2]}
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = keep
    res["what"] = ("oracle port (= the reference's torch ops) on cuda tensors, B=1 loop over the same batch, wall clock "
                   "incl. per-utterance D2H of the result; stock cuDNN/cuBLAS/cuFFT kernels")
    return res


def run_reference_arm(args, rank: int, world: int):
    """`--impl reference`: the reference's own CPU implementation of the path (oracle port; the
    reference checkout does not exist on the GPU box), all host threads, bounded sample per step."""
    if rank != 0:
        return
    tokens, durs, frames = workload(1)
    vals = []
    sample = ""
    # per-step sample sized so that the whole --steps K --warmup W run stays within ~2.5 minutes of CPU work
    budget = max(2.0, min(args.cpu_budget / 2, 150.0 / max(1, args.warmup + args.steps)))
    for s in range(args.warmup + args.steps):
        v, cores, sample = cpu_port_throughput(tokens, durs, frames, budget_s=budget, seed=s)
        if s >= args.warmup:
            vals.append(v)
    v = sum(vals) / len(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"BASELINE configs[1]: batch {BATCH} mixed 0.25-10 s utterances, B=1 loop (bounded sample)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


_JSON_FD = None


def claim_stdout():
    """stdout must carry exactly ONE JSON line, but libraries write there too (NCCL prints its version line to
    stdout whatever NCCL_DEBUG_FILE says): keep a private duplicate of fd 1 for the result and point fd 1 at
    stderr for everybody else."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    os.write(_JSON_FD if _JSON_FD is not None else 1, data)


def c1_latency(nets, dev, precision: str, iters: int = 40):
    """The metric's second half: p50 latency of ONE ~5 s utterance (BASELINE configs[0] shape: 120 phonemes,
    400 frames).  e2e = host call to host result: pinned tokens / durations in, CUDA-graph replay, waveform
    back to pinned host memory, stream synchronised.  device = CUDA events around the replay alone."""
    from stylish_tts_b200 import fixture
    from stylish_tts_b200.synth import Synthesizer
    syn = Synthesizer(nets, device=dev, precision=precision, use_graphs=True)
    g = torch.Generator().manual_seed(13)
    tok, dur, f = fixture.synth_tokens(120, g), fixture.synth_durations(120, 400, g), 400
    tok_h, dur_h = tok.pin_memory(), dur.to(torch.int32).pin_memory()
    host = torch.empty(300 * f, dtype=torch.float32).pin_memory()
    st = torch.cuda.current_stream(dev)
    for _ in range(4):
        syn.synthesize([tok_h], durations=[dur_h], frames=[f])
    syn.check()
    wall, devt = [], []
    for _ in range(iters):
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        out = syn.synthesize([tok_h], durations=[dur_h], frames=[f])
        e1.record()
        host.copy_(out["audio_flat"], non_blocking=True)
        st.synchronize()
        wall.append((time.perf_counter() - t0) * 1e3)
        devt.append(e0.elapsed_time(e1))
    wall.sort(); devt.sort()
    q = lambda v, p: v[min(len(v) - 1, int(p * len(v)))]
    return {"config": "C1: one 5 s utterance, 120 phonemes, 400 frames, CUDA graph replay", "precision": precision,
            "p50_ms": q(wall, 0.5), "p95_ms": q(wall, 0.95), "device_p50_ms": q(devt, 0.5), "iters": iters,
            "x_realtime_p50": 5000.0 / q(wall, 0.5),
            "what": "p50/p95: host call -> waveform in pinned host memory (H2D + replay + D2H + sync); device_p50: replay only"}


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="f16", choices=["f16", "fp32", "tf32x3"],
                    help="f16 = 16-bit mode (IEEE fp16 operand twins, fp32 accumulate)")
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the latency / modes / parity / torch-CUDA legs (kernel iteration runs)")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every step eagerly (no CUDA graph replay)")
    ap.add_argument("--lanes", type=int, default=3,
                    help="batches in flight per GPU (SynthesizerPool lanes: own context + CUDA graph + stream each)")
    ap.add_argument("--layers", default=None, help="write the per-layer GEMM table of the eager pass to this file")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else max(args.warmup, 0)

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference_arm(args, rank, world)
    if args.gpus > 1 and world == 1:
        raise SystemExit("launch multi-GPU runs with torch.distributed.run (one rank per GPU)")

    import torch.distributed as dist
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries exactly one JSON line
        dist.init_process_group("nccl", device_id=dev)

    from stylish_tts_b200 import fixture
    from stylish_tts_b200.config import ModelConfig
    from stylish_tts_b200.modules import build_model
    from stylish_tts_b200.shard import ShardPlan, WaveformGather
    from stylish_tts_b200.synth import SynthesizerPool

    sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
    nets = build_model(ModelConfig())
    for k, m in nets.items():
        m.load_state_dict(sds[k])
    pool = SynthesizerPool(nets, device=dev, precision=args.precision, lanes=max(1, args.lanes), use_graphs=False)
    syn = pool.lanes[0]
    eng = syn.engine

    # ---- one global batch, sharded by utterance (LPT on frames); this rank's shard
    g_tokens, g_durs, g_frames = workload(world)
    plan = ShardPlan(g_frames, world)
    mine = plan.local(rank)
    tokens, durs, frames = [g_tokens[i] for i in mine], [g_durs[i] for i in mine], [g_frames[i] for i in mine]
    audio_s = 300.0 * sum(g_frames) / 24000.0                     # whole job, per step
    tok_host = [t.pin_memory() for t in tokens]
    dur_host = [d.to(torch.int32).pin_memory() for d in durs]
    tok_dev = [t.to(dev) for t in tokens]
    dur_dev = [d.to(dev, torch.int32) for d in durs]
    gathers = [WaveformGather(plan, dev, dst=0) for _ in pool.lanes]
    host_elems = world * plan.cap if world > 1 else 300 * sum(frames)
    audio_host = [torch.empty(host_elems, dtype=torch.float32).pin_memory() for _ in pool.lanes] if rank == 0 else None
    flush_buf = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    use_graph = not args.no_graph

    # One step = the whole batch once.  Steps are submitted round-robin to the pool's lanes: batch i + 1 may
    # start while batch i is still running (independent batches; every one of the K steps runs to completion
    # inside the timed region, which ends with pool.drain() + a device synchronisation).
    def flush(lane):
        if flush_buf is not None:
            with torch.cuda.stream(pool.streams[lane]):
                flush_buf.fill_(1)

    def after_resident(lane, out):                # the final gather (world > 1): the path's only collective
        if world > 1:
            gathers[lane](out["audio_flat"])

    def after_e2e(lane, out):                     # gather, then the whole batch to pinned host memory on rank 0
        if world > 1:
            gathers[lane](out["audio_flat"])
            if rank == 0:
                audio_host[lane].copy_(gathers[lane].recv.view(-1), non_blocking=True)
        else:
            audio_host[lane].copy_(out["audio_flat"], non_blocking=True)

    def step_resident():
        flush(pool._next)
        return pool.submit(tok_dev, durations=dur_dev, frames=frames, graph=use_graph, after=after_resident)

    def step_eager():
        if flush_buf is not None:
            flush_buf.fill_(1)
        return syn.synthesize(tok_dev, durations=dur_dev, frames=frames, graph=False)

    def step_e2e():   # inputs from pinned host memory, waveform back to pinned host memory, every step
        flush(pool._next)
        return pool.submit(tok_host, durations=dur_host, frames=frames, graph=use_graph, after=after_e2e)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        pool.drain()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def warm(fn, n):
        for _ in range(n):
            fn()
        pool.synchronize()

    warm(step_resident, max(args.warmup, len(pool.lanes)))       # every lane warms up (and captures its graph)
    pool.check()
    ws_bytes = eng.workspace_bytes()              # of the timed workload's plan (later legs re-plan the context)
    # clocks / throttle reasons: nvidia-smi every 20 ms from two more untimed steps of the same load (the sampler
    # needs ~0.1 s to start; the timed region itself is ~0.1 s at the default K) through the timed region
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    warm(step_resident, 2 * len(pool.lanes))
    ms_total = timed(step_resident, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    # Per-GEMM CUDA events cannot be recorded inside a graph replay: the roofline pass times the same
    # K steps again, launched eagerly (same kernels, same inputs), with an event pair around every
    # conv-as-GEMM launch.  Kernel launches are counted on this eager pass (one step = the same
    # launches the graph replays).
    for _ in range(2):
        step_eager()
    ms_eager = timed(step_eager, args.steps)            # plain eager step (CPU launch rate shows for the small kernels)
    # Profiled pass: a spin kernel holds the stream for ~30 ms at the head of every step while the host enqueues the
    # whole step behind it, so the GPU then runs the launches back to back and an event pair brackets the KERNEL, not
    # the host's launch latency (a 10 us token-rate GEMM otherwise reads as ~27 us).
    spin_cycles = int(30e-3 * 1.9e9)

    def step_profiled():
        # the three token-rate chains normally run on three streams; here they share one, so that an event pair
        # times its kernel alone and not the kernels of the other two chains it shares the SMs with
        side = syn._side_streams
        cur = torch.cuda.current_stream(dev)
        syn._side_streams = (cur, cur)
        try:
            torch.cuda._sleep(spin_cycles)
            return step_eager()
        finally:
            syn._side_streams = side

    eng.set_profiling(True)
    l0 = eng.launch_count()
    timed(step_profiled, args.steps)
    launches = eng.launch_count() - l0
    prof = eng.profile()
    eng.set_profiling(False)

    warm(step_e2e, 2 * len(pool.lanes))
    ms_e2e = timed(step_e2e, args.steps)
    pool.synchronize()

    # ---- the other precision modes on the same workload (fp32-grade tf32x3, CUDA-core fp32)
    modes = {}
    if not args.no_extras:
        for mode, k in (("tf32x3", min(args.steps, 5)), ("fp32", min(args.steps, 3))):
            if mode == args.precision:
                continue
            for lane in pool.lanes:
                lane.set_precision(mode)
            warm(step_resident, len(pool.lanes) + 1)
            m_res = timed(step_resident, k)
            warm(step_e2e, len(pool.lanes))
            m_e2e = timed(step_e2e, k)
            pool.synchronize()
            modes[mode] = {"value": audio_s * k / (m_res / 1e3), "ms_per_step": m_res / k,
                           "e2e": audio_s * k / (m_e2e / 1e3), "steps": k, "unit": UNIT}
        for lane in pool.lanes:
            lane.set_precision(args.precision)

    tot_launch = torch.tensor([float(launches)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot_launch)
    value = audio_s * args.steps / (ms_total / 1e3)
    e2e_value = audio_s * args.steps / (ms_e2e / 1e3)

    if rank == 0:
        pk = peaks()
        n_tok = sum(int(t.numel()) for t in g_tokens)
        # ---- roofline: every GEMM-class launch of rank 0's step.  Engine backends: 0 = CUDA-core fp32, 1 = one-pass
        # fp16 twins, 2 = tf32x3, 3 = f16x3, 4 = one-pass fp16 islands (heads / prior convs / iSTFT), 5 = fused
        # WaveNet layer.  FLOPs are ALGORITHMIC (2 * valid rows * N * un-padded K; split modes count once).
        gem_ms = sum(p["ms"] for p in prof)
        gem_fl = sum(p["flops"] for p in prof)
        dft = [p for p in prof if p["name"].startswith(("stft", "istft"))]
        tokr = [p for p in prof if p["name"].startswith(("te.", "pr.", "se.", "pe.cross"))]
        top = max(prof, key=lambda p: p["ms"]) if prof else None
        if args.layers:
            with open(args.layers, "w") as f:
                f.write("| layer | backend | launches/step | ms/step | TFLOP/s (algorithmic) |\n|---|---|---|---|---|\n")
                for p in sorted(prof, key=lambda p: -p["ms"]):
                    f.write(f"| {p['name']} | {p['backend']} | {p['launches'] / args.steps:.0f} | {p['ms'] / args.steps:.4f} | "
                            f"{p['flops'] / max(p['ms'], 1e-9) / 1e9:.1f} |\n")
        peak_tf = pk["bf16_sustained"]
        achieved = (gem_fl / (gem_ms / 1e3) / 1e12) if gem_ms > 0 else 0.0
        my_audio_s = 300.0 * sum(frames) / 24000.0
        survey_flops = GFLOP_PER_AUDIO_S * 1e9 * audio_s                       # whole job
        exact_flops = (gem_fl - sum(p["flops"] for p in dft)) / args.steps + attention_flops(tokens, frames)   # rank 0
        tr = traffic_record()
        roofline = {
            "bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
            "frac": achieved / peak_tf if peak_tf else None,
            "kernel": ("all conv-as-GEMM launches of the step (tcgen05/TMA implicit-GEMM conv1d incl. one-pass islands "
                       "and fused WaveNet layers), CUDA events per launch on an eager pass whose launches are queued behind a spin kernel on ONE stream (GPU-side duration of each kernel alone)") if args.precision != "fp32"
                      else "conv_gemm_simt_kernel",
            "traffic": (tr["dram_bytes_read"] + tr["dram_bytes_write"]) if tr and args.precision == "f16" else None,
            "traffic_kernel": tr["kernel"] if tr else None,
            "traffic_source": tr["source"] if tr else None,
            "peak_source": pk["source"] + " bf16_tflops_sustained",
            # whole step against the same peak: SURVEY's algorithmic FLOPs / device time of the step
            "step_frac": survey_flops * args.steps / (ms_total / 1e3) / 1e12 / peak_tf / world,
            "step_tflops_algorithmic": survey_flops * args.steps / (ms_total / 1e3) / 1e12 / world,
            "step_flops_survey": survey_flops / world, "step_flops_exact_rank0": exact_flops,
            "dft_as_gemm_flops_rank0": sum(p["flops"] for p in dft) / args.steps,
            "gemm_ms_per_step": gem_ms / args.steps, "gemm_share_of_eager_step": gem_ms / ms_eager,
            "eager_ms_per_step": ms_eager / args.steps,
            "token_rate_gemms": {"ms_per_step": sum(p["ms"] for p in tokr) / args.steps,
                                 "tflops": (sum(p["flops"] for p in tokr) / max(sum(p["ms"] for p in tokr), 1e-9) / 1e9),
                                 "launches_per_step": sum(p["launches"] for p in tokr) / args.steps},
            "top_layer": {"name": top["name"], "ms_per_step": top["ms"] / args.steps,
                          "tflops": top["flops"] / (top["ms"] / 1e3) / 1e12} if top else None,
        }
        cpu = latency = parity = tcb = None
        if not args.no_extras:
            latency = c1_latency(nets, dev, args.precision)
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample = cpu_port_throughput(g_tokens, g_durs, g_frames, budget_s=args.cpu_budget)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
            if not args.no_extras:
                parity = parity_against_oracle(syn, g_tokens, g_durs, g_frames,
                                               [args.precision] + [m for m in ("tf32x3",) if m != args.precision])
                tcb = torch_cuda_baseline(g_tokens, g_durs, g_frames, dev)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": {"f16": "f16", "fp32": "f32", "tf32x3": "tf32x3"}[args.precision],
            "data": "synthetic",
            "config": {"workload": (f"one seeded batch of {BATCH} x {world} = {len(g_frames)} mixed-length utterances "
                                    f"0.25-10 s ({audio_s:.1f} audio-s, {n_tok} phonemes; "
                                    + ("BASELINE configs[1]" if world == 1 else
                                       ("BASELINE configs[3]" if world == 8 else "configs[3] shape at this N"))
                                    + "), default model.yml, fixture (seeded random-init) weights, synthetic durations"),
                       "precision": args.precision,
                       "launch": ("CUDA graph replay per step" if use_graph else "eager") +
                                 f", {len(pool.lanes)} batches in flight (round-robin lanes, own context + stream each)",
                       "parallelism": (f"utterances LPT-sharded by frame count over {world} GPU(s) "
                                       f"(partition balance {plan.balance():.4f}), no collective in the forward"
                                       + (", one NCCL gather of the waveforms to rank 0 per step" if world > 1 else "")),
                       "shard_audio_s": [round(300.0 * l / 24000.0, 1) for l in plan.loads],
                       "l2": ("256 MiB flush write on the lane's stream before every step; "
                              f"{len(pool.lanes)} batches in flight, each with a working set (GBs) >> the 126 MB L2")
                             if flush_buf is not None else "no flush; per-step working set (GBs) >> the 126 MB L2",
                       "parity_target": "audio_x0 (reference with the exact frame-0 STFT phase), oracle pitch injected",
                       "workspace_bytes": ws_bytes},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(sum(t.numel() * 8 for t in g_tokens) + sum(d.numel() * 4 for d in g_durs)),
                    "d2h_bytes_per_step": int(host_elems * 4),
                    "what": "pinned host tokens / durations in, synthesis, "
                            + ("NCCL gather to rank 0, " if world > 1 else "") + "whole batch's waveform to pinned host memory"},
            "gpu_launches": int(tot_launch.item()),
            "roofline": roofline,
            "cpu_baseline": cpu,
            "latency": latency,
            "modes": modes or None,
            "parity": parity,
            "torch_cuda_baseline": tcb,
        }
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
