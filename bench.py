#!/usr/bin/env python
"""Headline benchmark: audio-seconds synthesized per wall-second (1/RTF) of the Stylish-TTS
synthesis forward path on B200 (BASELINE.json metric), workload = BASELINE configs[1]
(batch 32 mixed-length utterances, 0.25-10 s, random-init fixture weights, synthetic phonemes).

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
BATCH = 32
FLOP_PER_F4_FRAME = 74.07e6          # frame-rate conv + linear FLOPs per hop-75 frame
FLOP_PER_F_FRAME = 8.0e6             # pitch / energy towers per hop-300 frame
FLOP_PER_TOKEN = 35e6 * 3            # three text encoders + predictors, per token (approx., 8d)


def workload(rank: int, batch: int = BATCH):
    """BASELINE configs[1]: `batch` utterances, lengths uniform in [0.25, 10] s, seeded per rank."""
    from stylish_tts_b200 import fixture
    secs = fixture.mixed_lengths(batch, seed=rank)
    items = fixture.synth_batch(secs, seed=rank)
    tokens = [it[0] for it in items]
    durs = [it[1] for it in items]
    frames = [it[2] for it in items]
    return tokens, durs, frames


def algorithmic_flops(tokens, frames) -> float:
    return (sum(frames) * 4 * FLOP_PER_F4_FRAME + sum(frames) * FLOP_PER_F_FRAME
            + sum(int(t.numel()) for t in tokens) * FLOP_PER_TOKEN)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "bf16_burst": p["bf16_tflops"], "bf16_sustained": p["bf16_tflops_sustained"],
                "source": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


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
                                          "--format=csv,noheader,nounits", "-lms", "100"],
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
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_throughput(tokens, durs, frames, budget_s: float, seed: int = 0):
    """The reference's CPU path (oracle port = pinned restatement of the reference PyTorch forward,
    B = 1 loop exactly like stage_type.py:491-512) on a bounded sample of the same workload."""
    from oracle import stylish_port as port
    from stylish_tts_b200 import fixture
    sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
    torch.set_num_threads(os.cpu_count() or 1)
    order = sorted(range(len(frames)), key=lambda i: frames[i])
    pick = [order[len(order) // 2], order[len(order) // 4], order[(3 * len(order)) // 4], order[-1], order[0]]
    done_audio, t_total, used = 0.0, 0.0, []
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        # warm-up on the smallest utterance (thread pools, allocator)
        i0 = order[0]
        port.synthesize(sds, tokens[i0][None], torch.tensor([tokens[i0].numel()]), fixture.draw_noise(frames[i0], g),
                        duration=durs[i0].float())
        # utterances in a length-stratified order, then the rest of the batch, cycling until the
        # budget is spent (so the sample is ~budget_s of CPU work whatever the host)
        seq = pick + [i for i in range(len(frames)) if i not in pick]
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


def run_reference_arm(args, rank: int, world: int):
    """`--impl reference`: the reference's own CPU implementation of the path (oracle port; the
    reference checkout does not exist on the GPU box), all host threads, bounded sample per step."""
    if rank != 0:
        return
    tokens, durs, frames = workload(0)
    vals = []
    sample = ""
    for s in range(args.warmup + args.steps):
        v, cores, sample = cpu_port_throughput(tokens, durs, frames, budget_s=args.cpu_budget / 2, seed=s)
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


# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the largest GEMM of the step (the forward STFT,
# "stft16") from the `ncu --set full` capture summarised in profiles/r1_full_stft16.txt: 285.9 MB read + 205.7 MB
# written, against 272 MB (fp16 frame operand, hi + lo) + 215 MB (fp16 twins of |X| and phase) algorithmic
TRAFFIC_TOP_KERNEL = 285.899776e6 + 205.674496e6


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


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="f16", choices=["f16", "bf16", "fp32", "tf32x3"],
                    help="f16 = 16-bit mode (fp16 operand twins, fp32 accumulate; \"bf16\" is an alias)")
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    from stylish_tts_b200.synth import SynthesizerPool

    sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
    nets = build_model(ModelConfig())
    for k, m in nets.items():
        m.load_state_dict(sds[k])
    pool = SynthesizerPool(nets, device=dev, precision=args.precision, lanes=max(1, args.lanes), use_graphs=False)
    syn = pool.lanes[0]
    eng = syn.engine

    tokens, durs, frames = workload(rank)
    audio_s = 300.0 * sum(frames) / 24000.0
    tok_host = [t.pin_memory() for t in tokens]
    dur_host = [d.to(torch.int32).pin_memory() for d in durs]
    tok_dev = [t.to(dev) for t in tokens]
    dur_dev = [d.to(dev, torch.int32) for d in durs]
    audio_host = [torch.empty(300 * sum(frames), dtype=torch.float32).pin_memory() for _ in pool.lanes]
    flush_buf = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    use_graph = not args.no_graph

    # One step = one batch of the workload.  Steps are submitted round-robin to the pool's lanes: batch
    # i + 1 may start while batch i is still running (independent batches; every one of the K steps runs
    # to completion inside the timed region, which ends with pool.drain() + a device synchronisation).
    def flush(lane):
        if flush_buf is not None:
            with torch.cuda.stream(pool.streams[lane]):
                flush_buf.fill_(1)

    def step_resident():
        flush(pool._next)
        return pool.submit(tok_dev, durations=dur_dev, frames=frames, graph=use_graph)

    def step_eager():
        if flush_buf is not None:
            flush_buf.fill_(1)
        return syn.synthesize(tok_dev, durations=dur_dev, frames=frames, graph=False)

    def step_e2e():   # inputs from pinned host memory, waveform back to pinned host memory, every step
        lane = pool._next
        flush(lane)
        return pool.submit(tok_host, durations=dur_host, frames=frames, out_host=audio_host[lane], graph=use_graph)

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

    for _ in range(max(args.warmup, len(pool.lanes))):     # every lane warms up (and captures its graph)
        step_resident()
    pool.synchronize()
    pool.check()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_total = timed(step_resident, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    # Per-GEMM CUDA events cannot be recorded inside a graph replay: the roofline pass times the same
    # K steps again, launched eagerly (same kernels, same inputs), with an event pair around every
    # conv-as-GEMM launch.  Kernel launches are counted on this eager pass (one step = the same
    # launches the graph replays).
    for _ in range(2):
        step_eager()
    eng.set_profiling(True)
    l0 = eng.launch_count()
    ms_eager = timed(step_eager, args.steps)
    launches = eng.launch_count() - l0
    prof = eng.profile()
    eng.set_profiling(False)

    for _ in range(2 * len(pool.lanes)):
        step_e2e()
    pool.synchronize()
    ms_e2e = timed(step_e2e, args.steps)
    pool.synchronize()

    tot_audio = torch.tensor([audio_s], device=dev, dtype=torch.float64)
    tot_launch = torch.tensor([float(launches)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot_audio)
        dist.all_reduce(tot_launch)
    value = float(tot_audio.item()) * args.steps / (ms_total / 1e3)
    e2e_value = float(tot_audio.item()) * args.steps / (ms_e2e / 1e3)

    if rank == 0:
        pk = peaks()
        # every tcgen05 conv-as-GEMM launch of the timed region (bf16, bf16x3 and tf32x3 backends; the
        # split modes count their ALGORITHMIC FLOPs once, not per pass); fp32 mode = the CUDA-core kernel
        tc = [p for p in prof if (p["backend"] in (1, 2, 3) if args.precision != "fp32" else p["backend"] == 0)]
        tc_ms = sum(p["ms"] for p in tc)
        tc_fl = sum(p["flops"] for p in tc)
        top = max(prof, key=lambda p: p["ms"]) if prof else None
        if args.layers:
            with open(args.layers, "w") as f:
                f.write("| layer | backend | ms/step | TFLOP/s (algorithmic) |\n|---|---|---|---|\n")
                for p in sorted(prof, key=lambda p: -p["ms"]):
                    f.write(f"| {p['name']} | {p['backend']} | {p['ms'] / args.steps:.4f} | "
                            f"{p['flops'] / max(p['ms'], 1e-9) / 1e9:.1f} |\n")
        peak_tf = pk["bf16_sustained"]
        achieved = (tc_fl / (tc_ms / 1e3) / 1e12) if tc_ms > 0 else 0.0
        roofline = {
            "bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
            "frac": achieved / peak_tf if peak_tf else None,
            # dram__bytes_read.sum + dram__bytes_write.sum of the largest GEMM launch of the step (bytes per launch)
            "traffic": TRAFFIC_TOP_KERNEL if args.precision in ("f16", "bf16") else None,
            "traffic_kernel": "stft16, the largest launch (algorithmic 272 MB in + 215 MB out), profiles/r1_full_stft16.txt",
            "kernel": "conv_gemm_tc_kernel (tcgen05/TMA implicit-GEMM conv1d), all launches of the timed region"
                      if args.precision != "fp32" else "conv_gemm_simt_kernel",
            "peak_source": pk["source"] + " bf16_tflops_sustained",
            "gemm_ms_per_step": sum(p["ms"] for p in prof) / args.steps,
            "gemm_share_of_step": sum(p["ms"] for p in prof) / ms_eager,
            "eager_ms_per_step": ms_eager / args.steps,
            "top_layer": {"name": top["name"], "ms_per_step": top["ms"] / args.steps,
                          "tflops": top["flops"] / (top["ms"] / 1e3) / 1e12} if top else None,
            "step_tflops_algorithmic": algorithmic_flops(tokens, frames) * args.steps / (ms_total / 1e3) / 1e12,
        }
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample = cpu_port_throughput(tokens, durs, frames, budget_s=args.cpu_budget)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": {"f16": "f16", "bf16": "f16", "fp32": "f32", "tf32x3": "tf32x3"}[args.precision],
            "data": "synthetic",
            "config": {"workload": f"BASELINE configs[1]: batch {BATCH} mixed-length utterances 0.25-10 s per GPU "
                                   f"({audio_s:.1f} audio-s, {sum(int(t.numel()) for t in tokens)} phonemes), default model.yml, "
                                   "fixture (seeded random-init) weights, synthetic durations",
                       "precision": args.precision, "launch": ("CUDA graph replay per step" if use_graph else "eager") +
                                 f", {len(pool.lanes)} batches in flight (round-robin lanes, own context + stream each)", "parallelism": f"utterance-sharded x{world}, no collective in the forward",
                       "l2": "256 MiB flush write before every step; per-step working set >> 126 MB L2"
                             if flush_buf is not None else "no flush",
                       "workspace_bytes": eng.workspace_bytes()},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(sum(t.numel() * 8 for t in tokens) + sum(d.numel() * 4 for d in durs)),
                    "d2h_bytes_per_step": int(audio_host[0].numel() * 4)},
            "gpu_launches": int(tot_launch.item()),
            "roofline": roofline,
            "cpu_baseline": cpu,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
