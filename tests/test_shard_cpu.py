"""CPU, world_size 2 over gloo: utterance sharding + final ragged gather (the N > 1 host path of
BASELINE configs[3]; the same code runs over NCCL in tests/test_shard_gpu.py and bench.py)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stylish_tts_b200.shard import ShardPlan, WaveformGather, gather_waveforms, lpt_shards, synthesize_sharded

FRAMES = [37, 400, 20, 123, 800, 55, 56, 310, 91, 640, 23, 77]


def _fake_audio(i, f):
    return torch.arange(300 * f, dtype=torch.float32) * 1e-6 + i


class _FakeSynth:
    """Stands in for Synthesizer.synthesize on CPU: utterance i -> a waveform that encodes (i, length)."""

    def synthesize(self, tokens, durations=None, frames=None, **kw):
        audio = [_fake_audio(int(t[0]), f) for t, f in zip(tokens, frames)]
        return {"audio": audio, "audio_flat": torch.cat(audio), "frames": list(frames)}


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shards = lpt_shards(FRAMES, world)
    mine = shards[rank]
    local = torch.cat([_fake_audio(i, FRAMES[i]) for i in mine])
    full = gather_waveforms(local, mine, FRAMES, shards)
    ok = all(torch.equal(full[i], _fake_audio(i, FRAMES[i])) for i in range(len(FRAMES)))
    # gather to rank 0 only (the product path: synthesize_sharded)
    tokens = [torch.tensor([i, 0, 0]) for i in range(len(FRAMES))]
    durs = [torch.tensor([f]) for f in FRAMES]
    out, got = synthesize_sharded(_FakeSynth(), tokens, durs, FRAMES, dst=0)
    if rank == 0:
        ok = ok and all(torch.equal(got[i], _fake_audio(i, FRAMES[i])) for i in range(len(FRAMES)))
    else:
        ok = ok and got is None
    ok = ok and out["frames"] == [FRAMES[i] for i in mine]
    # a reused gather object (bench.py: one per lane) gives the same result twice
    g = WaveformGather(ShardPlan(FRAMES, world), "cpu", dst=0)
    for _ in range(2):
        _, got = synthesize_sharded(_FakeSynth(), tokens, durs, FRAMES, gather=g)
        if rank == 0:
            ok = ok and all(torch.equal(got[i], _fake_audio(i, FRAMES[i])) for i in range(len(FRAMES)))
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_lpt_is_a_balanced_partition():
    for n in (1, 2, 4, 8):
        sh = lpt_shards(FRAMES, n)
        assert sorted(i for s in sh for i in s) == list(range(len(FRAMES)))
        loads = [sum(FRAMES[i] for i in s) for s in sh]
        assert max(loads) - min(loads) <= max(FRAMES)
        assert lpt_shards(FRAMES, n) == sh            # deterministic on every rank
        plan = ShardPlan(FRAMES, n)
        assert plan.shards == sh and plan.cap == 300 * max(loads)
        assert abs(plan.balance() - sum(loads) / (n * max(loads))) < 1e-12


def test_bench_workload_is_balanced_at_every_n():
    """The scaling bench shards ONE seeded batch of 32 x N mixed-length utterances: the partition itself
    must not cost more than 3 % (round 1 lost 10 % to per-rank random batches)."""
    from bench import workload
    for n in (1, 2, 4, 8):
        _, _, frames = workload(n)
        assert len(frames) == 32 * n
        assert ShardPlan(frames, n).balance() >= 0.97, (n, ShardPlan(frames, n).balance())


def test_single_rank_plan_needs_no_process_group():
    plan = ShardPlan(FRAMES, 1)
    local = torch.cat([_fake_audio(i, f) for i, f in enumerate(FRAMES)])
    full = WaveformGather(plan, "cpu")(local)
    assert all(torch.equal(full[i], _fake_audio(i, FRAMES[i])) for i in range(len(FRAMES)))


def test_gather_two_ranks_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]
