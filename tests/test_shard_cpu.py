"""CPU, world_size 2 over gloo: utterance sharding + final ragged gather (the N > 1 host path)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stylish_tts_b200.shard import gather_waveforms, lpt_shards

FRAMES = [37, 400, 20, 123, 800, 55, 56, 310, 91, 640, 23, 77]


def _fake_audio(i, f):
    return torch.arange(300 * f, dtype=torch.float32) * 1e-6 + i


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shards = lpt_shards(FRAMES, world)
    mine = shards[rank]
    local = torch.cat([_fake_audio(i, FRAMES[i]) for i in mine])
    full = gather_waveforms(local, mine, FRAMES, shards)
    ok = all(torch.equal(full[i], _fake_audio(i, FRAMES[i])) for i in range(len(FRAMES)))
    q.put((rank, ok))
    dist.destroy_process_group()


def test_lpt_is_a_balanced_partition():
    for n in (1, 2, 4, 8):
        sh = lpt_shards(FRAMES, n)
        assert sorted(i for s in sh for i in s) == list(range(len(FRAMES)))
        loads = [sum(FRAMES[i] for i in s) for s in sh]
        assert max(loads) - min(loads) <= max(FRAMES)
        assert lpt_shards(FRAMES, n) == sh            # deterministic on every rank


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
