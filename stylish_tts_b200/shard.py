"""Multi-GPU = utterance sharding (BASELINE.json configs[3]): utterances are
independent, so each rank synthesises its own sub-batch with no collective inside the forward; the
only exchange is the final ragged gather of waveforms (north_star "only a final gather").

  plan = ShardPlan(frames, world)               LPT bin packing on frame counts, identical on every rank
  out, full = synthesize_sharded(syn, tokens, durations, frames)     one process per GPU, NCCL gather
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


def lpt_shards(frames: Sequence[int], n_ranks: int) -> List[List[int]]:
    """Longest-processing-time bin packing of utterances by frame count: cost of an utterance is
    linear in its frames (all frame-rate operators are convolutions / per-row maps)."""
    loads = [0] * n_ranks
    shards: List[List[int]] = [[] for _ in range(n_ranks)]
    for i in sorted(range(len(frames)), key=lambda j: (-frames[j], j)):
        r = min(range(n_ranks), key=lambda k: (loads[k], k))
        shards[r].append(i)
        loads[r] += frames[i]
    return [sorted(s) for s in shards]


class ShardPlan:
    """Which utterance runs where, and where its samples land in the gathered buffer.  Depends only on
    (frames, world), so every rank computes the same plan without communication."""

    def __init__(self, frames: Sequence[int], world: int):
        self.frames = [int(f) for f in frames]
        self.world = int(world)
        self.shards = lpt_shards(self.frames, self.world)
        self.sizes = [300 * sum(self.frames[i] for i in s) for s in self.shards]      # samples per rank
        self.cap = max(self.sizes) if self.sizes else 0                               # padded slot per rank
        self.loads = [sum(self.frames[i] for i in s) for s in self.shards]

    def local(self, rank: int) -> List[int]:
        return self.shards[rank]

    def balance(self) -> float:
        """sum(load) / (world * max(load)): the scaling efficiency the partition itself allows."""
        return sum(self.loads) / (self.world * max(self.loads)) if self.loads and max(self.loads) else 1.0

    def split(self, slots: Sequence[torch.Tensor]) -> List[torch.Tensor]:
        """slots[r] = rank r's padded slot -> waveforms in utterance order (views, no copy)."""
        result: List[Optional[torch.Tensor]] = [None] * len(self.frames)
        for r, s in enumerate(self.shards):
            off = 0
            for i in s:
                n = 300 * self.frames[i]
                result[i] = slots[r][off: off + n]
                off += n
        return result  # type: ignore


class WaveformGather:
    """The path's one collective: rank r's concatenated waveform -> slot r of a [world, cap] buffer on
    rank ``dst`` (``dist.gather``: ncclSend / ncclRecv over NVLink on GPUs, gloo on CPU), or on every
    rank (``dst=None``: all_gather).  Buffers are allocated once per plan; the call enqueues on the
    current stream and does not synchronise the host."""

    def __init__(self, plan: ShardPlan, device, dtype=torch.float32, dst: Optional[int] = 0, group=None):
        self.plan, self.dst, self.group = plan, dst, group
        self.rank = dist.get_rank(group) if plan.world > 1 else 0
        self.send = torch.zeros(plan.cap, dtype=dtype, device=device)
        self.recv = None
        if plan.world > 1 and (dst is None or self.rank == dst):
            self.recv = torch.zeros((plan.world, plan.cap), dtype=dtype, device=device)

    def __call__(self, local_audio: torch.Tensor) -> Optional[List[torch.Tensor]]:
        """local_audio: this rank's waveforms concatenated in ``plan.local(rank)`` order.  Returns the
        waveforms of the whole batch in utterance order on the destination rank(s), else None."""
        n = self.plan.sizes[self.rank]
        if local_audio.numel() != n:
            raise ValueError(f"rank {self.rank}: expected {n} samples for its shard, got {local_audio.numel()}")
        if self.plan.world == 1:
            return self.plan.split([local_audio])
        self.send[:n].copy_(local_audio, non_blocking=True)
        if self.dst is None:
            dist.all_gather_into_tensor(self.recv.view(-1), self.send, group=self.group)
        elif self.rank == self.dst:
            dist.gather(self.send, list(self.recv.unbind(0)), dst=self.dst, group=self.group)
        else:
            dist.gather(self.send, None, dst=self.dst, group=self.group)
        return self.plan.split(list(self.recv.unbind(0))) if self.recv is not None else None


def gather_waveforms(local_audio: torch.Tensor, local_ids: Sequence[int], frames: Sequence[int],
                     shards: Sequence[Sequence[int]], group=None) -> List[torch.Tensor]:
    """All ranks end up with the full list of waveforms in utterance order (one all_gather on buffers
    padded to the largest shard; sizes are known on every rank from ``frames``)."""
    world = dist.get_world_size(group)
    plan = ShardPlan(frames, world)
    assert [list(s) for s in shards] == plan.shards and list(local_ids) == plan.local(dist.get_rank(group))
    return WaveformGather(plan, local_audio.device, local_audio.dtype, dst=None, group=group)(local_audio)


def synthesize_sharded(syn, tokens: Sequence[torch.Tensor], durations: Sequence[torch.Tensor],
                       frames: Optional[Sequence[int]] = None, dst: Optional[int] = 0, group=None,
                       gather: Optional[WaveformGather] = None, **kw):
    """BASELINE.json configs[3]: one batch (every rank passes the SAME tokens / durations) sharded by
    utterance over the ranks of ``group``.  Each rank synthesises ``ShardPlan.local(rank)`` with
    ``syn`` (a Synthesizer, or anything with its ``synthesize`` signature) and the waveforms are
    gathered once.  Returns (this rank's result dict, waveforms of the whole batch in utterance
    order on rank ``dst`` -- on every rank if ``dst`` is None -- else None)."""
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank(group) if world > 1 else 0
    if frames is None:
        frames = [int(d.sum()) for d in durations]
    plan = gather.plan if gather is not None else ShardPlan(frames, world)
    mine = plan.local(rank)
    out = syn.synthesize([tokens[i] for i in mine], durations=[durations[i] for i in mine],
                         frames=[plan.frames[i] for i in mine], **kw)
    if gather is None:
        gather = WaveformGather(plan, out["audio_flat"].device, out["audio_flat"].dtype, dst=dst, group=group)
    return out, gather(out["audio_flat"])
