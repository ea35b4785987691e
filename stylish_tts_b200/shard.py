"""Multi-GPU = utterance sharding: utterances are independent, so each
rank synthesises its own sub-batch with no collective inside the forward; the only exchange is
the final ragged gather of waveforms (BASELINE.json north_star "only a final gather")."""
from __future__ import annotations

from typing import List, Sequence

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


def gather_waveforms(local_audio: torch.Tensor, local_ids: Sequence[int], frames: Sequence[int],
                     shards: Sequence[Sequence[int]], group=None) -> List[torch.Tensor]:
    """All ranks end up with the full list of waveforms in utterance order.

    local_audio: the rank's waveforms concatenated in ``local_ids`` order (300 samples per frame).
    One all_gather on buffers padded to the largest shard (NCCL on GPUs, gloo in the CPU tests);
    sizes are known on every rank from ``frames``, so no size exchange is needed."""
    world = dist.get_world_size(group)
    sizes = [300 * sum(frames[i] for i in s) for s in shards]
    cap = max(sizes)
    buf = torch.zeros(cap, dtype=local_audio.dtype, device=local_audio.device)
    buf[: local_audio.numel()] = local_audio
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    result: List[torch.Tensor] = [None] * len(frames)  # type: ignore
    for r, s in enumerate(shards):
        off = 0
        for i in s:
            n = 300 * frames[i]
            result[i] = out[r][off: off + n]
            off += n
    return result
