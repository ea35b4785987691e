"""GPU: stts_frame_index_static, the allocation-free variant of stts_frame_index (its frame-offset scratch is a
file-scope `static __device__` array inside the library, csrc/length_reg.cu), called through the C-ABI and compared
with the existing entry point.  Integer work: the bar is bit equality.  Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _engine():
    from stylish_tts_b200.modules.frame_rate import default_engine
    return default_engine(torch.device(DEV))


def _batch(sizes, seed):
    gen = torch.Generator().manual_seed(seed)
    durs = [torch.randint(1, 47, (n,), generator=gen, dtype=torch.int32) for n in sizes]
    cu_tok = torch.tensor([0] + torch.cumsum(torch.tensor(sizes), 0).tolist(), dtype=torch.int32)
    total = int(sum(int(d.sum()) for d in durs))
    return torch.cat(durs).to(DEV), cu_tok.to(DEV), total


def test_static_scratch_frame_index_equals_frame_index_twice_in_a_row():
    eng = _engine()
    # second call is SMALLER than the first: it must not see the first call's frame offsets in the shared array
    calls = [_batch((3, 510, 1, 77, 129, 46, 255, 12), 11), _batch((64, 5, 300), 12)]
    # outputs are the caller's, allocated up front; the entry point itself allocates nothing
    outs = [(torch.full((tot,), -1, dtype=torch.int32, device=DEV), torch.zeros(1, dtype=torch.int32, device=DEV))
            for _, _, tot in calls]
    import gc
    gc.collect()                                          # tensors of earlier tests must not be freed mid-measurement
    torch.cuda.synchronize()
    for k, ((dur, cu_tok, tot), (idx, status)) in enumerate(zip(calls, outs)):
        reserved, allocated = torch.cuda.memory_reserved(), torch.cuda.memory_allocated()
        free_before = torch.cuda.mem_get_info()[0]
        eng.frame_index_static(dur, cu_tok, idx, status)
        torch.cuda.synchronize()
        # nothing allocated through torch (a decrease would only mean the garbage collector ran in between) ...
        assert torch.cuda.memory_reserved() <= reserved and torch.cuda.memory_allocated() <= allocated
        if k > 0:       # ... nor behind its back (the first call may load the module lazily; later ones add nothing)
            assert torch.cuda.mem_get_info()[0] >= free_before
        _, want_idx, want_status = eng.frame_index(dur, cu_tok, tot)
        assert int(status.item()) == int(want_status.item()) == 0
        assert torch.equal(idx, want_idx), k
        want = torch.repeat_interleave(torch.cat([torch.arange(int(b - a)) for a, b in
                                                  zip(cu_tok.cpu()[:-1], cu_tok.cpu()[1:])]), dur.cpu().long())
        assert torch.equal(idx.cpu(), want.to(torch.int32))


def test_static_scratch_frame_index_overflow_status_and_limit():
    from stylish_tts_b200.engine import SttsError
    eng = _engine()
    dur, cu_tok, tot = _batch((9, 200, 31), 13)
    idx = torch.zeros(tot - 1, dtype=torch.int32, device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    eng.frame_index_static(dur, cu_tok, idx, status)
    _, want_idx, want_status = eng.frame_index(dur, cu_tok, tot - 1)
    assert int(status.item()) == int(want_status.item()) == 1
    assert torch.equal(idx, want_idx)
    # beyond the static array's capacity: a clear error, nothing launched
    n = 1025
    dur = torch.ones(n, dtype=torch.int32, device=DEV)
    cu_tok = torch.arange(n + 1, dtype=torch.int32, device=DEV)
    idx = torch.zeros(n, dtype=torch.int32, device=DEV)
    status.zero_()
    with pytest.raises(SttsError, match="1024"):
        eng.frame_index_static(dur, cu_tok, idx, status)
    # the largest legal call
    n = 1024
    eng.frame_index_static(dur[:n].contiguous(), cu_tok[: n + 1].contiguous(), idx[:n], status)
    assert int(status.item()) == 0 and torch.equal(idx[:n].cpu(), torch.zeros(n, dtype=torch.int32))
