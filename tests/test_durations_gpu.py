"""GPU: length regulation (stts_durations / stts_frame_index) against golden vectors produced by the REAL
reference DurationProcessor (oracle/make_golden_durations.py, oracle/make_golden.py).  Integer work: bit-exact.

The one implementation-defined spot: `round(sum softmax * table)` when the exact value lies
within a few fp32 ulps of x.5 -- there the reference's own result depends on the last ulp of its libm's expf
(its CPU and CUDA builds disagree on the same rows).  The 2^20-row case therefore allows a mismatch ONLY on the
rows whose fp64-exact soft duration is within 1e-5 of a tie (11 rows of 2^20; 2 of them differ with a correctly
rounded exp), and the adversarial file covers the ties themselves with rows whose fp32 result is fixed by IEEE
arithmetic alone (exact x.5, x.5 -+ k ulp)."""
import json
import os

import pytest
import torch

from tests.util import ROOT, gold

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _engine():
    from stylish_tts_b200.modules.frame_rate import default_engine
    return default_engine(torch.device(DEV))


def test_golden_8192_rows_bit_exact():
    g = gold("durations")
    d = _engine().durations(g["logits"].to(DEV))
    assert torch.equal(d.cpu(), g["duration"])


def test_adversarial_tie_rows_bit_exact():
    g = gold("durations_adv")
    d = _engine().durations(g["logits"].to(DEV))
    bad = torch.nonzero(d.cpu() != g["duration"]).flatten().tolist()
    assert not bad, (bad[:8], g["logits"][bad[:2]])


def test_million_rows_against_reference():
    from oracle.durations_cases import random_logits
    g = gold("durations_1m")
    logits = random_logits(seed=int(g["seed"]))
    assert logits.shape[0] == g["duration"].numel() == 1 << 20
    d = _engine().durations(logits.to(DEV)).cpu()
    bad = torch.nonzero(d != g["duration"].to(torch.int32)).flatten()
    near = set(g["near_tie_rows"].tolist())
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "durations_1m.json"), "w") as f:
        json.dump({"rows": 1 << 20, "mismatch_rows": bad.tolist(), "near_tie_rows": sorted(near)}, f)
    assert all(int(i) in near for i in bad), [int(i) for i in bad if int(i) not in near][:8]
    assert bad.numel() <= len(near)
    assert int((d - g["duration"].to(torch.int32)).abs().max()) <= 1          # a tie moves the result by one frame


@pytest.mark.parametrize("case", ["all1", "all46", "max510", "ragged"])
def test_frame_index_edges_equal_repeat_interleave(case):
    """duration_to_alignment (utils.py:476-489) as indices: all-1 and all-46 durations, the longest legal
    utterance (510 tokens -> 4 800+ frames), a ragged batch; plus the overflow status when the cap is short."""
    eng = _engine()
    gen = torch.Generator().manual_seed(5)
    if case == "all1":
        durs = [torch.ones(37, dtype=torch.int32)]
    elif case == "all46":
        durs = [torch.full((29,), 46, dtype=torch.int32)]
    elif case == "max510":
        durs = [torch.randint(1, 47, (510,), generator=gen, dtype=torch.int32)]
    else:
        durs = [torch.randint(1, 47, (n,), generator=gen, dtype=torch.int32) for n in (3, 510, 1, 77, 129)]
    cu_tok = torch.tensor([0] + torch.cumsum(torch.tensor([d.numel() for d in durs]), 0).tolist(), dtype=torch.int32)
    total = int(sum(int(d.sum()) for d in durs))
    cu_f, idx, status = eng.frame_index(torch.cat(durs).to(DEV), cu_tok.to(DEV), total)
    want = torch.cat([torch.repeat_interleave(torch.arange(d.numel()), d.long()) for d in durs]).to(torch.int32)
    assert int(status.item()) == 0
    assert torch.equal(idx.cpu(), want)
    assert cu_f.cpu().tolist() == [0] + torch.cumsum(torch.tensor([int(d.sum()) for d in durs]), 0).tolist()
    _, _, status = eng.frame_index(torch.cat(durs).to(DEV), cu_tok.to(DEV), total - 1)
    assert int(status.item()) == 1


def test_mismatched_durations_raise_like_the_reference_shape_error():
    """host-supplied frame counts that disagree with the durations must not be silently truncated."""
    from tests.util import synthesizer
    syn = synthesizer("fp32")
    g = gold("tiny")
    tokens, dur = g["tokens"], g["duration_in"].to(torch.int64)
    syn.synthesize([tokens], durations=[dur], frames=[int(dur.sum()) + 3], graph=False)
    with pytest.raises(ValueError):
        syn.check()
    with pytest.raises(IndexError):
        syn.synthesize([torch.tensor([0, 5, 999, 0])])
