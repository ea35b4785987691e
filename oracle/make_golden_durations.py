"""Golden vectors for DurationProcessor.prediction_to_duration (train/utils.py:468-474) from the REAL
reference, at the size SURVEY 7.3 item 6 asks for.  Run here: ``python oracle/make_golden_durations.py``.
TEST INFRASTRUCTURE ONLY.  Writes:
  tests/golden/durations_1m.safetensors   2^20 seeded random logit rows -> reference durations (uint8); the rows
                                          themselves are regenerated from the seed by the tests
                                          (oracle.durations_cases.random_logits), plus the indices of the rows
                                          whose exact (fp64) soft duration lies within 1e-5 of a rounding tie
  tests/golden/durations_adv.safetensors  adversarial rows: exact x.5 ties (round-half-even), x.5 -+ k ulp,
                                          saturated / flat / all-equal rows, every (argmax < 7) class
"""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import stylish_port as port  # noqa: E402
from oracle.durations_cases import adversarial_logits, random_logits, tie_distance  # noqa: E402
from oracle.ref_import import import_reference  # noqa: E402
from safetensors.torch import save_file  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    torch.set_num_threads(8)
    build_model, load_cfg, DurationProcessor, yml = import_reference()
    mc = load_cfg(open(yml))
    proc = DurationProcessor(mc.duration_predictor.duration_classes, mc.duration_predictor.max_duration)
    with torch.no_grad():
        logits = random_logits()
        d = proc.prediction_to_duration(logits, None)
        assert torch.equal(d, port.prediction_to_duration(logits)), "port != reference on the 1M rows"
        assert float(d.min()) >= 1 and float(d.max()) <= 46
        dist, soft_path = tie_distance(logits)
        near = torch.nonzero((dist < 1e-5) & soft_path).flatten().to(torch.int32)
        save_file({"duration": d.to(torch.uint8), "near_tie_rows": near, "seed": torch.tensor([77])},
                  os.path.join(GOLD, "durations_1m.safetensors"))
        print(f"durations_1m: {logits.shape[0]} rows, {near.numel()} within 1e-5 of a tie, "
              f"soft path share {float(soft_path.float().mean()):.3f}, max {int(d.max())}")
        adv = adversarial_logits()
        da = proc.prediction_to_duration(adv, None)
        assert torch.equal(da, port.prediction_to_duration(adv)), "port != reference on the adversarial rows"
        save_file({"logits": adv, "duration": da.to(torch.int32)}, os.path.join(GOLD, "durations_adv.safetensors"))
        print(f"durations_adv: {adv.shape[0]} rows, histogram {torch.bincount(da.to(torch.int64)).tolist()}")


if __name__ == "__main__":
    main()
