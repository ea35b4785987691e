"""C1 latency only (bench.py's `latency` leg) for quick A/B runs: python tools/latency_c1.py [precision]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import c1_latency  # noqa: E402
from stylish_tts_b200 import fixture  # noqa: E402
from stylish_tts_b200.config import ModelConfig  # noqa: E402
from stylish_tts_b200.modules import build_model  # noqa: E402

dev = torch.device("cuda:0")
sds = fixture.fixture_nets(fixture.load_manifest(), seed=0)
nets = build_model(ModelConfig())
for k, m in nets.items():
    m.load_state_dict(sds[k])
    m.eval()
print(json.dumps(c1_latency(nets, dev, sys.argv[1] if len(sys.argv) > 1 else "f16", iters=int(os.environ.get("ITERS", "60")))))
