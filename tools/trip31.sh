mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
python - > gpurun_out/pcm16.log 2>&1 <<'PY'
import torch, numpy as np
from stylish_tts_b200.engine import Engine
e = Engine(torch.device("cuda:0"))
g = torch.Generator().manual_seed(3)
for n in (0, 1, 7, 8, 9, 1000003):
    x = (torch.rand(n, generator=g) * 2.2 - 1.1)
    if n > 8:
        x[:8] = torch.tensor([0.0, -1.0, 0.99999994, -0.99999994, 1/32768, -1/32768, 1.0, 0.5/32768])
    y = e.pcm16(x.cuda()).cpu().numpy()
    ref = np.clip(np.trunc(x.numpy() * np.float32(32768.0)), -32768, 32767).astype(np.int16)
    assert (y == ref).all(), (n, np.flatnonzero(y != ref)[:5])
    z = e.pcm16(x.cuda()[1:]).cpu().numpy() if n > 1 else None      # unaligned input
    assert z is None or (z == ref[1:]).all()
print("pcm16 ok")
PY
echo "pcm16 rc=$?" >> gpurun_out/rc.txt
L=flow3.wn1
STTS_PROFILE_LAYER=$L timeout 600 ncu --set full --import-source on --clock-control none --profile-from-start off -c 1 -o /tmp/r1d_$L python tools/profile_step.py f16 > gpurun_out/ncu_$L.log 2>&1; echo "ncu_$L rc=$?" >> gpurun_out/rc.txt
ncu -i /tmp/r1d_$L.ncu-rep --page raw --csv > gpurun_out/r1d_$L.raw.csv 2>/dev/null
ncu -i /tmp/r1d_$L.ncu-rep --page details > gpurun_out/r1d_$L.details.txt 2>/dev/null
ncu -i /tmp/r1d_$L.ncu-rep --page source --csv > gpurun_out/r1d_$L.source.csv 2>/dev/null
cat gpurun_out/rc.txt; cat gpurun_out/pcm16.log | tail -3
