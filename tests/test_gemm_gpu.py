"""GPU parity of the conv-as-GEMM kernels (CUDA-core fp32, tcgen05 bf16, tcgen05 tf32x3) through
the C ABI (stts_test_conv) against torch conv1d in float64 on the same packed ragged input."""
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = [  # (lens, cin, taps, pad, n_out)
    ([200, 77, 130], 130, 3, 1, 512),     # decoder encode conv1 shape, ragged + guard-only tile
    ([300], 578, 3, 1, 512),              # decoder block conv1
    ([128, 256], 128, 5, 2, 256),         # WN in_layer; segment ends exactly on a tile boundary
    ([90, 90], 512, 1, 0, 1536),          # pwconv1
    ([150], 768, 7, 3, 1025),             # output head: N not a multiple of 128
    ([64, 33], 64, 1, 0, 128),            # flow pre
]


def _reference(x_list, w, bias, pad):
    outs = []
    for x in x_list:  # x: [L, cin]
        y = torch.nn.functional.conv1d(x.t()[None].double(), w.double(), bias.double(), padding=pad)
        outs.append(y[0].t())
    return outs


def _run(backend, case, seed=0):
    from stylish_tts_b200.engine import Engine
    lens, cin, taps, pad, n_out = case
    g = torch.Generator().manual_seed(seed)
    eng = Engine("cuda:0")
    eng.set_layout(None, [max(l, 14) for l in lens])
    rows, starts = eng.layout_info(4)
    ldx = (cin + 63) // 64 * 64
    xs = [torch.randn(l, cin, generator=g) for l in lens]
    w = torch.randn(n_out, cin, taps, generator=g) / (cin * taps) ** 0.5
    bias = torch.randn(n_out, generator=g)
    x = torch.zeros(rows, ldx)
    for s, xi in zip(starts, xs):
        x[s: s + xi.shape[0], :cin] = xi
    ldy = (n_out + 63) // 64 * 64
    y = eng.test_conv(backend, x.cuda(), cin, w, bias, pad, ldy).cpu()
    ref = _reference([x[s: s + max(l, 14), :cin] for s, l in zip(starts, lens)], w, bias, pad)
    err, scale = 0.0, 0.0
    for s, l, r in zip(starts, lens, ref):
        L = max(l, 14)
        err = max(err, float((y[s: s + L, :n_out].double() - r).abs().max()))
        scale = max(scale, float(r.abs().max()))
    # guard rows and pad columns must stay untouched (zero)
    mask = torch.ones(rows, dtype=torch.bool)
    for s, l in zip(starts, lens):
        mask[s: s + max(l, 14)] = False
    assert float(y[mask].abs().max()) == 0.0, "kernel wrote a guard row"
    if ldy > n_out:
        assert float(y[:, n_out:].abs().max()) == 0.0, "kernel wrote a pad column"
    return err / scale


@pytest.mark.parametrize("case", CASES)
def test_simt_fp32(case):
    assert _run("fp32", case) <= 2e-6


@pytest.mark.parametrize("case", CASES)
def test_tc_tf32x3(case):
    assert _run("tf32x3", case) <= 5e-5   # fp32 accumulation inside the tensor core grows with K


@pytest.mark.parametrize("case", CASES)
def test_tc_f16(case):
    assert _run("f16", case) <= 2e-3      # one pass, fp16 operands (2^-12 rounding), fp32 accumulate
