"""GPU: SURVEY.md section 8 (f) rank 2 through the C-ABI -- stts_source_sinegen (SineGenerator,
cfm/cfm_mel_decoder.py:16-101) and stts_generator_block_snake (Snake AdaptiveGeneratorBlock, ada_norm.py:11-127)
against the golden vectors of the REAL reference classes and, at sizes the goldens do not reach, against the CPU oracle
(oracle/variants_port.py, pinned bit-exactly to those classes) run on the box.

Bars: the SineGenerator phase (two wrap-corrected fp64 cumulative sums, rounded to fp32 per sample) is integer-like
bookkeeping and must be BIT-EQUAL; its output is fp32 sin / dot / tanh: <= 1e-6.  The block is held to the waveform
tolerances of north_star relative to the block's output range: fp32 2e-5, tf32x3 1e-4, 16-bit mode 5e-3 (relative to
max |reference|)."""
import json
import os

import pytest
import torch

from oracle import variants_cases as vc
from tests.util import ROOT, gold

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SNAKE_REL = {"fp32": 2e-5, "tf32x3": 1e-4, "f16": 5e-3}


def _log(name, val):
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "variants.jsonl"), "a") as f:
        f.write(json.dumps({name: val}) + "\n")


def _sinegen_module(case):
    from stylish_tts_b200.modules import SineGenerator
    sg = SineGenerator(vc.SAMP_RATE, harmonic_num=vc.HARMONICS).to(DEV)
    with torch.no_grad():
        sg.merge[0].weight.copy_(case["merge_weight"])
    return sg


def test_sinegen_golden_phase_bit_equal_and_output():
    g = gold("variants")
    case = vc.sinegen_case(vc.SINEGEN_GOLDEN["lengths"], vc.SINEGEN_GOLDEN["seed"])
    sg = _sinegen_module(case)
    outs, phase = sg.forward_ragged(case["f0"], torch.stack(case["rand_ini"]), case["noise"], want_phase=True)
    for b in range(len(outs)):
        assert torch.equal(phase[b][::16].cpu(), g[f"sinegen.{b}.phase_s16"]), b
        sines = torch.sin(phase[b][::16].cpu() * 2 * torch.pi)
        assert float((sines - g[f"sinegen.{b}.sines_s16"]).abs().max()) <= 2.5e-7
        err = float((outs[b].cpu() - g[f"sinegen.{b}.out"]).abs().max())
        _log(f"sinegen.golden.{b}.out_max_abs", err)
        assert err <= 1e-6, err
    # the reference's own call shape: [B, L, 1] -> [B, L, 1]
    f0 = case["f0"][1][None, :, None].to(DEV)
    one = sg(f0, case["rand_ini"][1][None], case["noise"][1][None])
    assert one.shape == (1, f0.shape[1], 1)
    assert float((one.reshape(-1).cpu() - g["sinegen.1.out"]).abs().max()) <= 1e-6


def test_sinegen_ragged_long_batch_against_oracle():
    """60 s (1.44 M samples, 704 scan chunks), chunk-boundary lengths, a 7-sample utterance, an all-unvoiced one."""
    from oracle import variants_port as vp
    lengths = (1_440_000, 2048, 2049, 7, 100_001, 4096)
    case = vc.sinegen_case(lengths, 21)
    case["f0"][5] = torch.zeros(4096)                                 # all unvoiced: noise only
    sg = _sinegen_module(case)
    outs, phase = sg.forward_ragged(case["f0"], torch.stack(case["rand_ini"]), case["noise"], want_phase=True)
    torch.cuda.synchronize()
    worst = 0.0
    for b, n in enumerate(lengths):
        want = vp.sinegen(case["f0"][b].view(1, -1, 1), case["rand_ini"][b].view(1, -1), case["noise"][b][None],
                          case["merge_weight"], vc.SAMP_RATE, vc.HARMONICS)
        assert torch.equal(phase[b].cpu(), want["c2"][0]), (b, n)
        err = float((outs[b].cpu() - want["out"].reshape(-1)).abs().max())
        worst = max(worst, err)
        assert err <= 1e-6, (b, n, err)
    _log("sinegen.long.out_max_abs", worst)
    # second call on the same context (scratch reused) gives the same bits
    outs2 = sg.forward_ragged(case["f0"], torch.stack(case["rand_ini"]), case["noise"])
    assert all(torch.equal(a, b) for a, b in zip(outs, outs2))


def _block(channels, k, dil, seed, legacy=False):
    from stylish_tts_b200.modules import AdaptiveGeneratorBlock
    blk = AdaptiveGeneratorBlock(channels, kernel_size=k, dilation=dil, style_dim=vc.STYLE_DIM)
    if legacy:
        from stylish_tts_b200.modules.variants import variant_engine
        blk.to(DEV)
        eng = variant_engine(torch.device(DEV))
        eng.load_state(blk._name, vc.snake_state_dict(channels, k, seed, legacy=True))
        object.__setattr__(blk, "_loaded", eng)
        return blk
    blk.load_state_dict(vc.snake_state_dict(channels, k, seed))
    return blk.to(DEV)


@pytest.mark.parametrize("mode", ["fp32", "tf32x3", "f16"])
def test_snake_block_goldens(mode):
    from stylish_tts_b200.modules.variants import variant_engine
    g = gold("variants")
    variant_engine(torch.device(DEV)).set_precision(mode)
    for channels, k, dil, lengths, seed in vc.SNAKE_GOLDEN:
        blk = _block(channels, k, dil, seed)
        xs, s = vc.snake_case(channels, lengths, seed)
        outs = blk.forward_ragged([x.to(DEV) for x in xs], s.to(DEV))          # ragged: both utterances in one call
        for b, o in enumerate(outs):
            want = g[f"snake.k{k}.{b}"]
            rel = float((o.cpu() - want).abs().max() / want.abs().max())
            _log(f"snake.{mode}.k{k}.{b}.rel", rel)
            assert rel <= SNAKE_REL[mode], (mode, k, b, rel)
    variant_engine(torch.device(DEV)).set_precision("fp32")


def test_snake_block_larger_ragged_batch_and_module_call_against_oracle():
    """C = 256 (two 128-column GEMM tiles), k = 7, utterances crossing the 128-row tiles, a 2-frame utterance; the
    legacy weight_g / weight_v key dialect; the reference's dense [B, C, L] call."""
    from oracle import variants_port as vp
    from stylish_tts_b200.modules.variants import variant_engine
    variant_engine(torch.device(DEV)).set_precision("fp32")
    channels, k, dil, seed = 256, 7, (1, 3, 5), 31
    sd = vc.snake_state_dict(channels, k, seed)
    lengths = (700, 33, 129, 2, 128)
    xs, s = vc.snake_case(channels, lengths, seed)
    for legacy in (False, True):
        blk = _block(channels, k, dil, seed, legacy=legacy)
        outs = blk.forward_ragged([x.to(DEV) for x in xs], s.to(DEV))
        for b, o in enumerate(outs):
            with torch.no_grad():
                want = vp.snake_block(sd, xs[b][None], s[b: b + 1], k, dil)[0]
            rel = float((o.cpu() - want).abs().max() / want.abs().max())
            _log(f"snake.fp32.c256.legacy{int(legacy)}.{b}.rel", rel)
            # InstanceNorm over 2 frames maps every channel's pair of values to +-1 whatever their distance: rounding
            # differences are amplified by 1 / |y0 - y1| in each of the six norms (measured 5e-4; the reference's own
            # fp32 result is no better conditioned).  Every other length meets the fp32 bar (measured 2e-6).
            assert rel <= (5e-3 if lengths[b] < 4 else SNAKE_REL["fp32"]), (legacy, b, rel)
    x = torch.stack([xs[0][:, :96], xs[0][:, 100:196]])
    with torch.no_grad():
        want = vp.snake_block(sd, x, s[:2], k, dil)
    got = blk(x.to(DEV), s[:2].to(DEV))
    assert got.shape == want.shape
    assert float((got.cpu() - want).abs().max() / want.abs().max()) <= SNAKE_REL["fp32"]


STFT_REL = {"fp32": 5e-6, "tf32x3": 2e-5}      # relative to max |reference|; measured 1e-6 / 6e-6 (the fp32-grade bar is 1e-4)


@pytest.mark.parametrize("mode", ["fp32", "tf32x3"])
def test_stft_conv_module_goldens_and_oracle(mode):
    """STFT.transform / inverse (train/models/stft.py) for the module's default sizes and the export's (2048 / 300 /
    1200): ragged batch in one call against the reference goldens (every 4th bin) and the full oracle tensors; the
    inverse is fed the reference's own (magnitude, x, y)."""
    from oracle import variants_port as vp
    from stylish_tts_b200.modules import STFT
    from stylish_tts_b200.modules.variants import variant_engine
    g = gold("variants")
    variant_engine(torch.device(DEV)).set_precision(mode)
    for n_fft, hop, win, lengths, seed in vc.STFT_GOLDEN:
        m = STFT(filter_length=n_fft, hop_length=hop, win_length=win).to(DEV)
        waves = vc.wave_case(lengths, seed)
        mag, cx, cy = m.transform_ragged([w.to(DEV) for w in waves])
        ref = [vp.stft_conv_transform(w[None], n_fft, hop, win) for w in waves]
        for b, w in enumerate(waves):
            tag = f"stft.{n_fft}.{hop}.{win}.{b}"
            rmag, rx, ry, rre, rim = (t[0] for t in ref[b])
            scale = float(rmag.max())
            assert mag[b].shape == rmag.shape
            e_mag = float((mag[b].cpu() - rmag).abs().max()) / scale
            e_re = float((mag[b].cpu() * cx[b].cpu() - rre).abs().max()) / scale
            e_im = float((mag[b].cpu() * cy[b].cpu() - rim).abs().max()) / scale
            _log(f"stft.{mode}.{tag}.rel", [e_mag, e_re, e_im])
            assert max(e_mag, e_re, e_im) <= STFT_REL[mode], (tag, e_mag, e_re, e_im)
            assert float((mag[b].cpu()[::4] - g[tag + ".mag_s4"]).abs().max()) / scale <= STFT_REL[mode]
            # the unit vectors are well conditioned only where the bin carries energy
            strong = rmag > 1e-3 * scale
            assert float(((cx[b].cpu() - rx)[strong]).abs().max()) <= 1e-3
            assert float(((cy[b].cpu() - ry)[strong]).abs().max()) <= 1e-3
        outs = m.inverse_ragged([r[0][0].to(DEV) for r in ref], [r[1][0].to(DEV) for r in ref],
                                [r[2][0].to(DEV) for r in ref])
        for b, o in enumerate(outs):
            want = g[f"stft.{n_fft}.{hop}.{win}.{b}.inverse"]
            assert o.numel() == want.numel()
            e = float((o.cpu() - want).abs().max()) / float(want.abs().max())
            _log(f"stft.{mode}.{n_fft}.{hop}.{win}.{b}.inverse_rel", e)
            assert e <= STFT_REL[mode] * 5, (b, e)
        # the reference's dense call shapes, with `length`
        x = torch.stack([waves[0][:2000], waves[0][500:2500]])
        mg, xx, yy = m.transform(x.to(DEV))
        assert mg.shape == (2, n_fft // 2 + 1, 2000 // hop + 1)
        want = vp.stft_conv_inverse(*vp.stft_conv_transform(x, n_fft, hop, win)[:3], n_fft, hop, win, length=1500)
        got = m.inverse(mg, xx, yy, length=1500)
        assert got.shape == want.shape == (2, 1, 1500)
        assert float((got.cpu() - want).abs().max()) / float(want.abs().max()) <= STFT_REL[mode] * 20
    variant_engine(torch.device(DEV)).set_precision("fp32")
