"""CPU: the oracle restatements of SineGenerator / Snake AdaptiveGeneratorBlock (oracle/variants_port.py) against the
golden vectors produced by the REAL reference classes (oracle/make_golden_variants.py), and -- when /root/reference is
present -- against the live classes on a fresh case."""
import pytest
import torch

from oracle import variants_cases as vc
from oracle import variants_port as vp
from oracle.ref_import import reference_available
from tests.util import gold


def test_sinegen_port_reproduces_the_reference_goldens_bit_exactly():
    g = gold("variants")
    case = vc.sinegen_case(vc.SINEGEN_GOLDEN["lengths"], vc.SINEGEN_GOLDEN["seed"])
    for b, f0 in enumerate(case["f0"]):
        got = vp.sinegen(f0.view(1, -1, 1), case["rand_ini"][b].view(1, -1), case["noise"][b][None],
                         case["merge_weight"], vc.SAMP_RATE, vc.HARMONICS)
        assert torch.equal(got["out"].reshape(-1), g[f"sinegen.{b}.out"])
        assert torch.equal(got["sines"][0, ::16, :], g[f"sinegen.{b}.sines_s16"])
        assert torch.equal(got["c2"][0, ::16, :], g[f"sinegen.{b}.phase_s16"])
        # the track has voiced and unvoiced stretches, and the wrap correction keeps the phase bounded
        assert 0.05 < float((f0 > 0).float().mean()) < 0.98
        assert float(got["c2"].abs().max()) < 3.0


def test_snake_block_port_reproduces_the_reference_goldens_bit_exactly():
    g = gold("variants")
    for channels, k, dil, lengths, seed in vc.SNAKE_GOLDEN:
        sd = vc.snake_state_dict(channels, k, seed)
        xs, s = vc.snake_case(channels, lengths, seed)
        for b, x in enumerate(xs):
            with torch.no_grad():
                got = vp.snake_block(sd, x[None], s[b: b + 1], k, dil)[0]
            assert torch.equal(got, g[f"snake.k{k}.{b}"]), (k, b)


@pytest.mark.skipif(not reference_available(), reason="needs /root/reference (build container only)")
def test_ports_against_the_live_reference_on_a_fresh_case():
    from unittest import mock

    from oracle.ref_import import import_reference
    import_reference()
    from stylish_tts.train.models.ada_norm import AdaptiveGeneratorBlock
    from stylish_tts.train.models.cfm.cfm_mel_decoder import SineGenerator
    case = vc.sinegen_case((3001,), 77)
    ref = SineGenerator(vc.SAMP_RATE, harmonic_num=vc.HARMONICS)
    with torch.no_grad():
        ref.merge[0].weight.copy_(case["merge_weight"])
    x, ini, nz = case["f0"][0].view(1, -1, 1), case["rand_ini"][0].view(1, -1), case["noise"][0][None]
    with mock.patch.object(torch, "rand", lambda *a, **k: ini.clone()), \
            mock.patch.object(torch, "randn_like", lambda t, **k: nz.clone()), torch.no_grad():
        want = ref(x)
    assert torch.equal(vp.sinegen(x, ini, nz, case["merge_weight"], vc.SAMP_RATE, vc.HARMONICS)["out"], want)
    sd = vc.snake_state_dict(64, 5, 99)
    blk = AdaptiveGeneratorBlock(64, kernel_size=5, dilation=(1, 2, 4), style_dim=vc.STYLE_DIM).eval()
    blk.load_state_dict(sd)
    xs, s = vc.snake_case(64, (41,), 99)
    with torch.no_grad():
        assert torch.equal(vp.snake_block(sd, xs[0][None], s, 5, (1, 2, 4)), blk(xs[0][None], s))


def test_mirror_modules_have_the_reference_state_dict_keys():
    from stylish_tts_b200.modules import AdaptiveGeneratorBlock, SineGenerator
    blk = AdaptiveGeneratorBlock(128, kernel_size=7, dilation=(1, 3, 5), style_dim=64)
    sd = vc.snake_state_dict(128, 7, 1)
    assert set(blk.state_dict().keys()) == set(sd.keys())
    blk.load_state_dict(sd)
    sg = SineGenerator(vc.SAMP_RATE, harmonic_num=vc.HARMONICS)
    assert list(sg.state_dict().keys()) == ["merge.0.weight"] and sg.merge[0].weight.shape == (1, 9)
    with pytest.raises(RuntimeError):                               # no CPU fallback
        blk(torch.zeros(1, 128, 8), torch.zeros(1, 64))


def test_stft_conv_port_reproduces_the_reference_goldens_bit_exactly():
    g = gold("variants")
    for n_fft, hop, win, lengths, seed in vc.STFT_GOLDEN:
        for b, x in enumerate(vc.wave_case(lengths, seed)):
            mag, cx, cy, re, im = vp.stft_conv_transform(x[None], n_fft, hop, win)
            tag = f"stft.{n_fft}.{hop}.{win}.{b}"
            assert mag.shape == (1, n_fft // 2 + 1, x.numel() // hop + 1)
            assert torch.equal(mag[0, ::4], g[tag + ".mag_s4"])
            assert torch.equal((mag * cx)[0, ::4], g[tag + ".re_s4"])
            assert torch.equal((mag * cy)[0, ::4], g[tag + ".im_s4"])
            wave = vp.stft_conv_inverse(mag, cx, cy, n_fft, hop, win)
            assert torch.equal(wave.reshape(-1), g[tag + ".inverse"])
            assert wave.shape[-1] == (mag.shape[2] - 1) * hop


def test_stft_mirror_module_has_the_reference_buffers():
    from stylish_tts_b200.modules import STFT
    m = STFT(filter_length=2048, hop_length=300, win_length=1200)
    w = vp.stft_conv_weights(2048, 1200)
    assert list(m.state_dict().keys()) == ["window", "weight_forward_real", "weight_forward_imag", "weight_backward_real",
                                           "weight_backward_imag"]
    for k_mod, k_port in (("window", "window"), ("weight_forward_real", "fwd_real"), ("weight_forward_imag", "fwd_imag"),
                          ("weight_backward_real", "bwd_real"), ("weight_backward_imag", "bwd_imag")):
        assert torch.equal(getattr(m, k_mod), w[k_port]), k_mod
    with pytest.raises(RuntimeError):                               # no CPU fallback
        m.transform(torch.zeros(1, 4000))
