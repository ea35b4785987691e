"""Golden vectors for SURVEY.md section 8(f) rank 2 from the REAL reference classes (run in the build container, where
/root/reference exists):   python -m oracle.make_golden_variants

  * SineGenerator            /root/reference/src/stylish_tts/train/models/cfm/cfm_mel_decoder.py:16-101
  * AdaptiveGeneratorBlock   /root/reference/src/stylish_tts/train/models/ada_norm.py:11-127

For every case of oracle/variants_cases.py the reference module is built, given the seeded weights, and run; the
restatement oracle/variants_port.py must reproduce it BIT-EXACTLY (the script aborts otherwise: "PIN" lines), and the
reference outputs go to tests/golden/variants.safetensors.  The reference draws its random numbers inside forward
(torch.rand, torch.randn_like): they are replaced for the duration of the call by functions that hand out the case's
seeded draws, so reference and restatement see the same numbers.
"""
from __future__ import annotations

import os
from unittest import mock

import torch
from safetensors.torch import save_file

from oracle import variants_cases as vc
from oracle import variants_port as vp
from oracle.ref_import import import_reference

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pin(name, got, want):
    err = float((got.double() - want.double()).abs().max())
    print(f"PIN {name}: max|port - reference| = {err:.3e}, bit-equal = {torch.equal(got, want)}")
    if not torch.equal(got, want):
        raise SystemExit(f"oracle restatement differs from the reference on {name}")


def main():
    import_reference()
    from stylish_tts.train.models.ada_norm import AdaptiveGeneratorBlock
    from stylish_tts.train.models.cfm.cfm_mel_decoder import SineGenerator

    out = {}
    # ---------------------------------------------------------------- SineGenerator, one utterance at a time
    case = vc.sinegen_case(vc.SINEGEN_GOLDEN["lengths"], vc.SINEGEN_GOLDEN["seed"])
    ref = SineGenerator(vc.SAMP_RATE, harmonic_num=vc.HARMONICS)
    with torch.no_grad():
        ref.merge[0].weight.copy_(case["merge_weight"])
    for b, f0 in enumerate(case["f0"]):
        x = f0.view(1, -1, 1)
        ini, nz = case["rand_ini"][b].view(1, -1), case["noise"][b].unsqueeze(0)
        with mock.patch.object(torch, "rand", lambda *a, **k: ini.clone()), \
                mock.patch.object(torch, "randn_like", lambda t, **k: nz.clone()), torch.no_grad():
            want = ref(x)
            want_sines = ref._f02sine(torch.cat([x * (h + 1) for h in range(vc.HARMONICS + 1)], dim=2))
        got = vp.sinegen(x, ini, nz, case["merge_weight"], vc.SAMP_RATE, vc.HARMONICS)
        _pin(f"sinegen.{b}.out", got["out"], want)
        _pin(f"sinegen.{b}.sines", got["sines"], want_sines)
        out[f"sinegen.{b}.out"] = want.reshape(-1).contiguous()
        out[f"sinegen.{b}.sines_s16"] = want_sines[0, ::16, :].contiguous()
        # the phase in cycles (argument of the sine, :78, before * 2 pi) is an internal of _f02sine: taken from the
        # restatement, which the two PIN lines above tie bit-exactly to the reference's sines
        out[f"sinegen.{b}.phase_s16"] = got["c2"][0, ::16, :].contiguous()
    # ---------------------------------------------------------------- Snake AdaptiveGeneratorBlock
    for channels, k, dil, lengths, seed in vc.SNAKE_GOLDEN:
        sd = vc.snake_state_dict(channels, k, seed)
        ref = AdaptiveGeneratorBlock(channels, kernel_size=k, dilation=dil, style_dim=vc.STYLE_DIM).eval()
        missing = ref.load_state_dict(sd, strict=True)
        assert not missing.missing_keys and not missing.unexpected_keys
        xs, s = vc.snake_case(channels, lengths, seed)
        for b, x in enumerate(xs):
            with torch.no_grad():
                want = ref(x[None], s[b: b + 1])[0]
                got = vp.snake_block(sd, x[None], s[b: b + 1], k, dil)[0]
            _pin(f"snake.k{k}.{b}", got, want)
            out[f"snake.k{k}.{b}"] = want.contiguous()
        # the legacy weight_g / weight_v dialect must give the same weights
        sd_old = vc.snake_state_dict(channels, k, seed, legacy=True)
        with torch.no_grad():
            got = vp.snake_block(sd_old, xs[0][None], s[:1], k, dil)[0]
        _pin(f"snake.k{k}.legacy_keys", got, out[f"snake.k{k}.0"])
    # ---------------------------------------------------------------- conv-DFT STFT module (stft.py:6-195)
    from stylish_tts.train.models.stft import STFT
    for n_fft, hop, win, lengths, seed in vc.STFT_GOLDEN:
        ref = STFT(filter_length=n_fft, hop_length=hop, win_length=win).eval()
        for b, x in enumerate(vc.wave_case(lengths, seed)):
            with torch.no_grad():
                mag, cx, cy = ref.transform(x[None])
                wave = ref.inverse(mag, cx, cy)
                wave_len = ref.inverse(mag, cx, cy, length=x.numel() - 7)
            got = vp.stft_conv_transform(x[None], n_fft, hop, win)
            tag = f"stft.{n_fft}.{hop}.{win}.{b}"
            _pin(tag + ".mag", got[0], mag)
            _pin(tag + ".x", got[1], cx)
            _pin(tag + ".y", got[2], cy)
            _pin(tag + ".inverse", vp.stft_conv_inverse(mag, cx, cy, n_fft, hop, win), wave)
            _pin(tag + ".inverse_len", vp.stft_conv_inverse(mag, cx, cy, n_fft, hop, win, length=x.numel() - 7), wave_len)
            out[tag + ".mag_s4"] = mag[0, ::4].contiguous()
            out[tag + ".re_s4"] = (mag * cx)[0, ::4].contiguous()
            out[tag + ".im_s4"] = (mag * cy)[0, ::4].contiguous()
            out[tag + ".inverse"] = wave.reshape(-1).contiguous()
    path = os.path.join(ROOT, "tests", "golden", "variants.safetensors")
    save_file(out, path)
    print("wrote", path, {k: tuple(v.shape) for k, v in out.items()})


if __name__ == "__main__":
    main()
