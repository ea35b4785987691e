"""GPU: the BASELINE configs beyond the 5 s goldens (C2 batch 32 mixed, C3 60 s long form, C4 batch 256, C5 vocoder
only) against the CPU ORACLE (oracle/stylish_port.py, pinned bit-exact against the live reference), run on the box
at test time on the same seeded inputs and noise draws.

Conventions (stated next to every number):
  * target = the reference with the exact (real) frame-0 STFT phase, "audio_x0" (its raw frame-0 phase is FFT
    rounding noise, tests/test_oracle_cpu.py::test_frame0_stft_phase_of_the_reference_is_rounding_noise); the raw
    reference is matched beyond sample 7 500;
  * the oracle's pitch / energy are fed to the speech stage (the harmonic source integrates f0); the predicted
    curves are compared on their own;
  * the source STFT feeds atan2 phases straight into phase_prior_conv (generator.py:407-414).  A bin with Re < 0 and
    |Im / Re| below ~3e-7 -- about one per 3 M bins, i.e. one per ~10 s of audio -- has a phase of +pi or -pi decided by
    the rounding noise of the reference's own fp32 FFT (measured on the box: exact DFT of the reference's input
    Im = +3.7e-8, torch.stft fp32 Im = -1.7e-8, Re = -0.127; the reference's fp64 run sides with the exact value).
    The CUDA path computes the correctly rounded exact DFT (fp64 FFT, csrc/stft_fft.cu), so it differs from the fp32
    reference exactly on those bins, and one such bin moves ~50 output frames by up to 7e-3.  Max-abs is therefore
    reported, and the assertion bounds the mean, the 99th percentile and the share of samples above the tolerance
    (`robust_bounds`); utterances without such a bin (all goldens, the 60 s case) also meet the plain max-abs bar.
"""
import json
import os

import pytest
import torch

from oracle.metrics import maxabs
from tests.util import ROOT, fixture_sds, synthesizer

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

# (mean, 99th percentile, tolerance t, allowed share of samples above t) per precision mode; measured on B200
# (gpurun_out/parity_configs.jsonl, profiles/README.md): tf32x3 mean 0.7e-5 .. 2.0e-5, q99 2.3e-5 .. 1.3e-4 (with one
# +-pi event), f16 mean 4e-5 .. 1.1e-4, q99 1.3e-4 .. 3.7e-4
BOUNDS = {"fp32": (1e-5, 1e-4, 1e-4, 3e-2), "tf32x3": (4e-5, 2.5e-4, 1e-4, 3e-2), "f16": (2e-4, 8e-4, 2e-3, 5e-3)}


def _log(name, val):
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "parity_configs.jsonl"), "a") as f:
        f.write(json.dumps({name: val}) + "\n")


def robust_bounds(name, mode, got, want):
    e = (got.double().cpu().reshape(-1) - want.double().cpu().reshape(-1)).abs()
    mean_b, q99_b, tol, share_b = BOUNDS[mode]
    stats = {"max": float(e.max()), "mean": float(e.mean()), "q99": float(torch.quantile(e[::7].float(), 0.99)),
             "share_gt_tol": float((e > tol).double().mean()), "tol": tol, "n": int(e.numel())}
    _log(f"{mode}.{name}", stats)
    assert stats["mean"] <= mean_b, (name, mode, stats)
    assert stats["q99"] <= q99_b, (name, mode, stats)
    assert stats["share_gt_tol"] <= share_b, (name, mode, stats)
    return stats


def _oracle(tokens, dur, noise, **kw):
    from oracle import stylish_port as port
    with torch.no_grad():
        return port.synthesize(fixture_sds(), tokens[None], torch.tensor([tokens.numel()]), noise,
                               duration=dur.float(), **kw)


def _pick(frames, n):
    order = sorted(range(len(frames)), key=lambda i: frames[i])
    want = [order[-1], order[0], order[len(order) // 2], order[len(order) // 4], order[(3 * len(order)) // 4]]
    return want[:n]


@pytest.mark.parametrize("mode", ["tf32x3", "f16"])
def test_c2_batch32_sample_against_oracle(mode):
    """BASELINE configs[1] at full size: the whole 32-utterance batch runs on the GPU; the longest (10 s), the
    shortest (0.25 s) and three more utterances are checked against the oracle."""
    from bench import workload
    from oracle.metrics import mel_l1
    from stylish_tts_b200 import fixture
    tokens, durs, frames = workload(1)
    g = torch.Generator().manual_seed(2025)
    noise = [fixture.draw_noise(f, g) for f in frames]
    pick = _pick(frames, 5)
    ref = {i: (_oracle(tokens[i], durs[i], noise[i]), _oracle(tokens[i], durs[i], noise[i], exact_frame0=True))
           for i in pick}
    syn = synthesizer(mode)
    pe = None
    first = syn.synthesize(tokens, durations=durs, noise=noise)          # predicted curves for everybody ...
    syn.check()
    pe = [(first["pitch"][i].clone(), first["energy"][i].clone()) for i in range(len(tokens))]
    pitch_err = max(maxabs(first["pitch"][i], ref[i][0]["pitch"]) for i in pick)
    dur_ok = all(torch.equal(first["duration_pred"][i].cpu().to(torch.int32), ref[i][0]["duration"].to(torch.int32))
                 for i in pick)
    for i in pick:                                                         # ... the oracle's for the checked ones
        pe[i] = (ref[i][0]["pitch"], ref[i][0]["energy"])
    out = syn.synthesize(tokens, durations=durs, noise=noise, pitch_energy=pe)
    syn.check(audit=True)
    _log(f"{mode}.c2.pitch_hz", pitch_err)
    assert dur_ok
    assert pitch_err <= (3e-2 if mode == "tf32x3" else 0.6), pitch_err     # on ~100-270 Hz; f16 measured 0.27 Hz
    for i in pick:
        robust_bounds(f"c2.utt{i}.{300 * frames[i] / 24000:.2f}s.x0", mode, out["audio"][i], ref[i][1]["audio"])
        if 300 * frames[i] > 7500:
            robust_bounds(f"c2.utt{i}.raw_after_7500", mode, out["audio"][i][7500:],
                          ref[i][0]["audio"].reshape(-1)[7500:])
        _log(f"{mode}.c2.utt{i}.mel_l1", mel_l1(out["audio"][i], ref[i][1]["audio"]))
        assert out["audio"][i].numel() == 300 * frames[i]


@pytest.mark.parametrize("mode", ["tf32x3", "f16"])
def test_c3_long_form_60s_against_oracle(mode):
    """BASELINE configs[2]: one 60 s utterance, 510 phonemes, 4 800 frames (the largest legal segment)."""
    from oracle.metrics import mel_l1
    from stylish_tts_b200 import fixture
    g = torch.Generator().manual_seed(14)
    tokens, dur, frames = fixture.synth_tokens(510, g), fixture.synth_durations(510, 4800, g), 4800
    noise = fixture.draw_noise(frames, g)
    raw = _oracle(tokens, dur, noise)
    x0 = _oracle(tokens, dur, noise, exact_frame0=True)
    syn = synthesizer(mode)
    out = syn.synthesize([tokens], durations=[dur], noise=[noise], pitch_energy=[(raw["pitch"], raw["energy"])],
                         want_spec=True)
    syn.check(audit=True)
    assert torch.equal(out["duration_pred"][0].cpu().to(torch.int32), raw["duration"].to(torch.int32))
    perr = maxabs(out["pitch"][0], raw["pitch"])
    _log(f"{mode}.c3.pitch_hz", perr)
    assert perr <= (3e-2 if mode == "tf32x3" else 0.6), perr      # 60 s, 510 tokens: tf32x3 measured 1.1e-2 Hz
    robust_bounds("c3.60s.x0", mode, out["audio"][0], x0["audio"])     # measured max-abs: 6.9e-5 (tf32x3), 3.7e-4 (f16)
    robust_bounds("c3.60s.raw_after_7500", mode, out["audio"][0][7500:], raw["audio"].reshape(-1)[7500:])
    _log(f"{mode}.c3.mel_l1", mel_l1(out["audio"][0], x0["audio"]))
    _log(f"{mode}.c3.logamp_l1", float((out["logamp"][0].cpu() - x0["logamp"][0]).abs().mean()))


def test_c4_batch256_on_one_gpu_sample_against_oracle():
    """BASELINE configs[3] shape (256 mixed utterances) as ONE packed batch on one GPU (16-bit mode, 47 GB
    workspace): three utterances against the oracle, and its 8-way LPT shards reproduce the same waveforms
    (sharding is a pure partition: utterances do not interact)."""
    from bench import workload
    from stylish_tts_b200 import fixture
    from stylish_tts_b200.shard import ShardPlan
    tokens, durs, frames = workload(8)
    assert len(tokens) == 256
    g = torch.Generator().manual_seed(404)
    noise = [fixture.draw_noise(f, g) for f in frames]
    pick = _pick(frames, 3)
    ref = {i: _oracle(tokens[i], durs[i], noise[i], exact_frame0=True) for i in pick}
    syn = synthesizer("f16")
    first = syn.synthesize(tokens, durations=durs, noise=noise)
    pe = [(first["pitch"][i].clone(), first["energy"][i].clone()) for i in range(256)]
    for i in pick:
        pe[i] = (ref[i]["pitch"], ref[i]["energy"])
    out = syn.synthesize(tokens, durations=durs, noise=noise, pitch_energy=pe)
    syn.check()
    full = [a.clone() for a in out["audio"]]
    for i in pick:
        robust_bounds(f"c4.utt{i}.x0", "f16", full[i], ref[i]["audio"])
    plan = ShardPlan(frames, 8)
    assert plan.balance() >= 0.97
    worst = 0.0
    for r in (0, 5):                                                       # two of the eight shards, on this GPU
        ids = plan.local(r)
        o = syn.synthesize([tokens[i] for i in ids], durations=[durs[i] for i in ids], noise=[noise[i] for i in ids],
                           pitch_energy=[pe[i] for i in ids])
        worst = max([worst] + [maxabs(o["audio"][k], full[i]) for k, i in enumerate(ids)])
    _log("f16.c4.shard_vs_full_batch", worst)
    assert worst <= 2e-5, worst


def _c5_inputs(n_frames4, seed, unvoiced_share=0.3):
    """SURVEY 8d C5 inputs for one item: asr ~ N(0,1) [128, L], style ~ N(0,1) [64], pitch = 120 * 2^U(-.5,.5) Hz
    smoothed over 20 frames with ~30 % unvoiced (0) spans, energy ~ U(0,1)."""
    g = torch.Generator().manual_seed(seed)
    asr = torch.randn((1, 128, n_frames4), generator=g)
    style = torch.randn((1, 64), generator=g)
    raw = 120.0 * torch.pow(2.0, torch.rand(n_frames4 + 19, generator=g) - 0.5)
    pitch = torch.nn.functional.avg_pool1d(raw[None, None], 20, 1)[0, 0][:n_frames4].clone()
    pos = 0
    while pos < n_frames4:                                   # alternating voiced / unvoiced spans
        v = int(torch.randint(40, 200, (1,), generator=g))
        u = int(v * unvoiced_share / (1 - unvoiced_share))
        pitch[pos + v: pos + v + u] = 0.0
        pos += v + u
    energy = torch.rand(n_frames4, generator=g)
    noise = {"prior": torch.randn((1, 128, n_frames4), generator=g),
             "source": torch.randn((1, 1, 75 * n_frames4), generator=g), "phase": torch.rand((1, 1), generator=g)}
    return asr, style, pitch[None], energy[None], noise


def _port_vocoder(asr, style, pitch, energy, noise, exact_frame0):
    from oracle import stylish_port as port
    import torch.nn.functional as F
    sd = fixture_sds()["speech_predictor"]
    with torch.no_grad():
        x = port.decoder(sd, "decoder.", asr, pitch, energy, style)
        z, _, _ = port.prior_encoder(sd, "prior_encoder.", x, noise["prior"])
        z2 = port.flow_reverse(sd, "flow.", z, style.unsqueeze(-1))
        mel = F.linear(z2.mT, sd["post_flow.weight"], sd["post_flow.bias"]).mT
        audio, la, ph = port.generator(sd, "generator.", mel, style, pitch, noise["source"], noise["phase"],
                                       exact_frame0=exact_frame0)
    return dict(dec=x, z_prior=z, z_flow=z2, mel=mel, audio=audio, logamp=la, phase=ph)


@pytest.mark.parametrize("mode", ["tf32x3", "f16"])
def test_c5_vocoder_only_10s_against_port(mode):
    """BASELINE configs[4] ("SineGen + AdaIN conv decoder + iSTFT" = Decoder -> PriorEncoder -> flow^-1 -> post_flow ->
    Generator): one 10 s item (3 200 hop-75 frames) through the module-level C-ABI entries stts_decoder ->
    stts_prior_flow -> stts_generator, every stage fed the CUDA output of the previous one, against the port."""
    from tests.util import gpu_nets
    L = 3200
    asr, style, pitch, energy, noise = _c5_inputs(L, 55)
    raw = _port_vocoder(asr, style, pitch, energy, noise, False)
    x0 = _port_vocoder(asr, style, pitch, energy, noise, True)
    eng = gpu_nets()["speech_predictor"].engine()
    eng.set_precision(mode)
    eng.set_layout(None, [L])
    d = lambda t: t.to(DEV).contiguous().view(-1)
    dec = eng.decoder(d(asr), d(pitch), d(energy), d(style))
    mel, zp, zf = eng.prior_flow(dec, d(noise["prior"]), d(style), want_z=True)
    audio, la, ph = eng.generator(mel, d(style), d(pitch), d(noise["source"]), d(noise["phase"]).reshape(-1)[:1])
    eng.check_source_status()
    rel = {"fp32": 1.0, "tf32x3": 2.0, "f16": 400.0}[mode]
    for name, got, want, tol in (("dec", dec, raw["dec"], 2e-5), ("z_prior", zp, raw["z_prior"], 2e-5),
                                 ("z_flow", zf, raw["z_flow"], 1e-4), ("mel", mel, raw["mel"], 1e-4)):
        r = maxabs(got, want) / float(want.abs().max())
        _log(f"{mode}.c5.{name}.rel", r)
        assert r <= tol * rel, (name, r)
    robust_bounds("c5.10s.x0", mode, audio, x0["audio"])
    robust_bounds("c5.10s.raw_after_7500", mode, audio[7500:], raw["audio"].reshape(-1)[7500:])
    _log(f"{mode}.c5.logamp_l1", float((la.view(1025, L + 1).cpu() - x0["logamp"][0]).abs().mean()))


# ------------------------------------------------------------------------------------------ SURVEY 8c edge fixtures
@pytest.mark.parametrize("case", ["f0_above_750", "single_voiced_frame", "f0_at_thresholds"])
def test_source_edge_fixtures_against_port(case):
    """generate_pcph (generator.py:247-315) edge cases: f0 > 750 Hz (max_h = int(12000 / min f0) < 16 and per-harmonic
    Nyquist masks), ONE voiced frame in the utterance, f0 straddling the 10 Hz (voiced) / 20 Hz (min-f0) thresholds."""
    from oracle import stylish_port as port
    from stylish_tts_b200.engine import Engine
    L = 80
    g = torch.Generator().manual_seed(3)
    f0 = torch.zeros(1, 1, L)
    if case == "f0_above_750":
        f0[0, 0, 10:60] = 760.0 + 3000.0 * torch.rand(50, generator=g)         # up to 3 760 Hz: 3 .. 15 harmonics
    elif case == "single_voiced_frame":
        f0[0, 0, 37] = 212.5
    else:
        f0[0, 0, 5:30] = torch.linspace(9.0, 25.0, 25)                          # <= 10 unvoiced, (10, 20] voiced
        f0[0, 0, 40:50] = 20.5                                                  # min f0 just above 20 -> max_h 16
    noise = torch.randn((1, 1, 75 * L), generator=g)
    init = torch.rand((1, 1), generator=g)
    want = port.generate_pcph(f0, (f0 > 10.0).float(), noise, init)
    eng = Engine(DEV)
    eng.set_layout(None, [L])
    got = eng.source_pcph(f0.view(-1).to(DEV), noise.view(-1).to(DEV), init.view(-1).to(DEV))
    eng.check_source_status()
    err = maxabs(got, want)
    _log(f"source.{case}", err)
    assert err <= 2e-6, (case, err)


@pytest.mark.parametrize("case", ["all1", "all46"])
def test_extreme_duration_utterances_against_oracle(case):
    """Durations of all-1 (F = T: every frame its own token) and all-46 (the table maximum) through the whole path."""
    from stylish_tts_b200 import fixture
    g = torch.Generator().manual_seed(9)
    T = 24 if case == "all1" else 6
    tokens = fixture.synth_tokens(T, g)
    dur = torch.full((T,), 1 if case == "all1" else 46, dtype=torch.int64)
    noise = fixture.draw_noise(int(dur.sum()), g)
    x0 = _oracle(tokens, dur, noise, exact_frame0=True)
    for mode, tol in (("fp32", 1e-4), ("tf32x3", 1e-4), ("f16", 1e-3)):
        syn = synthesizer(mode)
        out = syn.synthesize([tokens], durations=[dur], noise=[noise], pitch_energy=[(x0["pitch"], x0["energy"])])
        syn.check()
        err = maxabs(out["audio"][0], x0["audio"])
        _log(f"{mode}.dur_{case}", {"audio_x0": err, "pitch": maxabs(out["pitch"][0], x0["pitch"])})
        assert err <= tol, (case, mode, err)
        assert torch.equal(out["duration_pred"][0].cpu().to(torch.int32), x0["duration"].to(torch.int32))


def test_cuda_graphs_survive_a_change_of_batch_shape():
    """the context holds ONE workspace plan, a captured graph bakes its pointers.  Alternating two
    shapes (A, B, A, B) in graph mode must give, every time, what the eager path gives for that shape."""
    g = torch.Generator().manual_seed(77)
    from stylish_tts_b200 import fixture
    shapes = {"A": [(9, 31), (14, 50)], "B": [(30, 120)]}
    items = {k: [(fixture.synth_tokens(t, g), fixture.synth_durations(t, f, g), f) for t, f in v]
             for k, v in shapes.items()}
    syn = synthesizer("fp32")
    want = {}
    for k, it in items.items():
        o = syn.synthesize([x[0] for x in it], durations=[x[1] for x in it], frames=[x[2] for x in it], graph=False)
        want[k] = ([p.clone() for p in o["pitch"]], [e.clone() for e in o["energy"]])
    for k in ("A", "B", "A", "B", "B", "A"):
        it = items[k]
        o = syn.synthesize([x[0] for x in it], durations=[x[1] for x in it], frames=[x[2] for x in it], graph=True)
        syn.check()
        for i in range(len(it)):
            assert torch.equal(o["pitch"][i], want[k][0][i]), (k, i)
            assert torch.equal(o["energy"][i], want[k][1][i]), (k, i)
            assert o["audio"][i].numel() == 300 * it[i][2] and bool(torch.isfinite(o["audio"][i]).all())
            assert float(o["audio"][i].abs().max()) <= 1.0


def test_saturation_audit_reports_clipped_twins():
    """fp16 twins saturate at 65504.  With fixture weights nothing saturates; with the decoder's
    input scaled far outside the validated range the audit must say so (check(audit=True) raises OverflowError)."""
    from tests.util import gpu_nets
    eng = gpu_nets()["speech_predictor"].engine()
    eng.set_precision("f16")
    L = 200
    asr, style, pitch, energy, noise = _c5_inputs(L, 56)
    eng.set_layout(None, [L])
    d = lambda t: t.to(DEV).contiguous().view(-1)
    eng.decoder(d(asr), d(pitch), d(energy), d(style))
    assert eng.saturation_count() == 0
    eng.decoder(d(asr * 1e6), d(pitch), d(energy), d(style))       # X0's twin of the raw input clips
    assert eng.saturation_count() > 0
    eng.set_precision("fp32")
