"""CPU tests: the oracle port against the golden vectors generated from the live reference
(oracle/make_golden.py), and -- when /root/reference is present -- against the reference itself."""
import json
import os

import pytest
import torch

from tests.util import GOLD, case_noise, fixture_sds, gold, maxabs
from oracle import stylish_port as port


def _run_case(g, seed, prefix=""):
    tokens = g[prefix + "tokens"]
    dur = g[prefix + "duration_in"].to(torch.float32)
    noise = case_noise(int(dur.sum()), seed)
    taps = {}
    with torch.no_grad():
        out = port.synthesize(fixture_sds(), tokens[None], torch.tensor([tokens.numel()]), noise, duration=dur,
                              taps=taps)
    return out, taps, noise


def test_pin_file_records_bit_exact_port():
    pin = json.load(open(os.path.join(GOLD, "PIN.json")))
    assert pin["worst_rel"] <= 2e-5


def test_tiny_all_intermediates():
    g = gold("tiny")
    out, taps, noise = _run_case(g, int(g["seed"]))
    assert torch.equal(out["duration"].to(torch.int32), g["duration_pred"])
    for k in ("audio", "logamp", "phase", "pitch", "energy"):
        assert maxabs(out[k], g["out." + k]) <= 2e-5 * max(1.0, float(g["out." + k].abs().max())), k
    for k, v in g.items():
        if k.startswith("tap.") and k[4:] in taps:
            assert maxabs(taps[k[4:]], v) <= 2e-5 * max(1.0, float(v.abs().max())), k
    for k in ("prior", "source", "phase"):
        assert torch.equal(noise[k], g["noise." + k])


def test_short_and_batch3():
    g = gold("short")
    out, _, _ = _run_case(g, int(g["seed"]))
    assert maxabs(out["audio"], g["out.audio"]) <= 2e-6
    g = gold("batch3")
    for i in range(3):
        out, _, _ = _run_case(g, int(g[f"{i}.seed"]), prefix=f"{i}.")
        assert maxabs(out["audio"], g[f"{i}.out.audio"]) <= 2e-6
        assert out["audio"].shape[-1] == 300 * int(g[f"{i}.duration_in"].sum())


def test_exact_frame0_variant_matches_reference_with_real_frame0():
    g = gold("short")
    tokens = g["tokens"]
    dur = g["duration_in"].to(torch.float32)
    noise = case_noise(int(dur.sum()), int(g["seed"]))
    with torch.no_grad():
        out = port.synthesize(fixture_sds(), tokens[None], torch.tensor([tokens.numel()]), noise, duration=dur,
                              exact_frame0=True)
    assert maxabs(out["audio"], g["out.audio_x0"]) <= 2e-6
    # the raw reference differs from it only through frame-0 rounding noise: large near the start
    assert maxabs(g["out.audio"], g["out.audio_x0"]) > 1e-3


def test_c1_audio():
    g = gold("c1")
    out, _, _ = _run_case(g, int(g["seed"]))
    assert maxabs(out["audio"], g["out.audio"]) <= 2e-6
    assert maxabs(out["logamp"][:, :, ::16], g["out.logamp_s16"]) <= 2e-5


def test_durations_bit_exact():
    g = gold("durations")
    d = port.prediction_to_duration(g["logits"])
    assert torch.equal(d.to(torch.int32), g["duration"])
    idx = port.duration_to_indices(d[:64])
    al = port.duration_to_alignment(d[:64])
    assert torch.equal(al.argmax(dim=0), idx)
    assert al.shape == (64, int(d[:64].sum())) and float(al.sum()) == al.shape[1]


def test_durations_million_rows_and_adversarial_ties():
    """The reference DurationProcessor's own results (oracle/make_golden_durations.py) on 2^20 seeded rows and
    on the exact-tie / tie -+ k ulp rows: the port must reproduce them bit for bit on this host."""
    from oracle.durations_cases import adversarial_logits, random_logits, tie_distance
    g = gold("durations_adv")
    assert torch.equal(adversarial_logits(), g["logits"])
    assert torch.equal(port.prediction_to_duration(g["logits"]).to(torch.int32), g["duration"])
    # exact ties round half to even: classes (7, 12) -> 9.5 -> 10, (9, 12) -> 10.5 -> 10
    row = torch.full((2, 16), -1e4)
    row[0, 6] = row[0, 8] = 0.0
    row[1, 7] = row[1, 8] = 0.0
    assert port.prediction_to_duration(row).tolist() == [10.0, 10.0]
    g = gold("durations_1m")
    logits = random_logits(seed=int(g["seed"]))
    d = port.prediction_to_duration(logits)
    bad = torch.nonzero(d.to(torch.uint8) != g["duration"]).flatten()
    near = set(g["near_tie_rows"].tolist())        # another host's libm may move these (and only these) rows
    assert all(int(i) in near for i in bad)
    dist, soft_path = tie_distance(logits)
    assert sorted(near) == torch.nonzero((dist < 1e-5) & soft_path).flatten().tolist()


def test_frame0_stft_phase_of_the_reference_is_rounding_noise():
    """Pins the "defined deviation" of DESIGN.md section 4: torch.stft(center=True, reflect) centres frame 0 on
    sample 0, the windowed frame is even, X_k is exactly real -- so the reference's atan2 phase there is the sign
    of FFT rounding noise.  Evidence: the SAME torch.stft in fp32 and in fp64 disagrees by +-pi sign flips on a
    large share of frame-0 bins and on none of the other frames' bins; the exact value (Im = 0) is what the CUDA
    path returns and what the "audio_x0" target encodes."""
    g = gold("tiny")
    src = g["tap.source"]
    win32, win64 = torch.hann_window(1200), torch.hann_window(1200, dtype=torch.float64)
    s32 = torch.stft(src, 2048, 75, 1200, window=win32, return_complex=True)[:, :, :-1]
    s64 = torch.stft(src.double(), 2048, 75, 1200, window=win64, return_complex=True)[:, :, :-1]
    p32, p64 = torch.angle(s32).double(), torch.angle(s64)
    flips = ((p32 - p64).abs() > 3.0)[0]                     # [1025 bins, frames]
    assert int(flips[:, 1:].sum()) == 0                       # no other frame flips
    assert int(flips[:, 0].sum()) >= 50                       # frame 0: a large share of the 1025 bins
    # frame 0 is numerically real: |Im| is at rounding level of |Re| in fp64
    assert float((s64.imag[0, :, 0].abs() / (s64.real[0, :, 0].abs() + 1e-30)).median()) < 1e-9
    # the golden phase of the reference shows both signs of pi in frame 0
    f0 = g["tap.har_phase"][0, :, 0]
    assert int((f0 > 3.0).sum()) > 20 and int((f0 < -3.0).sum()) > 20
    # the exact-frame-0 variant differs from the raw reference only in that frame
    x0 = port.exact_frame0_phase(g["tap.har_phase"])
    assert torch.equal(x0[:, :, 1:], g["tap.har_phase"][:, :, 1:])
    assert set(x0[0, :, 0].unique().tolist()) <= {0.0, float(torch.tensor(3.141592653589793, dtype=torch.float32))}


def test_known_answers():
    # facts usable without the oracle
    assert port.CLASS_TO_DUR == [1, 2, 3, 4, 5, 6, 7, 9, 12, 15, 18, 22, 27, 32, 38, 46]
    up = port.upsample4(torch.tensor([[0.0, 4.0, 8.0]]))[0]
    assert torch.allclose(up, torch.tensor([0, 0, .5, 1.5, 2.5, 3.5, 4.5, 5.5, 6.5, 7.5, 8, 8.0]))
    al = torch.zeros(1, 40, 3)
    al[0, 20, :] = 1
    m = port.band_mask(al, torch.zeros(1, 40, dtype=torch.bool), 5)[0, 0, 0]
    assert (~m).nonzero().flatten().tolist() == list(range(15, 26))  # inverted polarity: band is False


def test_unvoiced_and_error_paths():
    f0 = torch.zeros(1, 1, 20)
    n = torch.randn(1, 1, 1500)
    out = port.generate_pcph(f0, (f0 > 10).float(), n, torch.zeros(1, 1))
    assert torch.equal(out, 0.01 * n)                       # all-unvoiced early-out
    f0[..., 3] = 15.0                                       # voiced, but nothing above 20 Hz
    with pytest.raises(RuntimeError):
        port.generate_pcph(f0, (f0 > 10).float(), n, torch.zeros(1, 1))


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/stylish_tts"), reason="reference checkout absent")
def test_port_matches_live_reference_module():
    from oracle.ref_import import import_reference
    build_model, load_cfg, DurationProcessor, yml = import_reference()
    nets = build_model(load_cfg(open(yml)))
    sds = fixture_sds()
    dp = nets.duration_predictor
    dp.load_state_dict(sds["duration_predictor"])
    dp.eval()
    tok = torch.randint(1, 178, (1, 17))
    ln = torch.tensor([17])
    with torch.no_grad():
        assert maxabs(dp(tok, ln), port.duration_predictor(sds["duration_predictor"], tok, ln)) == 0.0
