"""Generate tests/golden/* from the REAL reference and pin the oracle port against it.

Run here (the container that has /root/reference):  ``python oracle/make_golden.py``
TEST INFRASTRUCTURE ONLY.  Writes:
  tests/golden/manifest.json          state_dict layout + default-init statistics (fixture.py)
  tests/golden/tiny.safetensors       T=12, F=40: inputs, noise and EVERY intermediate of SURVEY 8a
  tests/golden/short.safetensors      T=5,  F=20 (shortest legal utterance): audio only
  tests/golden/c1.safetensors         T=120, F=400 (BASELINE configs[0]): audio, pitch, energy,
                                      durations, logamp/phase every 16th frame
  tests/golden/batch3.safetensors     three mixed-length utterances as a B=1 loop
  tests/golden/durations.safetensors  8192 random + adversarial logit rows -> durations
  tests/golden/PIN.json               max-abs(port - reference) for every recorded tensor
The reference forward is run exactly as stage_type.py:483-523 / export_model.py:35-45 do, with
the three torch RNG draws replayed from ``fixture.draw_noise``.
"""
from __future__ import annotations

import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import stylish_port as port  # noqa: E402
from oracle.ref_import import import_reference  # noqa: E402
from stylish_tts_b200 import fixture  # noqa: E402
from safetensors.torch import save_file  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


class ReplayRNG:
    """Replays the three draws of one reference forward (flow.py:314, generator.py:275-277,306)."""

    def __init__(self, noise):
        self.noise = noise
        self.saved = None

    def __enter__(self):
        self.saved = (torch.randn_like, torch.randn, torch.rand)
        n = self.noise

        def randn_like(t, *a, **k):
            assert tuple(t.shape) == tuple(n["prior"].shape), (t.shape, n["prior"].shape)
            return n["prior"].clone()

        def randn(shape, *a, **k):
            assert tuple(shape) == tuple(n["source"].shape), (shape, n["source"].shape)
            return n["source"].clone()

        def rand(shape, *a, **k):
            assert tuple(shape) == (1, 1)
            return n["phase"].clone()

        torch.randn_like, torch.randn, torch.rand = randn_like, randn, rand
        return self

    def __exit__(self, *exc):
        torch.randn_like, torch.randn, torch.rand = self.saved


def ref_synthesize(nets, proc, tokens, lengths, noise, duration, taps, exact_frame0=False):
    """stage_type.py:483-523 for one utterance, with hooks recording the SURVEY 8a intermediates."""
    sp = nets.speech_predictor
    pe = nets.pitch_energy_predictor
    hooks = []

    def rec(name, pick=lambda o: o):
        def f(mod, inp, out):
            taps[name] = pick(out).detach().clone()
        return f

    def rec_in(name, idx=0):
        def f(mod, inp):
            taps[name] = inp[idx].detach().clone()
        return f

    hooks.append(sp.text_encoder.register_forward_hook(rec("sp_text", lambda o: o[0])))
    hooks.append(sp.style_encoder.register_forward_hook(rec("sp_style")))
    hooks.append(sp.decoder.register_forward_hook(rec("dec_out", lambda o: o[0])))
    hooks.append(sp.prior_encoder.register_forward_hook(rec("z_prior", lambda o: o[0])))
    hooks.append(sp.flow.register_forward_hook(rec("z_flow", lambda o: o[0])))
    hooks.append(sp.post_flow.register_forward_hook(rec("mel", lambda o: o.mT)))
    hooks.append(sp.generator.amp_prior_conv.register_forward_pre_hook(rec_in("har_spec")))
    hooks.append(sp.generator.phase_prior_conv.register_forward_pre_hook(rec_in("har_phase")))
    if exact_frame0:  # the real reference with the exact (real-valued) frame-0 spectrum phase
        hooks.append(sp.generator.phase_prior_conv.register_forward_pre_hook(
            lambda m, i: (port.exact_frame0_phase(i[0]),)))
    hooks.append(sp.generator.projector.register_forward_hook(rec("gen_proj")))
    hooks.append(sp.generator.convnext[3].register_forward_hook(rec("gen_convnext")))
    hooks.append(pe.prosody_encoder.register_forward_hook(rec("pe_prosody")))
    hooks.append(pe.F0[0].register_forward_pre_hook(rec_in("pe_cross")))
    hooks.append(sp.decoder.register_forward_pre_hook(
        lambda m, i: taps.update(pitch4=i[1].detach().clone(), energy4=i[2].detach().clone())))
    orig_prior = sp.generator.prior_generator

    def prior_rec(*a, **k):
        out = orig_prior(*a, **k)
        taps["source"] = out.squeeze(1).detach().clone()
        return out

    sp.generator.prior_generator = prior_rec
    try:
        with torch.no_grad(), ReplayRNG(noise):
            logits = nets.duration_predictor(tokens, lengths)
            dur_pred = proc.prediction_to_duration(logits[0], lengths[0])[: int(lengths[0])]
            dur = dur_pred if duration is None else duration.to(torch.float32)
            alignment = proc.duration_to_alignment(dur).unsqueeze(0)
            pe_enc, _, _ = nets.pe_text_encoder(tokens, lengths)
            pe_style = nets.pe_text_style_encoder(pe_enc, lengths)
            pitch, energy = pe(pe_enc, lengths, alignment, pe_style)
            pred = sp(tokens, lengths, alignment, pitch, energy)
    finally:
        sp.generator.prior_generator = orig_prior
        for h in hooks:
            h.remove()
    taps.update(dur_logits=logits, pe_text=pe_enc, pe_style=pe_style)
    return dict(audio=pred.audio, logamp=pred.magnitude, phase=pred.phase, duration=dur_pred,
                pitch=pitch, energy=energy)


def compare(name, a, b, pin, tol=2e-6):
    d = float((a.double() - b.double()).abs().max()) if a.numel() else 0.0
    scale = float(b.double().abs().max()) if b.numel() else 0.0
    pin[name] = {"max_abs": d, "ref_absmax": scale}
    flag = "" if d <= tol * max(1.0, scale) else "   <-- LARGE"
    print(f"    {name:24s} max|port-ref| = {d:.3e}  (|ref|max {scale:.3e}){flag}")
    return d


def run_case(name, nets, sds, proc, tokens, dur, seed, pin, keep_taps):
    gen = torch.Generator().manual_seed(seed)
    n_frames = int(dur.sum())
    noise = fixture.draw_noise(n_frames, gen)
    tk = tokens.unsqueeze(0)
    ln = torch.tensor([tokens.shape[0]], dtype=torch.int64)
    rtaps, ptaps = {}, {}
    ref = ref_synthesize(nets, proc, tk, ln, noise, dur, rtaps)
    with torch.no_grad():
        got = port.synthesize(sds, tk, ln, noise, duration=dur.to(torch.float32), taps=ptaps)
    ref_x0 = ref_synthesize(nets, proc, tk, ln, noise, dur, {}, exact_frame0=True)
    with torch.no_grad():
        got_x0 = port.synthesize(sds, tk, ln, noise, duration=dur.to(torch.float32), exact_frame0=True)
    compare(f"{name}.audio_x0", got_x0["audio"], ref_x0["audio"], pin)
    ref["audio_x0"] = ref_x0["audio"]
    ref["logamp_x0"] = ref_x0["logamp"]
    d0 = (ref_x0["audio"] - ref["audio"]).abs()[0, 0]
    print(f"    reference raw vs exact-frame-0: max {float(d0.max()):.3e}; beyond sample 7500: {float(d0[7500:].max()) if d0.numel() > 7500 else 0:.3e}")
    print(f"  case {name}: T={tokens.shape[0]} F={n_frames} audio absmax={float(ref['audio'].abs().max()):.4f} "
          f"voiced frac={float((rtaps['pitch4'] > 10).float().mean()):.2f}")
    for k in ("audio", "logamp", "phase", "pitch", "energy"):
        compare(f"{name}.{k}", got[k], ref[k], pin)
    assert torch.equal(got["duration"], ref["duration"]), "duration mismatch port vs reference"
    for k in sorted(rtaps):
        if k in ptaps:
            compare(f"{name}.tap.{k}", ptaps[k], rtaps[k], pin)
    out = {"tokens": tokens, "duration_in": dur.to(torch.int32), "duration_pred": ref["duration"].to(torch.int32),
           "seed": torch.tensor([seed])}
    return ref, rtaps, noise, out


def main():
    torch.set_num_threads(8)
    build_model, load_cfg, DurationProcessor, yml = import_reference()
    mc = load_cfg(open(yml))
    torch.manual_seed(0)
    nets = build_model(mc)
    proc = DurationProcessor(mc.duration_predictor.duration_classes, mc.duration_predictor.max_duration)

    # ---- manifest
    manifest = {}
    for m in fixture.MODULES:
        manifest[m] = {}
        for k, v in nets[m].state_dict().items():
            vf = v.detach().float()
            std = float(vf.std(unbiased=False)) if vf.numel() > 1 else 0.0
            manifest[m][k] = {"shape": list(v.shape), "mean": float(vf.mean()), "std": std}
    os.makedirs(GOLD, exist_ok=True)
    with open(os.path.join(GOLD, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=0, sort_keys=True)
    print("manifest:", {m: len(manifest[m]) for m in manifest})

    # ---- fixture weights into the reference modules
    sds = fixture.fixture_nets(manifest, seed=0)
    for m in fixture.MODULES:
        missing, unexpected = nets[m].load_state_dict(sds[m], strict=False)
        assert not unexpected, unexpected
        assert all(k.startswith("posterior_encoder.") for k in missing), missing
        nets[m].eval()

    pin = {"torch": torch.__version__, "threads": torch.get_num_threads()}

    # ---- tiny: all intermediates
    g = torch.Generator().manual_seed(11)
    tokens = fixture.synth_tokens(12, g)
    dur = fixture.synth_durations(12, 40, g)
    ref, taps, noise, out = run_case("tiny", nets, sds, proc, tokens, dur, 101, pin, True)
    for k, v in ref.items():
        out[f"out.{k}"] = v.float().contiguous()
    del out["out.logamp_x0"]
    for k, v in taps.items():
        out[f"tap.{k}"] = v.float().contiguous()
    for k, v in noise.items():
        out[f"noise.{k}"] = v.contiguous()
    save_file({k: v.contiguous() for k, v in out.items()}, os.path.join(GOLD, "tiny.safetensors"))

    # ---- short: shortest legal utterance
    g = torch.Generator().manual_seed(12)
    tokens = fixture.synth_tokens(5, g)
    dur = fixture.synth_durations(5, 20, g)
    ref, taps, noise, out = run_case("short", nets, sds, proc, tokens, dur, 102, pin, False)
    out["out.audio"] = ref["audio"].contiguous()
    out["out.audio_x0"] = ref["audio_x0"].contiguous()
    out["out.pitch"] = ref["pitch"].contiguous()
    save_file(out, os.path.join(GOLD, "short.safetensors"))

    # ---- c1: BASELINE configs[0]
    g = torch.Generator().manual_seed(13)
    tokens = fixture.synth_tokens(120, g)
    dur = fixture.synth_durations(120, 400, g)
    ref, taps, noise, out = run_case("c1", nets, sds, proc, tokens, dur, 103, pin, False)
    out["out.audio"] = ref["audio"].contiguous()
    out["out.pitch"] = ref["pitch"].contiguous()
    out["out.energy"] = ref["energy"].contiguous()
    out["out.audio_x0"] = ref["audio_x0"].contiguous()
    out["out.logamp_x0_s16"] = ref["logamp_x0"][:, :, ::16].contiguous()
    out["out.logamp_s16"] = ref["logamp"][:, :, ::16].contiguous()
    out["out.phase_s16"] = ref["phase"][:, :, ::16].contiguous()
    out["tap.mel_s16"] = taps["mel"][:, :, ::16].contiguous()
    out["tap.dec_out_s16"] = taps["dec_out"][:, :, ::16].contiguous()
    save_file(out, os.path.join(GOLD, "c1.safetensors"))

    # ---- batch3: mixed lengths, B = 1 loop
    out = {}
    for i, (t, f_) in enumerate([(7, 23), (16, 57), (9, 31)]):
        g = torch.Generator().manual_seed(20 + i)
        tokens = fixture.synth_tokens(t, g)
        dur = fixture.synth_durations(t, f_, g)
        ref, taps, noise, o = run_case(f"batch3.{i}", nets, sds, proc, tokens, dur, 200 + i, pin, False)
        for k, v in o.items():
            out[f"{i}.{k}"] = v
        out[f"{i}.out.audio"] = ref["audio"].contiguous()
        out[f"{i}.out.audio_x0"] = ref["audio_x0"].contiguous()
        out[f"{i}.out.pitch"] = ref["pitch"].contiguous()
        out[f"{i}.out.energy"] = ref["energy"].contiguous()
    save_file(out, os.path.join(GOLD, "batch3.safetensors"))

    # ---- durations: reference DurationProcessor on random + adversarial rows
    g = torch.Generator().manual_seed(31)
    logits = torch.randn((8192, 16), generator=g) * 3.0
    logits[4096:] *= 0.3  # flatter rows -> soft path with larger durations
    logits[:16] = 0.0
    logits[:16, :] += torch.eye(16) * 20.0  # one row per class, saturated argmax
    with torch.no_grad():
        dref = proc.prediction_to_duration(logits, None)
        dport = port.prediction_to_duration(logits)
        assert torch.equal(dref, dport)
        al = proc.duration_to_alignment(dref[:64])
        assert torch.equal(al, port.duration_to_alignment(dref[:64]))
        assert torch.equal(al.argmax(dim=0), port.duration_to_indices(dref[:64]))
    save_file({"logits": logits, "duration": dref.to(torch.int32)}, os.path.join(GOLD, "durations.safetensors"))
    print("  durations: 8192 rows, port == reference bit-exact; histogram max", int(dref.max()))

    worst = max(v["max_abs"] / max(1.0, v["ref_absmax"]) for k, v in pin.items() if isinstance(v, dict))
    pin["worst_rel"] = worst
    with open(os.path.join(GOLD, "PIN.json"), "w") as f:
        json.dump(pin, f, indent=1, sort_keys=True)
    print(f"worst normalised deviation port vs reference: {worst:.3e}")
    assert worst <= 2e-5, "oracle port does not reproduce the reference"


if __name__ == "__main__":
    main()
