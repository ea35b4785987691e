"""Stage-by-stage parity of the CUDA path against the golden intermediates of the live reference
(tests/golden/tiny.safetensors).  Each stage is fed the *reference's* inputs, so errors do not
accumulate and a failing stage is pinpointed.  Used by tests/test_stages_gpu.py and runnable as
``python -m tests.stage_runner [precision]`` (prints every error, writes gpurun_out/stages_*.json)."""
import json
import math
import os
import sys

import torch

from tests.util import ROOT, fixture_sds, gold, gpu_nets, maxabs


def _wrapdiff(a, b):
    d = (a.double().cpu() - b.double().cpu()).abs()
    return torch.minimum(d, (d - 2 * math.pi).abs())


def oracle_case(tokens, dur, seed):
    """A golden-style dict for an arbitrary case, produced by the CPU oracle port at run time
    (used for sizes whose intermediates are too large to commit, e.g. the 5 s C1 utterance)."""
    from oracle import stylish_port as port
    from stylish_tts_b200 import fixture
    from tests.util import case_noise
    noise = case_noise(int(dur.sum()), seed)
    sds = fixture_sds()
    tk, ln = tokens[None], torch.tensor([tokens.numel()])
    taps = {}
    with torch.no_grad():
        ref = port.synthesize(sds, tk, ln, noise, duration=dur.float(), taps=taps)
        ref_x0 = port.synthesize(sds, tk, ln, noise, duration=dur.float(), exact_frame0=True)
    g = {"tokens": tokens, "duration_in": dur.to(torch.int32), "duration_pred": ref["duration"].to(torch.int32)}
    for k in ("audio", "logamp", "phase", "pitch", "energy"):
        g["out." + k] = ref[k].float().contiguous()
    g["out.audio_x0"] = ref_x0["audio"].contiguous()
    for k, v in taps.items():
        g["tap." + k] = v.float().contiguous()
    for k, v in noise.items():
        g["noise." + k] = v.contiguous()
    return g


def run(precision="fp32", g=None, tag=""):
    g = gold("tiny") if g is None else g
    dev = "cuda:0"
    nets = gpu_nets()
    sp, pep = nets["speech_predictor"], nets["pitch_energy_predictor"]
    eng = sp.engine()
    eng.set_precision(precision)
    peng = pep.engine()
    peng.set_precision(precision)
    res = {}

    def rec(name, got, want, wrap=False):
        want = want.reshape(got.shape) if want.numel() == got.numel() else want
        d = _wrapdiff(got, want) if wrap else (got.double().cpu() - want.double().cpu()).abs()
        res[name] = {"max_abs": float(d.max()), "ref_absmax": float(want.abs().max()),
                     "rel": float(d.max()) / max(1e-12, float(want.abs().max())),
                     "frac_gt_1e-3": float((d > 1e-3).double().mean())}

    dur = g["duration_in"].to(torch.int64)
    F_ = int(dur.sum())
    F4 = 4 * F_
    tokens = g["tokens"].to(dev)[None]
    lens = torch.tensor([tokens.shape[1]], device=dev)
    idx = torch.repeat_interleave(torch.arange(dur.numel()), dur).to(dev)
    alignment = torch.zeros(1, dur.numel(), F_, device=dev)
    alignment[0, idx, torch.arange(F_, device=dev)] = 1

    style = g["tap.sp_style"].to(dev).contiguous()
    init = g["noise.phase"].to(dev).reshape(-1)[:1].contiguous()

    def sec_length_regulation():  # length regulation
        from stylish_tts_b200.modules import DurationProcessor
        proc = DurationProcessor(16, 50).to(dev)
        d_pred = proc.prediction_to_duration(g["tap.dur_logits"][0].to(dev))[: tokens.shape[1]]
        res["duration_pred"] = {"max_abs": float((d_pred.cpu().to(torch.int32) - g["duration_pred"]).abs().max()),
                                "ref_absmax": float(g["duration_pred"].max()), "rel": 0.0, "frac_gt_1e-3": 0.0}
        al = proc.duration_to_alignment(dur.to(dev).float())
        res["alignment"] = {"max_abs": float((al - alignment[0]).abs().max()), "ref_absmax": 1.0, "rel": 0.0,
                            "frac_gt_1e-3": 0.0}

    def sec_text_encoders():  # native token-rate text encoders (stts_text_encoder) + duration logits
        sp.text_encoder.attach_native(eng, "speech_predictor")
        enc, _, _ = sp.text_encoder(tokens, lens)
        rec("sp_text", enc, g["tap.sp_text"])
        pe_enc, _, _ = nets["pe_text_encoder"](tokens, lens)
        rec("pe_text", pe_enc, g["tap.pe_text"])
        rec("dur_logits", nets["duration_predictor"](tokens, lens), g["tap.dur_logits"])
        pep.prosody_encoder.attach_native(peng, "pitch_energy_predictor")
        rec("pe_prosody", pep.prosody_encoder(g["tap.pe_text"].to(dev), g["tap.pe_style"].to(dev), lens),
            g["tap.pe_prosody"])
        # native style encoders (stts_text_style_encoder) on the reference's text encodings
        sp.style_encoder.attach_native(eng, "speech_predictor")
        rec("sp_style", sp.style_encoder(g["tap.sp_text"].to(dev), lens), g["tap.sp_style"])
        rec("pe_style", nets["pe_text_style_encoder"](g["tap.pe_text"].to(dev), lens), g["tap.pe_style"])
        # native cross-attention block (stts_pe_cross) on the reference's prosody encoding
        peng.set_token_layout([tokens.shape[1]])
        peng.set_layout([F_], None)
        cu_f = torch.tensor([0, F_], dtype=torch.int32, device=dev)
        pros_tok = g["tap.pe_prosody"].to(dev).reshape(tokens.shape[1], -1).contiguous()
        xc = peng.pe_cross(pros_tok, g["tap.pe_style"].to(dev), idx.to(torch.int32).contiguous(), cu_f,
                           want_dense=True)
        rec("pe_cross", xc, g["tap.pe_cross"])

    def sec_pitch___energy_towers_on_the_reference_s():  # pitch / energy towers on the reference's cross-attended frames
        peng.set_layout([F_], None)
        pitch, energy = peng.pe_towers(g["tap.pe_cross"].to(dev).contiguous().view(-1), g["tap.pe_style"].to(dev))
        rec("pe_towers.pitch", pitch, g["out.pitch"])
        rec("pe_towers.energy", energy, g["out.energy"])
        # full module (token-rate torch + native towers)
        p2, e2 = pep(g["tap.pe_text"].to(dev), lens, alignment, g["tap.pe_style"].to(dev))
        rec("PitchEnergyPredictor.pitch", p2, g["out.pitch"])

    def sec_decoder():  # decoder
        asr = (g["tap.sp_text"].to(dev) @ alignment.repeat_interleave(4, dim=2)).contiguous()
        x, _ = sp.decoder(asr, g["tap.pitch4"].to(dev), g["tap.energy4"].to(dev), g["tap.sp_style"].to(dev))
        rec("decoder", x, g["tap.dec_out"])

    def sec_prior___flow___post_flow():  # prior + flow + post_flow
        eng.set_layout(None, [F4])
        mel, zp, zf = eng.prior_flow(g["tap.dec_out"].to(dev).contiguous().view(-1),
                                     g["noise.prior"].to(dev).contiguous().view(-1), style.view(-1), want_z=True)
        rec("z_prior", zp, g["tap.z_prior"])
        rec("z_flow", zf, g["tap.z_flow"])
        rec("mel", mel, g["tap.mel"])

    def sec_source():  # source
        src = eng.source_pcph(g["tap.pitch4"].to(dev).contiguous().view(-1),
                              g["noise.source"].to(dev).contiguous().view(-1), init)
        rec("source", src, g["tap.source"])

    def sec_generator__fed_the_reference_mel_():  # generator (fed the reference mel)
        audio, la, ph = eng.generator(g["tap.mel"].to(dev).contiguous().view(-1), style.view(-1),
                                      g["tap.pitch4"].to(dev).contiguous().view(-1),
                                      g["noise.source"].to(dev).contiguous().view(-1), init)
        rows, starts = eng.layout_info(4)
        s0 = starts[0]
        hs = eng.debug_read("HS")[s0: s0 + F4, :1025].t()
        hp = eng.debug_read("HP")[s0: s0 + F4, :1025].t()
        catp = eng.debug_read("CATP")[s0: s0 + F4]
        sd = {k: v.to(dev).double() for k, v in fixture_sds()["speech_predictor"].items() if k.startswith("generator.")}
        F = torch.nn.functional
        lp = F.conv1d(g["tap.har_spec"].to(dev).double(), sd["generator.amp_prior_conv.weight"],
                      sd["generator.amp_prior_conv.bias"], padding=3)[0]
        pp = F.conv1d(g["tap.har_phase"].to(dev).double(), sd["generator.phase_prior_conv.weight"],
                      sd["generator.phase_prior_conv.bias"], padding=3)[0]
        rec("logamp_prior", catp[:, 512:768].t(), lp)
        rec("phase_prior", catp[:, 768:1024].t(), pp)
        # frame 0 of the reference's phase is FFT rounding noise; the k7 conv spreads it over
        # output frames 0..3 -- from frame 4 on the prior must match
        rec("phase_prior_from4", catp[4:, 768:1024].t(), pp[:, 4:])
        d_un = (hp.double().cpu() - g["tap.har_phase"][0].double()).abs()
        res["har_phase.unwrapped_flips"] = {"max_abs": float((d_un > 3.0).sum()), "ref_absmax": float(d_un.numel()),
                                            "rel": float((d_un > 3.0).double().mean()), "frac_gt_1e-3": 0.0}
        rec("har_spec", hs, g["tap.har_spec"][0])
        rec("har_phase(wrapped)", hp, g["tap.har_phase"][0], wrap=True)
        rec("gen_convnext", eng.debug_read("XG")[s0: s0 + F4, :512].t(), g["tap.gen_convnext"][0])
        rec("gen.logamp", la.view(1025, F4 + 1), g["out.logamp"][0])
        rec("gen.phase", ph.view(1025, F4 + 1), g["out.phase"][0])
        rec("gen.audio", audio, g["out.audio"])
        rec("gen.audio_x0", audio, g["out.audio_x0"])
        rec("gen.audio_after_7500", audio[7500:], g["out.audio"].reshape(-1)[7500:])

    def sec_generator_from_reference_stft():  # strict: generator fed the reference's own |X| / phase
        audio, la, ph = eng.generator_from_stft(g["tap.mel"].to(dev).contiguous().view(-1), style.view(-1),
                                                g["tap.har_spec"].to(dev).contiguous().view(-1),
                                                g["tap.har_phase"].to(dev).contiguous().view(-1))
        rows, starts = eng.layout_info(4)
        s0 = starts[0]
        rec("strict.gen_convnext", eng.debug_read("XG")[s0: s0 + F4, :512].t(), g["tap.gen_convnext"][0])
        rec("strict.logamp", la.view(1025, F4 + 1), g["out.logamp"][0])
        rec("strict.phase", ph.view(1025, F4 + 1), g["out.phase"][0])
        rec("strict.audio", audio, g["out.audio"])

    def sec_SpeechPredictor_module_end_to_end__nativ():  # SpeechPredictor module end to end (native frame-rate path, reference pitch/energy)
        sp.noise_override = {k[6:]: v for k, v in g.items() if k.startswith("noise.")}
        pred = sp(tokens, lens, alignment, g["out.pitch"].to(dev), g["out.energy"].to(dev))
        sp.noise_override = None
        rec("SpeechPredictor.audio", pred.audio, g["out.audio"])
        rec("SpeechPredictor.audio_x0", pred.audio, g["out.audio_x0"])
        rec("SpeechPredictor.logamp", pred.magnitude, g["out.logamp"])

    with torch.no_grad():
        for fn in (sec_length_regulation, sec_text_encoders, sec_pitch___energy_towers_on_the_reference_s, sec_decoder,
                   sec_prior___flow___post_flow, sec_source, sec_generator__fed_the_reference_mel_,
                   sec_generator_from_reference_stft,
                   sec_SpeechPredictor_module_end_to_end__nativ):
            try:
                fn()
            except Exception as e:  # keep going: one trip to the GPU box should report every stage
                res[fn.__name__ + ".EXCEPTION"] = {"max_abs": float("inf"), "ref_absmax": 0.0, "rel": float("inf"),
                                                   "frac_gt_1e-3": 1.0, "error": repr(e)[:500]}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"stages_{precision}{tag}.json"), "w") as f:
        json.dump(res, f, indent=1)
    return res


if __name__ == "__main__":
    prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
    if len(sys.argv) > 2:          # python -m tests.stage_runner fp32 c1  -> oracle-generated C1 intermediates
        gc = gold(sys.argv[2])
        out = run(prec, oracle_case(gc["tokens"], gc["duration_in"].to(torch.int64), int(gc["seed"])), "_" + sys.argv[2])
    else:
        out = run(prec)
    for k, v in out.items():
        if "error" in v:
            print(prec, k, v["error"])
            continue
        print(f"{prec:7s} {k:32s} max_abs {v['max_abs']:.3e}  rel {v['rel']:.3e}  |ref| {v['ref_absmax']:.3e}  "
              f"frac>1e-3 {v['frac_gt_1e-3']:.2e}")
