"""Batched, mixed-length synthesis: the user-facing call of this package.

The reference synthesises one utterance at a time (validate_duration, train/stage_type.py:483-523;
ExportModel, train/models/export_model.py:35-45).  ``Synthesizer.synthesize`` runs the same chain
for a ragged batch with *per-utterance* semantics (every norm / GRN / STFT edge sees only its own
utterance): token-rate encoders on a padded batch with exact masking, the
frame-rate path on the packed layout of the C library.  One context holds both weight sets.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
from torch import nn

from .config import ModelConfig
from .engine import Engine
from .modules.frame_rate import build_model


class Synthesizer:
    def __init__(self, nets: Optional[Dict[str, nn.Module]] = None, device="cuda:0", precision: str = "fp32",
                 model_config=None):
        self.device = torch.device(device)
        self.nets = nets if nets is not None else build_model(model_config or ModelConfig())
        for m in self.nets.values():
            m.to(self.device).eval()
            for p in m.parameters():
                p.requires_grad_(False)
        self.engine = Engine(self.device)
        self.reload_weights()
        self.engine.set_precision(precision)

    def reload_weights(self):
        self.engine.load_state("speech_predictor", self.nets["speech_predictor"].state_dict())
        self.engine.load_state("pitch_energy_predictor", self.nets["pitch_energy_predictor"].state_dict())
        self.engine.load_state("pe_text_encoder", self.nets["pe_text_encoder"].state_dict())
        self.engine.load_state("duration_predictor", self.nets["duration_predictor"].state_dict())
        self.engine.finalize()
        self.nets["pe_text_encoder"].attach_native(self.engine, "pe_text_encoder")
        self.nets["duration_predictor"].attach_engine(self.engine)
        self.nets["speech_predictor"].attach_engine(self.engine)
        self.nets["pitch_energy_predictor"].attach_engine(self.engine)

    def set_precision(self, precision: str):
        self.engine.set_precision(precision)

    # ------------------------------------------------------------------ the call
    @torch.no_grad()
    def synthesize(self, tokens: Sequence[torch.Tensor], durations: Optional[Sequence[torch.Tensor]] = None,
                   noise: Optional[Sequence[Dict[str, torch.Tensor]]] = None, want_spec: bool = False,
                   want_intermediates: bool = False, frames: Optional[Sequence[int]] = None,
                   pitch_energy: Optional[Sequence] = None) -> Dict[str, object]:
        """tokens[b]: int64 [T_b] (host or device).  durations[b]: optional integer durations
        overriding the predicted ones (they are still predicted and returned).  noise[b]:
        optional dict(prior [1,128,4F], source [1,1,300F], phase [1,1]) -- the three reference
        draws; drawn with torch on the device otherwise.  pitch_energy[b] = (pitch [F], energy [F]):
        optional curves fed to the speech stage instead of the predicted ones (the predicted ones are
        still computed and returned) -- the harmonic source integrates f0 over the whole utterance, so
        waveform parity is only defined for identical pitch input (DESIGN.md "conditioning").
        Returns dict(audio=[...[300 F_b]], duration_pred=[...], pitch=[...], energy=[...], ...)."""
        dev, eng, nets = self.device, self.engine, self.nets
        B = len(tokens)
        T = [int(t.numel()) for t in tokens]
        Tmax = max(T)
        tok = torch.zeros((B, Tmax), dtype=torch.int64, device=dev)
        for b, t in enumerate(tokens):
            tok[b, : T[b]] = t.to(dev, non_blocking=True)
        lens = torch.tensor(T, dtype=torch.int64, device=dev)

        # ---- durations (duration_predictor + prediction_to_duration)
        logits = nets["duration_predictor"](tok, lens, host_lengths=T)                      # [B, Tmax, 16]
        dur_all = eng.durations(logits.reshape(B * Tmax, 16)).view(B, Tmax)  # int32
        valid = torch.arange(Tmax, device=dev)[None, :] < lens[:, None]
        dur_pred = dur_all * valid
        if durations is not None:
            dur_used = torch.zeros((B, Tmax), dtype=torch.int32, device=dev)
            for b, d in enumerate(durations):
                dur_used[b, : T[b]] = d.to(dev, torch.int32, non_blocking=True)
            # host-known frame counts: no device sync (pass `frames` when the durations live on the device)
            F_list = [int(f) for f in frames] if frames is not None else [int(d.sum()) for d in durations]
        else:
            dur_used = dur_pred
            F_list = [int(v) for v in dur_pred.sum(dim=1).tolist()]          # the one unavoidable D2H
        F_tot = sum(F_list)
        cu_tok = torch.arange(B + 1, device=dev, dtype=torch.int32) * Tmax    # token offsets into the padded batch
        cu_frames, idx, _ = eng.frame_index(dur_used.contiguous().view(-1), cu_tok, F_tot)
        eng.set_layout(F_list, [4 * f for f in F_list])

        # ---- pitch / energy
        pe_enc, _, _ = nets["pe_text_encoder"](tok, lens, host_lengths=T)
        pe_style = nets["pe_text_style_encoder"](pe_enc, lens)
        pep = nets["pitch_energy_predictor"]
        pep.prosody_encoder.attach_native(eng, "pitch_energy_predictor")
        prosody = pep.prosody_encoder(pe_enc, pe_style, lens, host_lengths=T)
        Fmax = max(F_list)
        f_lens = torch.tensor(F_list, dtype=torch.int64, device=dev)
        # padded [B, Fmax] frame -> token map from the packed index vector
        pos = torch.arange(Fmax, device=dev)[None, :]
        gather = (cu_frames[:-1].to(torch.int64)[:, None] + pos).clamp_(max=max(F_tot - 1, 0))
        frame_token = torch.where(pos < f_lens[:, None], idx[:F_tot].to(torch.int64)[gather], torch.zeros_like(gather))
        x = pep.compute_cross(prosody, frame_token, pe_style, lens, frame_lengths=f_lens)   # [B, 320, Fmax]
        x_dense = torch.cat([x[b, :, : F_list[b]].reshape(-1) for b in range(B)])
        pitch, energy = eng.pe_towers(x_dense, pe_style.contiguous())
        pitch_in, energy_in = pitch, energy
        if pitch_energy is not None:
            pitch_in = torch.cat([pe[0].reshape(-1) for pe in pitch_energy]).to(dev, torch.float32).contiguous()
            energy_in = torch.cat([pe[1].reshape(-1) for pe in pitch_energy]).to(dev, torch.float32).contiguous()

        # ---- speech
        sp = nets["speech_predictor"]
        sp.text_encoder.attach_native(eng, "speech_predictor")
        enc, _, _ = sp.text_encoder(tok, lens, host_lengths=T)
        style = sp.style_encoder(enc, lens).contiguous()
        enc_tok = enc.transpose(1, 2).contiguous().view(B * Tmax, 128)
        if noise is not None:
            n_prior = torch.cat([n["prior"].reshape(-1) for n in noise]).to(dev)
            n_src = torch.cat([n["source"].reshape(-1) for n in noise]).to(dev)
            n_phase = torch.cat([n["phase"].reshape(-1)[:1] for n in noise]).to(dev).contiguous()
        else:
            n_prior = torch.randn(128 * 4 * F_tot, device=dev)
            n_src = torch.randn(300 * F_tot, device=dev)
            n_phase = torch.rand(B, device=dev)
        audio, la, ph = eng.speech_frames(enc_tok, cu_tok, idx, cu_frames, pitch_in, energy_in, style, n_prior, n_src,
                                          n_phase, want_spec=want_spec)
        out: Dict[str, object] = {}
        a_off = [0]
        for f in F_list:
            a_off.append(a_off[-1] + 300 * f)
        out["audio_flat"] = audio
        out["audio_offsets"] = a_off
        out["audio"] = [audio[a_off[b]: a_off[b + 1]] for b in range(B)]
        out["frames"] = F_list
        out["duration_pred"] = [dur_pred[b, : T[b]] for b in range(B)]
        f_off = [0]
        for f in F_list:
            f_off.append(f_off[-1] + f)
        out["pitch"] = [pitch[f_off[b]: f_off[b + 1]] for b in range(B)]
        out["energy"] = [energy[f_off[b]: f_off[b + 1]] for b in range(B)]
        if want_spec:
            s_off = [0]
            for f in F_list:
                s_off.append(s_off[-1] + 1025 * (4 * f + 1))
            out["logamp"] = [la[s_off[b]: s_off[b + 1]].view(1025, 4 * F_list[b] + 1) for b in range(B)]
            out["phase"] = [ph[s_off[b]: s_off[b + 1]].view(1025, 4 * F_list[b] + 1) for b in range(B)]
        if want_intermediates:
            out["pe_cross"] = x
            out["pe_style"], out["sp_style"], out["sp_text"], out["pe_text"] = pe_style, style, enc, pe_enc
            out["dur_logits"] = logits
        return out

    def check(self):
        """Raises RuntimeError if an utterance hit the reference's torch.min-of-empty error path."""
        self.engine.check_source_status()
