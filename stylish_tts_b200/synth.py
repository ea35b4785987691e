"""Batched, mixed-length synthesis: the user-facing call of this package.

The reference synthesises one utterance at a time (validate_duration, train/stage_type.py:483-523;
ExportModel, train/models/export_model.py:35-45).  ``Synthesizer.synthesize`` runs the same chain
for a ragged batch with *per-utterance* semantics (every norm / GRN / STFT edge sees only its own
utterance): the token-rate encoders on the packed token layout and the
frame-rate path on the packed frame layouts of the C library.  One context holds every weight set;
between the input tokens and the output waveform every kernel is a libstts_b200 launch (torch
only draws the three noise tensors).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence

import torch
from torch import nn

from .config import ModelConfig
from .engine import Engine
from .modules.frame_rate import build_model
from .modules.token_rate import TokenPack


_CHAIN_ORDER = tuple(os.environ.get("STTS_CHAIN_ORDER", "pe,dp,sp").split(","))


class Synthesizer:
    def __init__(self, nets: Optional[Dict[str, nn.Module]] = None, device="cuda:0", precision: str = "fp32",
                 model_config=None, use_graphs: bool = False):
        self.device = torch.device(device)
        self.nets = nets if nets is not None else build_model(model_config or ModelConfig())
        for m in self.nets.values():
            m.to(self.device).eval()
            for p in m.parameters():
                p.requires_grad_(False)
        self.engine = Engine(self.device)
        self._side_streams = (torch.cuda.Stream(self.device), torch.cuda.Stream(self.device))
        self.use_graphs = use_graphs
        self.model_config = model_config or ModelConfig()
        self._graphs, self._packs, self._finfo = {}, {}, {}
        self._layout_key = None        # (token counts, frame counts) the engine workspace is planned for
        self._last_index = None        # (status, cu_frames, frame counts) of the last eager frame_index call
        self.reload_weights()
        self.engine.set_precision(precision)

    def reload_weights(self):
        self._graphs.clear() if hasattr(self, "_graphs") else None
        self.engine.load_state("speech_predictor", self.nets["speech_predictor"].state_dict())
        self.engine.load_state("pitch_energy_predictor", self.nets["pitch_energy_predictor"].state_dict())
        self.engine.load_state("pe_text_encoder", self.nets["pe_text_encoder"].state_dict())
        self.engine.load_state("pe_text_style_encoder", self.nets["pe_text_style_encoder"].state_dict())
        self.engine.load_state("duration_predictor", self.nets["duration_predictor"].state_dict())
        self.engine.finalize()
        self.nets["pe_text_encoder"].attach_native(self.engine, "pe_text_encoder")
        self.nets["pe_text_style_encoder"].attach_native(self.engine, "pe_text_style_encoder")
        self.nets["duration_predictor"].attach_engine(self.engine)
        self.nets["speech_predictor"].attach_engine(self.engine)
        self.nets["pitch_energy_predictor"].attach_engine(self.engine)

    def set_precision(self, precision: str):
        self.engine.set_precision(precision)
        self._graphs.clear()          # captured graphs bake the kernel variants of the old precision

    # ------------------------------------------------------------------ the call
    @torch.no_grad()
    def synthesize(self, tokens: Sequence[torch.Tensor], durations: Optional[Sequence[torch.Tensor]] = None,
                   noise: Optional[Sequence[Dict[str, torch.Tensor]]] = None, want_spec: bool = False,
                   want_intermediates: bool = False, frames: Optional[Sequence[int]] = None,
                   pitch_energy: Optional[Sequence] = None, graph: Optional[bool] = None) -> Dict[str, object]:
        """tokens[b]: int64 [T_b] (host or device).  durations[b]: optional integer durations
        overriding the predicted ones (they are still predicted and returned).  noise[b]:
        optional dict(prior [1,128,4F], source [1,1,300F], phase [1,1]) -- the three reference
        draws; drawn with torch on the device otherwise.  pitch_energy[b] = (pitch [F], energy [F]):
        optional curves fed to the speech stage instead of the predicted ones (the predicted ones are
        still computed and returned) -- the harmonic source integrates f0 over the whole utterance, so
        waveform parity is only defined for identical pitch input (DESIGN.md "conditioning").
        graph: replay a CUDA graph captured for this batch shape (token counts, frame counts).  Needs
        host-known frame counts (``durations`` + ``frames``) and torch-drawn noise; default = the
        Synthesizer's ``use_graphs`` setting.  Outputs of a graph call live in the graph's static
        buffers: consume (or clone) them before the next call with the same shape.
        Returns dict(audio=[...[300 F_b]], duration_pred=[...], pitch=[...], energy=[...], ...)."""
        dev = self.device
        T = [int(t.numel()) for t in tokens]
        n_vocab = int(self.model_config.text_encoder.tokens)
        for b, t in enumerate(tokens):           # nn.Embedding raises IndexError on an out-of-range id
            if not t.is_cuda and t.numel() and (int(t.min()) < 0 or int(t.max()) >= n_vocab):
                raise IndexError(f"utterance {b}: token id outside [0, {n_vocab})")
        pack = self._token_pack(T, max(T))
        use_graph = self.use_graphs if graph is None else graph
        use_graph = use_graph and durations is not None and noise is None and pitch_energy is None and not want_intermediates
        F_list = None
        if durations is not None:
            # host-known frame counts: no device sync (pass `frames` when the durations live on the device)
            F_list = [int(f) for f in frames] if frames is not None else [int(d.sum()) for d in durations]
        if use_graph and not any(t.is_cuda for t in tokens) and not any(d.is_cuda for d in durations):
            # inputs on the host: they are written into the graph's pinned staging buffers and the H2D copies are
            # nodes of the graph -- no device work outside the replay
            res = self._graphed(None, pack, None, F_list, want_spec, host=(tokens, durations))
            return self._package(res, T, want_spec)
        # one H2D copy of the concatenated tokens: every stage works on the packed [sum T] token order
        tok = torch.cat([t.reshape(-1) for t in tokens]).to(dev, torch.int32, non_blocking=True)
        dur_used = None
        if durations is not None:
            dur_used = torch.cat([d.reshape(-1) for d in durations]).to(dev, torch.int32, non_blocking=True)
        if use_graph:
            res = self._graphed(tok, pack, dur_used, F_list, want_spec)
        else:
            res = self._core(tok, pack, dur_used, F_list, noise, pitch_energy, want_spec, want_intermediates)
        return self._package(res, T, want_spec)

    @torch.no_grad()
    def predict_durations(self, tokens: Sequence[torch.Tensor]) -> List[torch.Tensor]:
        """DurationPredictor.forward + DurationProcessor.prediction_to_duration (duration_predictor.py:30-36,
        utils.py:468-474) for a ragged batch: int32 durations per utterance, on the device.  The sharded
        path (shard.synthesize_sharded) needs the frame counts before it can partition the batch."""
        T = [int(t.numel()) for t in tokens]
        pack = self._token_pack(T, max(T))
        tok = torch.cat([t.reshape(-1) for t in tokens]).to(self.device, torch.int32, non_blocking=True)
        eng, mc = self.engine, self.model_config
        eng.set_token_layout(pack.lengths)
        if self._layout_key is not None and self._layout_key[0] != tuple(pack.lengths):
            self._layout_key = None
            self._graphs.clear()
        dp_mu, _ = eng.text_encoder("duration_predictor", tok, mc.inter_dim, want_x=False)
        dp_style = eng.text_style_encoder("duration_predictor", None)
        _, logits = eng.prosody_encoder("duration_predictor", dp_mu, dp_style, mc.inter_dim + mc.style_dim,
                                        want_logits=True, want_out=False)
        dur = eng.durations(logits)
        off = [0]
        for n in T:
            off.append(off[-1] + n)
        return [dur[off[b]: off[b + 1]] for b in range(len(T))]

    def _token_pack(self, T, Tmax) -> TokenPack:
        key = (tuple(T), Tmax)
        if key not in self._packs:
            if len(self._packs) > 64:
                self._packs.clear()
            pk = TokenPack(T, Tmax, self.device)
            pk.lens = torch.tensor(T, dtype=torch.int64).to(self.device)
            cu = [0]
            for n in T:
                cu.append(cu[-1] + n)
            pk.cu = torch.tensor(cu, dtype=torch.int32).to(self.device)          # token offsets of the packed order
            pk.seg_id = torch.repeat_interleave(torch.arange(len(T)), torch.tensor(T)).to(self.device)
            self._packs[key] = pk
        return self._packs[key]

    def _frame_info(self, F_list):
        key = tuple(F_list)
        if key not in self._finfo:
            if len(self._finfo) > 64:
                self._finfo.clear()
            self._finfo[key] = torch.tensor(F_list, dtype=torch.int64).to(self.device)
        return self._finfo[key]

    def _graphed(self, tok, pack, dur_used, F_list, want_spec, host=None):
        """One CUDA graph per batch shape: the whole step (three token-rate chains on forked streams,
        length regulation, frame-rate path) is replayed with no host work between kernels.
        host = (tokens, durations) on the CPU: the graph starts with the H2D copies from its own pinned staging
        buffers (rewritten here once the previous replay of this graph has finished)."""
        key = (tuple(pack.lengths), tuple(F_list), bool(want_spec), host is not None)
        # A captured graph bakes the arena pointers, layout tables and TMA maps of the workspace plan it was
        # captured under, and the context holds ONE plan: any call with another shape (graph or eager)
        # re-plans it.  Graphs of other shapes are therefore dropped as soon as the plan moves on.
        if self._layout_key != key[:2]:
            self._graphs.clear()
        ent = self._graphs.get(key)
        if ent is None:
            if len(self._graphs) >= 8:
                self._graphs.clear()
            if host is not None:
                n_tok = sum(pack.lengths)
                h_tok = torch.empty(n_tok, dtype=torch.int32).pin_memory()
                h_dur = torch.empty(n_tok, dtype=torch.int32).pin_memory()
                self._stage_host(h_tok, h_dur, host)
                s_tok, s_dur = h_tok.to(self.device), h_dur.to(self.device)
            else:
                s_tok, s_dur = tok.clone(), dur_used.clone()
            self._frame_info(F_list)
            cur = torch.cuda.current_stream(self.device)
            warm = torch.cuda.Stream(self.device)
            warm.wait_stream(cur)
            with torch.cuda.stream(warm):        # allocations, layouts, lazy attribute setup happen here
                for _ in range(2):
                    self._core(s_tok, pack, s_dur, F_list, None, None, want_spec, False)
            cur.wait_stream(warm)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                if host is not None:
                    s_tok.copy_(h_tok, non_blocking=True)
                    s_dur.copy_(h_dur, non_blocking=True)
                res = self._core(s_tok, pack, s_dur, F_list, None, None, want_spec, False)
            ent = (g, s_tok, s_dur, res) if host is None else (g, h_tok, h_dur, res, [torch.cuda.Event(), False])
            self._graphs[key] = ent
        if host is not None:
            g, h_tok, h_dur, res, busy = ent
            if busy[1]:
                busy[0].synchronize()            # the previous replay has read the staging buffers
            self._stage_host(h_tok, h_dur, host)
            g.replay()
            busy[0].record()
            busy[1] = True
            return res
        g, s_tok, s_dur, res = ent
        s_tok.copy_(tok)
        s_dur.copy_(dur_used)
        g.replay()
        return res

    @staticmethod
    def _stage_host(h_tok, h_dur, host):
        tokens, durations = host
        if len(tokens) == 1:
            h_tok.copy_(tokens[0].reshape(-1))
            h_dur.copy_(durations[0].reshape(-1))
        else:
            h_tok.copy_(torch.cat([t.reshape(-1) for t in tokens]))
            h_dur.copy_(torch.cat([d.reshape(-1) for d in durations]))

    def _core(self, tok, pack, dur_used, F_list, noise, pitch_energy, want_spec, want_intermediates):
        """The device work of one step: every stage is a libstts_b200 call on the packed layouts.
        tok / dur_used: int32 [sum T] in packed token order.  Free of host synchronisation when
        F_list is given."""
        dev, eng = self.device, self.engine
        mc = self.model_config
        B = len(pack.lengths)
        capturing = torch.cuda.is_current_stream_capturing()
        eng.set_token_layout(pack.lengths)
        # The duration, pitch/energy and speech chains only share the tokens: enqueue them on three
        # streams (their kernels are small and latency-bound at token rate) and join before the
        # frame-rate path.  The C library keeps one scratch set per chain (token_pipeline.cu).
        cur = torch.cuda.current_stream(dev)
        s_dp, s_sp = self._side_streams
        s_dp.wait_stream(cur)
        s_sp.wait_stream(cur)

        pe_inter = mc.pitch_energy_predictor.inter_dim
        out = {}

        def chain_dp():      # durations: text encoder -> style -> prosody + duration_proj -> argmax rule   [stream s_dp]
            with torch.cuda.stream(s_dp):
                dp_mu, _ = eng.text_encoder("duration_predictor", tok, mc.inter_dim, want_x=False)
                dp_style = eng.text_style_encoder("duration_predictor", None)
                _, logits = eng.prosody_encoder("duration_predictor", dp_mu, dp_style, mc.inter_dim + mc.style_dim,
                                                want_logits=True, want_out=False)
                out["dp"] = (dp_mu, dp_style, logits, eng.durations(logits))        # durations: int32 [sum T]

        def chain_sp():      # speech text / style encoders                                                [stream s_sp]
            with torch.cuda.stream(s_sp):
                enc_tok, _ = eng.text_encoder("speech_predictor", tok, mc.inter_dim, want_x=False)
                out["sp"] = (enc_tok, eng.text_style_encoder("speech_predictor", None))

        def chain_pe():      # pitch / energy token-rate part                                            [current stream]
            pe_mu, _ = eng.text_encoder("pe_text_encoder", tok, pe_inter, want_x=False)
            pe_style = eng.text_style_encoder("pe_text_style_encoder", None)
            prosody, _ = eng.prosody_encoder("pitch_energy_predictor", pe_mu, pe_style, pe_inter + mc.style_dim,
                                             want_out=want_intermediates)
            out["pe"] = (pe_mu, pe_style, prosody)

        # Enqueue order = node order of a captured graph, and a replay releases the branches in that order, tens of
        # microseconds apart: the two chains the length regulation waits for go first, the speech encoders (needed
        # only after the pitch towers) last.
        for name in _CHAIN_ORDER:
            {"dp": chain_dp, "sp": chain_sp, "pe": chain_pe}[name]()
        dp_mu, dp_style, logits, dur_pred = out["dp"]
        enc_tok, style = out["sp"]
        pe_mu, pe_style, prosody = out["pe"]

        cur.wait_stream(s_dp)
        if not capturing:
            for t_ in (dp_mu, dp_style, logits, dur_pred):
                t_.record_stream(cur)
        if dur_used is None:
            dur_used = dur_pred
            per_utt = torch.zeros(B, dtype=torch.int64, device=dev).index_add_(0, pack.seg_id, dur_pred.to(torch.int64))
            F_list = [int(v) for v in per_utt.tolist()]                         # the one unavoidable D2H
        F_tot = sum(F_list)
        cu_frames, idx, status = eng.frame_index(dur_used, pack.cu, F_tot)
        if not capturing:
            self._last_index = (status, cu_frames, list(F_list))
        eng.set_layout(F_list, [4 * f for f in F_list])
        if self._layout_key != (tuple(pack.lengths), tuple(F_list)):
            self._layout_key = (tuple(pack.lengths), tuple(F_list))
            if not capturing:
                self._graphs.clear()

        # ---- pitch / energy frame-rate part: cross attention over the alignment, then the two towers
        x = eng.pe_cross(None, pe_style, idx, cu_frames, want_dense=want_intermediates)
        pitch, energy = eng.pe_towers(None, pe_style)
        pitch_in, energy_in = pitch, energy
        if pitch_energy is not None:
            pitch_in = torch.cat([pe[0].reshape(-1).to(dev, torch.float32) for pe in pitch_energy]).contiguous()
            energy_in = torch.cat([pe[1].reshape(-1).to(dev, torch.float32) for pe in pitch_energy]).contiguous()

        # ---- speech frame-rate path
        cur.wait_stream(s_sp)
        if not capturing:
            for t_ in (enc_tok, style):
                t_.record_stream(cur)
        if noise is not None:
            n_prior = torch.cat([n["prior"].reshape(-1) for n in noise]).to(dev)
            n_src = torch.cat([n["source"].reshape(-1) for n in noise]).to(dev)
            n_phase = torch.cat([n["phase"].reshape(-1)[:1] for n in noise]).to(dev).contiguous()
        else:
            n_prior = torch.randn(128 * 4 * F_tot, device=dev)
            n_src = torch.randn(300 * F_tot, device=dev)
            n_phase = torch.rand(B, device=dev)
        audio, la, ph = eng.speech_frames(enc_tok, pack.cu, idx, cu_frames, pitch_in, energy_in, style, n_prior, n_src,
                                          n_phase, want_spec=want_spec)
        res = dict(audio=audio, la=la, ph=ph, dur_pred=dur_pred, pitch=pitch, energy=energy, frames=F_list)
        if want_intermediates:   # token-major [sum T, C] / flat dense [320, F_b] per utterance
            res.update(pe_cross=x, pe_style=pe_style, sp_style=style, sp_text=enc_tok, pe_text=pe_mu,
                       pe_prosody=prosody, dur_logits=logits)
        return res

    @staticmethod
    def _package(res, T, want_spec) -> Dict[str, object]:
        B, F_list = len(T), res["frames"]
        audio, la, ph, dur_pred, pitch, energy = (res[k] for k in ("audio", "la", "ph", "dur_pred", "pitch", "energy"))
        out: Dict[str, object] = {}
        a_off = [0]
        for f in F_list:
            a_off.append(a_off[-1] + 300 * f)
        out["audio_flat"] = audio
        out["audio_offsets"] = a_off
        out["audio"] = [audio[a_off[b]: a_off[b + 1]] for b in range(B)]
        out["frames"] = F_list
        t_off = [0]
        for n in T:
            t_off.append(t_off[-1] + n)
        out["duration_pred"] = [dur_pred[t_off[b]: t_off[b + 1]] for b in range(B)]
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
        for k in ("pe_cross", "pe_style", "sp_style", "sp_text", "pe_text", "pe_prosody", "dur_logits"):
            if k in res:
                out[k] = res[k]
        return out

    def check(self, audit: bool = False):
        """Synchronising post-call validation.  ``audit`` (16-bit mode): also raise OverflowError if any fp16
        operand twin of the last run saturated at +-65504 (a checkpoint outside the validated range).  Raises RuntimeError if an utterance hit the reference's
        torch.min-of-empty error path, ValueError if the durations given to the last eager call do not
        sum to the frame counts it was laid out for (the reference's `text_encoding @ alignment` shape
        error)."""
        self.engine.check_source_status()
        if self._last_index is not None:
            status, cu_frames, F_list = self._last_index
            self._last_index = None
            got = cu_frames.cpu().tolist()
            want = [0]
            for f in F_list:
                want.append(want[-1] + f)
            if int(status.item()) != 0 or got != want:
                raise ValueError(f"durations do not sum to the given frame counts: cumulative frames {got} vs {want}")
        if audit and self.engine.precision == "f16":
            n = self.engine.saturation_count()
            if n:
                raise OverflowError(f"{n} fp16 operand values saturated at +-65504: use precision='tf32x3' for this "
                                    "checkpoint")


class SynthesizerPool:
    """Throughput mode: ``lanes`` independent Synthesizers (own C context, workspace, CUDA graphs and
    streams; the nn.Modules are shared) fed round-robin.  Batches are independent, so the token-rate
    chains of batch i + 1 (small, latency-bound kernels) overlap the frame-rate GEMMs of batch i on the
    same GPU.  ``submit`` enqueues one batch and returns immediately; a lane is reused only after its
    previous batch (including the optional device -> host copy of the waveform) has finished."""

    def __init__(self, nets=None, device="cuda:0", precision: str = "fp32", model_config=None, lanes: int = 2,
                 use_graphs: bool = True):
        self.device = torch.device(device)
        first = Synthesizer(nets, device, precision, model_config, use_graphs)
        self.lanes = [first] + [Synthesizer(first.nets, device, precision, model_config, use_graphs)
                                for _ in range(lanes - 1)]
        self.streams = [torch.cuda.Stream(self.device) for _ in range(lanes)]
        if lanes > 1:       # concurrent lanes: look-ahead CTAs of one lane would take SM slots from the others
            for lane in self.lanes:
                lane.engine.set_pdl(False)
                lane._graphs.clear()
        self._done = [None] * lanes
        self._next = 0

    def submit(self, tokens, durations=None, frames=None, out_host: Optional[torch.Tensor] = None, after=None, **kw):
        """Enqueue one batch on the next lane.  ``out_host``: pinned [sum 300 F] buffer that receives the
        concatenated waveform.  ``after(lane, out)``: extra work enqueued on the lane's stream behind the batch
        (the sharded path's waveform gather).  Returns (result dict, event recorded after the batch's last
        operation)."""
        i = self._next
        self._next = (i + 1) % len(self.lanes)
        if self._done[i] is not None:
            self._done[i].synchronize()          # the lane's static graph buffers / host buffer are free again
        s = self.streams[i]
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            out = self.lanes[i].synthesize(tokens, durations=durations, frames=frames, **kw)
            if out_host is not None:
                out_host.copy_(out["audio_flat"], non_blocking=True)
            if after is not None:
                after(i, out)
            ev = torch.cuda.Event()
            ev.record(s)
        self._done[i] = ev
        return out, ev

    def drain(self):
        """Make the current stream wait for every lane (no host synchronisation)."""
        cur = torch.cuda.current_stream(self.device)
        for s in self.streams:
            cur.wait_stream(s)

    def synchronize(self):
        for ev in self._done:
            if ev is not None:
                ev.synchronize()

    def check(self):
        for lane in self.lanes:
            lane.check()
