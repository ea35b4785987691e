"""Thin torch-facing wrapper over the C ABI: torch supplies device memory and streams, the
library does the work.  Every method enqueues on torch's current stream."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Iterable, Optional, Sequence

import torch

from . import _lib

# "f16": the 16-bit mode -- one-pass tensor-core GEMMs on IEEE fp16 operand twins, fp32 accumulate.  It is
# what the north star calls "bf16 mode" (same tcgen05 kind::f16 rate, 8x finer rounding), but it is NOT
# bfloat16: fp16 saturates at 65504 where bf16 keeps fp32's range, so the name "bf16" is rejected instead
# of being silently aliased (stts_saturation_count / Synthesizer.check(audit=True) report clipped twins).
PRECISIONS = {"fp32": 0, "f16": 1, "tf32x3": 2}


class SttsError(RuntimeError):
    pass


def _ptr(t: Optional[torch.Tensor], dtype=torch.float32) -> Optional[int]:
    if t is None:
        return None
    if not t.is_cuda:
        raise SttsError("expected a CUDA tensor")
    if t.dtype != dtype:
        raise SttsError(f"expected dtype {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise SttsError("expected a contiguous tensor")
    return t.data_ptr()


class Engine:
    """One context per device: prepacked weights + workspace for one packed layout at a time."""

    def __init__(self, device: Optional[torch.device] = None):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise SttsError("no CUDA device: the synthesis path has no CPU fallback")
        self.device = torch.device(device if device is not None else "cuda:0")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        h = C.c_void_p()
        rc = self.lib.stts_create(C.byref(h), self.device.index)
        if rc != 0:
            raise SttsError(f"stts_create failed ({rc}): needs an sm_100 (B200) device")
        self.h = h
        self.n_seg = 0
        self.len_f: Optional[Sequence[int]] = None
        self.len_f4: Optional[Sequence[int]] = None
        self.precision = "fp32"
        self.len_t: Sequence[int] = []

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.lib.stts_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # ------------------------------------------------------------------ helpers
    def _check(self, rc: int):
        if rc < 0:
            msg = self.lib.stts_last_error(self.h).decode()
            if rc == -2:
                raise ValueError(msg)
            if rc == -7:
                raise RuntimeError(msg)  # same type torch.min raises on an empty tensor
            raise SttsError(f"stts error {rc}: {msg}")
        return rc

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    # ------------------------------------------------------------------ weights
    def load_state(self, module: str, sd: Dict[str, torch.Tensor], prefix: str = ""):
        for k, v in sd.items():
            t = v.detach().to("cpu", torch.float32).contiguous()
            shape = (C.c_int64 * max(t.dim(), 1))(*t.shape)
            self._check(self.lib.stts_load_tensor(self.h, module.encode(), (prefix + k).encode(), t.data_ptr(), shape,
                                                  t.dim()))

    def finalize(self):
        # the analysis window of TorchSTFT is a non-persistent buffer of the reference module (generator.py:30), so it
        # is not in any state_dict: hand over torch's own hann_window values (the forward STFT multiplies by them)
        self.load_state("speech_predictor", {"generator.stft.window": torch.hann_window(1200, dtype=torch.float32)})
        self._check(self.lib.stts_finalize_weights(self.h))
        self.n_seg = 0

    def set_precision(self, name: str):
        if name not in PRECISIONS:
            raise ValueError(f"unknown precision {name!r}: choose one of {sorted(PRECISIONS)} "
                             "(the 16-bit mode is 'f16' = IEEE fp16 operand twins; there is no bfloat16 mode)")
        self._check(self.lib.stts_set_precision(self.h, PRECISIONS[name]))
        self.precision = name

    def set_pdl(self, on: bool):
        """Programmatic dependent launch of this context's kernels (default on; see stts_set_pdl)."""
        self._check(self.lib.stts_set_pdl(self.h, 1 if on else 0))

    def set_small_batch_shapes(self, on: bool):
        """Small-batch kernel shapes of this context (default on; see stts_set_small_batch_shapes)."""
        self._check(self.lib.stts_set_small_batch_shapes(self.h, 1 if on else 0))

    # ------------------------------------------------------------------ layout
    def set_layout(self, len_f: Optional[Iterable[int]], len_f4: Optional[Iterable[int]]):
        lf = [int(x) for x in len_f] if len_f is not None else None
        l4 = [int(x) for x in len_f4] if len_f4 is not None else None
        n = len(lf if lf is not None else l4)
        a = (C.c_int * n)(*lf) if lf is not None else None
        b = (C.c_int * n)(*l4) if l4 is not None else None
        with torch.cuda.device(self.device):
            self._check(self.lib.stts_set_layout(self.h, n, a, b))
        self.n_seg, self.len_f, self.len_f4 = n, lf, l4

    def set_token_layout(self, len_t: Iterable[int]):
        lt = [int(x) for x in len_t]
        with torch.cuda.device(self.device):
            self._check(self.lib.stts_set_token_layout(self.h, len(lt), (C.c_int * len(lt))(*lt)))
        self.len_t = lt

    # ------------------------------------------------------------------ token-rate encoders
    def text_encoder(self, which: str, tokens: torch.Tensor, inter_dim: int, want_mu: bool = True, want_x: bool = True):
        """TextEncoder.forward (text_encoder.py:433-462) on the current token layout.  tokens: int32
        [sum T] concatenated.  Returns token-major (mu [sum T, inter_dim], x [sum T, 128])."""
        n = int(tokens.numel())
        if n != sum(self.len_t):
            raise ValueError("token count does not match the token layout")
        tokens = tokens.to(torch.int32).contiguous()
        mu = torch.empty((n, inter_dim), dtype=torch.float32, device=self.device) if want_mu else None
        x = torch.empty((n, 128), dtype=torch.float32, device=self.device) if want_x else None
        self._check(self.lib.stts_text_encoder(self.h, which.encode(), _ptr(tokens, torch.int32), _ptr(mu), _ptr(x),
                                               self._stream()))
        return mu, x

    def prosody_encoder(self, which: str, enc_tok: torch.Tensor, style: torch.Tensor, hid: int,
                        want_logits: bool = False, want_out: bool = True):
        """ProsodyEncoder.forward (prosody_encoder.py:63-81) on the current token layout.  enc_tok
        token-major [sum T, d_model]; style [B, 64].  Returns (out [sum T, hid], logits [sum T, 16] | None)."""
        n = int(enc_tok.shape[0])
        if n != sum(self.len_t):
            raise ValueError("token count does not match the token layout")
        enc_tok = enc_tok.to(torch.float32).contiguous()
        style = style.to(torch.float32).contiguous()
        out = torch.empty((n, hid), dtype=torch.float32, device=self.device) if want_out else None
        logits = torch.empty((n, 16), dtype=torch.float32, device=self.device) if want_logits else None
        self._check(self.lib.stts_prosody_encoder(self.h, which.encode(), _ptr(enc_tok), _ptr(style), _ptr(out),
                                                  _ptr(logits), self._stream()))
        return out, logits

    def text_style_encoder(self, which: str, enc_tok: Optional[torch.Tensor]) -> torch.Tensor:
        """TextStyleEncoder.forward (text_style_encoder.py:20-26) on the current token layout.
        enc_tok token-major [sum T, inter_dim], or None = the mu rows the text encoder of the same
        chain just left in the workspace.  Returns style [B, 64]."""
        if enc_tok is not None:
            if int(enc_tok.shape[0]) != sum(self.len_t):
                raise ValueError("token count does not match the token layout")
            enc_tok = enc_tok.to(torch.float32).contiguous()
        style = torch.empty((len(self.len_t), 64), dtype=torch.float32, device=self.device)
        self._check(self.lib.stts_text_style_encoder(self.h, which.encode(), _ptr(enc_tok), _ptr(style),
                                                     self._stream()))
        return style

    def pe_cross(self, prosody_tok: Optional[torch.Tensor], style: torch.Tensor, idx: torch.Tensor,
                 cu_frames: torch.Tensor, want_dense: bool = False) -> Optional[torch.Tensor]:
        """PitchEnergyPredictor.compute_cross (pitch_energy_predictor.py:83-102) on the current token
        and hop-300 layouts.  The result stays in the workspace for ``pe_towers(None, style)``;
        ``want_dense`` also returns it as flat dense [320, F_b] per utterance."""
        if prosody_tok is not None:
            prosody_tok = prosody_tok.to(torch.float32).contiguous()
        style = style.to(torch.float32).contiguous()
        x = torch.empty(320 * sum(self.len_f), device=self.device) if want_dense else None
        self._check(self.lib.stts_pe_cross(self.h, _ptr(prosody_tok), _ptr(style), _ptr(idx, torch.int32),
                                           _ptr(cu_frames, torch.int32), _ptr(x), self._stream()))
        return x

    def workspace_bytes(self) -> int:
        return int(self.lib.stts_workspace_bytes(self.h))

    def launch_count(self) -> int:
        return int(self.lib.stts_launch_count(self.h))

    # ------------------------------------------------------------------ length regulation
    def durations(self, logits: torch.Tensor) -> torch.Tensor:
        """[n_tokens, 16] logits -> int32 durations (utils.py:468-474)."""
        logits = logits.contiguous()
        n = logits.shape[0]
        out = torch.empty(n, dtype=torch.int32, device=self.device)
        self._check(self.lib.stts_durations(self.h, _ptr(logits), n, _ptr(out, torch.int32), self._stream()))
        return out

    def frame_index(self, dur: torch.Tensor, cu_tok: torch.Tensor, cap: int):
        n_seg = cu_tok.numel() - 1
        cu_frames = torch.empty(n_seg + 1, dtype=torch.int32, device=self.device)
        # zero-filled: if the durations sum to less than `cap` the tail still names a valid token
        idx = torch.zeros(max(cap, 1), dtype=torch.int32, device=self.device)
        status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._check(self.lib.stts_frame_index(self.h, _ptr(dur, torch.int32), _ptr(cu_tok, torch.int32), n_seg,
                                              _ptr(cu_frames, torch.int32), _ptr(idx, torch.int32), cap,
                                              _ptr(status, torch.int32), self._stream()))
        return cu_frames, idx, status

    def frame_index_static(self, dur: torch.Tensor, cu_tok: torch.Tensor, idx: torch.Tensor, status: torch.Tensor):
        """stts_frame_index_static: same indices as ``frame_index`` written into the caller's ``idx`` / ``status``
        (``status`` zeroed by the caller); the frame-offset scratch is a static device array inside the library,
        so neither this method nor the library allocates.  At most 1024 utterances per call; calls that overlap on
        the GPU share the array and race."""
        self._check(self.lib.stts_frame_index_static(self.h, _ptr(dur, torch.int32), _ptr(cu_tok, torch.int32),
                                                     cu_tok.numel() - 1, _ptr(idx, torch.int32), idx.numel(),
                                                     _ptr(status, torch.int32), self._stream()))

    # ------------------------------------------------------------------ variants outside the default model.yml
    def sinegen(self, f0: torch.Tensor, lengths: Sequence[int], rand_ini: torch.Tensor, noise: torch.Tensor,
                merge_weight: torch.Tensor, harmonic_num: int, samp_rate: float, sine_amp: float = 0.1,
                noise_std: float = 0.003, voiced_threshold: float = 0.0, want_phase: bool = False):
        """SineGenerator.forward (cfm/cfm_mel_decoder.py:85-101) on a ragged batch.  f0 [sum L] per-sample Hz, the
        utterances back to back; rand_ini [B, dim]; noise [sum L, dim].  Returns (out [sum L], phase [sum L, dim] | None),
        phase = the wrap-corrected cumulative phase in cycles (cfm_mel_decoder.py:78 before * 2 pi)."""
        lens = [int(x) for x in lengths]
        dim = harmonic_num + 1
        tot = sum(lens)
        if f0.numel() != tot or noise.numel() != tot * dim or rand_ini.numel() != len(lens) * dim:
            raise ValueError("sinegen: f0 / noise / rand_ini do not match the lengths")
        w = merge_weight.detach().to("cpu", torch.float32).reshape(-1).contiguous()
        if w.numel() != dim:
            raise ValueError("sinegen: merge weight must have harmonic_num + 1 entries")
        out = torch.empty(tot, dtype=torch.float32, device=self.device)
        phase = torch.empty((tot, dim), dtype=torch.float32, device=self.device) if want_phase else None
        self._check(self.lib.stts_source_sinegen(
            self.h, _ptr(f0), (C.c_int * len(lens))(*lens), len(lens), _ptr(rand_ini), _ptr(noise),
            (C.c_float * dim)(*w.tolist()), harmonic_num, float(samp_rate), float(sine_amp), float(noise_std),
            float(voiced_threshold), _ptr(out), _ptr(phase), self._stream()))
        return out, phase

    def snake_block(self, module: str, channels: int, kernel_size: int, dilation: Sequence[int], x: torch.Tensor,
                    lengths: Sequence[int], style: torch.Tensor):
        """AdaptiveGeneratorBlock.forward (ada_norm.py:109-120) on a ragged batch; weights loaded under ``module`` with
        ``load_state``.  x: the utterances' dense [channels, L_b] tensors back to back; style [B, 64]."""
        lens = [int(v) for v in lengths]
        if x.numel() != channels * sum(lens) or style.numel() != 64 * len(lens):
            raise ValueError("snake_block: x / style do not match the lengths")
        out = torch.empty_like(x)
        with torch.cuda.device(self.device):
            self._check(self.lib.stts_generator_block_snake(
                self.h, module.encode(), channels, kernel_size, (C.c_int * 3)(*[int(d) for d in dilation]), len(lens),
                (C.c_int * len(lens))(*lens), _ptr(x), _ptr(style), _ptr(out), self._stream()))
        return out

    def stft_conv_transform(self, module: str, n_fft: int, hop: int, win: int, wave: torch.Tensor,
                            lengths: Sequence[int], pad_mode: str = "replicate"):
        """STFT.transform (train/models/stft.py:98-136) on a ragged batch; returns (mag, x, y) flat, per utterance a
        dense [n_fft // 2 + 1, len // hop + 1] block."""
        lens = [int(v) for v in lengths]
        if wave.numel() != sum(lens):
            raise ValueError("stft_conv_transform: waveform does not match the lengths")
        if pad_mode not in ("replicate", "constant"):
            raise ValueError("pad_mode must be 'replicate' or 'constant'")
        bins = n_fft // 2 + 1
        tot = sum(bins * (n // hop + 1) for n in lens)
        mag, x, y = (torch.empty(tot, dtype=torch.float32, device=self.device) for _ in range(3))
        with torch.cuda.device(self.device):
            self._check(self.lib.stts_stft_conv_transform(
                self.h, module.encode(), n_fft, hop, win, 0 if pad_mode == "replicate" else 1, len(lens),
                (C.c_int * len(lens))(*lens), _ptr(wave), _ptr(mag), _ptr(x), _ptr(y), self._stream()))
        return mag, x, y

    def stft_conv_inverse(self, module: str, n_fft: int, hop: int, win: int, mag: torch.Tensor, x: torch.Tensor,
                          y: torch.Tensor, frames: Sequence[int], out_len: Optional[Sequence[int]] = None):
        """STFT.inverse (train/models/stft.py:138-186): flat [bins, frames_b] blocks in, waveforms back to back out."""
        fr = [int(v) for v in frames]
        bins = n_fft // 2 + 1
        if not (mag.numel() == x.numel() == y.numel() == bins * sum(fr)):
            raise ValueError("stft_conv_inverse: inputs do not match the frame counts")
        ol = [int(v) for v in out_len] if out_len is not None else [(f - 1) * hop for f in fr]
        wave = torch.empty(max(sum(ol), 1), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self._check(self.lib.stts_stft_conv_inverse(
                self.h, module.encode(), n_fft, hop, win, len(fr), (C.c_int * len(fr))(*fr), (C.c_int * len(fr))(*ol),
                _ptr(mag), _ptr(x), _ptr(y), _ptr(wave), self._stream()))
        return wave[: sum(ol)], ol

    # ------------------------------------------------------------------ stages (dense tensors, see header)
    def pe_towers(self, x: torch.Tensor, style: torch.Tensor):
        tot = sum(self.len_f)
        pitch = torch.empty(tot, device=self.device)
        energy = torch.empty(tot, device=self.device)
        self._check(self.lib.stts_pe_towers(self.h, _ptr(x), _ptr(style), _ptr(pitch), _ptr(energy), self._stream()))
        return pitch, energy

    def decoder(self, asr, f0, n, style):
        out = torch.empty(512 * sum(self.len_f4), device=self.device)
        self._check(self.lib.stts_decoder(self.h, _ptr(asr), _ptr(f0), _ptr(n), _ptr(style), _ptr(out), self._stream()))
        return out

    def prior_flow(self, x, noise, style, want_z: bool = False):
        tot = sum(self.len_f4)
        mel = torch.empty(512 * tot, device=self.device)
        zp = torch.empty(128 * tot, device=self.device) if want_z else None
        zf = torch.empty(128 * tot, device=self.device) if want_z else None
        self._check(self.lib.stts_prior_flow(self.h, _ptr(x), _ptr(noise), _ptr(style), _ptr(mel), _ptr(zp), _ptr(zf),
                                             self._stream()))
        return mel, zp, zf

    def source_pcph(self, pitch4, noise, init_phase):
        out = torch.empty(75 * sum(self.len_f4), device=self.device)
        self._check(self.lib.stts_source_pcph(self.h, _ptr(pitch4), _ptr(noise), _ptr(init_phase), _ptr(out),
                                              self._stream()))
        return out

    def generator(self, mel, style, pitch4, noise, init_phase, want_spec: bool = True):
        tot = sum(self.len_f4)
        audio = torch.empty(75 * tot, device=self.device)
        la = torch.empty(1025 * (tot + self.n_seg), device=self.device) if want_spec else None
        ph = torch.empty(1025 * (tot + self.n_seg), device=self.device) if want_spec else None
        self._check(self.lib.stts_generator(self.h, _ptr(mel), _ptr(style), _ptr(pitch4), _ptr(noise),
                                            _ptr(init_phase), _ptr(audio), _ptr(la), _ptr(ph), self._stream()))
        return audio, la, ph

    def generator_from_stft(self, mel, style, har_spec, har_phase):
        tot = sum(self.len_f4)
        audio = torch.empty(75 * tot, device=self.device)
        la = torch.empty(1025 * (tot + self.n_seg), device=self.device)
        ph = torch.empty(1025 * (tot + self.n_seg), device=self.device)
        self._check(self.lib.stts_generator_from_stft(self.h, _ptr(mel), _ptr(style), _ptr(har_spec), _ptr(har_phase),
                                                      _ptr(audio), _ptr(la), _ptr(ph), self._stream()))
        return audio, la, ph

    def speech_frames(self, enc_tok, cu_tok, idx, cu_frames, pitch, energy, style, noise_prior, noise_source,
                      init_phase, want_spec: bool = False):
        tot = sum(self.len_f4)
        audio = torch.empty(75 * tot, device=self.device)
        la = torch.empty(1025 * (tot + self.n_seg), device=self.device) if want_spec else None
        ph = torch.empty(1025 * (tot + self.n_seg), device=self.device) if want_spec else None
        self._check(self.lib.stts_speech_frames(
            self.h, _ptr(enc_tok), _ptr(cu_tok, torch.int32), _ptr(idx, torch.int32), _ptr(cu_frames, torch.int32),
            _ptr(pitch), _ptr(energy), _ptr(style), _ptr(noise_prior), _ptr(noise_source), _ptr(init_phase),
            _ptr(audio), _ptr(la), _ptr(ph), self._stream()))
        return audio, la, ph

    def check_source_status(self):
        self._check(self.lib.stts_check_source_status(self.h, self._stream()))

    def saturation_count(self) -> int:
        """Elements of the fp16 operand twins at +-65504 after the last run (synchronises)."""
        return int(self._check(self.lib.stts_saturation_count(self.h, self._stream())))

    def pcm16(self, audio: torch.Tensor) -> torch.Tensor:
        """fp32 device waveform -> int16 PCM on the device (train/test_onnx.py:79)."""
        a = audio.contiguous()
        out = torch.empty(a.shape, dtype=torch.int16, device=self.device)
        self._check(self.lib.stts_pcm16(self.h, _ptr(a), a.numel(), out.data_ptr(), self._stream()))
        return out

    def set_profiling(self, on: bool):
        self._check(self.lib.stts_set_profiling(self.h, 1 if on else 0))

    def profile(self):
        """Per-layer GEMM timings recorded since set_profiling(True) (synchronises)."""
        import json
        return json.loads(self.lib.stts_profile_json(self.h).decode())

    # ------------------------------------------------------------------ test hooks
    def test_conv(self, backend: str, x: torch.Tensor, cin: int, w: torch.Tensor, bias: Optional[torch.Tensor],
                  pad: int, ldy: int) -> torch.Tensor:
        """x: packed [R, ldx]; w: [N, cin, taps] (CPU or CUDA); returns packed [R, ldy]."""
        n_out, _, taps = w.shape
        wc = w.detach().to("cpu", torch.float32).contiguous()
        bc = bias.detach().to("cpu", torch.float32).contiguous() if bias is not None else None
        y = torch.zeros((x.shape[0], ldy), device=self.device)
        self._check(self.lib.stts_test_conv(self.h, PRECISIONS[backend], _ptr(x), x.shape[1], cin, taps, pad,
                                            wc.data_ptr(), bc.data_ptr() if bc is not None else None, n_out, _ptr(y),
                                            ldy, self._stream()))
        return y

    def debug_read(self, name: str) -> torch.Tensor:
        rows, ld = C.c_int(), C.c_int()
        n = self._check(self.lib.stts_debug_read(self.h, name.encode(), None, 0, C.byref(rows), C.byref(ld), None))
        out = torch.empty(n, device=self.device)
        self._check(self.lib.stts_debug_read(self.h, name.encode(), _ptr(out), n, C.byref(rows), C.byref(ld),
                                             self._stream()))
        return out.view(rows.value, ld.value) if ld.value > 1 else out

    def layout_info(self, rate: int):
        n = self.n_seg
        rows = C.c_int()
        starts = (C.c_int * max(n, 1))()
        self._check(self.lib.stts_layout_info(self.h, rate, C.byref(rows), starts))
        return rows.value, list(starts)[:n]
