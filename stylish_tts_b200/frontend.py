"""Host-side inference front / back end around the synthesis kernels.

What the reference does around its model call when it synthesises text:
  * ``TextCleaner`` -- phoneme string -> token ids (lib/text_utils.py:8-41, symbols from model.yml:81-85),
  * pad id 0 on both ends of every utterance (train/test_onnx.py:49-53, train/dataloader.py:178-179),
  * int16 PCM: ``np.multiply(audio, 32768).astype(np.int16)`` and a 24 kHz WAV (train/test_onnx.py:79-92),
  * long-form text: one model call per line with at most 510 symbols, leading / trailing silence trimmed
    by an amplitude threshold and the pieces joined (legacy logic, tts/ttab/inference.py:180-200,541-578).
Everything here is host code over ``Synthesizer``; the waveform -> int16 conversion of a batch runs on
the device (``stts_pcm16``) so that the device -> host copy moves 2 bytes per sample.
"""
from __future__ import annotations

import logging
import struct
from dataclasses import dataclass
from typing import Iterable, List, Sequence

import numpy as np
import torch

logger = logging.getLogger(__name__)

MAX_SYMBOLS = 510           # longest model input in tokens, pad tokens included (the encoders' segment cap)


@dataclass
class SymbolConfig:
    """model.yml:81-85 (lib/config_loader.py:163-180)."""
    pad: str = "$"
    punctuation: str = ";:,.!?¡¿—…\"()“” "
    letters: str = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz"
    letters_ipa: str = ("ɑɐɒæɓʙβɔɕçɗɖðʤəɘɚɛɜɝɞɟʄɡɠɢʛɦɧħɥʜɨɪʝɭɬɫɮʟɱɯɰŋɳɲɴøɵɸθœɶʘɹɺɾɻʀʁɽʂʃʈʧʉʊʋⱱʌɣɤʍχʎʏʑʐʒʔʡʕʢǀǁᵊǃˈˌːˑʼʴʰʱʲʷˠˤ˞"
                        "↓↑→↗↘'̩'ᵻ")


class TextCleaner:
    """Same constructor, call and ``word_index_dictionary`` as lib/text_utils.py:8-41: symbols are
    numbered in the order pad, punctuation, letters, IPA letters (a repeated symbol keeps its LAST
    index, as the reference's dict construction does); unknown characters are logged and dropped."""

    def __init__(self, symbols=None):
        symbols = symbols or SymbolConfig()
        self._pad = symbols.pad
        self._punctuation = symbols.punctuation
        self._letters = symbols.letters
        self._letters_ipa = symbols.letters_ipa
        self.word_index_dictionary = self.build_text_cleaner()

    def build_text_cleaner(self):
        table = [self._pad] + list(self._punctuation) + list(self._letters) + list(self._letters_ipa)
        return {sym: i for i, sym in enumerate(table)}

    def __call__(self, text: str) -> List[int]:
        out = []
        for ch in text:
            idx = self.word_index_dictionary.get(ch)
            if idx is None:
                logger.error("Meld " + ch + ": " + text)
            else:
                out.append(idx)
        return out


def wrap_tokens(ids: Sequence[int]) -> torch.Tensor:
    """ids -> int64 [T + 2] with pad id 0 at both ends (train/test_onnx.py:50-53)."""
    t = torch.zeros(len(ids) + 2, dtype=torch.int64)
    if len(ids):
        t[1:-1] = torch.as_tensor(list(ids), dtype=torch.int64)
    return t


_BREAKS = (".!?…", ";:—", ",", " ")     # preferred split points, strongest first


def split_long_text(text: str, max_symbols: int = MAX_SYMBOLS - 2) -> List[str]:
    """Long-form input -> pieces of at most ``max_symbols`` characters (so that piece + 2 pad tokens fits
    the 510-token segment).  A piece ends after the last sentence mark inside the window, else after the
    last clause mark, comma or space; a window without any break is cut hard.  The reference's legacy
    front end feeds one line per call under the same cap (tts/ttab/inference.py:180-186)."""
    if max_symbols < 1:
        raise ValueError("max_symbols must be >= 1")
    pieces, rest = [], text.strip()
    while len(rest) > max_symbols:
        window = rest[:max_symbols]
        cut = -1
        for marks in _BREAKS:
            cut = max(window.rfind(m) for m in marks)
            if cut > 0:
                break
        cut = cut + 1 if cut > 0 else max_symbols
        pieces.append(rest[:cut].strip())
        rest = rest[cut:].strip()
    if rest:
        pieces.append(rest)
    return [p for p in pieces if p]


def trim_silence(audio: np.ndarray, threshold: float = 1000.0 / 32768.0, tail: int = 9000) -> np.ndarray:
    """Leading / trailing samples below ``threshold`` removed, ``tail`` samples of the trailing part
    kept (tts/ttab/inference.py:190-199; same index arithmetic, vectorised)."""
    a = np.asarray(audio)
    loud = np.flatnonzero(np.abs(a) >= threshold)
    n = a.shape[0]
    i = int(loud[0]) if loud.size else n
    j = int(loud[-1]) if loud.size else 0
    j = min(n, j + tail)
    return a[i:j]


def blend_together(loose_end: np.ndarray, start_next: np.ndarray) -> np.ndarray:
    """Linear cross-fade of two overlapping pieces, the shorter one zero-padded (at its start for the
    incoming piece, at its end for the outgoing one) -- tts/ttab/inference.py:563-570."""
    diff = loose_end.shape[0] - start_next.shape[0]
    if diff > 0:
        start_next = np.concatenate([np.zeros([diff], dtype=start_next.dtype), start_next])
    elif diff < 0:
        loose_end = np.concatenate([loose_end, np.zeros([-diff], dtype=loose_end.dtype)])
    w = np.linspace(0, 1, loose_end.shape[0], dtype=loose_end.dtype)
    return w * start_next + (1 - w) * loose_end


def to_int16(audio) -> np.ndarray:
    """``np.multiply(audio, 32768).astype(np.int16)`` (train/test_onnx.py:79): truncation toward zero.
    The reference's cast is undefined for the single value +1.0 (32768 does not fit); it is clamped to
    32767 here -- every other input gives the reference's result."""
    a = np.asarray(audio.detach().cpu() if isinstance(audio, torch.Tensor) else audio, dtype=np.float32)
    return np.clip(np.trunc(a * np.float32(32768.0)), -32768, 32767).astype(np.int16)


def wav_bytes(samples: np.ndarray, sample_rate: int = 24000) -> bytes:
    """Canonical 44-byte-header mono PCM-16 RIFF/WAVE -- what ``scipy.io.wavfile.write(path, 24000,
    int16_array)`` produces (train/test_onnx.py:84-92)."""
    pcm = np.ascontiguousarray(samples, dtype="<i2")
    data = pcm.tobytes()
    hdr = b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVE" + b"fmt " + struct.pack(
        "<IHHIIHH", 16, 1, 1, sample_rate, sample_rate * 2, 2, 16) + b"data" + struct.pack("<I", len(data))
    return hdr + data


def write_wav(path: str, samples: np.ndarray, sample_rate: int = 24000) -> None:
    with open(path, "wb") as fh:
        fh.write(wav_bytes(samples, sample_rate))


class TextToSpeech:
    """phoneme strings in, int16 PCM out: TextCleaner -> pad wrap -> one ragged ``Synthesizer`` batch
    (durations predicted on the device) -> device int16 conversion -> host.  Mirrors the loop of
    train/test_onnx.py:49-92 with the batch replacing the per-text model calls."""

    def __init__(self, synthesizer, symbols=None):
        self.syn = synthesizer
        self.cleaner = TextCleaner(symbols)

    def tokenize(self, texts: Iterable[str]) -> List[torch.Tensor]:
        toks = []
        for t in texts:
            ids = self.cleaner(t)
            if len(ids) + 2 > MAX_SYMBOLS:
                raise ValueError(f"utterance of {len(ids)} symbols exceeds the {MAX_SYMBOLS - 2}-symbol segment; "
                                 "split it with split_long_text()")
            toks.append(wrap_tokens(ids))
        return toks

    def synthesize(self, texts: Sequence[str], combine: bool = False):
        """-> list of int16 arrays (or one concatenated array, the reference's ``combine=True``)."""
        tokens = self.tokenize(texts)
        out = self.syn.synthesize(tokens)
        flat = self.syn.engine.pcm16(out["audio_flat"]).cpu().numpy()      # one kernel, one 2-byte/sample D2H copy
        if combine:
            return flat
        off = out["audio_offsets"]
        return [flat[off[b]: off[b + 1]] for b in range(len(tokens))]

    def synthesize_long(self, text: str, trim: bool = True) -> np.ndarray:
        """Long-form text: split under the segment cap, synthesise the pieces as ONE batch, trim each
        piece's edge silence (threshold 1000 / 32768, 9000-sample tail) and join."""
        pieces = split_long_text(text)
        tokens = self.tokenize(pieces)
        out = self.syn.synthesize(tokens)
        parts = []
        for a in out["audio"]:
            x = a.float().cpu().numpy()
            parts.append(trim_silence(x) if trim else x)
        joined = np.concatenate(parts) if parts else np.zeros(0, np.float32)
        return to_int16(joined)
