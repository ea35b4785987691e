"""CPU tests of the host front / back end (stylish_tts_b200/frontend.py) against golden vectors
generated from the live reference (oracle/make_frontend_golden.py -> tests/golden/frontend.json)."""
import io
import json
import os

import numpy as np
import pytest

from tests.util import ROOT
from stylish_tts_b200 import frontend as fe


@pytest.fixture(scope="module")
def gold():
    return json.load(open(os.path.join(ROOT, "tests", "golden", "frontend.json")))


def test_symbol_defaults_are_the_reference_model_yml(gold):
    s = fe.SymbolConfig()
    for k in ("pad", "punctuation", "letters", "letters_ipa"):
        assert getattr(s, k) == gold["symbols"][k], k


def test_text_cleaner_table_and_tokenisation(gold):
    c = fe.TextCleaner()
    assert c.word_index_dictionary == gold["table"]
    assert max(c.word_index_dictionary.values()) < 178           # text_encoder.tokens of model.yml
    for text, ids in zip(gold["texts"], gold["ids"]):
        assert c(text) == ids                                     # unknown characters are dropped


def test_wrap_tokens_pads_both_ends():
    t = fe.wrap_tokens([5, 6, 7])
    assert t.tolist() == [0, 5, 6, 7, 0] and str(t.dtype) == "torch.int64"
    assert fe.wrap_tokens([]).tolist() == [0, 0]


def test_blend_together_matches_reference(gold):
    a, b = np.array(gold["blend"]["a"]), np.array(gold["blend"]["b"])
    np.testing.assert_array_equal(fe.blend_together(a, b), np.array(gold["blend"]["ab"]))
    np.testing.assert_array_equal(fe.blend_together(b, a), np.array(gold["blend"]["ba"]))


def test_trim_silence_matches_reference_loop(gold):
    rng = np.random.default_rng(gold["trim"]["signal_seed"])
    rng.standard_normal(37), rng.standard_normal(52)             # the generator's earlier draws
    sig = np.concatenate([np.zeros(100), 0.02 * rng.standard_normal(50), 0.5 * rng.standard_normal(400),
                          0.01 * rng.standard_normal(80), np.zeros(20000)]).astype(np.float32)
    out = fe.trim_silence(sig)
    np.testing.assert_array_equal(out, sig[gold["trim"]["i"]:gold["trim"]["j"]])
    assert fe.trim_silence(np.zeros(50, np.float32)).shape[0] == 0    # all silent: reference gives audio[n:n]


def test_int16_conversion_matches_numpy_cast(gold):
    x = np.array(gold["pcm"]["x"], dtype=np.float32)
    np.testing.assert_array_equal(fe.to_int16(x), np.array(gold["pcm"]["i16"], dtype=np.int16))
    assert fe.to_int16(np.array([1.0, 2.0, -2.0], np.float32)).tolist() == [32767, 32767, -32768]


def test_wav_bytes_round_trip():
    from scipy.io import wavfile
    pcm = (np.arange(-500, 500) * 7).astype(np.int16)
    blob = fe.wav_bytes(pcm, 24000)
    ref = io.BytesIO()
    wavfile.write(ref, 24000, pcm)                               # what the reference calls (test_onnx.py:84-92)
    assert blob == ref.getvalue()
    sr, back = wavfile.read(io.BytesIO(blob))
    assert sr == 24000 and (back == pcm).all()


def test_split_long_text_respects_cap_and_loses_nothing():
    sent = "ðɪs ɪz ɐ tˈɛst sˈɛntəns, wɪð ɐ klˈɔːz; ænd mˈɔːɹ. "
    text = sent * 40
    parts = fe.split_long_text(text)
    assert all(0 < len(p) <= fe.MAX_SYMBOLS - 2 for p in parts)
    assert "".join(p.replace(" ", "") for p in parts) == text.replace(" ", "")
    assert all(p[-1] in ".;," for p in parts[:-1])               # cuts fall on punctuation when there is some
    hard = "a" * 1300                                            # no break anywhere: hard cuts
    assert [len(p) for p in fe.split_long_text(hard)] == [508, 508, 284]
    assert fe.split_long_text("  ") == []
    with pytest.raises(ValueError):
        fe.split_long_text("abc", 0)


def test_tokenize_rejects_over_long_utterance():
    tts = fe.TextToSpeech(synthesizer=None)
    with pytest.raises(ValueError):
        tts.tokenize(["a" * 509])
    assert tts.tokenize(["a" * 508])[0].numel() == 510
