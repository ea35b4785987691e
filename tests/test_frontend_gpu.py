"""GPU: device int16 conversion (stts_pcm16) and the text front end over the Synthesizer."""
import numpy as np
import pytest
import torch

from tests.util import synthesizer
from stylish_tts_b200 import frontend as fe

pytestmark = pytest.mark.gpu


def test_pcm16_is_the_numpy_cast_bit_exact():
    eng = synthesizer("fp32").engine
    g = torch.Generator().manual_seed(3)
    edge = torch.tensor([0.0, -1.0, 0.99999994, -0.99999994, 1 / 32768, -1 / 32768, 1.0, 0.5 / 32768, 3.0, -3.0])
    for n in (0, 1, 7, 8, 9, 75 * 1601, 1_000_003):
        x = torch.rand(n, generator=g) * 2.2 - 1.1
        if n >= edge.numel():
            x[:edge.numel()] = edge
        want = fe.to_int16(x.numpy())
        got = eng.pcm16(x.cuda()).cpu().numpy()
        assert got.dtype == np.int16 and (got == want).all(), n
        if n > 1:                                   # 4-byte-aligned but not 16-byte-aligned input
            assert (eng.pcm16(x.cuda()[1:]).cpu().numpy() == want[1:]).all(), n


def test_text_to_speech_batch_equals_per_call_pipeline():
    syn = synthesizer("fp32")
    tts = fe.TextToSpeech(syn)
    texts = ["ðə kwˈɪk bɹˈaʊn fˈɑːks.", "hɛlˈoʊ, wˈɜːld!"]
    torch.manual_seed(11)
    torch.cuda.manual_seed(11)
    pcm = tts.synthesize(texts)
    syn.check()
    torch.manual_seed(11)
    torch.cuda.manual_seed(11)
    ref = syn.synthesize(tts.tokenize(texts))
    assert len(pcm) == 2
    for p, a, d in zip(pcm, ref["audio"], ref["duration_pred"]):
        assert p.dtype == np.int16 and p.shape[0] == a.numel() == 300 * int(d.sum())
        assert (p == fe.to_int16(a)).all()
    one = tts.synthesize(texts, combine=True)
    assert one.shape[0] == sum(p.shape[0] for p in pcm)


def test_long_form_text_is_split_and_joined():
    syn = synthesizer("fp32")
    tts = fe.TextToSpeech(syn)
    text = "ðɪs ɪz ɐ tˈɛst sˈɛntəns, wɪð ɐ klˈɔːz; ænd mˈɔːɹ. " * 12      # > 510 symbols
    assert len(text) > fe.MAX_SYMBOLS
    pcm = tts.synthesize_long(text, trim=False)
    syn.check()
    pieces = fe.split_long_text(text)
    assert len(pieces) >= 2 and pcm.dtype == np.int16 and pcm.shape[0] % 300 == 0 and pcm.shape[0] > 0
