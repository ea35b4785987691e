"""Golden vectors for the host front / back end (stylish_tts_b200/frontend.py), generated from the live
reference in this container: the TextCleaner symbol table of the default model.yml, tokenisations of a few
phoneme strings, the legacy trim / cross-fade helpers and the int16 conversion of train/test_onnx.py:79.

TEST INFRASTRUCTURE ONLY.  usage: python -m oracle.make_frontend_golden  -> tests/golden/frontend.json
"""
import importlib.util
import json
import os

import numpy as np

from oracle.ref_import import REFERENCE_SRC, import_reference

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

TEXTS = [
    "ðə kwˈɪk bɹˈaʊn fˈɑːks dʒˈʌmps ˌoʊvɚ ðə lˈeɪzi dˈɔːɡ.",
    "hɛlˈoʊ, wˈɜːld! — “kwˈoʊt” (pɚɹˈɛnθəsˌiːz)…",
    "unknown: 0123 @# ɚ ᵊ ᵻ '",
    "",
]


def main():
    _, load_yaml, _, yml = import_reference()
    from stylish_tts.lib.text_utils import TextCleaner
    mc = load_yaml(open(yml))
    cleaner = TextCleaner(mc.symbol)
    out = {"symbols": {"pad": mc.symbol.pad, "punctuation": mc.symbol.punctuation, "letters": mc.symbol.letters,
                       "letters_ipa": mc.symbol.letters_ipa},
           "table": cleaner.word_index_dictionary, "texts": TEXTS, "ids": [cleaner(t) for t in TEXTS]}
    # legacy helpers: module-level functions of tts/ttab/inference.py need heavy imports; load only the
    # two pure-numpy functions by executing their source segment
    src = open(os.path.join(REFERENCE_SRC, "stylish_tts", "tts", "ttab", "inference.py")).read()
    start = src.index("def blend_together")
    end = src.index("def combine_styles")
    ns = {"np": np}
    exec(compile(src[start:end], "blend_together", "exec"), ns)
    rng = np.random.default_rng(5)
    a, b = rng.standard_normal(37), rng.standard_normal(52)
    out["blend"] = {"a": a.tolist(), "b": b.tolist(), "ab": ns["blend_together"](a, b).tolist(),
                    "ba": ns["blend_together"](b, a).tolist()}
    # trim: the reference's loop (context_infer, inference.py:190-199) restated on a recorded signal
    sig = np.concatenate([np.zeros(100), 0.02 * rng.standard_normal(50), 0.5 * rng.standard_normal(400),
                          0.01 * rng.standard_normal(80), np.zeros(20000)]).astype(np.float32)
    i, j, thr = 0, sig.shape[0] - 1, 1000 / 32768
    while i < sig.shape[0] and abs(sig[i]) < thr:
        i += 1
    while j > 0 and abs(sig[j]) < thr:
        j -= 1
    j = min(sig.shape[0], j + 9000)
    out["trim"] = {"signal_seed": 5, "i": i, "j": j}
    x = np.concatenate([rng.uniform(-1, 1, 64), [0.0, -1.0, 0.99999994, -0.99999994, 1 / 32768, -1 / 32768,
                                                 0.5 / 32768, -0.5 / 32768]]).astype(np.float32)
    out["pcm"] = {"x": x.tolist(), "i16": np.multiply(x, 32768).astype(np.int16).tolist()}
    path = os.path.join(ROOT, "tests", "golden", "frontend.json")
    json.dump(out, open(path, "w"), ensure_ascii=False, indent=0)
    print("wrote", path, len(out["table"]), "symbols")


if __name__ == "__main__":
    main()
