"""Import the *real* reference (read-only, /root/reference) in this container.

TEST INFRASTRUCTURE ONLY (see oracle/stylish_port.py).  The reference needs two modules that
are absent here and never called on the hot path: ``munch`` (attribute
dict, train/utils.py:4, models/models.py:20) and ``matplotlib`` (train/utils.py:3).  They are
stubbed in ``sys.modules`` before the import.  Nothing is copied from the reference.
"""
from __future__ import annotations

import os
import sys
import types
import warnings

REFERENCE_SRC = "/root/reference/src"


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "stylish_tts"))


def import_reference():
    """Returns (build_model, load_model_config_yaml, DurationProcessor, model_yml_path)."""
    if not reference_available():
        raise RuntimeError("reference checkout not present at /root/reference (expected on the GPU box)")
    if "munch" not in sys.modules:
        class Munch(dict):
            __getattr__ = dict.__getitem__
            __setattr__ = dict.__setitem__
        m = types.ModuleType("munch")
        m.Munch = Munch
        sys.modules["munch"] = m
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    warnings.filterwarnings("ignore", category=FutureWarning)
    from stylish_tts.lib.config_loader import load_model_config_yaml
    from stylish_tts.train.models.models import build_model
    from stylish_tts.train.utils import DurationProcessor
    yml = os.path.join(REFERENCE_SRC, "stylish_tts", "train", "config", "model.yml")
    return build_model, load_model_config_yaml, DurationProcessor, yml
