"""Shadows the reference's `src/model.py` on sys.path: `from model import M2FNet` in the UNCHANGED src/train.py:11 and
src/test.py:5 then resolves to the B200-native drop-in.  Nothing else lives here."""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module("multimodal-emotion-recognition_b200")

M2FNet = _pkg.M2FNet
FusionAttentionModule = _pkg.model.FusionAttentionModule
