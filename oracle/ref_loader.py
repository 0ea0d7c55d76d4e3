"""Loads the UNMODIFIED reference module /root/reference/models/utils.py (container only).

The file imports ``skimage.io`` and ``decord`` (absent here, unused by the Warper path), so empty
stand-ins are placed in ``sys.modules`` first.  Used only by oracle/gen_golden.py
and by the container-only tests that diff the C oracle against the live reference; never on the
GPU box (where /root/reference does not exist) and never by the product.
"""
import importlib.util
import os
import sys
import types

REF_UTILS = "/root/reference/models/utils.py"


def available() -> bool:
    return os.path.exists(REF_UTILS)


def load():
    for name in ("skimage", "skimage.io", "decord"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            if name == "decord":
                m.VideoReader = object
                m.cpu = lambda *a, **k: None
            sys.modules[name] = m
    sys.modules["skimage"].io = sys.modules["skimage.io"]
    spec = importlib.util.spec_from_file_location("_tc_reference_utils", REF_UTILS)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
