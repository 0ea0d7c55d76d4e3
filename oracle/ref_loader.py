"""Loads the UNMODIFIED reference module models/utils.py: from /root/reference in the build container, else from the
verbatim run-time copy oracle/build_ref.py leaves in oracle/_ref/ (git-ignored; it travels to the GPU box).

The file imports ``skimage.io`` and ``decord`` (absent here, unused by the Warper path), so empty
stand-ins are placed in ``sys.modules`` first.  Used only by oracle/gen_golden*.py, by the tests that
diff the C oracle against the live reference, and by bench.py's CPU-baseline legs; never by the product.
"""
import importlib.util
import os
import sys
import types

_CANDIDATES = ("/root/reference/models/utils.py",
               os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "reference_models_utils.py"))


def path() -> str:
    for p in _CANDIDATES:
        if os.path.exists(p):
            return p
    return ""


def available() -> bool:
    return bool(path())


def load():
    for name in ("skimage", "skimage.io", "decord"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            if name == "decord":
                m.VideoReader = object
                m.cpu = lambda *a, **k: None
            sys.modules[name] = m
    sys.modules["skimage"].io = sys.modules["skimage.io"]
    spec = importlib.util.spec_from_file_location("_tc_reference_utils", path())
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
