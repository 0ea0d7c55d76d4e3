"""CPU: the C-ABI library builds, loads and exports every symbol include/tcwarp.h declares
(no compute calls — there is no GPU in the build container)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "tcwarp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tcw_[a-z_]+)\s*\(", text)))


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for name in ("tcw_create", "tcw_destroy", "tcw_forward_warp", "tcw_forward_warp_host", "tcw_splat",
                 "tcw_project", "tcw_clean_points", "tcw_workspace_bytes", "tcw_last_error", "tcw_version",
                 "tcw_launch_count"):
        assert name in syms


def test_library_builds_and_exports_every_declared_symbol():
    from trajectorycrafter_b200.build import build_library
    path = build_library()
    lib = ctypes.CDLL(path)
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} is declared in include/tcwarp.h but not exported"
    lib.tcw_version.restype = ctypes.c_int
    assert lib.tcw_version() == 200


def test_python_binding_covers_the_header():
    from trajectorycrafter_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    _lib.load()


def test_argument_errors_do_not_need_a_gpu():
    from trajectorycrafter_b200 import _lib
    lib = _lib.load()
    need = ctypes.c_size_t(0)
    assert lib.tcw_workspace_bytes(None, 1, 8, 8, 0, ctypes.byref(need)) == -1
    assert b"bad argument" in lib.tcw_last_error()
    assert lib.tcw_forward_warp(None, None, None, None, None, None, 1, 1, 8, 8, 0,
                                None, None, None, None, None, None, 0, None) == -1


def test_no_cpu_fallback():
    import torch
    from trajectorycrafter_b200 import Warper
    with pytest.raises(ValueError):
        Warper(device="cpu")
    if not torch.cuda.is_available():
        w = Warper(device="cuda:0")
        with pytest.raises(RuntimeError):
            w.forward_warp(torch.zeros(1, 3, 8, 8), None, torch.ones(1, 1, 8, 8), torch.eye(4)[None],
                           torch.eye(4)[None], torch.eye(3)[None], None)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "trajectorycrafter_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "liborc" not in text, f
