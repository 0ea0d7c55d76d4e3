"""Shared helpers for the parity tests: golden-case loading and the parity definitions
(SURVEY.md §8c): masks / flow12 / projected points bit-exact; colour |d| <= 1e-5*max(1,|ref|);
depth |d| <= 1e-5*|ref| on valid pixels."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
COLOUR_TOL = 1e-5
DEPTH_RTOL = 1e-5


def load_cases():
    z = np.load(os.path.join(GOLDEN_DIR, "warp_cases.npz"))
    cases = {}
    for key in z.files:
        name, field = key.split("/", 1)
        cases.setdefault(name, {})[field] = z[key]
    return cases


def warp_case_names(twice=None, clean=None):
    names = []
    for name, c in load_cases().items():
        if "flags" not in c:
            continue
        is_clean, is_twice = bool(c["flags"][0]), bool(c["flags"][1])
        if twice is not None and is_twice != twice:
            continue
        if clean is not None and is_clean != clean:
            continue
        names.append(name)
    return names


def bit_equal(a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    assert a.shape == b.shape and a.dtype == b.dtype, (a.shape, b.shape, a.dtype, b.dtype)
    # +0.0 and -0.0 compare equal on purpose (torch.where/clamp do not preserve the sign of zero)
    return bool(np.array_equal(a, b))


def assert_colour_close(got, ref, what="colour", tol=COLOUR_TOL):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    err = np.abs(got - ref)
    bound = tol * np.maximum(1.0, np.abs(ref))
    bad = err > bound
    assert not bad.any(), f"{what}: {int(bad.sum())} px out of tolerance, max |d|={err.max():.3e}"
    return float(err.max())


def assert_depth_close(got, ref, valid, what="depth", rtol=DEPTH_RTOL):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    valid = np.asarray(valid) > 0
    assert np.array_equal(got[~valid], ref[~valid]), f"{what}: hole values differ"
    if not valid.any():
        return 0.0
    rel = np.abs(got[valid] - ref[valid]) / np.maximum(np.abs(ref[valid]), 1e-30)
    assert rel.max() <= rtol, f"{what}: max rel err {rel.max():.3e} > {rtol}"
    return float(rel.max())
