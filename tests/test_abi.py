"""CPU: the C-ABI library builds, loads and exports every symbol include/mantis_cuda.h declares.
No compute calls (there is no GPU here and the library has no CPU fallback)."""
import ctypes
import os
import re

import pytest

import mantis_b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "mantis_cuda.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mantis_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    path = mantis_b200.library_path()
    if not os.path.exists(path):
        mantis_b200.build_library()
    lib = ctypes.CDLL(path)
    syms = declared_symbols()
    assert "mantis_create" in syms and "mantis_edges_batch" in syms
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in mantis_cuda.h but not exported"


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(mantis_b200.MantisError, match="no CUDA device"):
        mantis_b200.Context(0)


def test_strerror():
    lib = mantis_b200.load_library()
    assert lib.mantis_strerror(0) == b"ok"
    assert b"fallback" in lib.mantis_strerror(-3)


def test_product_does_not_import_oracle():
    for dp, _, fns in os.walk(os.path.join(ROOT, "mantis_b200")):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dp, fn)).read()
                for pat in (r"^\s*(from|import)\s+oracle", r"#include\s*[<\"].*oracle", r"libmantis_oracle", r"pyoracle"):
                    assert not re.search(pat, text, flags=re.M), f"{fn} uses the oracle ({pat})"


def test_create_rejects_bad_config_before_touching_a_device():
    """mantis_create validates mantis_config up front: limits the kernels are built for and
    non-finite / non-positive tunables come back as MANTIS_ERR_INVALID (-1), never as a CUDA error later."""
    lib = mantis_b200.load_library()
    bad = []
    for field, value in (("particles", 97), ("particles", -1), ("iterations", -3), ("iterations", 1 << 30), ("grid_size", 10),
                         ("grid_size", 0), ("grid_spacing", 0.0), ("grid_spacing", float("nan")), ("projection_bias", -1.0),
                         ("color_projection_bias", float("inf")), ("rot_sigma", -0.1), ("canny_hysteresis", 100000)):
        cfg = mantis_b200.binding.default_config()
        setattr(cfg, field, value)
        h = ctypes.c_void_p()
        bad.append((field, value, lib.mantis_create(ctypes.byref(cfg), 0, ctypes.byref(h)), h.value))
    assert all(rc == -1 and handle is None for _, _, rc, handle in bad), bad
