"""CPU: the C-ABI library loads and exports every symbol include/dirt_b200.h declares; argument
validation works without a GPU; the product package never touches oracle/."""
import ctypes
import os
import re

import pytest

from dirt_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "dirt_b200.h")).read()
    return sorted(set(re.findall(r"DIRT_API[^;(]*?\b(dirt_\w+)\s*\(", text)))


def test_header_and_binding_agree():
    declared = _declared_symbols()
    assert len(declared) >= 9
    assert declared == sorted(_lib.EXPORTS)


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    for name in _declared_symbols():
        assert getattr(lib, name) is not None
    assert lib.dirt_abi_version() == _lib.ABI_VERSION


def test_workspace_bytes_and_shape_validation():
    lib = _lib.load()
    ok = _lib.DirtShape(4, 64, 64, 3, 100, 200)
    n = lib.dirt_workspace_bytes(ctypes.byref(ok))
    assert n >= 4 * 64 * 64 * 8
    for bad in (_lib.DirtShape(1, 0, 64, 3, 1, 1), _lib.DirtShape(1, 64, 64, 0, 1, 1),
                _lib.DirtShape(1, 64, 64, 17, 1, 1), _lib.DirtShape(-1, 64, 64, 3, 1, 1)):
        assert lib.dirt_workspace_bytes(ctypes.byref(bad)) == 0


def test_forward_rejects_bad_arguments_without_touching_the_gpu():
    lib = _lib.load()
    assert lib.dirt_rasterise_forward(None, None) == -1
    args = _lib.DirtForwardArgs()
    args.shape = _lib.DirtShape(1, 8, 8, 3, 3, 1)
    args.vertices_bstride = 5            # neither 0 nor >= V*4
    assert lib.dirt_rasterise_forward(ctypes.byref(args), None) == -1
    assert b"vertices" in lib.dirt_last_error()
    with pytest.raises(_lib.InvalidArgumentError):
        _lib.check(-1)


def test_product_package_never_references_the_oracle():
    pkg = os.path.join(ROOT, "dirt_b200")
    for dirpath, _, files in os.walk(pkg):
        if os.path.basename(dirpath) == "build":
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower() or f == "rasterise_ops.py" and False, os.path.join(dirpath, f)
                assert "golden" not in text.lower(), os.path.join(dirpath, f)


def test_python_surface_matches_reference_signatures():
    import inspect
    import dirt_b200
    # dirt/rasterise_ops.py:10, :57, :91
    assert list(inspect.signature(dirt_b200.rasterise).parameters)[:9] == [
        "background", "vertices", "vertex_colors", "faces", "camera_pos", "height", "width", "channels", "name"]
    assert list(inspect.signature(dirt_b200.rasterise_batch).parameters) == [
        "background", "vertices", "vertex_colors", "faces", "height", "width", "channels", "name"]
    assert list(inspect.signature(dirt_b200.rasterise_grad).parameters)[:9] == [
        "background", "vertices", "vertex_colors", "faces", "camera_pos", "height", "width", "channels", "name"]
