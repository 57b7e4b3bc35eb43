"""samples/c_caller.c: the header is valid C99, a plain C program links against libdirt_b200.so with no
CUDA headers, and the host-buffer entry points either do the work (GPU box: the README square, bit
for bit, and its colour gradient) or say in words why not (no GPU here: there is no CPU path)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    from dirt_b200 import build
    lib = build.build()
    exe = str(tmp_path / "c_caller")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "samples", "c_caller.c"), "-L" + os.path.dirname(lib), "-ldirt_b200",
           "-Wl,-rpath," + os.path.dirname(lib), "-lm", "-o", exe]
    subprocess.check_call(cmd)
    return exe


@pytest.mark.skipif(shutil.which("gcc") is None, reason="no gcc")
def test_c_caller_compiles_links_and_fails_loudly_without_a_gpu(tmp_path):
    import torch
    exe = _build(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: covered by the gpu test")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 1
    assert "no usable CUDA device" in r.stderr and "no CPU path" in r.stderr


@pytest.mark.gpu
@pytest.mark.skipif(shutil.which("gcc") is None, reason="no gcc")
def test_c_caller_renders_the_readme_square(tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "covered pixels: 256 (expected 256), mismatches: 0" in r.stdout
    assert "stateless backward matches" in r.stdout
