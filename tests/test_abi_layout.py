"""The ctypes mirrors in dirt_b200/_lib.py against the C structs of include/dirt_b200.h, as a C compiler lays them
out: a small C99 program prints sizeof / offsetof of every field, and each must equal what ctypes computed.  A field
added, reordered or retyped on one side only would otherwise pass garbage across the drop-in boundary."""
import ctypes
import os
import shutil
import subprocess

import pytest

from dirt_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STRUCTS = [_lib.DirtShape, _lib.DirtForwardArgs, _lib.DirtBackwardArgs, _lib.DirtSharedBackwardArgs, _lib.DirtPeerAllreduceArgs,
           _lib.DirtVertexShadeArgs, _lib.DirtSceneArgs]


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs a C compiler")
def test_ctypes_structs_have_the_layout_the_c_compiler_gives_the_header(tmp_path):
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "dirt_b200.h"', 'int main(void) {']
    for st in STRUCTS:
        name = st.__name__
        lines.append('    printf("%s . %%zu\\n", sizeof(%s));' % (name, name))
        for field, _ in st._fields_:
            lines.append('    printf("%s %s %%zu\\n", offsetof(%s, %s));' % (name, field, name, field))
    lines += ['    printf("ABI %d\\n", DIRT_B200_ABI_VERSION);', '    return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines) + "\n")
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    got = {}
    for ln in out.splitlines():
        parts = ln.split()
        if len(parts) == 3:
            got[(parts[0], parts[1])] = int(parts[2])
    for st in STRUCTS:
        name = st.__name__
        assert got[(name, ".")] == ctypes.sizeof(st), name
        for field, _ in st._fields_:
            assert got[(name, field)] == getattr(st, field).offset, "%s.%s" % (name, field)
    assert ("ABI %d" % _lib.ABI_VERSION) in out
