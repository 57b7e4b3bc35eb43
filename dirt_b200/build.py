"""Builds libdirt_b200.so in-tree with nvcc for sm_100a (no torch, no TF dependency).

    python -m dirt_b200.build [--force] [--verbose]

The library links the CUDA runtime statically, so it loads in any process (torch, TF, plain C).
-fmad=false: nvcc must not contract a*b+c into FMAs; coverage and depth keys are specified
as individually rounded IEEE operations (docs/RASTER_SPEC.md).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdirt_b200.so")
SOURCES = ["setup_raster.cu", "depth_order.cu", "forward.cu", "backward.cu", "dilate.cu", "geometry.cu", "lighting.cu", "peer_allreduce.cu", "scene_oceanic.cu", "scene_shaders.cu", "capi.cu", "capi_host.cu", "capi_noalloc.cu"]
HEADERS = ["raster_core.cuh", "forward_common.cuh", "kernels.h", os.path.join("..", "..", "include", "dirt_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-cudart", "static",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, defines=(), out=None):
    """`defines` / `out` build a tuning variant (e.g. defines=["KCTAS=8"], out="libvar.so")."""
    if out is None and not force and not needs_build():
        return LIB
    objdir = os.path.join(HERE, "build" if out is None else "build_" + os.path.basename(out))
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()
    procs = []
    objs = []
    for s in SOURCES:
        obj = os.path.join(objdir, s.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, s), "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        log, _ = p.communicate()
        if verbose or p.returncode != 0:
            sys.stderr.write(log)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    target = LIB if out is None else os.path.join(HERE, out)
    cmd = [nvcc, "-shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a", "-o", target] + objs
    subprocess.check_call(cmd)
    return target


if __name__ == "__main__":
    defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    outs = [a[6:] for a in sys.argv[1:] if a.startswith("--out=")]
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, defines=defs, out=outs[0] if outs else None))
