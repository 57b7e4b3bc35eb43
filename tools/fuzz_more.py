"""development aid: runs the fuzz families of tests/test_fuzz_gpu.py over seeds the committed tests do not use.
    python tools/fuzz_more.py FIRST_SEED COUNT"""
import os, sys, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import test_fuzz_gpu as t

first, count = int(sys.argv[1]), int(sys.argv[2])
families = [("forward_backward", t.test_fuzz_forward_backward, 1.0), ("depth_ordered", t.test_fuzz_depth_ordered_path, 0.25),
            ("extreme", t.test_fuzz_extreme_geometry, 0.5), ("spiky", t.test_fuzz_spiky_spheres_with_and_without_a_hint, 0.25), ("shared", t.test_fuzz_shared_mesh_backward, 0.2), ("dilate", t.test_fuzz_dilate, 0.2)]
bad = 0
for name, fn, share in families:
    n = int(count * share)
    for seed in range(first, first + n):
        try:
            fn(seed)
        except Exception as e:
            bad += 1
            print("FAIL", name, seed, type(e).__name__, str(e).splitlines()[0][:200] if str(e) else "", flush=True)
    print(name, "done", n, flush=True)
print("failures:", bad)
