"""Diagnostic (GPU): per accumulator, the largest difference between a solve with option "deterministic" and a default
solve, relative to the plane maximum and per element — the numbers behind the dynamic-range limit stated in DESIGN.md
section 4.  Run on the GPU box from the repo root: python tools/det_diag.py"""
import sys
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from dartray_b200 import scenes, abi
from oracle import oracle as O
from test_gpu_zedge import _full_run
def report(name, sc):
    a = _full_run(sc, (("deterministic", 1),))
    b = _full_run(sc, ())
    for k in a:
        x, y = np.asarray(a[k], dtype=float), np.asarray(b[k], dtype=float)
        if x.ndim == 2 and k in ("u_fest", "u_final", "lumcell"):
            pm = np.abs(y).max(axis=0, keepdims=True)
        else:
            pm = np.abs(y).max()
        pm = np.where(pm == 0, 1.0, pm)
        d = np.abs(x - y)
        rel = (d / np.maximum(np.abs(y), 1e-300)).max()
        print(f"{name} {k}: max abs/planemax {np.max(d / pm):.3e}  max rel {rel:.3e}", flush=True)
report("galaxy", scenes.galaxy_scene(max_lvl=2, min_lvl=2, lnum=4, name="galaxy_tiny"))
report("trusti", scenes.trusti_scene(tau_z=1.0, base=(10, 2), max_lvl=2, lnum=3, name="trustI_tiny"))
report("magtar", scenes.magtar_scene(max_lvl=2, lnum=4, name="magtar_small"))
