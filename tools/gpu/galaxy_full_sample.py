"""ncu target: the deposit kernel on accumulator rows that do NOT fit L2 (galaxy-shaped grid at the EX1 depth: 165 943 cells x 829
doubles = 1.1 GB).  A whole solve of this grid takes 157 s; this script keeps the luminosity of every 50th emitting cell only, so that
the direct-light pass finishes in seconds while every deposit still lands in the full-size accumulator array (the memory regime of
the full solve).  Prints the pass counters and the device time of the k_deposit launches."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from dartray_b200 import abi, engine  # noqa: E402

scene, desc = bench.load_scene("galaxy_full")
emit = np.nonzero(scene.lumcell.sum(axis=1) > 0)[0]
keep = emit[::50]
lum = np.zeros_like(scene.lumcell)
lum[keep] = scene.lumcell[keep]
eng = engine.Engine(scene, device=0)
eng.set_sources(lum)
for rt in (abi.RTT_PRECALC_CELL, abi.RTT_DIR_CELL):
    c = eng.rt_loop(rt)
    print(rt, "sources", len(keep), "rays", c.rays, "intersections", c.intersections, "deposits", c.deposits,
          "pass_ms", round(c.pass_ms, 2), "k_deposit_ms", round(c.kernel_ms, 2), "launches", c.trace_launches,
          "algorithmic GB/s of k_deposit", round(bench.deposit_bytes(c) / max(c.kernel_ms, 1e-9) / 1e6, 1) if rt == abi.RTT_DIR_CELL else "")
eng.close()
