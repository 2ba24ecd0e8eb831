"""GPU parity at the FULL size of the BASELINE configs: the CUDA engine through the C ABI against the
CPU oracle on the same inputs, default (non-deterministic) accumulation mode.

Per pass: every event counter and the order-independent intersection checksum bit for bit (the checksum covers
unit, step, ray ordinal, intersection ordinal, cell, exit face and the bits of psel of EVERY ray-cell intersection);
u_fest / u_final / every scaspe plane / scaspe_tot / lum_lost / lumcell / tot_rad_en / psel_av / i_obs / SED within 1e-5
relative per element.  What these sizes exercise that the reduced scenes of test_gpu_parity.py do not: accumulator rows
of 830-1040 doubles staged in two chunks with all 32 lanes live, four tree levels, direction tables up to nside 512.

The file sorts before test_gpu_parity.py / test_gpu_zedge.py on purpose (pytest -x stops at the first failure).
The oracle runs on all host cores: C1 ~15 s, the galaxy bench workload ~1.5 min, C2 ~2.5 min on 16 cores."""
import os

import numpy as np
import pytest

from test_gpu_parity import run_sequence

pytestmark = pytest.mark.gpu

NTH = os.cpu_count() or 8


def test_c1_trusti_full_size():
    """C1 = TRUSTI_GRIDS/input_trustI_tau0.01_LD.in: 21 801 cells, 32 wavelengths, 1 point source, z wall; the whole
    solve (precalc_src, dir_src, i_obs_dir_src, every scattering iteration up to convergence, i_obs, SED)."""
    from dartray_b200 import scenes
    sc = scenes.trusti_scene()
    assert sc.grid.ncell == 21801 and sc.lnum == 32
    run_sequence(sc, record=0, iterations=6, nthreads=NTH, maps=False)


def test_c3_galaxy_bench_workload_full_size():
    """The galaxy-shaped emitting-cell grid bench.py keeps in `also` (input_galaxy_example1.in parameters, max_lvl 3):
    6 994 cells, 5 733 emitting leaves x 48 sectors, 16 wavelengths: precalc_cell, dir_cell, i_obs_dir_cell, scattering
    iteration 1, i_obs."""
    from dartray_b200 import scenes
    sc = scenes.galaxy_scene(max_lvl=3, min_lvl=2)
    assert sc.lnum == 16
    run_sequence(sc, record=0, iterations=1, nthreads=NTH, maps=False)


def test_c2_magtar_full_size():
    """C2 = MAGTAR_GRIDS/input_shell.in: 16 337 cells, 28 wavelengths, point source at the origin (on a cell corner),
    isotropic + kp 0 + kp 1 wavelengths: precalc_src, dir_src (nside up to 512), i_obs_dir_src, scattering iteration 1, i_obs."""
    from dartray_b200 import scenes
    sc = scenes.magtar_scene()
    assert sc.lnum == 28
    run_sequence(sc, record=0, iterations=1, nthreads=NTH, maps=False)


def test_c4_2dto3d_full_size():
    """C4 = 2DTO3D_GRIDS/input_2dto3d_example.in at its own depth (max_lvl 4: 433 783 cells, 387 099 emitting leaves, 9 wavelengths):
    rt_loop_iobs towards the internal observer at (8000, 0, 0) pc, then the kp_maps = 8 map (HEALPix nside 256)."""
    from dartray_b200 import scenes
    from test_gpu_parity import run_iobs_only
    sc = scenes.twod3d_scene()
    assert sc.grid.ncell == 433783
    run_iobs_only(sc, kp_maps=8)
