"""GPU parity on edge cases of the path (through the C ABI, against the CPU oracle): wavelength counts that leave
lanes ragged or need several wavelengths per lane, both source kinds in one model, and no luminous source at all.
Same bar as test_gpu_parity.py: counters and the intersection checksum bit for bit, fields within 1e-5 relative."""
import numpy as np
import pytest

import edge_scenes as E
from test_gpu_parity import _engine, _oracle, assert_counters_equal, run_sequence

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("lnum", [1, 33, 40, 70])
def test_wavelength_counts(tiny_trusti, lnum):
    """lnum = 1 (31 idle lanes), 33 and 40 (two wavelengths per lane, second word ragged), 70 (three words, run by the
    four-per-lane instantiation of the ray march): precalc, direct light, i_obs, SED, maps, one scattering iteration.
    The intersection list itself is compared for lnum = 33 (the first count past one word)."""
    run_sequence(E.wide(tiny_trusti, lnum), record=4_000_000 if lnum == 33 else 0, iterations=1)


def test_point_source_in_emitting_dusty_grid(tiny_galaxy):
    """Emitting cells and a point source in one model; the source sits in a dusty cell of the disk, so the host-cell
    formulas (quirks Q4 and Q10: rt_routines.f90:2409-2416, 2519-2525, 2552) run with tau > 0."""
    sc = E.with_point_source(tiny_galaxy, (1000.0, 500.0, 20.0))
    assert sc.dens_arr[sc.cell_src[0]].min() > 0
    run_sequence(sc, record=3_000_000, iterations=1)


@pytest.mark.parametrize("fixture", ["tiny_trusti", "tiny_galaxy"])
def test_no_luminous_source(fixture, request):
    """Empty cell list (luminosity gate rt_routines.f90:1016 rejects every cell / the point source has zero
    luminosity): every pass returns at once with zero counters and leaves the accumulators untouched."""
    from dartray_b200 import abi
    from oracle import oracle as O
    sc = E.dark(request.getfixturevalue(fixture))
    eng = _engine(sc)
    orc, st = _oracle(sc)
    gst = O.State(sc)
    for rt in (abi.RTT_PRECALC_CELL, abi.RTT_PRECALC_SRC, abi.RTT_DIR_CELL, abi.RTT_DIR_SRC):
        if rt in (abi.RTT_PRECALC_SRC, abi.RTT_DIR_SRC) and sc.nsrc == 0:
            continue
        c_o = orc.run_pass(rt, 0, st)
        c_g = eng.rt_loop(rt, 0)
        assert_counters_equal(c_g, c_o)
        assert c_g.rays == 0 and c_g.intersections == 0
    eng.download_state(gst)
    assert not gst.u_final.any() and not gst.u_fest.any() and not gst.lum_lost.any()
    assert all(not p.any() for p in gst.scaspe)
    eng.close()


OPTION_SETS = {
    "unit_batches": (("batch_units", 480),),            # 11 088 units in 24 batches: lists rebuilt per batch
    "phase_computed": (("ftab_kmax", 2),),              # phase-function rows computed per ray above nside 4 (and Q2's pixel-1 row)
    "no_phase_tables": (("ftab_kmax", 0),),
    "lane_atomics": (("use_bulk", 0),),                 # per-lane fp64 reductions instead of the bulk reduction
    "two_slot_ring": (("ring_slots", 2),),
    "stage_quarters": (("stage_split", 4), ("warps", 8)),
    "whole_row_stage": (("stage_split", 1), ("warps", 4)),
    "scan_three_kernels": (("scan_chained", 0),),
    "one_ray_chunks": (("chunk", 1), ("ctas_per_sm", 2)),
}


@pytest.mark.parametrize("variant", sorted(OPTION_SETS))
def test_tuning_options_do_not_change_results(tiny_galaxy, variant):
    """Every dre_set_option knob selects between implementations of the same arithmetic (A/B switches of
    profiles/r01_sweeps.md, the unit-batch size, the computed phase function used beyond the tables): counters and the
    intersection checksum stay bit-exact, fields within 1e-5 of the oracle, whichever is chosen."""
    run_sequence(tiny_galaxy, record=0, iterations=1, options=OPTION_SETS[variant])


def test_unit_batches_point_source(small_trusti):
    """One point source = 48 units: batches of 48 units exactly, and a batch size that does not divide 48."""
    run_sequence(small_trusti, record=0, iterations=1, options=(("batch_units", 48),))


def test_direction_table_limit_is_reported(tiny_trusti):
    """A ray tree that needs a finer HEALPix resolution than the direction tables hold ends the pass with an error
    (no silent truncation): tables up to nside 8, the slab needs more."""
    from dartray_b200 import abi, engine
    eng = _engine(tiny_trusti, (("dir_kmax", 3),))
    eng.rt_loop(abi.RTT_PRECALC_SRC, 0)
    with pytest.raises(engine.EngineError, match="direction tables"):
        eng.rt_loop(abi.RTT_DIR_SRC, 0)
    eng.close()


GRIDS = {
    # six levels with base 2 (the smallest subdivision factor): 1 609 cells, leaves at levels 3-6
    "base2_six_levels": dict(base=(2, 2), max_lvl=6, min_lvl=3, lnum=3, max_dtau=1.0, max_dlum=1e-2, name="galaxy_b2_deep"),
    # different subdivision factors at level 1 and below (base(1) = 5, base(2) = 2), leaves at every level 1-4
    "base5_2": dict(base=(5, 2), max_lvl=4, min_lvl=1, lnum=3, max_dtau=1.0, max_dlum=1e-2, name="galaxy_b52"),
    # no refinement at all: 27 level-1 cells (every neighbour step is a table hit or leaves the model)
    "uniform_27_cells": dict(base=(3, 3), max_lvl=1, min_lvl=1, lnum=3, name="galaxy_uniform27"),
}


@pytest.mark.parametrize("grid", sorted(GRIDS))
def test_grid_shapes(grid):
    """Tree shapes the standard scenes do not have (set_base / increment / ccindd_to_cc, smooth_grid_routines.f90:
    622-649, 751-769, 820-852): intersection lists bit for bit, fields within 1e-5."""
    from dartray_b200 import scenes
    run_sequence(scenes.galaxy_scene(**GRIDS[grid]), record=0 if grid == "base2_six_levels" else 12_000_000, iterations=1)


def test_uniform_grid_point_source():
    """TRUST-I slab on the unrefined 10 x 10 x 10 grid (max_lvl = 1)."""
    from dartray_b200 import scenes
    run_sequence(scenes.trusti_scene(tau_z=1.0, base=(10, 2), max_lvl=1, lnum=3, name="trusti_uniform"), record=1_000_000, iterations=2)


@pytest.mark.parametrize("kp_sca_max", [0, 2])
def test_scattering_resolution(kp_sca_max):
    """kp_sca_max = 0 (12 HEALPix directions per anisotropic wavelength) and 2 (192): other row lengths and phase tables
    than the kp_sca_max = 1 of every shipped input file (set_kp_sca / create_scaspe, rt_routines.f90:2784-2852)."""
    from dartray_b200 import scenes
    sc = scenes.trusti_scene(tau_z=1.0, base=(10, 2), max_lvl=2, lnum=3, kp_sca_max=kp_sca_max, name=f"trusti_kp{kp_sca_max}")
    assert int(sc.npix.max()) == 12 * 4 ** kp_sca_max + sc.ndir
    run_sequence(sc, record=0, iterations=2)


def test_three_point_sources_and_walls_on_every_axis(tiny_galaxy):
    """Several point sources (prepare_p_src rt_routines.f90:535-601) in an emitting grid, and walls on x, y and z
    (check_wall_hit :2572-2596)."""
    import copy
    from dartray_b200 import scenes
    sc = copy.copy(tiny_galaxy)
    sc.name = "galaxy_tiny_3src_walls"
    pos, sc.src_cell, sc.cell_src = scenes.prepare_p_src(sc.grid, [[4000.0, 500.0, 20.0], [-6000.0, 2500.0, -300.0], [0.0, 0.0, 0.0]])
    sc.ccoord_p_src = pos
    tot = tiny_galaxy.lumcell.sum(axis=0)
    sc.lum_p_src = np.ascontiguousarray(np.outer([0.05, 0.02, 0.1], tot))
    ms = sc.grid.modelsize
    sc.wall_on = (1, 1, 1)
    sc.wall_coord = (-0.3 * ms, 0.3 * ms, -0.25 * ms, 0.35 * ms, -0.2 * ms, 0.2 * ms)
    assert len(set(sc.cell_src.tolist())) == 3
    run_sequence(sc, record=0, iterations=1)
