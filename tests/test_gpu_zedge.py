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
    "fused_march": (("split", 0),),                     # geometry and deposits in one warp-per-ray kernel (the round-1 form) instead of k_geom + k_deposit
    "coop_lists": (("coop", 1),),                       # the list side of a step as two cooperative launches (grid-wide barriers) instead of graph-replayed kernels
    "no_coop": (("coop", 0),),                          # small steps through the graph-replayed kernels too
    "no_graphs": (("graphs", 0), ("coop", 0)),                      # the kernels of a step launched one by one instead of replayed from CUDA graphs
    "refused_steps": (("cap0", 64), ("batch_units", 960)),   # list buffers start at 64 entries: steps are refused on the device and repeated after growth
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


def test_direction_tables_grow_on_demand(tiny_trusti):
    """The direction tables are built up to nside 2^dir_kmax at the start of a pass and extended when a ray tree refines further
    (the reference allows nside <= 8192): with tables up to nside 8 only, the slab's direct light still matches the oracle."""
    run_sequence(tiny_trusti, record=0, iterations=1, options=(("dir_kmax", 3),))


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


# ---- order-independent accumulation (option "deterministic") ------------------------------------------------------------
def _full_run(scene, options, iterations=2):
    """precalc -> direct light -> scattering iterations on the device; returns every accumulator of the path."""
    from dartray_b200 import abi
    from oracle import oracle as O
    eng = _engine(scene, options)
    st = O.State(scene)
    has_src = scene.nsrc > 0
    has_cell = bool(np.any(scene.lumcell))
    out = {}
    for rt in ([abi.RTT_PRECALC_CELL] if has_cell else []) + ([abi.RTT_PRECALC_SRC] if has_src else []) + \
              ([abi.RTT_DIR_CELL] if has_cell else []) + ([abi.RTT_DIR_SRC] if has_src else []):
        eng.rt_loop(rt, 0)
    psel = np.zeros(scene.grid.ncell + scene.nsrc)
    out["psel_av_direct"] = eng.download(abi.ARR_PSEL_AV, psel).copy()
    tot = np.zeros(scene.lnum)
    for it in range(1, iterations + 1):
        if eng.rt_prepare_scatt(it, tot):
            break
        eng.rt_loop(abi.RTT_SCATT, it)
    eng.download_state(st)
    out["psel_av_scatt"] = eng.download(abi.ARR_PSEL_AV, psel).copy()
    out.update(u_fest=st.u_fest, u_final=st.u_final, lum_lost=st.lum_lost, lumcell=st.lumcell, tot_rad_en=tot)
    for l in range(scene.lnum):
        out[f"scaspe{l}"] = st.scaspe[l]; out[f"scaspe_tot{l}"] = st.scaspe_tot[l]
    eng.close()
    return out


SCHEDULES = (
    (("warps", 16), ("chunk", 8)),
    (("warps", 5), ("chunk", 1), ("ctas_per_sm", 2), ("batch_units", 336)),
    (("use_bulk", 0), ("warps", 9), ("chunk", 3), ("batch_units", 2000), ("stage_split", 3)),
)


@pytest.mark.parametrize("fixture", ["tiny_galaxy", "tiny_trusti"])
def test_deterministic_mode_is_bit_reproducible(fixture, request):
    """With option "deterministic" every contribution is rounded to a per-wavelength grid before it is added, so all
    partial sums are exact and the accumulators cannot depend on the order of the reductions: three runs that schedule
    the rays differently (warps per CTA, rays per chunk, unit batches, bulk reductions or per-lane atomics) give the same
    bits in u_fest, u_final, every scaspe / scaspe_tot plane, lum_lost, lumcell, tot_rad_en and psel_av."""
    scene = request.getfixturevalue(fixture)
    runs = [_full_run(scene, (("deterministic", 1),) + s) for s in SCHEDULES]
    assert runs[0]["u_final"].max() > 0 and runs[0]["scaspe0"].max() > 0
    for other in runs[1:]:
        for k, v in runs[0].items():
            assert np.array_equal(v, other[k]), f"{k} differs between two deterministic runs"


DET_K = 2048    # allowance in quanta: ~sqrt(n)/2 for up to n = 1.6e7 contributions per element


@pytest.mark.parametrize("fixture", ["tiny_galaxy", "tiny_trusti", "small_magtar"])
def test_deterministic_mode_parity(fixture, request):
    """The rounding changes no decision of the ray trees (counters and the intersection checksum stay bit-exact) and every
    accumulator stays within 1e-5 relative + DET_K quanta of the oracle, the quanta being those the engine reports
    (2^-52 of the a-priori bound of each wavelength: dre_get_quanta)."""
    from dartray_b200 import abi
    from oracle import oracle as O
    from test_gpu_parity import assert_fields_close
    scene = request.getfixturevalue(fixture)
    eng = _engine(scene, (("deterministic", 1),))
    orc, st = _oracle(scene)
    gst = O.State(scene)
    has_src = scene.nsrc > 0
    has_cell = bool(np.any(scene.lumcell))
    L = scene.lnum
    qs_max = np.zeros(L)

    def compare(tag, scatt):
        nonlocal qs_max
        eng.download_state(gst)
        qs, qu, ql = eng.quanta()
        assert (qu > 0).all() and (qs > 0).all()
        qs_max = np.maximum(qs_max, qs)
        assert_fields_close(gst.u_fest, st.u_fest, f"u_fest {tag}", atol=DET_K * qu[None, :])
        assert_fields_close(gst.u_final, st.u_final, f"u_final {tag}", atol=DET_K * qu[None, :])
        assert_fields_close(gst.lum_lost, st.lum_lost, f"lum_lost {tag}", atol=DET_K * ql)
        for l in range(L):
            assert_fields_close(gst.scaspe[l], st.scaspe[l], f"scaspe[{l}] {tag}", atol=DET_K * qs[l])
            if scatt:
                assert_fields_close(gst.scaspe_tot[l], st.scaspe_tot[l], f"scaspe_tot[{l}] {tag}", atol=DET_K * qs_max[l])

    for rt in ([abi.RTT_PRECALC_CELL] if has_cell else []) + ([abi.RTT_PRECALC_SRC] if has_src else []) + \
              ([abi.RTT_DIR_CELL] if has_cell else []) + ([abi.RTT_DIR_SRC] if has_src else []):
        c_o = orc.run_pass(rt, 0, st)
        c_g = eng.rt_loop(rt, 0)
        assert_counters_equal(c_g, c_o)
        compare(f"after pass {rt}", False)
    tot_g = np.zeros(L)
    for it in (1, 2):
        f_o = orc.prepare_scatt(it, st)
        f_g = eng.rt_prepare_scatt(it, tot_g)
        assert f_o == f_g
        # tot_rad_en(l) sums N x npix_hp(l) elements of the direct-light scaspe planes, each on the grid qs_max(l)
        assert_fields_close(tot_g, st.tot_rad_en, "tot_rad_en",
                            atol=DET_K * qs_max * np.sqrt(scene.grid.ncell * scene.npix_hp.astype(float)))
        if f_o:
            break
        orc.set_sources(st.lumcell)
        c_o = orc.run_pass(abi.RTT_SCATT, it, st)
        c_g = eng.rt_loop(abi.RTT_SCATT, it)
        assert_counters_equal(c_g, c_o)
        compare(f"after scattering iteration {it}", True)
    eng.close()


@pytest.mark.parametrize("kind", ["galaxy", "trusti"])
def test_two_gpus_bit_identical_in_deterministic_mode(kind):
    """Sums of grid values are exact, so the NCCL all-reduce (rt_mpi) cannot change them either: two GPUs sharing the
    units give the same bits as one GPU with another schedule."""
    import os, subprocess, sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29519", os.path.join(root, "tests", "multigpu_check.py"), kind, "det"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "multigpu_check ok" in r.stdout and "det=1" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_trusti_full_size_properties():
    """C1 at its full size (21 801 cells, 32 wavelengths, 181-MB accumulator rows), through properties that need no
    oracle run:
    (a) a checksum linking the two accumulators: the light scattered out of the rays, sum over cells and HEALPix
        directions of the direct-light scaspe (= tot_rad_en, rt_routines.f90:690-698), equals ksca_norm x the energy
        removed from the rays, sum_c u_final dens cs V (deposit :2499-2525, with sum_pix ffn = 1 :2258-2260);
    (b) linearity, bit for bit: with order-independent accumulation a solve with twice the source luminosity gives
        exactly twice every accumulator (all operations on the path are linear in the luminosity, x2 is exact in fp64,
        the rounding grids scale with it) and the same ray trees."""
    import copy
    from dartray_b200 import abi, scenes
    from oracle import oracle as O
    sc = scenes.trusti_scene()
    assert sc.grid.ncell == 21801 and sc.lnum == 32
    sc2 = copy.copy(sc)
    sc2.lum_p_src = 2.0 * sc.lum_p_src
    res = []
    for s in (sc, sc2):
        eng = _engine(s, (("deterministic", 1),))
        cnt = [eng.rt_loop(abi.RTT_PRECALC_SRC, 0), eng.rt_loop(abi.RTT_DIR_SRC, 0)]
        st = O.State(s)
        eng.download_state(st)
        u_dir, lost_dir = st.u_final.copy(), st.lum_lost.copy()
        tot = np.zeros(s.lnum)
        assert not eng.rt_prepare_scatt(1, tot)
        cnt.append(eng.rt_loop(abi.RTT_SCATT, 1))
        eng.download_state(st)
        res.append((cnt, u_dir, lost_dir, tot, st))
        eng.close()
    # (a) in the default mode (the grid of the deterministic mode is too coarse for the far infrared, where 1e-10 of
    # the luminosity is scattered: DESIGN.md section 4)
    eng = _engine(sc)
    eng.rt_loop(abi.RTT_PRECALC_SRC, 0)
    c_dir = eng.rt_loop(abi.RTT_DIR_SRC, 0)
    sd = O.State(sc)
    eng.download_state(sd)
    tot_d = np.zeros(sc.lnum)
    eng.rt_prepare_scatt(1, tot_d)
    eng.close()
    cs = abi.make_params(sc, abi.RTT_DIR_SRC).cs
    removed = (sd.u_final * sc.dens_arr * (cs * sc.grid.csize ** 3)[:, None]).sum(axis=0)
    assert tot_d.min() > 0
    assert np.allclose(tot_d, sc.ksca_norm * removed, rtol=1e-9, atol=0.0), np.abs(tot_d / (sc.ksca_norm * removed) - 1).max()
    assert np.all(tot_d + sd.lum_lost < sc.lum_p_src[0])      # scattered + lost light cannot exceed what was emitted
    # (b)
    cnt, u_dir, lost_dir, tot, st = res[0]
    cnt2, u_dir2, lost_dir2, tot2, st2 = res[1]
    assert_counters_equal(c_dir, cnt[1])                      # the rounding changes no ray tree
    for a, b in zip(cnt, cnt2):
        assert_counters_equal(a, b)
    assert cnt[1].intersections > 500_000 and cnt[2].intersections > cnt[1].intersections
    assert np.array_equal(u_dir2, 2.0 * u_dir) and np.array_equal(tot2, 2.0 * tot) and np.array_equal(lost_dir2, 2.0 * lost_dir)
    assert np.array_equal(st2.u_final, 2.0 * st.u_final) and np.array_equal(st2.lumcell, 2.0 * st.lumcell)
    assert np.array_equal(st2.lum_lost, 2.0 * st.lum_lost)
    for l in range(sc.lnum):
        assert np.array_equal(st2.scaspe[l], 2.0 * st.scaspe[l]) and np.array_equal(st2.scaspe_tot[l], 2.0 * st.scaspe_tot[l])
