"""N5: the 2D mode (rt_loop_2D rt_routines.f90:1061-1140, skip_cell_2D_rt_loop :1174-1187, find_linked_cells :3364-3578,
fix_symmetry / fix_symmetry_part1-3 :3605-4025, calc_scaspe_indices :4083-4217).  Oracle pins on the CPU, GPU parity below."""
import copy

import numpy as np
import pytest

FLIP = [(1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 0), (1, 0, 1), (0, 1, 1), (1, 1, 1)]


def sym_scene(sc):
    """The 2D mode computes no i_obs (dartray_hub.f90:69) and its scratch arrays conform only without observer directions."""
    s = copy.copy(sc)
    s.dir_i_out = np.zeros((2, 0))
    return s


def test_linked_cells_are_the_mirror_images(tiny_galaxy):
    """check_grid_symmetry (rt_routines.f90:3283-3359): the seven linked cells sit at the mirrored coordinates, same level."""
    from oracle import oracle as O
    sc = sym_scene(tiny_galaxy)
    orc = O.Oracle(sc)
    g = sc.grid
    pos = [i for i in np.nonzero(g.cchild == -1)[0] if (g.ccoord[i] >= 0).all()]
    assert len(pos) > 20
    for i in pos:
        ll = orc.linked_cells(i)
        for q, f in enumerate(FLIP):
            want = g.ccoord[i] * np.where(np.array(f) == 1, -1.0, 1.0)
            assert np.array_equal(g.ccoord[ll[q]], want) and g.lvl[ll[q]] == g.lvl[i] and g.cchild[ll[q]] == -1


@pytest.mark.parametrize("kp", [0, 1, 2])
def test_scaspe_indices_are_the_mirrored_pixels(kp, tiny_galaxy):
    """calc_scaspe_indices: every index array is an involution that sends a HEALPix pixel to the pixel whose centre is the mirror
    image of its own (to rounding of the centre angles)."""
    from oracle import oracle as O
    orc = O.Oracle(tiny_galaxy)
    n = 12 * 4 ** kp
    v = orc.pix2vec(2 ** kp, np.arange(n))
    for q, f in enumerate(FLIP):
        ix = orc.scaspe_symmetry_index(kp, q)
        assert np.array_equal(ix[ix], np.arange(n)), (kp, q)
        want = v * np.where(np.array(f) == 1, -1.0, 1.0)
        assert np.abs(v[ix] - want).max() < 1e-6, (kp, q)


def solve_2d_and_main(sc, orc, O, abi, iterations=1):
    """precalc_cell, dir_cell and scattering iterations in the 2D mode and in the main mode on the same (symmetric) scene."""
    st2, st3 = O.State(sc), O.State(sc)
    n2 = n3 = 0
    for rt in (abi.RTT_PRECALC_CELL, abi.RTT_DIR_CELL):
        n2 += sum(c.intersections for c in orc.run_pass_2d(rt, 0, st2))
        n3 += orc.run_pass(rt, 0, st3).intersections
    for it in range(1, iterations + 1):
        f2 = orc.prepare_scatt(it, st2)
        lum2 = st2.lumcell.copy()
        f3 = orc.prepare_scatt(it, st3)
        assert f2 == f3
        if f2:
            break
        orc.set_sources(lum2)
        n2 += sum(c.intersections for c in orc.run_pass_2d(abi.RTT_SCATT, it, st2))
        orc.set_sources(st3.lumcell)
        n3 += orc.run_pass(abi.RTT_SCATT, it, st3).intersections
    return st2, st3, n2, n3


def test_2d_mode_equals_main_mode_on_a_symmetric_model(tiny_galaxy):
    """The pin of the whole 2D machinery: on a model with the three mirror symmetries the 2D mode (one octant of the sources,
    symmetrised, plus the sources on the planes) gives the fields of the main mode -- u_fest, u_final, every scaspe plane, lum_lost --
    with an eighth of the off-plane ray tracing."""
    from dartray_b200 import abi
    from oracle import oracle as O
    sc = sym_scene(tiny_galaxy)
    orc = O.Oracle(sc, nthreads=8)
    st2, st3, n2, n3 = solve_2d_and_main(sc, orc, O, abi, iterations=1)
    assert n2 < 0.8 * n3        # 9 cells per axis: a third of the leaves sit on the Cartesian planes and are traced in full

    def close(a, b, name, tol=1e-9):
        scale = np.abs(b).max()
        assert scale > 0 and np.abs(a - b).max() <= tol * scale, (name, np.abs(a - b).max() / scale)

    close(st2.u_fest, st3.u_fest, "u_fest")
    # fix_symmetry_part3 adds u_final - u_fest of the mirror cells (rt_routines.f90:3841): the difference of two nearly equal
    # numbers carries their rounding, ~1e-16 u / (delta u), into the sum -- 4e-8 of the scale here, inherent to the 2D formula
    close(st2.u_final, st3.u_final, "u_final", tol=1e-6)
    for l in range(sc.lnum):
        close(st2.scaspe[l], st3.scaspe[l], f"scaspe[{l}]")
    close(st2.lum_lost, st3.lum_lost, "lum_lost")


@pytest.mark.gpu
def test_gpu_symmetry_tables_match_calc_scaspe_indices(tiny_galaxy):
    """The engine reflects the pixel centre and locates it with ang2pix_nest; the oracle restates the reference's ring search."""
    from dartray_b200 import engine
    from oracle import oracle as O
    sc = sym_scene(tiny_galaxy)
    eng, orc = engine.Engine(sc), O.Oracle(sc)
    for kp in (0, 1, 2):
        for w in range(7):
            assert np.array_equal(eng.scaspe_symmetry_index(kp, w), orc.scaspe_symmetry_index(kp, w)), (kp, w)
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["galaxy_base3", "galaxy_base2", "galaxy_isotropic"])
def test_gpu_2d_mode_matches_oracle(kind, tiny_galaxy):
    """dre_run_pass_2d against the oracle's rt_loop_2D: precalculation, direct light and two scattering iterations on a symmetric
    model; counters and the intersection checksum of both loops bit for bit, fields within 1e-5.  base (3,3) has cells on the
    Cartesian planes (second loop, links that coincide with the cell); base (2,2) has none; the third case makes the last wavelength
    isotropic."""
    from dartray_b200 import abi, engine, scenes
    from oracle import oracle as O
    from test_gpu_parity import assert_counters_equal, assert_fields_close
    sc = sym_scene(tiny_galaxy if kind != "galaxy_base2" else scenes.galaxy_scene(max_lvl=3, min_lvl=3, lnum=3, base=(2, 2), name="galaxy_b2"))
    if kind == "galaxy_isotropic":       # the isotropic branches of fix_symmetry_part2 / part3 (incl. the stale-`cc` sum of part2)
        k = sc.kp_sca.copy(); k[-1] = -1; sc.kp_sca = k
    eng = engine.Engine(sc)
    eng.set_option("hash", 1)
    orc, st = O.Oracle(sc, nthreads=8), O.State(sc)
    gst = O.State(sc)

    def total(cs):
        t = abi.Counters()
        for k, _ in abi.Counters._fields_:
            setattr(t, k, sum(getattr(c, k) for c in cs))
        return t

    def compare(tag):
        eng.download_state(gst)
        assert_fields_close(gst.u_fest, st.u_fest, f"u_fest {tag}")
        assert_fields_close(gst.u_final, st.u_final, f"u_final {tag}")
        for l in range(sc.lnum):
            assert_fields_close(gst.scaspe[l], st.scaspe[l], f"scaspe[{l}] {tag}")
        assert_fields_close(gst.lum_lost, st.lum_lost, f"lum_lost {tag}")

    for rt in (abi.RTT_PRECALC_CELL, abi.RTT_DIR_CELL):
        c_o = total(orc.run_pass_2d(rt, 0, st))
        c_g = eng.rt_loop_2d(rt, 0)
        assert_counters_equal(c_g, c_o)
        compare(f"after 2D pass {rt}")
    tot_g = np.zeros(sc.lnum)
    for it in (1, 2):
        f_o = orc.prepare_scatt(it, st)
        assert f_o == eng.rt_prepare_scatt(it, tot_g)
        if f_o:
            break
        orc.set_sources(st.lumcell)
        c_o = total(orc.run_pass_2d(abi.RTT_SCATT, it, st))
        c_g = eng.rt_loop_2d(abi.RTT_SCATT, it)
        assert_counters_equal(c_g, c_o)
        compare(f"after 2D scattering iteration {it}")
    eng.close()
