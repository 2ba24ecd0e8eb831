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
