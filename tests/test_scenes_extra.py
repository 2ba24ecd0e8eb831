"""Harness checks of the C4 (2-D tables on the 3-D grid) and C5 (particles on the 3-D grid) scene builders: the arrays have the
invariants the hot path relies on and the cell values follow the reference's averaging rules."""
import numpy as np


def _tree_ok(sc):
    from oracle import oracle as O
    orc = O.Oracle(sc)
    g = sc.grid
    leaves = np.nonzero(g.cchild == -1)[0]
    rng = np.random.default_rng(1)
    for c in rng.choice(leaves, 200, replace=False):
        cell, lvl = orc.cindex_roundtrip(int(c))
        assert cell == c and lvl == g.lvl[c]
    # leaves tile the model exactly
    assert abs((g.csize[leaves] ** 3).sum() / g.modelsize ** 3 - 1) < 1e-12
    # brute-force point location agrees with find_cc_new2
    pts = rng.uniform(-0.49, 0.49, (300, 3)) * g.modelsize
    got = orc.locate(pts, np.full(len(pts), leaves[0], dtype=np.int32))      # cc_old is always a leaf in the path
    for p, c in zip(pts, got):
        assert g.cchild[c] == -1 and np.all(np.abs(p - g.ccoord[c]) <= g.csize[c] / 2 * (1 + 1e-6))


def test_c4_scene_follows_av_dens_2dto3d():
    """av_dens_2dto3d (user_routines_2dto3d.f90:360-475): the separable evaluation equals the point-by-point triple loop, the model
    is symmetric, and the luminosity of the leaves adds up to the volume integral of the tables inside the model cube."""
    from dartray_b200 import scenes
    sc = scenes.twod3d_scene(max_lvl=3)
    _tree_ok(sc)
    tab = scenes._Table2D(sc.grid.modelsize)
    x, y, z, cs = 1200.0, -3100.0, 173.0, sc.grid.modelsize / 27
    j, k = tab.cell_average(x, y, z, cs)
    j2, _ = tab.cell_average(-y, x, -z, cs)
    assert np.array_equal(j, j2) and np.all(k == 0) and np.all(j > 0)
    npt = min(max(int(cs / (sc.grid.modelsize / 1000.0)), 2), 200)
    step = cs / (npt - 1)
    R, Z = tab.R, tab.Z
    tot = np.zeros(sc.lnum)
    for ix in range(npt - 1):
        for iy in range(npt - 1):
            for iz in range(npt - 1):
                xp, yp, zp = x - cs / 2 + (ix + .5) * step, y - cs / 2 + (iy + .5) * step, z - cs / 2 + (iz + .5) * step
                rp = (xp * xp + yp * yp) ** .5
                ir = len(R) - 1
                while ir >= 1 and not rp > R[ir]:
                    ir -= 1
                i2 = len(Z) - 1
                while i2 >= 1 and not abs(zp) > Z[i2]:
                    i2 -= 1
                if rp > R[-1] or abs(zp) > Z[-1]:
                    continue
                f = tab.F[:sc.lnum]
                d = [(R[ir + 1] - rp) * (Z[i2 + 1] - abs(zp)), (R[ir + 1] - rp) * (abs(zp) - Z[i2]), (rp - R[ir]) * (Z[i2 + 1] - abs(zp)),
                     (rp - R[ir]) * (abs(zp) - Z[i2]), (R[ir + 1] - R[ir]) * (Z[i2 + 1] - Z[i2])]
                tot += (f[:, ir, i2] * d[0] + f[:, ir, i2 + 1] * d[1] + f[:, ir + 1, i2] * d[2] + f[:, ir + 1, i2 + 1] * d[3]) / d[4]
    assert np.abs(j / (tot * step ** 3 / cs ** 3) - 1).max() < 1e-12
    assert sc.ccoord_obs.shape == (1, 3) and sc.ccoord_obs[0, 0] == 8000.0


def test_c5_scene_bins_every_particle_once():
    """av_galaxy (user_routines_Nbody_SPH.f90:403-649): a cell holds the sums of the particles inside its half-open box; leaves tile
    the model, so the leaf luminosities add up to the luminosity of the particles inside the cube, whatever the refinement."""
    from dartray_b200 import scenes
    sc = scenes.nbody_scene(max_lvl=3, lnum=4, max_dlum=1e-3)
    _tree_ok(sc)
    star, l_old, l_young, gas, dust = scenes.nbody_particles()
    inside = (np.abs(star) < sc.grid.modelsize / 2).all(axis=1)
    g, w_old, w_young = scenes.nbody_grid(max_lvl=3, max_dlum=1e-3)
    leaf = g.cchild == -1
    assert abs((w_old[leaf] * g.csize[leaf] ** 3).sum() / (0.7 * l_old[inside].sum()) - 1) < 1e-12
    assert abs((w_young[leaf] * g.csize[leaf] ** 3).sum() / (0.3 * l_young[inside].sum()) - 1) < 1e-12
    # same seed, same particles
    s2 = scenes.nbody_particles()[0]
    assert np.array_equal(star, s2)
    # the refinement criterion holds on every leaf that is not at the deepest level
    coarse = leaf & (g.lvl < g.max_lvl)
    assert np.all((w_old + w_young)[coarse] * g.csize[coarse] ** 3 <= 1e-3 * (1 + 1e-12))


def test_grid_files_round_trip(tmp_path, tiny_galaxy):
    """The reference's grid datasets (io_routines.f90:189-197, 397-498) in the raw-array container of dartray_b200/gridfile.py:
    main grid + one lambda grid per wavelength, written and read back bit for bit; csize_arr rebuilt by the make_csize_arr rule."""
    from dartray_b200 import gridfile
    sc = tiny_galaxy
    gridfile.write_scene_grids(str(tmp_path), "m", sc)
    g, dens_arr, lumcell = gridfile.read_scene_grids(str(tmp_path), "m", sc.lambda_um)
    for k in ("cchild", "lvl", "cindex", "ccoord", "csize", "dens", "dens_stars", "csize_arr", "carea_arr", "cvol_arr"):
        assert np.array_equal(getattr(g, k), getattr(sc.grid, k)), k
    assert g.base == sc.grid.base and g.max_lvl == sc.grid.max_lvl and g.modelsize == sc.grid.modelsize
    assert np.array_equal(dens_arr, sc.dens_arr)
    assert np.abs(lumcell - sc.lumcell).max() <= 1e-15 * np.abs(sc.lumcell).max()      # x / V * V
    import pytest
    with pytest.raises(ValueError):
        gridfile.read_lambda_grid(str(tmp_path / gridfile.lambda_grid_name("m", float(sc.lambda_um[0]))), g.ncell + 1)
