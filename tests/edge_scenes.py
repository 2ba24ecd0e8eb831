"""Edge-case scenes for the parity tests, derived from the standard test scenes (conftest.py).

wide(scene, L)          the same grid with L wavelengths (per-wavelength tables resampled over the wavelength index):
                        L = 1, L = 33 / 40 (two wavelengths per lane, ragged second word), L = 70 (three words: the
                        engine's 4-per-lane ray-march instantiation with one word empty).
with_point_source(...)  a point source inside a dusty cell of a grid whose cells emit too: both source kinds in one
                        model, and the host-cell formulas Q4 / Q10 with tau > 0.
dark(scene)             no luminous source at all (empty cell list, rt_routines.f90:1016).
"""
import copy

import numpy as np


def _resample_cols(a, L):
    a = np.asarray(a, dtype=np.float64)
    n = a.shape[-1]
    if n == 1:
        return np.repeat(a, L, axis=-1)
    x = np.linspace(0.0, n - 1.0, L)
    i0 = np.minimum(np.floor(x).astype(int), n - 2)
    w = x - i0
    return np.ascontiguousarray(a[..., i0] * (1.0 - w) + a[..., i0 + 1] * w)


def wide(scene, L, name=None):
    from dartray_b200 import scenes
    sc = copy.copy(scene)
    sc.name = name or f"{scene.name}_L{L}"
    sc.lambda_um = _resample_cols(scene.lambda_um, L)
    sc.dens_arr = _resample_cols(scene.dens_arr, L)
    sc.ksca_norm = _resample_cols(scene.ksca_norm, L)
    sc.gsca = _resample_cols(scene.gsca, L)
    sc.kp_sca = scenes.set_kp_sca(sc.gsca, 1)
    sc.lumcell = _resample_cols(scene.lumcell, L)
    sc.lum_p_src = _resample_cols(scene.lum_p_src, L) if scene.nsrc else np.zeros((0, L))
    return sc


def with_point_source(scene, pos, lum_scale=0.05, name=None):
    """Adds one point source at `pos` whose luminosity per wavelength is lum_scale x the total cell luminosity."""
    from dartray_b200 import scenes
    sc = copy.copy(scene)
    sc.name = name or f"{scene.name}_psrc"
    p, src_cell, cell_src = scenes.prepare_p_src(scene.grid, [list(pos)])
    sc.ccoord_p_src = p
    sc.src_cell = src_cell
    sc.cell_src = cell_src
    sc.lum_p_src = np.ascontiguousarray(lum_scale * scene.lumcell.sum(axis=0)[None, :])
    return sc


def dark(scene, name=None):
    sc = copy.copy(scene)
    sc.name = name or f"{scene.name}_dark"
    sc.lumcell = np.zeros_like(scene.lumcell)
    sc.lum_p_src = np.zeros_like(scene.lum_p_src)
    return sc
