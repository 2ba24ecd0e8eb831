"""Synthetic inputs for the ray-tracing path: adaptive Cartesian grids + opacities + sources.

The reference ships no grid files, so the bench/test inputs are generated
here.  The builder follows the structure of the reference grid programs
(create_adap_grid_trustI.f90:157-395: level-by-level subdivision, children appended as contiguous
blocks of base**3 cells with x fastest, 2:1 face balance through `subdivide_neighbour_cells`,
`cindex` packing at :344-348) so that the arrays have exactly the invariants the hot path relies
on.  Grid creation itself is OUT of the hot-path scope; this is harness code.
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import dataclass, field

import numpy as np

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")

# single-precision pi of the reference (smooth_grid_routines.f90:311-315)
HALFPI = float(np.float32(np.arcsin(np.float32(1.0))))
PI = 2 * HALFPI
PARSEC = float(np.float32(3.08567758e16))
CSPEED = float(np.float32(2.99792458e8))


def _powi(x: float, n: int) -> float:
    """libgcc __powidf2 (what gfortran emits for real**integer)."""
    m = abs(n)
    r = x if (m & 1) else 1.0
    m >>= 1
    while m:
        x = x * x
        if m & 1:
            r = r * x
        m >>= 1
    return r if n >= 0 else 1.0 / r


@dataclass
class Grid:
    modelsize: float
    base: tuple
    max_lvl: int
    cchild: np.ndarray
    lvl: np.ndarray
    cindex: np.ndarray
    ccoord: np.ndarray  # (N,3) C order == Fortran (3,N)
    csize: np.ndarray
    dens: np.ndarray  # extinction coefficient at the reference wavelength [1/pc]
    dens_stars: np.ndarray
    csize_arr: np.ndarray = field(default=None)
    carea_arr: np.ndarray = field(default=None)
    cvol_arr: np.ndarray = field(default=None)

    @property
    def ncell(self):
        return int(self.cchild.shape[0])

    @property
    def leaves(self):
        return np.nonzero(self.cchild == -1)[0]


class _Builder:
    """Restatement of the grid-creation loop (create_adap_grid_trustI.f90:157-395 and the
    tree helpers of smooth_grid_routines.f90:608-852)."""

    def __init__(self, modelsize, base, max_lvl, calc_dens, subdivision, max_ncell=4_000_000):
        self.modelsize = float(modelsize)
        self.base = (int(base[0]), int(base[1]))
        self.max_lvl = int(max_lvl)
        self.calc_dens = calc_dens
        self.subdivision = subdivision
        self.max_ncell = max_ncell
        self.baseinv = (1.0 / self.base[0], 1.0 / self.base[1])
        self.basediv = []
        for b in self.base:  # set_base smooth_grid_routines.f90:622-649
            j = 0
            while not (b ** 3 < 2 ** j):
                j += 1
            self.basediv.append(2 ** j)
        self.basemask = [d - 1 for d in self.basediv]
        self.cchild = [-1]
        self.lvl = [0]
        self.cindex = [0]
        self.ccoord = [(0.0, 0.0, 0.0)]
        self.csize = [self.modelsize]
        d, s = calc_dens(0.0, 0.0, 0.0, self.modelsize)
        self.dens = [d]
        self.dens_stars = [s]

    def cellsize(self, clvl):  # calc_cellsize :608-619
        if clvl > 0:
            return self.modelsize * self.baseinv[0] * _powi(self.baseinv[1], clvl - 1)
        return self.modelsize

    def cindex_to_ccindd(self, nc, clvl):  # :793-813
        cc = [[0, 0, 0] for _ in range(self.max_lvl + 1)]
        ci = self.cindex[nc]
        for i in range(1, clvl + 1):
            if i > 1:
                ib = 1
                j = ((ci // (self.basediv[0] * self.basediv[1] ** (i - 2))) & self.basemask[1]) - 1
            else:
                ib = 0
                j = (ci & self.basemask[0]) - 1
            b = self.base[ib]
            cc[i] = [j % b, (j // b) % b, (j // (b * b)) % b]
        return cc

    def ccindd_to_cc(self, cc):  # :820-852
        l = 1
        for i in range(1, self.max_lvl + 1):
            b = self.base[1 if i > 1 else 0]
            k = (cc[i][2] * b + cc[i][1]) * b + cc[i][0] + 1
            l = l + k - 1
            if self.cchild[l] == -1:
                return l, i
            l = self.cchild[l]
        return -1, -1

    def find_neighbour(self, nc, isel, inc):  # find_neighbours :656-743
        clvl = self.lvl[nc]
        cc = self.cindex_to_ccindd(nc, clvl)
        for a in range(3):
            if a == isel:
                k = inc
                for j in range(clvl, 0, -1):
                    b = self.base[1 if j > 1 else 0]
                    v = cc[j][a] + k
                    if v >= b:
                        v = 0
                    elif v < 0:
                        v = b - 1
                    else:
                        k = 0
                    cc[j][a] = v
                if k != 0:
                    return -1
                for j in range(clvl + 1, self.max_lvl + 1):
                    b = self.base[1 if j > 1 else 0]
                    cc[j][a] = 0 if inc == 1 else b - 1
            else:
                bcoord = 0.5 * self.cellsize(clvl)
                for j in range(clvl + 1, self.max_lvl + 1):
                    cs = self.cellsize(j)
                    cc[j][a] = int(math.floor(bcoord / cs))
                    bcoord = math.fmod(bcoord, cs)
        cell, _ = self.ccindd_to_cc(cc)
        return cell

    def subdivide_cell(self, cc, nlevel_cc):  # :318-395
        self.cchild[cc] = len(self.cchild)
        ib = 1 if nlevel_cc > 0 else 0
        b = self.base[ib]
        if len(self.cchild) + b ** 3 > self.max_ncell:
            raise RuntimeError("too many cells")
        cellsize = self.cellsize(nlevel_cc + 1)
        x0, y0, z0 = self.ccoord[cc]
        par = (x0, y0, z0)
        for iz in range(b):
            for iy in range(b):
                for ix in range(b):
                    code = (iz * b + iy) * b + ix + 1
                    if nlevel_cc > 0:
                        ci = self.cindex[cc] | (code * self.basediv[0] * self.basediv[1] ** (nlevel_cc - 1))
                    else:
                        ci = self.cindex[cc] | code
                    ii = (ix, iy, iz)
                    if b % 2 == 1:
                        pos = tuple(par[j] + cellsize * float(ii[j] - b // 2) for j in range(3))
                    else:
                        pos = tuple(par[j] + cellsize * float(ii[j] - b // 2 + 0.5) for j in range(3))
                    d, s = self.calc_dens(pos[0], pos[1], pos[2], cellsize)
                    self.cindex.append(ci)
                    self.csize.append(cellsize)
                    self.lvl.append(nlevel_cc + 1)
                    self.ccoord.append(pos)
                    self.dens.append(d)
                    self.dens_stars.append(s)
                    self.cchild.append(-1)

    def subdivide_neighbours(self, ncurr, nlevel):  # :287-316 + check_level_jump :774-787
        for isel in range(3):
            for inc in (1, -1):
                cc = self.find_neighbour(ncurr, isel, inc)
                if cc < 0:
                    continue
                lc, ln = self.lvl[cc], self.lvl[ncurr]
                jump = (ln == nlevel and ln > lc) or (ln < nlevel and ln > lc + 1)
                if jump:
                    self.subdivide_cell(cc, lc)

    def run(self):  # subdivision_loop :157-283
        nlevel = 0
        nstart = [0, 1]
        while True:
            for ncurr in range(nstart[nlevel], nstart[nlevel + 1]):
                x, y, z = self.ccoord[ncurr]
                cellsize = self.cellsize(nlevel)
                user = nlevel < self.max_lvl and self.subdivision(x, y, z, cellsize, nlevel, self.dens[ncurr], self.dens_stars[ncurr])
                if user or self.lvl[ncurr] < nlevel:
                    if self.lvl[ncurr] > 1:
                        self.subdivide_neighbours(ncurr, nlevel)
                    if self.lvl[ncurr] == nlevel and self.cchild[ncurr] == -1:
                        self.subdivide_cell(ncurr, nlevel)
            nlevel += 1
            nstart.append(len(self.cchild))
            if nlevel < self.max_lvl:
                continue
            break
        a, b = nstart[nlevel], nstart[nlevel + 1]
        while True:
            for ncurr in range(a, b):
                if self.lvl[ncurr] < nlevel and self.lvl[ncurr] > 1:
                    self.subdivide_neighbours(ncurr, nlevel)
            a, b = b, len(self.cchild)
            if b - a > 0:
                continue
            break
        return self.finish()

    def finish(self) -> Grid:
        g = Grid(
            modelsize=self.modelsize, base=self.base, max_lvl=self.max_lvl,
            cchild=np.asarray(self.cchild, dtype=np.int32), lvl=np.asarray(self.lvl, dtype=np.int32),
            cindex=np.asarray(self.cindex, dtype=np.int64),
            ccoord=np.ascontiguousarray(np.asarray(self.ccoord, dtype=np.float64)),
            csize=np.asarray(self.csize, dtype=np.float64), dens=np.asarray(self.dens, dtype=np.float64),
            dens_stars=np.asarray(self.dens_stars, dtype=np.float64))
        # make_csize_arr io_routines.f90:669-696
        cs = [self.modelsize]
        for i in range(1, self.max_lvl + 1):
            cs.append(self.modelsize * (self.baseinv[0] * _powi(self.baseinv[1], i - 1)))
        g.csize_arr = np.asarray(cs, dtype=np.float64)
        g.carea_arr = g.csize_arr * g.csize_arr
        g.cvol_arr = g.csize_arr * g.csize_arr * g.csize_arr
        return g


def build_grid(modelsize, base, max_lvl, calc_dens, subdivision, max_ncell=4_000_000) -> Grid:
    return _Builder(modelsize, base, max_lvl, calc_dens, subdivision, max_ncell).run()


# --------------------------------------------------------------------------------------
# per-wavelength scattering resolution: set_npix_arr rt_routines.f90:5235-5315
# --------------------------------------------------------------------------------------
def set_kp_sca(gsca, kp_sca_max):
    nmin = 5
    g_th = float(np.float32(2) * np.float32(1e-3))
    kp = np.zeros(len(gsca), dtype=np.int32)
    for i, gg in enumerate(gsca):
        g = abs(float(gg))
        if g < g_th:
            kp[i] = -1
            continue
        pmax = (1 - g ** 2) / (1 + g ** 2 - 2 * g) ** 1.5
        pmin = (1 - g ** 2) / (1 + g ** 2 + 2 * g) ** 1.5
        cos_hm = ((2 * (1 - g ** 2) / (pmax + pmin)) ** float(np.float32(2.0) / np.float32(3.0)) - (1 + g ** 2)) / (-2 * g)
        fwhm = 2 * math.acos(cos_hm)
        delta_min = fwhm / nmin
        k_min = int(1 / (2 * math.log10(2.0)) * math.log10(4 * PI / (12 * delta_min ** 2)))
        kp[i] = min(k_min, kp_sca_max)
    return kp


@dataclass
class Scene:
    name: str
    grid: Grid
    lambda_um: np.ndarray
    dens_arr: np.ndarray      # (N,L) C order == Fortran dens_arr(L,N)
    ksca_norm: np.ndarray
    gsca: np.ndarray
    kp_sca: np.ndarray
    dir_i_out: np.ndarray     # (2,ndir) C order == Fortran dir_i_out(ndir,2)
    lumcell: np.ndarray       # (N,L)
    ccoord_p_src: np.ndarray  # (S,3)
    lum_p_src: np.ndarray     # (S,L)
    src_cell: np.ndarray      # (N) int32, -1 = none
    cell_src: np.ndarray      # (S)
    bm_par: int = 5
    bm_par_sca: int = 5
    bm_par_max: int = 200
    rad_lim: float = 0.1
    accuracy: float = 0.1
    conv_en_lim: float = 0.01
    wall_on: tuple = (0, 0, 0)
    wall_coord: tuple = (0.0, 0.0, 0.0, 0.0, 0.0, 0.0)  # model units (after set_walls)
    ccoord_obs: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))

    @property
    def lnum(self):
        return int(self.dens_arr.shape[1])

    @property
    def ndir(self):
        return int(self.dir_i_out.shape[1])

    @property
    def nsrc(self):
        return int(self.ccoord_p_src.shape[0])

    @property
    def npix_hp(self):
        return np.where(self.kp_sca < 0, 1, 12 * 4 ** np.maximum(self.kp_sca, 0)).astype(np.int32)

    @property
    def npix(self):
        return (self.npix_hp + self.ndir).astype(np.int32)

    def en_lim(self):
        """set_en_lim rt_routines.f90:3137-3158"""
        tot_sources = int((self.grid.cchild == -1).sum()) + self.nsrc
        return self.accuracy / (tot_sources * 0.25)

    def cs(self):
        return CSPEED / PARSEC  # set_units rt_routines.f90:5192


def prepare_p_src(grid: Grid, pos):
    """prepare_p_src + fix_ccoord_p_src rt_routines.f90:535-601"""
    pos = np.array(pos, dtype=np.float64).reshape(-1, 3).copy()
    tol_p = float(np.float32(1e-6))
    src_cell = np.full(grid.ncell, -1, dtype=np.int32)
    cell_src = np.full(pos.shape[0], -1, dtype=np.int32)
    leaves = grid.leaves
    for i in range(pos.shape[0]):
        rel = pos[i][None, :] - grid.ccoord[leaves]
        inside = (np.abs(rel) < (1.0 + tol_p) * grid.csize[leaves][:, None] / 2.0).all(axis=1)
        idx = np.nonzero(inside)[0]
        if len(idx) == 0:
            raise RuntimeError("host cell not found")
        j = int(leaves[idx[0]])
        src_cell[j] = i
        cell_src[i] = j
        r = pos[i] - grid.ccoord[j]
        cs = grid.csize[j]
        for a in range(3):
            if abs(r[a]) >= cs / 2.0:
                pos[i, a] += -tol_p * cs if r[a] > 0 else tol_p * cs
    return pos, src_cell, cell_src


def _load_table(name):
    with open(os.path.join(_DATA, name + ".json")) as f:
        return json.load(f)


def _cache_path(tag):
    d = os.environ.get("DARTRAY_CACHE", "/tmp/dartray_b200_cache")
    os.makedirs(d, exist_ok=True)
    return os.path.join(d, tag + ".npz")


def _grid_cached(tag, make):
    p = _cache_path(tag)
    if os.path.exists(p):
        z = np.load(p)
        return Grid(modelsize=float(z["modelsize"]), base=tuple(int(x) for x in z["base"]), max_lvl=int(z["max_lvl"]),
                    cchild=z["cchild"], lvl=z["lvl"], cindex=z["cindex"], ccoord=z["ccoord"], csize=z["csize"],
                    dens=z["dens"], dens_stars=z["dens_stars"], csize_arr=z["csize_arr"], carea_arr=z["carea_arr"],
                    cvol_arr=z["cvol_arr"])
    g = make()
    np.savez(p, modelsize=g.modelsize, base=np.asarray(g.base), max_lvl=g.max_lvl, cchild=g.cchild, lvl=g.lvl,
             cindex=g.cindex, ccoord=g.ccoord, csize=g.csize, dens=g.dens, dens_stars=g.dens_stars,
             csize_arr=g.csize_arr, carea_arr=g.carea_arr, cvol_arr=g.cvol_arr)
    return g


def trusti_grid(tau_z=0.01, base=(10, 2), max_lvl=4, min_lvl=1, modelsize=10.0) -> Grid:
    """TRUST-I slab grid: create_adap_grid_trustI.f90:56-100 + user_routines_trustI.f90:167-221
    (slab x,y in [-5,5], z in [-5,-2]; `subdivision` 'standard' with min_lvl_in = 0)."""
    z0, z1 = -5.0, -2.0
    ndust = tau_z / 3.0

    def calc_dens(x, y, z, cs):
        inside = (-5.0 <= x <= 5.0) and (-5.0 <= y <= 5.0) and (z0 <= z <= z1)
        return (ndust if inside else 0.0), 0.0

    def crit(x, y, z, cs, nlevel, dens, dens_stars):
        return (abs(z + cs / 2.0 - z1) / abs(z1) < float(np.float32(1e-5))) or (nlevel < min_lvl)

    tag = f"trusti_t{tau_z}_b{base[0]}_{base[1]}_l{max_lvl}_m{min_lvl}"
    return _grid_cached(tag, lambda: build_grid(modelsize, base, max_lvl, calc_dens, crit))


def trusti_scene(tau_z=0.01, base=(10, 2), max_lvl=4, lnum=None, kp_sca_max=1, name="trustI") -> Scene:
    """C1: TRUSTI_GRIDS/input_trustI_tau0.01_LD.in — 1 point source at (0,0,4), 7 directions."""
    g = trusti_grid(tau_z, base, max_lvl)
    t = _load_table("trusti")
    L = len(t["lambda_um"]) if lnum is None else lnum
    sel = np.arange(len(t["lambda_um"]))
    if lnum is not None:  # sub-sample wavelengths evenly (small test cases)
        sel = np.unique(np.linspace(0, len(t["lambda_um"]) - 1, lnum).round().astype(int))
        L = len(sel)
    kscale = np.asarray(t["kext_scale"])[sel]
    dens_arr = np.ascontiguousarray(g.dens[:, None] * kscale[None, :])
    gsca = np.asarray(t["gsca"])[sel]
    theta = np.arange(7) * (math.pi / 6)
    theta = np.array([float("%.6g" % v) for v in theta])  # the file stores 6 significant digits
    dir_i_out = np.ascontiguousarray(np.stack([theta, np.zeros(7)]))
    pos, src_cell, cell_src = prepare_p_src(g, [[0.0, 0.0, 4.0]])
    ms = g.modelsize
    return Scene(name=name, grid=g, lambda_um=np.asarray(t["lambda_um"])[sel], dens_arr=dens_arr,
                 ksca_norm=np.asarray(t["ksca_norm"])[sel], gsca=gsca, kp_sca=set_kp_sca(gsca, kp_sca_max),
                 dir_i_out=dir_i_out, lumcell=np.zeros((g.ncell, L)), ccoord_p_src=pos,
                 lum_p_src=np.ascontiguousarray(np.asarray(t["lum_p_src"])[sel][None, :]),
                 src_cell=src_cell, cell_src=cell_src, bm_par=10, bm_par_sca=5, bm_par_max=200, rad_lim=0.1,
                 accuracy=0.1, conv_en_lim=0.01, wall_on=(0, 0, 1),
                 wall_coord=(0.0, 0.0, 0.0, 0.0, 0.0 * ms - ms / 2.0, 0.3 * ms - ms / 2.0))


# --------------------------------------------------------------------------------------
# C2: MAGTAR ellipsoidal dust shell with a central point source
# --------------------------------------------------------------------------------------
def magtar_grid(tau_z=0.48, base=(10, 3), max_lvl=3, min_lvl=1, max_dtau=0.048, modelsize=10.0,
                ax=1.09, by=2.18, cz=2.18, width=0.10) -> Grid:
    """MAGTAR_GRIDS/input_grid_shell.in: av_rho_dust_shell user_routines_magtar.f90:159-210 (9^3 sub-samples
    of the cell, ellipsoidal radius within `width` of 1) and the 'shell' criterion
    create_adap_grid_magtar.f90:80-106."""
    ax, by, cz, width = (float(np.float32(v)) for v in (ax, by, cz, width))  # real32 module variables (:17-23)
    nstep = 10
    frac = (np.arange(nstep - 1) + 0.5) / (nstep - 1)

    def calc_dens(x, y, z, cs):
        xp = (x - cs / 2.0) + frac * cs
        yp = (y - cs / 2.0) + frac * cs
        zp = (z - cs / 2.0) + frac * cs
        rad = np.sqrt((xp[:, None, None] / ax) ** 2 + (yp[None, :, None] / by) ** 2 + (zp[None, None, :] / cz) ** 2)
        n_in = int((np.abs(rad - 1.0) < width).sum())
        s = cs / (nstep - 1)
        return n_in * tau_z * (s * s * s) / cs ** 3, 0.0

    def crit(x, y, z, cs, nlevel, dens, dens_stars):
        return (dens * cs > max_dtau) or (nlevel < min_lvl)

    tag = f"magtar_t{tau_z}_b{base[0]}_{base[1]}_l{max_lvl}_m{min_lvl}_d{max_dtau}"
    return _grid_cached(tag, lambda: build_grid(modelsize, base, max_lvl, calc_dens, crit))


def magtar_scene(tau_z=0.48, base=(10, 3), max_lvl=3, lnum=None, kp_sca_max=1, max_dtau=0.048, name="magtar") -> Scene:
    """C2: MAGTAR_GRIDS/input_shell.in — point source at the origin (T_eff 3660 K Planck), 7 directions, no wall."""
    g = magtar_grid(tau_z, base, max_lvl, max_dtau=max_dtau)
    t = _load_table("magtar")
    sel = np.arange(len(t["lambda_um"]))
    if lnum is not None:
        sel = np.unique(np.linspace(0, len(sel) - 1, lnum).round().astype(int))
    L = len(sel)
    kscale = np.asarray(t["kext_scale"])[sel]
    gsca = np.asarray(t["gsca"])[sel]
    theta = np.array([float("%.6g" % v) for v in np.arange(7) * (math.pi / 6)])
    dir_i_out = np.ascontiguousarray(np.stack([theta, np.zeros(7)]))
    pos, src_cell, cell_src = prepare_p_src(g, [[0.0, 0.0, 0.0]])
    return Scene(name=name, grid=g, lambda_um=np.asarray(t["lambda_um"])[sel],
                 dens_arr=np.ascontiguousarray(g.dens[:, None] * kscale[None, :]),
                 ksca_norm=np.asarray(t["ksca_norm"])[sel], gsca=gsca, kp_sca=set_kp_sca(gsca, kp_sca_max),
                 dir_i_out=dir_i_out, lumcell=np.zeros((g.ncell, L)), ccoord_p_src=pos,
                 lum_p_src=np.ascontiguousarray(np.asarray(t["lum_p_src"])[sel][None, :]),
                 src_cell=src_cell, cell_src=cell_src, bm_par=10, bm_par_sca=5, bm_par_max=200, rad_lim=0.1,
                 accuracy=0.1, conv_en_lim=0.01)


# --------------------------------------------------------------------------------------
# C3: galaxy-shaped grid (disk + bulge), every luminous leaf cell emits
# --------------------------------------------------------------------------------------
def _zavg_exp(z, cs, zs):
    """mean of exp(-|z'|/zs) over z' in [z-cs/2, z+cs/2] (analytic vertical cell average)."""
    a, b = z - cs / 2.0, z + cs / 2.0
    if a >= 0 or b <= 0:
        lo, hi = (a, b) if a >= 0 else (-b, -a)
        return zs * (math.exp(-lo / zs) - math.exp(-hi / zs)) / cs
    return zs * ((1.0 - math.exp(a / zs)) + (1.0 - math.exp(-b / zs))) / cs


def galaxy_grid(base=(3, 3), max_lvl=4, min_lvl=2, max_dtau=0.01, max_dlum=0.5e-6, modelsize=28000.0,
                z_subd_lim=50.0, r_subd_lim=4000.0, tau_scale=1.0):
    """Galaxy-shaped stand-in for GALAXY_GRIDS/input_grid_galaxy_example1.in: geometry
    parameters from the input file (old/young stellar disks h=3200 pc, z=140/50 pc; dust disks
    tau_B=3.17/0.61, h=5200/3200 pc, z=140/50 pc; bulge r_eff=400 pc; truncation R<14000, |z|<4000),
    double-exponential profiles, 'standard' subdivision criterion create_adap_grid_galaxy.f90:270-296.
    Returns (grid, w_old, w_young): dens = B-band extinction coefficient [1/pc]; w_* = per-cell stellar
    weights (fraction of the component's total luminosity per unit volume)."""
    rtrun, max_z = 14000.0, 4000.0
    hs, zs_old, zs_young = 3200.0, 140.0, 50.0
    hd1, zd1, tau1 = 5200.0, 140.0, 3.17 * tau_scale
    hd2, zd2, tau2 = 3200.0, 50.0, 0.61 * tau_scale
    reff, bd = 400.0, 1.0
    norm_disk = lambda h, z: 4.0 * math.pi * h * h * z      # volume integral of exp(-R/h-|z|/z)
    norm_bulge = 8.0 * math.pi * (reff / 7.67 ** 4) ** 3 * 0.6 * math.factorial(11) / 4.0   # int exp(-7.67 (m/reff)^(1/4)) dV, q = 0.6

    def comps(x, y, z, cs):
        R = math.sqrt(x * x + y * y)
        if R - cs > rtrun or abs(z) - cs / 2.0 > max_z:
            return 0.0, 0.0, 0.0
        dust = tau1 / (2.0 * zd1) * math.exp(-R / hd1) * _zavg_exp(z, cs, zd1) + \
            tau2 / (2.0 * zd2) * math.exp(-R / hd2) * _zavg_exp(z, cs, zd2)
        m = math.sqrt(R * R + (z / 0.6) ** 2)
        old = math.exp(-R / hs) * _zavg_exp(z, cs, zs_old) / norm_disk(hs, zs_old) + \
            bd * math.exp(-7.67 * (max(m, 0.25 * cs) / reff) ** 0.25) / norm_bulge
        young = math.exp(-R / hs) * _zavg_exp(z, cs, zs_young) / norm_disk(hs, zs_young)
        return dust, old / (1.0 + bd), young

    def calc_dens(x, y, z, cs):
        d, o, yv = comps(x, y, z, cs)
        return d, o + yv

    def crit(x, y, z, cs, nlevel, dens, dens_stars):
        return (dens * cs > max_dtau) or (0.5 * dens_stars * cs ** 3 > max_dlum) or (nlevel < min_lvl) or \
            (abs(z) < z_subd_lim and math.sqrt(x * x + y * y) < r_subd_lim)

    tag = f"galaxy_b{base[0]}_{base[1]}_l{max_lvl}_m{min_lvl}_t{max_dtau}_d{max_dlum}_s{tau_scale}"
    g = _grid_cached(tag, lambda: build_grid(modelsize, base, max_lvl, calc_dens, crit))
    w_old = np.zeros(g.ncell); w_young = np.zeros(g.ncell)
    for i in np.nonzero(g.cchild == -1)[0]:
        _, o, yv = comps(g.ccoord[i, 0], g.ccoord[i, 1], g.ccoord[i, 2], g.csize[i])
        w_old[i], w_young[i] = o, yv
    return g, w_old, w_young


def galaxy_scene(max_lvl=4, min_lvl=2, lnum=None, kp_sca_max=1, max_dtau=0.01, max_dlum=0.5e-6, tau_scale=1.0,
                 name="galaxy", base=(3, 3)) -> Scene:
    """C3: GALAXY_GRIDS/input_galaxy_example1.in — no point sources, every leaf with stellar emissivity
    emits (rtt_precalc_cell / rtt_dir_cell), 8 directions theta = 0..pi/2, z wall [0.3, 0.7]."""
    g, w_old, w_young = galaxy_grid(base=base, max_lvl=max_lvl, min_lvl=min_lvl, max_dtau=max_dtau, max_dlum=max_dlum,
                                    tau_scale=tau_scale)
    t = _load_table("galaxy")
    sel = np.arange(len(t["lambda_um"]))
    if lnum is not None:
        sel = np.unique(np.linspace(0, len(sel) - 1, lnum).round().astype(int))
    L = len(sel)
    kscale = np.asarray(t["kext_scale"])[sel]
    gsca = np.asarray(t["gsca"])[sel]
    sed_old = np.asarray(t["sed_old"], dtype=np.float64); sed_young = np.asarray(t["sed_young"], dtype=np.float64)
    # the pt11 SED files stop at 2.2 um: continue with a Rayleigh-Jeans-like fall-off (synthetic)
    lam = np.asarray(t["lambda_um"])
    for sed in (sed_old, sed_young):
        last = int(np.nonzero(sed)[0][-1])
        for j in range(last + 1, len(sed)):
            sed[j] = sed[last] * (lam[last] / lam[j]) ** 2
    vol = g.csize ** 3
    lumcell = (w_old * vol)[:, None] * sed_old[sel][None, :] + (w_young * vol)[:, None] * sed_young[sel][None, :]
    lumcell[g.cchild != -1] = 0.0
    theta = np.array([0.0, 0.224399481500898, 0.448798963001796, 0.673198444502694, 0.897597926003592,
                      1.12199740750449, 1.34639688900539, 1.57079637050629])
    dir_i_out = np.ascontiguousarray(np.stack([theta, np.zeros(8)]))
    ms = g.modelsize
    return Scene(name=name, grid=g, lambda_um=lam[sel], dens_arr=np.ascontiguousarray(g.dens[:, None] * kscale[None, :]),
                 ksca_norm=np.asarray(t["ksca_norm"])[sel], gsca=gsca, kp_sca=set_kp_sca(gsca, kp_sca_max),
                 dir_i_out=dir_i_out, lumcell=np.ascontiguousarray(lumcell), ccoord_p_src=np.zeros((0, 3)),
                 lum_p_src=np.zeros((0, L)), src_cell=np.full(g.ncell, -1, dtype=np.int32),
                 cell_src=np.zeros(0, dtype=np.int32), bm_par=5, bm_par_sca=5, bm_par_max=200, rad_lim=0.1,
                 accuracy=0.3, conv_en_lim=0.01, wall_on=(0, 0, 1),
                 wall_coord=(0.0, 0.0, 0.0, 0.0, 0.3 * ms - ms / 2.0, 0.7 * ms - ms / 2.0))


def delta_lambda_bins(lambda_si):
    """delta_lambda_bin of set_lambda_arr_si (sed_routines.f90:974-978): bin edges at the logarithmic mid-points of the
    wavelength list, first / last bin closed at the first / last wavelength."""
    lam = np.asarray(lambda_si, dtype=np.float64)
    edges = 10.0 ** ((np.log10(lam[1:]) + np.log10(lam[:-1])) / 2.0)
    d = np.empty_like(lam)
    d[1:-1] = edges[1:] - edges[:-1]
    d[0] = edges[0] - lam[0]
    d[-1] = lam[-1] - edges[-1]
    return d


# --------------------------------------------------------------------------------------
# C4: 2DTO3D_GRIDS/input_2dto3d_example.in -- an axisymmetric (R, z) model on the 3-D grid
# --------------------------------------------------------------------------------------
class _Table2D:
    """The 2-D tables of the example (read_grid_2d user_routines_2dto3d.f90:168-323) and the cell average of a tabulated
    quantity, av_dens_2dto3d (:360-475): bilinear interpolation in (R, |z|) at the mid-points of an (np-1)^3 sub-grid of the cell,
    np = cellsize / (modelsize / 1000) clamped to [2, 200], zero beyond the last R / z of the table.  The bilinear weights separate
    in R and z, so the triple sum is evaluated as a_R^T F a_z for all tables at once (the reference adds the same terms point by
    point: agreement to rounding, this is input generation)."""

    def __init__(self, modelsize):
        t = _load_table("twod3d_ex1")
        R = np.asarray(t["R"]); z = np.asarray(t["z"])
        self.nz = int(np.argmax(R != R[0]))            # z varies fastest (:236-244)
        self.nr = len(R) // self.nz
        self.R = R[::self.nz].copy(); self.Z = z[:self.nz].copy()
        self.lambda_um = np.asarray(t["lambda_um"])
        self.F = np.stack([np.asarray(a).reshape(self.nr, self.nz) for a in t["j"]] +
                          [np.asarray(a).reshape(self.nr, self.nz) for a in t["krho"]])     # (2 L, nr, nz)
        self.pos_obs = np.asarray(t["pos_obs"], dtype=np.float64).reshape(-1, 3)
        self.modelsize = float(modelsize)
        self.memo = {}

    @staticmethod
    def _weights(v, grid):
        """Per-node sums of the linear-interpolation weights of the points v (those beyond the last node are skipped);
        the lower node is the largest index >= 0 with v > grid[index], searched as at :409-419."""
        v = v[v <= grid[-1]]
        i = np.clip(np.searchsorted(grid, v, side="left") - 1, 0, len(grid) - 2)
        d = grid[i + 1] - grid[i]
        w = np.zeros(len(grid))
        np.add.at(w, i, (grid[i + 1] - v) / d)
        np.add.at(w, i + 1, (v - grid[i]) / d)
        return w

    def cell_average(self, x, y, z, cs):
        """(j(L), krho(L)) averaged over the cell; the model is symmetric under x <-> -x, y <-> -y, z <-> -z and x <-> y."""
        ax, ay = sorted((abs(x), abs(y)))
        key = (round(ax, 6), round(ay, 6), round(abs(z), 6), round(cs, 6))
        hit = self.memo.get(key)
        if hit is not None:
            return hit
        npt = min(max(int(cs / (self.modelsize / 1000.0)), 2), 200)
        step = cs / (npt - 1)
        off = (np.arange(npt - 1) + 0.5) * step
        xp = (ax - cs / 2.0) + off; yp = (ay - cs / 2.0) + off; zp = (abs(z) - cs / 2.0) + off
        rp = np.sqrt(xp[:, None] ** 2 + yp[None, :] ** 2).ravel()
        a_r = self._weights(rp, self.R); a_z = self._weights(np.abs(zp), self.Z)
        v = np.einsum("r,qrz,z->q", a_r, self.F, a_z) * step ** 3 / cs ** 3
        L = len(self.lambda_um)
        out = (v[:L].copy(), v[L:].copy())
        self.memo[key] = out
        return out


def twod3d_grid(max_lvl=4, min_lvl=2, base=(3, 3), modelsize=28000.0, max_dtau=0.01, max_dlum=0.5e-6, z_subd_lim=50.0,
                r_subd_lim=4000.0):
    """The main grid of 2DTO3D_GRIDS/input_grid_2dto3d_example.in: dust density and emissivity of the reference wavelength
    (lambda_ref = 12 um, the first table) averaged per cell, subdivision criterion 'standard' of create_adap_grid_2dto3d.f90:163-185.
    lnu_tot in that criterion is a module variable of user_routines_2dto3d.f90 that nothing assigns (:19-20): it is 0 there, so
    every emitting cell is subdivided down to max_lvl."""
    tab = _Table2D(modelsize)
    lnu_tot = 0.0

    def calc_dens(x, y, z, cs):
        j, k = tab.cell_average(x, y, z, cs)
        return float(k[0]), float(j[0])

    def crit(x, y, z, cs, nlevel, dens, dens_stars):
        return (dens * cs > max_dtau) or (dens_stars * cs ** 3 > max_dlum * lnu_tot) or (nlevel < min_lvl) or \
            (abs(z) < z_subd_lim and math.sqrt(x * x + y * y) < r_subd_lim)

    tag = f"twod3d_b{base[0]}_{base[1]}_l{max_lvl}_m{min_lvl}"
    g = _grid_cached(tag, lambda: build_grid(modelsize, base, max_lvl, calc_dens, crit))
    return g, tab


def twod3d_scene(max_lvl=4, name="2dto3d") -> Scene:
    """C4: 2DTO3D_GRIDS/input_2dto3d_example.in -- rt_algorithm 'i_obs', only_direct_rt: the emission of the nine shipped tables
    (12 - 850 um, dust opacity zero in all of them) seen by one internal observer at (8000, 0, 0) pc, no external directions
    (use_dir_out false), use_lambda_grid (per-wavelength dens / dens_stars from the lambda grids), kp_maps = 8."""
    g, tab = twod3d_grid(max_lvl=max_lvl)
    L = len(tab.lambda_um)
    dens_arr = np.zeros((g.ncell, L)); lumcell = np.zeros((g.ncell, L))
    for i in np.nonzero(g.cchild == -1)[0]:
        j, k = tab.cell_average(g.ccoord[i, 0], g.ccoord[i, 1], g.ccoord[i, 2], g.csize[i])
        dens_arr[i] = k
        lumcell[i] = j * g.csize[i] ** 3          # calc_total_luminosity rt_routines.f90:605-660
    ms = g.modelsize
    return Scene(name=name, grid=g, lambda_um=tab.lambda_um, dens_arr=np.ascontiguousarray(dens_arr), ksca_norm=np.zeros(L),
                 gsca=np.zeros(L), kp_sca=np.full(L, -1, dtype=np.int32), dir_i_out=np.zeros((2, 0)),
                 lumcell=np.ascontiguousarray(lumcell), ccoord_p_src=np.zeros((0, 3)), lum_p_src=np.zeros((0, L)),
                 src_cell=np.full(g.ncell, -1, dtype=np.int32), cell_src=np.zeros(0, dtype=np.int32), bm_par=5, bm_par_sca=5,
                 bm_par_max=2000, rad_lim=0.1, accuracy=0.10, conv_en_lim=0.01, wall_on=(0, 0, 1),
                 wall_coord=(0.0, 0.0, 0.0, 0.0, 0.3 * ms - ms / 2.0, 0.7 * ms - ms / 2.0), ccoord_obs=tab.pos_obs.copy())


# --------------------------------------------------------------------------------------
# C5: NBODY_SPH_GRIDS/input_nbody_example1.in -- a particle snapshot binned onto the adaptive grid
# --------------------------------------------------------------------------------------
def nbody_particles(n_star=200_000, n_gas=200_000, seed=12345):
    """The reference's snapshot (run708mainDiff.01000_unbar.h5) is not shipped.  Synthetic stand-in: 2e5 star and
    2e5 gas particles from numpy's PCG64 generator, seed 12345 -- an exponential disc (h = 3 kpc, sech-like |z| scale 300 pc), a
    compact bulge, and 40 star-forming clumps that hold a tenth of the gas and all the young stars.  Returns
    (star xyz [pc], old-star luminosity, young-star luminosity, gas xyz, dust mass = metallicity-weighted gas mass)."""
    rng = np.random.default_rng(seed)

    def disc(n, h, zs):
        r = rng.gamma(2.0, h, n)                       # surface density ~ exp(-R/h)
        ph = rng.uniform(0.0, 2.0 * math.pi, n)
        z = rng.laplace(0.0, zs, n)
        return np.stack([r * np.cos(ph), r * np.sin(ph), z], axis=1)

    n_bulge = n_star // 5
    n_young = n_star // 10
    n_disc = n_star - n_bulge - n_young
    centres = disc(40, 3500.0, 60.0)
    def clumps(n):
        k = rng.integers(0, len(centres), n)
        return centres[k] + rng.normal(0.0, 120.0, (n, 3))
    star = np.concatenate([disc(n_disc, 3000.0, 300.0), rng.normal(0.0, 450.0, (n_bulge, 3)) * np.array([1.0, 1.0, 0.6]), clumps(n_young)])
    lum_old = np.concatenate([np.full(n_disc, 1.0), np.full(n_bulge, 1.5), np.zeros(n_young)])
    lum_young = np.concatenate([np.zeros(n_disc + n_bulge), rng.lognormal(0.0, 0.5, n_young)])
    n_cl = n_gas // 10
    gas = np.concatenate([disc(n_gas - n_cl, 4500.0, 120.0), clumps(n_cl)])
    feh = -0.2 * np.sqrt(gas[:, 0] ** 2 + gas[:, 1] ** 2) / 4500.0 + rng.normal(0.0, 0.05, n_gas)   # [Fe/H] gradient
    dust = 10.0 ** feh                                  # p_met * mgas with equal-mass gas particles (av_galaxy :562-564)
    return star, lum_old / lum_old.sum(), lum_young / max(lum_young.sum(), 1e-300), gas, dust / dust.sum()


def nbody_grid(base=(3, 3), max_lvl=4, min_lvl=1, max_dtau=0.01, max_dlum=0.10e-4, modelsize=20000.0, tau_face=1.5, young_frac=0.3):
    """Main grid of NBODY_SPH_GRIDS/input_grid_nbody_example1.in: every cell holds the sums of the particles inside its half-open
    box divided by its volume (av_galaxy user_routines_Nbody_SPH.f90:403-649), subdivision criterion 'standard' of
    create_adap_grid_Nbody_SPH.f90:194-212 (relative to the root cell: luminosity fraction > max_dlum, optical depth >
    max_dtau x the root's, level < min_lvl).  The particles are binned once on the finest level; a coarser cell is the sum of its
    block.  dens is scaled so that the face-on central B-band optical depth of the disc is tau_face."""
    star, l_old, l_young, gas, dust = nbody_particles()
    nf = base[0] * base[1] ** (max_lvl - 1)
    fine = modelsize / nf

    def hist(xyz, w):
        ijk = np.floor((xyz + modelsize / 2.0) / fine).astype(np.int64)
        ok = ((ijk >= 0) & (ijk < nf)).all(axis=1)
        flat = (ijk[ok, 2] * nf + ijk[ok, 1]) * nf + ijk[ok, 0]
        return np.bincount(flat, weights=w[ok], minlength=nf ** 3).reshape(nf, nf, nf)     # [z][y][x]

    H = {"old": hist(star, l_old), "young": hist(star, l_young), "dust": hist(gas, dust)}
    lv = {}
    for l in range(0, max_lvl + 1):
        n = 1 if l == 0 else base[0] * base[1] ** (l - 1)
        b = nf // n
        lv[l] = {k: v.reshape(n, b, n, b, n, b).sum(axis=(1, 3, 5)) for k, v in H.items()}
    # column through the centre, perpendicular to the disc: sum of dust mass / area  -> scale to tau_face
    col = lv[max_lvl]["dust"][:, nf // 2, nf // 2].sum() / fine ** 2
    kscale = tau_face / col

    def lookup(x, y, z, cs):
        l = int(round(math.log(modelsize / cs) / math.log(base[1]))) if cs < modelsize * 0.999 else 0
        n = 1 if l == 0 else base[0] * base[1] ** (l - 1)
        i = [min(max(int((c + modelsize / 2.0) / cs), 0), n - 1) for c in (x, y, z)]
        vol = cs ** 3
        return (lv[l]["dust"][i[2], i[1], i[0]] / vol, lv[l]["old"][i[2], i[1], i[0]] / vol, lv[l]["young"][i[2], i[1], i[0]] / vol)

    def calc_dens(x, y, z, cs):
        d, o, yv = lookup(x, y, z, cs)
        return d * kscale, (1.0 - young_frac) * o + young_frac * yv

    tot_dens0 = lv[0]["dust"][0, 0, 0] / modelsize ** 3 * kscale

    def crit(x, y, z, cs, nlevel, dens, dens_stars):
        # dens_stars(0) * modelsize^3 = 1 (luminosity fractions); tau relative to the root cell
        return (dens_stars * cs ** 3 > max_dlum) or (dens * cs > max_dtau * tot_dens0 * modelsize) or (nlevel < min_lvl)

    tag = f"nbody_b{base[0]}_{base[1]}_l{max_lvl}_m{min_lvl}_t{max_dtau}_d{max_dlum}_f{tau_face}_y{young_frac}"
    g = _grid_cached(tag, lambda: build_grid(modelsize, base, max_lvl, calc_dens, crit))
    w_old = np.zeros(g.ncell); w_young = np.zeros(g.ncell)
    for i in np.nonzero(g.cchild == -1)[0]:
        _, o, yv = lookup(g.ccoord[i, 0], g.ccoord[i, 1], g.ccoord[i, 2], g.csize[i])
        w_old[i], w_young[i] = (1.0 - young_frac) * o, young_frac * yv
    return g, w_old, w_young


def nbody_scene(max_lvl=4, max_dlum=0.10e-4, lnum=None, kp_sca_max=1, name="nbody") -> Scene:
    """C5: NBODY_SPH_GRIDS/input_nbody_example1.in -- rt_algorithm 'main', every luminous leaf emits, 28 stellar wavelengths of
    lambda_list_trustI_basic.dat (max_lambda_stars = 30) with the TRUST effective-grain opacities, 8 directions of
    dir_out_galaxy.dat, bm_par 5, accuracy 0.2, no wall.  Old and young stars radiate as 4 000 K and 15 000 K black bodies
    (the starburst99 library of the example is not shipped)."""
    g, w_old, w_young = nbody_grid(max_lvl=max_lvl, max_dlum=max_dlum)
    t = _load_table("magtar")
    sel = np.arange(len(t["lambda_um"]))
    if lnum is not None:
        sel = np.unique(np.linspace(0, len(sel) - 1, lnum).round().astype(int))
    L = len(sel)
    lam = np.asarray(t["lambda_um"])[sel]
    nu_shape = lambda T: (1.0 / lam ** 3) / np.expm1(14387.77 / (lam * T))      # B_nu shape, lambda in um
    sed_old, sed_young = nu_shape(4000.0), nu_shape(15000.0)
    sed_old = sed_old / sed_old.max(); sed_young = sed_young / sed_young.max()
    vol = g.csize ** 3
    lumcell = (w_old * vol)[:, None] * sed_old[None, :] + (w_young * vol)[:, None] * sed_young[None, :]
    lumcell[g.cchild != -1] = 0.0
    lumcell *= 1e30
    kscale = np.asarray(t["kext_scale"])[sel]
    kb = float(np.interp(0.44, np.asarray(t["lambda_um"]), np.asarray(t["kext_scale"])))
    gsca = np.asarray(t["gsca"])[sel]
    theta = np.array([0.0, 0.224399481500898, 0.448798963001796, 0.673198444502694, 0.897597926003592,
                      1.12199740750449, 1.34639688900539, 1.57079637050629])
    return Scene(name=name, grid=g, lambda_um=lam, dens_arr=np.ascontiguousarray(g.dens[:, None] * (kscale / kb)[None, :]),
                 ksca_norm=np.asarray(t["ksca_norm"])[sel], gsca=gsca, kp_sca=set_kp_sca(gsca, kp_sca_max),
                 dir_i_out=np.ascontiguousarray(np.stack([theta, np.zeros(8)])), lumcell=np.ascontiguousarray(lumcell),
                 ccoord_p_src=np.zeros((0, 3)), lum_p_src=np.zeros((0, L)), src_cell=np.full(g.ncell, -1, dtype=np.int32),
                 cell_src=np.zeros(0, dtype=np.int32), bm_par=5, bm_par_sca=5, bm_par_max=200, rad_lim=0.10, accuracy=0.20,
                 conv_en_lim=0.01)
