"""Grid files of the reference, as a raw-array container (N3).

The reference stores its grids in HDF5 (io_routines.f90:185-199 datasets, read_main_grid :297-393, read_lambda_grid :397-498);
this image has no HDF5 (no library, no h5py), so the same datasets -- same names, same shapes, same element types, Fortran
(column-major) memory order -- are written into one little-endian binary file behind a JSON header:

    b"DRGRID1\\n" | uint32 header length | JSON {"datasets": {name: {"dtype", "shape" (Fortran order), "offset", "nbytes"}}, "attrs": {...}} | padding to 64 B | data ...

Main grid  (`grid_<label>_main`):   ncell i32[N], ccoord f64 (3,N), cchild i32[N], cindex i64[N], lvl i32[N], csize f64[N],
                                    dens f64[N] (extinction at the reference wavelength), dens_stars f64[N], base i32[2]
Lambda grid (`grid_<label>_l<F9.3 lambda>um`, file name rule io_routines.f90:2327): dens f64[N], dens_stars f64[N]

A maintainer converts between the two with any HDF5 tool: every dataset is a contiguous array with the HDF5 dataset's name.
"""
from __future__ import annotations

import json
import os

import numpy as np

from .scenes import Grid, _powi

MAGIC = b"DRGRID1\n"
MAIN_DATASETS = ("ncell", "ccoord", "cchild", "cindex", "lvl", "csize", "dens", "dens_stars", "base")


def _write(path, arrays, attrs):
    """arrays: name -> (ndarray in C order whose memory image is the Fortran array, Fortran shape)."""
    meta, off = {}, 0
    blobs = []
    for name, (a, fshape) in arrays.items():
        a = np.ascontiguousarray(a)
        meta[name] = {"dtype": a.dtype.newbyteorder("<").str, "shape": list(fshape), "offset": off, "nbytes": int(a.nbytes)}
        blobs.append(a.astype(a.dtype.newbyteorder("<"), copy=False).tobytes())
        off += (a.nbytes + 63) // 64 * 64
    hdr = json.dumps({"datasets": meta, "attrs": attrs}).encode()
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(np.uint32(len(hdr)).tobytes())
        f.write(hdr)
        pos = len(MAGIC) + 4 + len(hdr)
        f.write(b"\0" * ((-pos) % 64))
        for b in blobs:
            f.write(b)
            f.write(b"\0" * ((-len(b)) % 64))


def _read(path):
    with open(path, "rb") as f:
        if f.read(len(MAGIC)) != MAGIC:
            raise ValueError(f"{path}: not a DRGRID1 file")
        n = int(np.frombuffer(f.read(4), dtype="<u4")[0])
        hdr = json.loads(f.read(n).decode())
        pos = len(MAGIC) + 4 + n
        base = pos + ((-pos) % 64)
        out = {}
        for name, m in hdr["datasets"].items():
            f.seek(base + m["offset"])
            out[name] = np.frombuffer(f.read(m["nbytes"]), dtype=np.dtype(m["dtype"])).copy()
            out[name + ".shape"] = tuple(m["shape"])
        return out, hdr["attrs"]


def lambda_grid_name(label, lambda_um):
    """grid_<label>_l<lambda, F9.3, left-adjusted>um (io_routines.f90:2327)."""
    return f"grid_{label}_l{lambda_um:.3f}um.drg"


def write_main_grid(path, g: Grid):
    """The nine datasets of the main grid file (io_routines.f90:189-197)."""
    N = g.ncell
    _write(path, {
        "ncell": (np.arange(N, dtype=np.int32), (N,)),
        "ccoord": (np.ascontiguousarray(g.ccoord, dtype=np.float64), (3, N)),       # C (N,3) == Fortran (3,N)
        "cchild": (g.cchild.astype(np.int32), (N,)),
        "cindex": (g.cindex.astype(np.int64), (N,)),
        "lvl": (g.lvl.astype(np.int32), (N,)),
        "csize": (g.csize.astype(np.float64), (N,)),
        "dens": (g.dens.astype(np.float64), (N,)),
        "dens_stars": (g.dens_stars.astype(np.float64), (N,)),
        "base": (np.asarray(g.base, dtype=np.int32), (2,)),
    }, {"modelsize": g.modelsize, "max_lvl": g.max_lvl})


def read_main_grid(path) -> Grid:
    """read_main_grid (io_routines.f90:297-393): the arrays as stored; max_lvl = maxval(lvl) (:389), modelsize = csize(0), and the
    per-level size tables of make_csize_arr (:669-696)."""
    d, attrs = _read(path)
    for k in MAIN_DATASETS:
        if k not in d:
            raise ValueError(f"{path}: dataset {k} missing")
    N = len(d["cchild"])
    base = tuple(int(x) for x in d["base"])
    max_lvl = int(d["lvl"].max())
    modelsize = float(d["csize"][0])
    g = Grid(modelsize=modelsize, base=base, max_lvl=max_lvl, cchild=d["cchild"].astype(np.int32), lvl=d["lvl"].astype(np.int32),
             cindex=d["cindex"].astype(np.int64), ccoord=np.ascontiguousarray(d["ccoord"].reshape(N, 3)),
             csize=d["csize"].astype(np.float64), dens=d["dens"].astype(np.float64), dens_stars=d["dens_stars"].astype(np.float64))
    baseinv = (1.0 / base[0], 1.0 / base[1])
    cs = [modelsize] + [modelsize * (baseinv[0] * _powi(baseinv[1], i - 1)) for i in range(1, max_lvl + 1)]
    g.csize_arr = np.asarray(cs, dtype=np.float64)
    g.carea_arr = g.csize_arr * g.csize_arr
    g.cvol_arr = g.csize_arr * g.csize_arr * g.csize_arr
    return g


def write_lambda_grid(path, dens, dens_stars):
    """One lambda grid: dens and dens_stars of every cell at one wavelength (io_routines.f90:397-498)."""
    N = len(dens)
    _write(path, {"dens": (np.asarray(dens, dtype=np.float64), (N,)), "dens_stars": (np.asarray(dens_stars, dtype=np.float64), (N,))}, {})


def read_lambda_grid(path, ncell):
    d, _ = _read(path)
    if len(d["dens"]) != ncell or len(d["dens_stars"]) != ncell:
        raise ValueError(f"{path}: dimensions of the lambda grid do not coincide with the main grid")   # io_routines.f90:444
    return d["dens"], d["dens_stars"]


def write_scene_grids(directory, label, scene):
    """Main grid + one lambda grid per wavelength of a scene (use_lambda_grid): dens = dens_arr(l,:), dens_stars = lumcell(l,:)/csize^3."""
    os.makedirs(directory, exist_ok=True)
    g = scene.grid
    write_main_grid(os.path.join(directory, f"grid_{label}_main.drg"), g)
    vol = g.csize ** 3
    for l, lam in enumerate(scene.lambda_um):
        write_lambda_grid(os.path.join(directory, lambda_grid_name(label, float(lam))), scene.dens_arr[:, l], scene.lumcell[:, l] / vol)


def read_scene_grids(directory, label, lambda_um):
    """(grid, dens_arr (N,L), lumcell (N,L)) from the files: lumcell = dens_stars * csize^3 for leaves (calc_total_luminosity
    rt_routines.f90:605-660)."""
    g = read_main_grid(os.path.join(directory, f"grid_{label}_main.drg"))
    N, L = g.ncell, len(lambda_um)
    dens_arr = np.zeros((N, L)); lumcell = np.zeros((N, L))
    vol = g.csize ** 3
    for l, lam in enumerate(lambda_um):
        d, s = read_lambda_grid(os.path.join(directory, lambda_grid_name(label, float(lam))), N)
        dens_arr[:, l] = d
        lumcell[:, l] = np.where(g.cchild == -1, s * vol, 0.0)
    return g, dens_arr, lumcell
