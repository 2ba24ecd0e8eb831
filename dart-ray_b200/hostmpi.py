"""Host-side mirror of the reference's multi-rank bookkeeping for the ray-tracing path.

* `my_units` — the cyclic shard of the (source, sector) work units of a pass
  (rt_routines.f90:1003-1005 assigns sources cyclically by rank; the engine shards the 48 sectors of
  every source as well because configs C1/C2 have a single source).
* `rt_mpi_host` — `rt_mpi` (io_routines.f90:3667-3705) on HOST arrays (the Fortran driver's
  `u_fest_arr`, `u_final_arr`, `scaspe_arr(:)%a`, `lum_lost`, `psel_av_arr`), with any
  `torch.distributed` backend as the transport.  It restates reduce_u_fest_arr (:3708-3730),
  reduce_u_final_arr incl. the delta trick (:3733-3762), reduce_scaspe_arr (:3765-3838) and
  reduce_lum_lost with its correction (:3841-3855).  The device-resident equivalent is
  `dre_allreduce` (NCCL, csrc/dre_engine.cu); this module is what a host that keeps the arrays on
  its side calls between `dre_run_pass` calls, and it is what the world_size-2 gloo tests exercise.
"""
from __future__ import annotations

import numpy as np

from . import abi


def my_units(n_sources: int, rank: int, nranks: int) -> np.ndarray:
    """Global unit ids (source_ordinal*48 + sector) processed by `rank`: gid % nranks == rank."""
    total = int(n_sources) * 48
    return np.arange(rank, total, nranks, dtype=np.int64)


def source_multiplier(n_sources: int) -> int:
    """Multiplier of the scrambled source order of the engine (dre_run_pass_resident): near nsrc / golden ratio, coprime to nsrc."""
    import math
    if n_sources <= 2:
        return 1
    m = max(1, int(n_sources * 0.6180339887))
    while math.gcd(m, n_sources) != 1:
        m += 1
    return m


def engine_units(n_sources: int, rank: int, nranks: int, sector_rotation: bool = True) -> np.ndarray:
    """The units the ENGINE gives to `rank`, as global unit ids source_ordinal*48 + sector, in processing order (k_seed,
    csrc/dre_kernels.cuh): position gq = rank + i*nranks; source ordinal = (gq // 48 * multiplier) mod nsrc; sector =
    (gq % 48 + gq // 48) mod 48 -- rotated by the source position, so that a rank does not trace the same sectors of every
    source when 48 % nranks == 0."""
    gq = my_units(n_sources, rank, nranks)
    pos = gq // 48
    sord = (pos * source_multiplier(n_sources)) % max(int(n_sources), 1)
    sector = (gq % 48 + pos) % 48 if sector_rotation else gq % 48
    return sord * 48 + sector


def unit_count(n_sources: int, rank: int, nranks: int) -> int:
    total = int(n_sources) * 48
    return (total - rank + nranks - 1) // nranks if total > rank else 0


def _allreduce_sum(arr: np.ndarray, group=None):
    """In-place sum over ranks of a contiguous float64/int64 numpy array."""
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(arr)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)


def rt_mpi_host(st, rt_type: int, nranks: int, allreduce=_allreduce_sum):
    """Combine the per-rank partial results of one pass, in place, exactly as `rt_mpi` does.

    `st` carries u_fest, u_final, scaspe (list of (N, P(l)) planes), lum_lost, lum_lost_prev,
    psel_sum / ipsel.  For rtt_scatt every rank starts the pass from the same u_final (= u_fest) and
    lum_lost (= lum_lost_prev), so only the increments are summed:
      u_final = allreduce(max(u_final - u_fest, 0)) + u_fest                    (:3743-3757)
      lum_lost = allreduce(lum_lost - (nranks-1)/nranks * lum_lost_prev)        (:3846-3848)
    """
    if nranks <= 1:
        return
    if rt_type in (abi.RTT_PRECALC_CELL, abi.RTT_PRECALC_SRC):
        allreduce(st.u_fest)
        return
    if rt_type in (abi.RTT_DIR_CELL, abi.RTT_DIR_SRC):
        allreduce(st.u_final)
        for p in st.scaspe:
            allreduce(p)
        allreduce(st.lum_lost)
    elif rt_type == abi.RTT_SCATT:
        d = st.u_final - st.u_fest
        np.maximum(d, 0.0, out=d)
        allreduce(d)
        st.u_final[...] = d + st.u_fest
        for p in st.scaspe:          # zeroed by rt_prepare before the pass: plain sum
            allreduce(p)
        st.lum_lost[...] = st.lum_lost - (nranks - 1) / float(nranks) * st.lum_lost_prev
        allreduce(st.lum_lost)
    else:
        return                       # i_obs passes are not split over ranks (rt_routines.f90:1498)
    if getattr(st, "psel_sum", None) is not None:
        allreduce(st.psel_sum)
        allreduce(st.ipsel)
