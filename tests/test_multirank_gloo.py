"""world_size-2 `gloo` test of the N>1 host logic (no GPU): every rank runs its cyclic shard of the
(source, sector) units on the CPU oracle, `hostmpi.rt_mpi_host` combines the partial arrays over
torch.distributed exactly as rt_mpi (io_routines.f90:3667-3881) does, and the result must equal the
one-rank solve: counters/checksums add up bit for bit, fields agree to summation order."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _solve(sc, rank, nranks, allreduce_ready):
    """precalc -> direct -> one scattering iteration, sharded over `nranks` (1 = reference run)."""
    from dartray_b200 import abi, hostmpi
    from oracle import oracle as O
    orc = O.Oracle(sc, nthreads=2)
    st = O.State(sc)
    tot = {"intersections": 0, "isect_hash": 0, "rays": 0}
    has_src = sc.nsrc > 0

    def run(rt, it=0):
        st.psel_sum[...] = 0; st.ipsel[...] = 0     # psel_av_arr(i0, iterations, :) is a fresh slice per pass
        c = orc.run_pass(rt, it, st, rank=rank, nranks=nranks)
        if allreduce_ready:
            hostmpi.rt_mpi_host(st, rt, nranks)
        tot["intersections"] += c.intersections
        tot["rays"] += c.rays
        tot["isect_hash"] = (tot["isect_hash"] + c.isect_hash) % (1 << 64)

    run(abi.RTT_PRECALC_SRC if has_src else abi.RTT_PRECALC_CELL)
    run(abi.RTT_DIR_SRC if has_src else abi.RTT_DIR_CELL)
    conv = orc.prepare_scatt(1, st)
    assert not conv
    orc.set_sources(st.lumcell)
    run(abi.RTT_SCATT, 1)
    return st, tot


def _worker(rank, world, port, scene_kind, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from dartray_b200 import scenes
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        sc = _scene(scenes, scene_kind)
        st, tot = _solve(sc, rank, world, True)
        cnt = torch.tensor([tot["intersections"], tot["rays"]], dtype=torch.int64)
        dist.all_reduce(cnt)
        hashes = [None] * world
        dist.all_gather_object(hashes, tot["isect_hash"])
        # every rank must hold identical combined arrays after rt_mpi
        chk = torch.tensor([float(st.u_final.sum()), float(st.u_fest.sum()), float(st.lum_lost.sum())], dtype=torch.float64)
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(torch.equal(lo, hi))
        if rank == 0:
            q.put({"u_final": st.u_final.copy(), "u_fest": st.u_fest.copy(), "lum_lost": st.lum_lost.copy(),
                   "scaspe": [p.copy() for p in st.scaspe], "psel_sum": st.psel_sum.copy(), "ipsel": st.ipsel.copy(),
                   "intersections": int(cnt[0]), "rays": int(cnt[1]), "hash": sum(hashes) % (1 << 64), "same": same})
    finally:
        dist.destroy_process_group()


def _scene(scenes, kind):
    if kind == "trusti":
        return scenes.trusti_scene(tau_z=1.0, base=(10, 2), max_lvl=2, lnum=3, name="trustI_tiny")
    return scenes.galaxy_scene(max_lvl=2, min_lvl=2, lnum=4, name="galaxy_tiny")


@pytest.mark.parametrize("kind", ["trusti", "galaxy"])
def test_two_ranks_equal_one_rank(kind):
    import torch.multiprocessing as mp
    sys.path.insert(0, ROOT)
    from dartray_b200 import scenes
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, kind, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    one, tot1 = _solve(_scene(scenes, kind), 0, 1, False)
    assert got["same"], "ranks disagree after rt_mpi"
    assert got["intersections"] == tot1["intersections"] and got["rays"] == tot1["rays"]
    assert got["hash"] == tot1["isect_hash"]
    assert np.array_equal(got["ipsel"], one.ipsel)
    for name in ("u_fest", "u_final", "lum_lost", "psel_sum"):
        a, b = got[name], getattr(one, name)
        assert np.allclose(a, b, rtol=1e-11, atol=1e-300), name
    for a, b in zip(got["scaspe"], one.scaspe):
        assert np.allclose(a, b, rtol=1e-11, atol=1e-300)


def test_unit_sharding_covers_every_unit_once():
    sys.path.insert(0, ROOT)
    from dartray_b200 import hostmpi
    for nsrc in (1, 7, 231):
        for n in (1, 2, 3, 8):
            parts = [hostmpi.my_units(nsrc, r, n) for r in range(n)]
            assert [len(p) for p in parts] == [hostmpi.unit_count(nsrc, r, n) for r in range(n)]
            allu = np.sort(np.concatenate(parts))
            assert np.array_equal(allu, np.arange(nsrc * 48))


def test_engine_unit_order_is_a_permutation_and_balances_sectors():
    """The engine's own unit order (scrambled sources, sectors rotated by the source position; k_seed) restated on the host:
    every (source, sector) pair exactly once over the ranks, and with 8 ranks every rank sees every sector equally often."""
    sys.path.insert(0, ROOT)
    from dartray_b200 import hostmpi
    for nsrc in (1, 2, 7, 48, 231):
        for n in (1, 2, 3, 8):
            for rot in (True, False):
                parts = [hostmpi.engine_units(nsrc, r, n, rot) for r in range(n)]
                assert np.array_equal(np.sort(np.concatenate(parts)), np.arange(nsrc * 48))
    per_rank = [np.bincount(hostmpi.engine_units(48 * 5, r, 8) % 48, minlength=48) for r in range(8)]
    assert all((c == per_rank[0][0]).all() for c in per_rank)                 # rotated: 30 units of every sector on every rank
    unrot = np.bincount(hostmpi.engine_units(48 * 5, 0, 8, False) % 48, minlength=48)
    assert (unrot == 0).sum() == 42                                           # not rotated: rank 0 only ever sees 6 sectors
