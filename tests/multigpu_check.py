"""Run under torchrun on N GPUs: the sharded solve (units cyclic over ranks + dre_allreduce = rt_mpi over
NCCL) must reproduce the one-GPU solve: event counters and the intersection checksum add up bit for
bit, every rank ends with the same fields, and they match the one-GPU fields to summation order.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/multigpu_check.py [scene] [det]

With `det` both solves run with option "deterministic" (order-independent accumulation) and the N-GPU fields must
equal the one-GPU fields bit for bit.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def solve(eng, sc, abi, iterations=2):
    for a in (abi.ARR_U_FEST, abi.ARR_U_FINAL, abi.ARR_SCASPE, abi.ARR_SCASPE_PREV, abi.ARR_SCASPE_TOT, abi.ARR_LUM_LOST):
        eng.zero(a)
    eng.set_sources(sc.lumcell)
    has_src, has_cell = sc.nsrc > 0, bool(np.any(sc.lumcell))
    tot = {"intersections": 0, "rays": 0, "deposits": 0, "isect_hash": 0}

    def acc(c):
        for k in ("intersections", "rays", "deposits"):
            tot[k] += getattr(c, k)
        tot["isect_hash"] = (tot["isect_hash"] + c.isect_hash) % (1 << 64)

    if has_cell:
        acc(eng.rt_loop(abi.RTT_PRECALC_CELL))
    if has_src:
        acc(eng.rt_loop(abi.RTT_PRECALC_SRC))
    eng.rt_mpi(abi.RTT_PRECALC_SRC)
    if has_cell:
        acc(eng.rt_loop(abi.RTT_DIR_CELL))
    if has_src:
        acc(eng.rt_loop(abi.RTT_DIR_SRC))
    eng.rt_mpi(abi.RTT_DIR_SRC)
    tr = np.zeros(sc.lnum)
    for it in range(1, iterations + 1):
        if eng.rt_prepare_scatt(it, tr):
            break
        acc(eng.rt_loop(abi.RTT_SCATT, it))
        eng.rt_mpi(abi.RTT_SCATT)
    return tot


def main():
    import torch
    import torch.distributed as dist
    from dartray_b200 import abi, engine, scenes
    from oracle import oracle as O   # State container only (host arrays); nothing is computed on the CPU
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    kind = sys.argv[1] if len(sys.argv) > 1 else "trusti"
    if kind == "galaxy":
        sc = scenes.galaxy_scene(max_lvl=2, min_lvl=2, lnum=4, name="galaxy_tiny")
    else:
        sc = scenes.trusti_scene(lnum=4, max_lvl=3)
    obj = [engine.Engine.get_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(obj, src=0)
    eng = engine.Engine(sc, device=lr, rank=rank, nranks=world, nccl_id=obj[0])
    eng.set_option("hash", 1)
    det = len(sys.argv) > 2 and sys.argv[2] == "det"
    if det:
        eng.set_option("deterministic", 1)
    tot = solve(eng, sc, abi)
    st = O.State(sc)
    eng.download_state(st)
    t = torch.tensor([tot["intersections"], tot["rays"], tot["deposits"]], dtype=torch.int64, device="cuda")
    dist.all_reduce(t)
    hashes = [None] * world
    dist.all_gather_object(hashes, tot["isect_hash"])
    # all ranks hold the same combined arrays
    chk = torch.tensor([st.u_final.sum(), st.u_fest.sum(), st.lum_lost.sum(), sum(float(p.sum()) for p in st.scaspe_tot)],
                       dtype=torch.float64, device="cuda")
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    ok = bool(torch.equal(lo, hi))
    eng.close()
    if rank == 0:
        one = engine.Engine(sc, device=lr)
        one.set_option("hash", 1)
        if det:
            one.set_option("deterministic", 1)
            one.set_option("warps", 7)          # a different schedule as well
            one.set_option("batch_units", 1000)
        tot1 = solve(one, sc, abi)
        s1 = O.State(sc)
        one.download_state(s1)
        one.close()
        assert ok, "ranks disagree after dre_allreduce"
        assert int(t[0]) == tot1["intersections"] and int(t[1]) == tot1["rays"] and int(t[2]) == tot1["deposits"], (t, tot1)
        assert sum(hashes) % (1 << 64) == tot1["isect_hash"]
        for name in ("u_fest", "u_final", "lum_lost", "lumcell"):
            a, b = getattr(st, name), getattr(s1, name)
            assert np.array_equal(a, b) if det else np.allclose(a, b, rtol=1e-10, atol=1e-300), name
        for a, b in zip(st.scaspe_tot + st.scaspe, s1.scaspe_tot + s1.scaspe):
            assert np.array_equal(a, b) if det else np.allclose(a, b, rtol=1e-10, atol=1e-300)
        assert st.u_final.max() > 0 and st.scaspe_tot[0].max() > 0
        print(f"multigpu_check ok: world={world} scene={kind} det={int(det)} intersections={int(t[0])}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
