#!/usr/bin/env python
"""Freeze the oracle's outputs on the tiny TRUST-I scene (regression fixture; the reference itself cannot
be run in this image).  Usage: python tests/golden/make_golden.py"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from dartray_b200 import abi, scenes  # noqa: E402
from oracle import oracle as O  # noqa: E402

KEYS = ("sources", "rays", "intersections", "deposits", "beam_tests", "splits", "merges", "sca_elems", "active_lambda",
        "steps", "isect_hash")


def main():
    sc = scenes.trusti_scene(tau_z=1.0, base=(10, 2), max_lvl=2, lnum=3, name="trustI_tiny")
    orc = O.Oracle(sc, nthreads=1)
    st = O.State(sc)
    out = {}
    for name, rt in (("precalc_src", abi.RTT_PRECALC_SRC), ("dir_src", abi.RTT_DIR_SRC)):
        c = orc.run_pass(rt, 0, st).as_dict()
        out[name] = {k: c[k] for k in KEYS}
    out["u_final_colsum"] = st.u_final.sum(axis=0).tolist()
    out["scaspe_sum"] = [float(s.sum()) for s in st.scaspe]
    orc.prepare_scatt(1, st)
    orc.set_sources(st.lumcell)
    c = orc.run_pass(abi.RTT_SCATT, 1, st).as_dict()
    out["scatt1"] = {k: c[k] for k in KEYS}
    out["lum_lost"] = st.lum_lost.tolist()
    json.dump(out, open(os.path.join(HERE, "trusti_tiny.json"), "w"), indent=1)
    print(json.dumps(out["scatt1"]))


if __name__ == "__main__":
    main()
