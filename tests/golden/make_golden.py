"""Generates the committed fixtures under tests/golden/.

  demo_data.json      the only series-shaped data in the reference tree
                      (examples/spring-boot-demo/src/main/resources/data{1,2}.txt),
                      read from /root/reference when this script is run in the
                      build container; the GPU box never reads the reference.
  oracle_records.npz  outputs of the CPU oracle on seeded synthetic batches
                      (include/fm_synth.h), frozen so that a change of the oracle
                      or of the generator is caught.

Run:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import oracle_lib as O  # noqa: E402
from foremast_1_b200 import _abi as A  # noqa: E402
from foremast_1_b200.engine import _as_matrix, synth_desc, synth_host  # noqa: E402
import ctypes as C  # noqa: E402

# name -> (config, S, n_cur, n_base, n_hist, pairwise mode, strategy)
CASES = {
    "c1_job": (1, 4, 60, 1440, 0, A.FM_PW_MANN_WHITE, A.FM_STRAT_CANARY),
    "c3_any": (3, 256, 60, 1440, 0, A.FM_PW_ANY, A.FM_STRAT_CANARY),
    "c2_hist": (2, 32, 60, 1440, 10080, A.FM_PW_MANN_WHITE, A.FM_STRAT_CANARY),
    "c5_ma": (5, 512, 60, 0, 240, A.FM_PW_NONE, A.FM_STRAT_HISTORICAL),
}


def case_inputs(name):
    cfg, S, nc, nb, nh, mode, strat = CASES[name]
    d = synth_desc(cfg, S, nc, nb, nh)
    cur, base, hist, mt = synth_host(d, n_threads=2)
    g, t = A.default_params()
    g.pairwise_mode = mode
    return g, t, strat, cur, base, hist, mt


def host_batch(strat, cur, base, hist, mt):
    b = A.Batch()
    b.struct_size = C.sizeof(A.Batch)
    b.strategy = strat
    b.n_series = cur.shape[0]
    b.cur, b.n_cur, b.cur_pitch = A.ptr(cur), cur.shape[1], cur.shape[1]
    if base is not None:
        b.base, b.n_base, b.base_pitch = A.ptr(base), base.shape[1], base.shape[1]
    if hist is not None:
        b.hist, b.n_hist, b.hist_pitch = A.ptr(hist), hist.shape[1], hist.shape[1]
    b.metric_type = A.ptr(mt)
    return b


def run_case(name):
    g, t, strat, cur, base, hist, mt = case_inputs(name)
    b = host_batch(strat, cur, base, hist, mt)
    out = O.score_batch(g, t, b, n_threads=2)
    return {"rec": out["rec"], "aux": out["aux"]}


def demo_data():
    res = "/root/reference/examples/spring-boot-demo/src/main/resources"
    out = {}
    for k in ("data1", "data2"):
        with open(os.path.join(res, k + ".txt")) as f:
            # lines are "YYYY-MM-DD hh:mm:ss,value"
            out[k] = [float(line.rsplit(",", 1)[1]) for line in f if "," in line]
    return out


if __name__ == "__main__":
    if os.path.isdir("/root/reference"):
        with open(os.path.join(HERE, "demo_data.json"), "w") as f:
            json.dump(demo_data(), f)
    arrs = {}
    for name in CASES:
        for k, v in run_case(name).items():
            arrs[f"{name}__{k}"] = v
    np.savez_compressed(os.path.join(HERE, "oracle_records.npz"), **arrs)
    print("wrote", sorted(arrs))
