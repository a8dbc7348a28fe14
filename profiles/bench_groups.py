"""Config 4 timing (diagnostics, not the bench.py contract): V services x 3..8 pod groups x 60 integer
counts, Kruskal-Wallis + per-group bounds against a 1440-point pooled history, device-resident rows."""
import ctypes as C
import json
import sys
import os

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_1_b200 import _abi as A  # noqa: E402
from foremast_1_b200.engine import Context, synth_desc  # noqa: E402

V = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
NH = int(sys.argv[2]) if len(sys.argv) > 2 else 1440          # history points (0: no model)
MODE = sys.argv[3] if len(sys.argv) > 3 else "KRUSKAL"     # KRUSKAL | NONE
PAD = int(sys.argv[4]) if len(sys.argv) > 4 else 0          # bytes allocated first (shifts every buffer: placement effects)
rng = np.random.default_rng(44)
K = rng.integers(3, 9, V)
off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
G = int(off[-1])
svc_of = np.repeat(np.arange(V, dtype=np.uint64), K)
ids = (svc_of << np.uint64(4)) + (np.arange(G, dtype=np.uint64) - off[:-1].astype(np.uint64)[svc_of.astype(np.int64)])
hids = np.arange(V, dtype=np.uint64) << np.uint64(4)
with Context(0) as ctx:
    g, t = A.default_params()
    g.pairwise_mode = A.FM_PW_KRUSKAL if MODE == "KRUSKAL" else A.FM_PW_NONE
    ctx.set_params(g, t)
    d_pad = ctx.dev_alloc(PAD) if PAD else None
    d_ids, d_hids = ctx.dev_alloc(ids.nbytes), ctx.dev_alloc(hids.nbytes)
    ctx.h2d(d_ids, ids); ctx.h2d(d_hids, hids)
    dg = synth_desc(4, G, 60, 0, 0, row_ids=d_ids)
    dh = synth_desc(4, V, 60, 0, max(NH, 4), row_ids=d_hids)
    d_cur, d_hist, d_mt = ctx.dev_alloc(G * dg.cur_pitch * 4), ctx.dev_alloc(V * dh.hist_pitch * 4), ctx.dev_alloc(V)
    ctx.synth_device(dg, d_cur, None, None, None)
    ctx.synth_device(dh, None, None, d_hist, d_mt)
    d_off = ctx.dev_alloc(off.nbytes); ctx.h2d(d_off, off)
    gb = A.GroupBatch()
    gb.struct_size = C.sizeof(A.GroupBatch)
    gb.n_services, gb.n_groups, gb.group_offsets = V, G, d_off
    gb.cur, gb.n_cur, gb.cur_pitch = d_cur, 60, dg.cur_pitch
    gb.hist, gb.n_hist, gb.hist_pitch = (d_hist if NH else None), (NH if NH else 0), dh.hist_pitch
    gb.metric_type = d_mt
    d_out, d_svc = ctx.dev_alloc(G * 96), ctx.dev_alloc(V * 16)
    tm = A.Timing()
    ms = []
    for i in range(8):
        ctx.score_groups_device(gb, d_out, d_svc, tm)
        if i >= 3:
            ms.append(tm.main_kernel_ms)
    bytes_alg = 4 * (G * 60 + V * NH) + G * 96 + V * 16
    m = float(np.median(ms))
    print(json.dumps({"workload": "config 4", "n_hist": NH, "mode": MODE, "services": V, "groups": G, "kernel_ms": m,
                      "services_per_s": V / (m / 1e3), "algorithmic_GBps": bytes_alg / (m / 1e3) / 1e9,
                      "frac_of_measured_hbm": bytes_alg / (m / 1e3) / 1e9 / 6557.8}))
