"""Config 3 by metric family (diagnostics): S series of ONE family (error counts / latency / cpu /
memory = row id mod 4 in include/fm_synth.h), 60 points against a 1440-point baseline, device-resident,
Mann-Whitney + Wilcoxon.  Shows what heavy ties (the count family) cost the rank kernel."""
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_1_b200 import _abi as A  # noqa: E402
from foremast_1_b200.engine import Context, synth_desc  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
MODE = sys.argv[2] if len(sys.argv) > 2 else "ANY"
with Context(0) as ctx:
    g, t = A.default_params()
    g.pairwise_mode = getattr(A, "FM_PW_" + MODE)
    ctx.set_params(g, t)
    out = {}
    for fam, name in enumerate(["count", "latency", "cpu", "memory", "mixed"]):
        ids = (np.arange(S, dtype=np.uint64) * (4 if fam < 4 else 1) + (fam if fam < 4 else 0))
        d_ids = ctx.dev_alloc(ids.nbytes); ctx.h2d(d_ids, ids)
        d = synth_desc(3, S, 60, 1440, 0, row_ids=d_ids)
        ptrs = [ctx.dev_alloc(S * d.cur_pitch * 4), ctx.dev_alloc(S * d.base_pitch * 4), None, ctx.dev_alloc(S)]
        ctx.synth_device(d, *ptrs)
        b = A.Batch()
        b.struct_size = C.sizeof(A.Batch)
        b.strategy, b.n_series = A.FM_STRAT_CANARY, S
        b.cur, b.n_cur, b.cur_pitch = ptrs[0], 60, d.cur_pitch
        b.base, b.n_base, b.base_pitch = ptrs[1], 1440, d.base_pitch
        b.metric_type = ptrs[3]
        b.record_kind = A.FM_REC_FULL
        d_out = ctx.dev_alloc(S * 96)
        tm = A.Timing()
        ms = []
        for i in range(7):
            ctx.score_device(b, d_out, None, None, None, timing=tm)
            if i >= 2:
                ms.append(tm.main_kernel_ms)
        out[name] = round(float(np.median(ms)), 4)
        for q in [d_ids, d_out] + [p for p in ptrs if p]:
            ctx.dev_free(q)
    print(json.dumps({"series": S, "pairwise": MODE, "ms": out}))
