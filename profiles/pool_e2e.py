"""End-to-end throughput of ONE process driving every visible GPU through fm_pool (config 3 rows in
pinned host memory): cold (every row crosses PCIe) and warm (second poll of the same row ids:
baselines resident in HBM, only the current rows cross).  Prints one JSON object.

    python profiles/pool_e2e.py [series per GPU] [steps]
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_1_b200 import _abi as A  # noqa: E402
from foremast_1_b200 import _ffi  # noqa: E402
from foremast_1_b200.engine import Pinned, Pool, synth_desc  # noqa: E402

per_gpu = int(sys.argv[1]) if len(sys.argv) > 1 else 500_000
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
with Pool() as pool:
    G = pool.size
    S = per_gpu * G
    d = synth_desc(3, S, 60, 1440, 0)
    p_cur, p_base = Pinned((S, d.cur_pitch), np.float32), Pinned((S, d.base_pitch), np.float32)
    p_mt, p_out, p_ids = Pinned((S,), np.uint8), Pinned((S,), A.SERIES_REC), Pinned((S,), np.uint64)
    assert _ffi.lib().fm_synth_fill_host(C.byref(d), p_cur.ptr, p_base.ptr, None, p_mt.ptr, 0) == 0
    p_ids.array[:] = (np.arange(S, dtype=np.uint64) + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15)
    b = A.Batch()
    b.struct_size = C.sizeof(A.Batch)
    b.strategy, b.n_series = A.FM_STRAT_CANARY, S
    b.cur, b.n_cur, b.cur_pitch = p_cur.ptr, 60, d.cur_pitch
    b.base, b.n_base, b.base_pitch = p_base.ptr, 1440, d.base_pitch
    b.metric_type = p_mt.ptr
    g, t = A.default_params()
    pool.set_params(g, t)

    def timed(row_ids):
        pool.score_host(b, p_out.ptr, row_ids=row_ids)
        t0 = time.perf_counter()
        for _ in range(steps):
            tm = pool.score_host(b, p_out.ptr, row_ids=row_ids)
        dt = (time.perf_counter() - t0) / steps
        return {"series_per_s": S / dt, "ms_per_step": 1e3 * dt,
                "h2d_bytes_per_step": int(sum(tm[i].h2d_bytes for i in range(G))),
                "kernel_ms_per_device": [round(float(tm[i].kernel_ms), 2) for i in range(G)]}

    cold = timed(None)
    first = p_out.array[:4096].copy()
    pool.cache_configure(per_gpu * d.base_pitch * 4 + (64 << 20))
    warm = timed(p_ids.ptr)
    same = bool((p_out.array[:4096] == first).all())
    print(json.dumps({"gpus": G, "series": S, "cold": cold, "warm": warm, "warm_equals_cold": same}))
