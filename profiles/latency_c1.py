"""Config 1 latency (diagnostics): one healthcheck job = 1 service x 4 metrics, 60-point canary vs
1440-point baseline, through fm_score_batch with host rows (H2D + kernels + D2H inside the call)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_1_b200 import _abi as A  # noqa: E402
from foremast_1_b200.engine import Context, synth_desc, synth_host  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 4
d = synth_desc(1, S, 60, 1440, 0)
cur, base, _, mt = synth_host(d)
with Context(0) as ctx:
    g, t = A.default_params()
    ctx.set_params(g, t)
    api = np.zeros(S, np.uint32)
    for _ in range(20):
        ctx.score(cur, base, None, metric_type=mt, api_id=api, n_apis=1)
    ts = []
    for _ in range(300):
        t0 = time.perf_counter()
        out = ctx.score(cur, base, None, metric_type=mt, api_id=api, n_apis=1)
        ts.append(time.perf_counter() - t0)
    ts = np.array(ts) * 1e6
    print(json.dumps({"workload": "config 1", "series": S, "call_us_median": float(np.median(ts)),
                      "call_us_p99": float(np.percentile(ts, 99)), "verdicts": out["rec"]["verdict"][:4].tolist()}))
