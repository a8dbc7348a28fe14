"""Host-side decode rate of Prometheus query_range bodies: json.loads + numpy against
the C scanner fm_decode_query_range.  CPU only."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foremast_1_b200 import brain  # noqa: E402

rng = np.random.default_rng(1)
S, N = 400, 1440
res = []
for s in range(S):
    v = rng.normal(250, 40, N)
    res.append({"metric": {"__name__": "namespace_pod:http_server_requests_latency", "pod": "p-%d" % s},
                "values": [[1554076800 + 60 * i, repr(float(x))] for i, x in enumerate(v)]})
body = json.dumps({"status": "success", "data": {"resultType": "matrix", "result": res}}).encode()


def best(fn, n=3):
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        out = fn(body)
        ts.append(time.perf_counter() - t0)
    return min(ts), out


t_py, a = best(brain.decode_query_range)
t_c, b = best(brain.decode_query_range_fast)
assert all(x[2].tobytes() == y[2].tobytes() for x, y in zip(a, b))
print(json.dumps({"body_MB": len(body) / 1e6, "samples": S * N, "python_s": t_py, "c_s": t_c,
                  "python_MBps": len(body) / 1e6 / t_py, "c_MBps": len(body) / 1e6 / t_c,
                  "c_Msamples_per_s": S * N / t_c / 1e6, "speedup": t_py / t_c}))
