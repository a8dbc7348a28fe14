"""Small batches through every kernel path (all FOREMAST_CUDA_PATH settings, ragged rows, roll-up,
groups, rank dump, window series).  Written for `compute-sanitizer --tool memcheck`; that tool is
closed on this pool (gpurun answers rc 86), so it only serves as a quick all-paths smoke run."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from foremast_1_b200 import _abi as A  # noqa: E402
from foremast_1_b200.engine import Context, synth_desc, synth_host  # noqa: E402

rng = np.random.default_rng(1)
for path in ("auto", "elect", "generic"):
    os.environ["FOREMAST_CUDA_PATH"] = path
    with Context(0) as ctx:
        g, t = A.default_params()
        ctx.set_params(g, t)
        for (S, nc, nb) in ((70, 60, 1440), (33, 64, 2048), (40, 7, 100), (9, 1, 1), (17, 33, 513)):
            cur, base, _, mt = synth_host(synth_desc(3, S, nc, nb, 0))
            cl = rng.integers(0, nc + 1, S).astype(np.uint32)
            bl = rng.integers(0, nb + 1, S).astype(np.uint32)
            ctx.score(cur, base, metric_type=mt, want_aux=True)
            ctx.score(cur, base, cur_len=cl, base_len=bl, metric_type=mt,
                      api_id=(np.arange(S) // 4).astype(np.uint32), api_service=np.zeros((S + 3) // 4, np.uint32),
                      n_apis=(S + 3) // 4, n_services=1)
        for nh in (60, 200, 700, 3000):
            cur, _, hist, mt = synth_host(synth_desc(5, 50, 60, 0, nh))
            ctx.score(cur, None, hist, strategy=A.FM_STRAT_HISTORICAL, metric_type=mt, record_kind=A.FM_REC_COMPACT,
                      hist_len=rng.integers(0, nh + 1, 50).astype(np.uint32))
        cur, base, hist, mt = synth_host(synth_desc(2, 20, 60, 1440, 10080))
        ctx.score(cur, base, hist, metric_type=mt)
        K = rng.integers(3, 9, 30)
        off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
        gc = rng.integers(0, 5, (int(off[-1]), 60)).astype(np.float32)
        ctx.score_groups(gc, off, rng.integers(0, 5, (30, 1440)).astype(np.float32))
        ctx.rank_dump(rng.normal(0, 1, 60), rng.normal(0, 1, 1440))
        ctx.ma_window_series(rng.normal(0, 1, 3000), 100)
print("sanitize_smoke: all paths ran")
