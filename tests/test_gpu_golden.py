"""GPU results against the committed golden fixtures (tests/golden/), through the C ABI."""
import os
import sys

import numpy as np
import pytest

from foremast_1_b200 import _abi as A
from foremast_1_b200.engine import Context
from parity import assert_records_match, bound_masks

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
pytestmark = pytest.mark.gpu


def test_gpu_matches_golden_records():
    from make_golden import CASES, case_inputs
    z = np.load(os.path.join(HERE, "golden", "oracle_records.npz"))
    with Context(0) as ctx:
        for name in CASES:
            g, t, strat, cur, base, hist, mt = case_inputs(name)
            ctx.set_params(g, t)
            out = ctx.score(cur, base, hist, strategy=strat, metric_type=mt, want_aux=True)
            nc = CASES[name][2]
            assert_records_match(out["rec"], z[f"{name}__rec"], cur[:, :nc], None,
                                 bound_masks(t, mt, cur.shape[0]), out["aux"], z[f"{name}__aux"])


def test_gpu_demo_fixture():
    import json
    with open(os.path.join(HERE, "golden", "demo_data.json")) as f:
        d = json.load(f)
    with Context(0) as ctx:
        g, t = A.default_params()
        ctx.set_params(g, t)
        out = ctx.score(np.array(d["data2"], np.float32), None, np.array(d["data1"], np.float32),
                        strategy=A.FM_STRAT_HISTORICAL, metric_type=np.array([1], np.uint8))
        r = out["rec"][0]
        spikes = [i for i, v in enumerate(d["data2"]) if v > 40]
        assert r["verdict"] == A.FM_UNHEALTHY and r["n_anom"] == 2 and r["first_idx"] == spikes[0]
        assert r["anom_mask"] == sum(1 << i for i in spikes)
