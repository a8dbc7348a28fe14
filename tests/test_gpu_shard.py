"""Sharding on the GPU: each "rank" is a ctx scoring its own range of series ids; the
concatenation equals the single-ctx result byte for byte.  With one visible
GPU the ranks run one after another on the same device."""
import numpy as np
import pytest

from foremast_1_b200 import _abi as A
from foremast_1_b200 import shard
from foremast_1_b200.engine import Context, synth_desc, synth_host

pytestmark = pytest.mark.gpu


def test_sharded_equals_whole():
    import torch
    n, ngpu = 6000, max(1, torch.cuda.device_count())
    g, t = A.default_params()
    svc = np.arange(n) // 32

    def score(dev, lo, hi):
        with Context(dev) as ctx:
            ctx.set_params(g, t)
            cur, base, _, mt = synth_host(synth_desc(3, hi - lo, 60, 1440, 0, first_series=lo))
            return ctx.score(cur, base, metric_type=mt)["rec"]

    whole = score(0, 0, n)
    for world in (2, 3, 8):
        bounds = shard.shard_bounds(n, world, svc)
        parts = [score(r % ngpu, lo, hi) for r, (lo, hi) in enumerate(bounds) if hi > lo]
        assert np.concatenate(parts).tobytes() == whole.tobytes()
