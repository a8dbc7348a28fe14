"""One process, several contexts driven concurrently from host threads (fm_pool):
the pooled result must equal the single-ctx result byte for byte — records, exact integers and
roll-ups — with the batch cut at service boundaries.  On a one-GPU box the pool lists that GPU
twice (two ctxs, two threads, six streams on one device); with more GPUs it spans them."""
import ctypes as C

import numpy as np
import pytest

from foremast_1_b200 import _abi as A
from foremast_1_b200 import _ffi
from foremast_1_b200.engine import Context, Pool, synth_desc, synth_host

pytestmark = pytest.mark.gpu


def host_batch(cur, base, mt, api=None, svc=None, n_svc=0):
    b = A.Batch()
    b.struct_size = C.sizeof(A.Batch)
    b.strategy, b.n_series = A.FM_STRAT_CANARY, cur.shape[0]
    b.cur, b.n_cur, b.cur_pitch = cur.ctypes.data, 60, cur.shape[1]
    b.base, b.n_base, b.base_pitch = base.ctypes.data, base.shape[1] if base.shape[1] % 4 == 0 else 0, base.shape[1]
    b.metric_type = mt.ctypes.data
    if api is not None:
        b.api_id, b.api_service, b.n_apis, b.n_services = api.ctypes.data, svc.ctypes.data, svc.size, n_svc
    return b


def device_lists():
    n = C.c_int32(0)
    _ffi.lib().fm_device_count(C.byref(n))
    lists = [[0, 0], [0, 0, 0]]
    if n.value >= 2:
        lists.append(list(range(n.value)))
    return lists


@pytest.mark.parametrize("devices", device_lists())
def test_pool_equals_single_context(devices):
    S = 20011                                            # not a multiple of anything
    d = synth_desc(3, S, 60, 1440, 0)
    cur, base, _, mt = synth_host(d)
    # services of uneven size: 1..40 series per API, 1..6 APIs per service
    rng = np.random.default_rng(3)
    api = np.repeat(np.arange(S), rng.integers(1, 41, S))[:S].astype(np.uint32)
    n_api = int(api[-1]) + 1
    svc = np.repeat(np.arange(n_api), rng.integers(1, 7, n_api))[:n_api].astype(np.uint32)
    n_svc = int(svc[-1]) + 1
    g, t = A.default_params()
    with Context(0) as ctx:
        ctx.set_params(g, t)
        ref = ctx.score(cur, base, metric_type=mt, want_aux=True, api_id=api, api_service=svc, n_apis=n_api,
                        n_services=n_svc)
    b = host_batch(cur, base, mt, api, svc, n_svc)
    b.n_base = 1440
    rec = np.zeros(S, A.SERIES_REC)
    aux = np.zeros(S, A.RANK_AUX)
    rapi = np.zeros(n_api, A.ROLLUP_REC)
    rsvc = np.zeros(n_svc, A.ROLLUP_REC)
    ids = (np.arange(S, dtype=np.uint64) + np.uint64(7)) * np.uint64(0x9E3779B97F4A7C15)
    with Pool(devices) as pool:
        assert pool.size == len(devices)
        pool.set_params(g, t)
        tm = pool.score_host(b, rec.ctypes.data, aux.ctypes.data, rapi.ctypes.data, rsvc.ctypes.data)
        assert rec.tobytes() == ref["rec"].tobytes() and aux.tobytes() == ref["aux"].tobytes()
        assert rapi.tobytes() == ref["api"].tobytes() and rsvc.tobytes() == ref["svc"].tobytes()
        assert all(tm[i].launches > 0 for i in range(pool.size))      # every ctx took part
        # the same through every ctx's row cache, cold and warm
        pool.cache_configure(256 << 20)
        for _ in range(2):
            rec[:] = 0; rapi[:] = 0; rsvc[:] = 0
            pool.score_host(b, rec.ctypes.data, aux.ctypes.data, rapi.ctypes.data, rsvc.ctypes.data, row_ids=ids.ctypes.data)
            assert rec.tobytes() == ref["rec"].tobytes()
            assert rapi.tobytes() == ref["api"].tobytes() and rsvc.tobytes() == ref["svc"].tobytes()
        # without a roll-up the cut is an equal share
        rec[:] = 0
        b2 = host_batch(cur, base, mt)
        b2.n_base = 1440
        pool.score_host(b2, rec.ctypes.data)
        assert rec.tobytes() == ref["rec"].tobytes()


def test_pool_errors_are_reported():
    with pytest.raises(_ffi.ForemastCudaError):
        Pool([99])
    with Pool([0]) as pool:
        b = A.Batch()
        b.struct_size = 8
        with pytest.raises(_ffi.ForemastCudaError):
            pool.score_host(b, 0)
