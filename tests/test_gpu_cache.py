"""Device-resident row cache (fm_score_batch_cached; MAX_CACHE_SIZE, foremast-brain/README.md:30):
a re-poll of the same jobs must give byte-identical records without re-reading baseline / history
rows from the host, evicted rows must come back correctly, and a batch larger than the cache must
still be scored."""
import numpy as np
import pytest

from foremast_1_b200 import _abi as A
from foremast_1_b200.engine import Context, synth_desc, synth_host

pytestmark = pytest.mark.gpu


def ids_of(lo, n, salt=0x9E3779B97F4A7C15):
    return ((np.arange(lo, lo + n, dtype=np.uint64) + np.uint64(1)) * np.uint64(salt)).astype(np.uint64)


def test_warm_equals_cold_and_skips_the_host_rows():
    S = 6000
    d = synth_desc(3, S, 60, 1440, 0)
    cur, base, _, mt = synth_host(d)
    ids = ids_of(0, S)
    with Context(0) as ctx:
        g, t = A.default_params()
        ctx.set_params(g, t)
        plain = ctx.score(cur, base, metric_type=mt, want_aux=True)
        ctx.cache_configure(64 << 20)
        cold = ctx.score(cur, base, metric_type=mt, want_aux=True, row_ids=ids)
        st = ctx.cache_stats()
        assert st["misses"] == S and st["hits"] == 0 and st["resident_rows"] == S
        assert cold["rec"].tobytes() == plain["rec"].tobytes() and cold["aux"].tobytes() == plain["aux"].tobytes()
        # second poll: same ids, the host baseline is garbage now — it must not be read
        junk = np.full_like(base, 1.0e30)
        warm = ctx.score(cur, junk, metric_type=mt, want_aux=True, row_ids=ids)
        st = ctx.cache_stats()
        assert st["hits"] == S and st["misses"] == S
        assert warm["rec"].tobytes() == plain["rec"].tobytes() and warm["aux"].tobytes() == plain["aux"].tobytes()
        assert warm["timing"].h2d_bytes < cold["timing"].h2d_bytes / 10          # 240 B of current row, not 6 KB
        assert ctx.cache_query(ids[:5]).all() and not ctx.cache_query(ids_of(10**6, 5)).any()
        # a new current window on resident baselines = the plain result for those rows
        cur2 = np.roll(cur, 7, axis=0)
        warm2 = ctx.score(cur2, junk, metric_type=mt, want_aux=True, row_ids=ids)
        plain2 = ctx.score(cur2, base, metric_type=mt, want_aux=True)
        assert warm2["rec"].tobytes() == plain2["rec"].tobytes()
        # a permuted re-poll (jobs finished / arrived in between) still finds every row
        perm = np.random.default_rng(0).permutation(S)
        warm3 = ctx.score(cur[perm], junk[perm], metric_type=mt[perm], want_aux=True, row_ids=ids[perm])
        assert warm3["rec"].tobytes() == plain["rec"][perm].tobytes()
        ctx.cache_configure(0)
        off = ctx.score(cur, base, metric_type=mt, want_aux=True, row_ids=ids)     # cache off: plain path
        assert off["rec"].tobytes() == plain["rec"].tobytes()


def test_eviction_and_batches_larger_than_the_cache():
    S = 1500
    d = synth_desc(3, 2 * S, 60, 500, 0)
    cur, base, _, mt = synth_host(d)
    row_bytes = base.shape[1] * 4
    with Context(0) as ctx:
        g, t = A.default_params()
        ctx.set_params(g, t)
        ref = ctx.score(cur, base, metric_type=mt)["rec"]
        ctx.cache_configure(1000 * row_bytes)                    # room for 1000 rows
        a, b = slice(0, 800), slice(S, S + 800)
        ra = ctx.score(cur[a], base[a], metric_type=mt[a], row_ids=ids_of(0, 800))["rec"]
        rb = ctx.score(cur[b], base[b], metric_type=mt[b], row_ids=ids_of(S, 800))["rec"]
        st = ctx.cache_stats()
        assert st["capacity_rows"] == 1000 and st["evictions"] == 600 and st["resident_rows"] == 1000
        assert ra.tobytes() == ref[a].tobytes() and rb.tobytes() == ref[b].tobytes()
        # the first job set again: 200 of its rows survived, 600 come back over PCIe
        ra2 = ctx.score(cur[a], base[a], metric_type=mt[a], row_ids=ids_of(0, 800))["rec"]
        st2 = ctx.cache_stats()
        assert ra2.tobytes() == ref[a].tobytes()
        assert st2["hits"] - st["hits"] == 200 and st2["misses"] - st["misses"] == 600
        # one call with more rows than the cache holds: the surplus is scored uncached
        big = ctx.score(cur[:S], base[:S], metric_type=mt[:S], row_ids=ids_of(5000, S))["rec"]
        assert big.tobytes() == ref[:S].tobytes()
        assert ctx.cache_stats()["uncached"] == S - 1000


def test_ragged_rows_history_and_layout_change():
    rng = np.random.default_rng(5)
    S = 900
    d = synth_desc(2, S, 60, 300, 700)
    cur, base, hist, mt = synth_host(d)
    bl = rng.integers(0, 301, S).astype(np.uint32)
    hl = rng.integers(0, 701, S).astype(np.uint32)
    ids = ids_of(0, S)
    with Context(0) as ctx:
        g, t = A.default_params()
        ctx.set_params(g, t)
        ref = ctx.score(cur, base, hist, metric_type=mt, base_len=bl, hist_len=hl)["rec"]
        ctx.cache_configure(32 << 20)
        for _ in range(2):       # cold, then warm with lengths the caller no longer knows
            got = ctx.score(cur, base, hist, metric_type=mt, base_len=bl, hist_len=hl, row_ids=ids)["rec"]
            assert got.tobytes() == ref.tobytes()
            bl, hl = np.zeros_like(bl), np.zeros_like(hl)
        # historical strategy (no baseline) is another layout: it gets the second cache, the canary
        # rows stay resident beside it (a poll of mixed jobs alternates between the two)
        ref2 = ctx.score(cur, None, hist, strategy=A.FM_STRAT_HISTORICAL, metric_type=mt)["rec"]
        for _ in range(2):
            got2 = ctx.score(cur, None, hist, strategy=A.FM_STRAT_HISTORICAL, metric_type=mt, row_ids=ids)["rec"]
            assert got2.tobytes() == ref2.tobytes()
        assert ctx.cache_stats()["resident_rows"] == 2 * S
        hits = ctx.cache_stats()["hits"]
        got = ctx.score(cur, base, hist, metric_type=mt, base_len=bl, hist_len=hl, row_ids=ids)["rec"]
        assert got.tobytes() == ref.tobytes() and ctx.cache_stats()["hits"] == hits + S
        # a third layout replaces the one used longest ago (the history-only one)
        ctx.score(cur, base[:, :100], None, metric_type=mt, row_ids=ids)
        assert ctx.cache_query(ids[:3]).all() and ctx.cache_stats()["resident_rows"] == 2 * S


def test_cached_multi_chunk_with_rollup():
    S = 30000
    d = synth_desc(3, S, 60, 1440, 0)
    cur, base, _, mt = synth_host(d)
    api = (np.arange(S) // 4).astype(np.uint32)
    svc = (np.arange(S // 4) // 8).astype(np.uint32)
    kw = dict(metric_type=mt, api_id=api, api_service=svc, n_apis=S // 4, n_services=(S // 4 + 7) // 8)
    with Context(0) as ctx:
        g, t = A.default_params()
        ctx.set_params(g, t)
        ref = ctx.score(cur, base, **kw)
        ctx.cache_configure(512 << 20)
        for _ in range(2):
            got = ctx.score(cur, base, row_ids=ids_of(0, S), **kw)
            assert got["rec"].tobytes() == ref["rec"].tobytes()
            assert (got["api"] == ref["api"]).all() and (got["svc"] == ref["svc"]).all()


@pytest.mark.parametrize("n_base,n_hist", [(3000, 0), (10080, 0), (1440, 10080), (0, 60), (0, 1440)])
def test_resident_rows_are_read_in_place(n_base, n_hist):
    """k_canary_team (baselines above 2048 points), k_canary_count with a separate history row and
    k_ma_only take a resident row straight from its cache slot; a cache half the size of the call
    mixes resident and uploaded rows in one launch, a permuted re-poll scatters the slots."""
    rng = np.random.default_rng(n_base + n_hist)
    S = 600
    d = synth_desc(2, S, 60, max(n_base, 4), max(n_hist, 4))
    cur, base, hist, mt = synth_host(d)
    base = base if n_base else None
    hist = hist if n_hist else None
    strat = A.FM_STRAT_CANARY if n_base else A.FM_STRAT_HISTORICAL
    ids = ids_of(7, S)
    with Context(0) as ctx:
        g, t = A.default_params()
        ctx.set_params(g, t)
        ref = ctx.score(cur, base, hist, strategy=strat, metric_type=mt, want_aux=True)
        row_bytes = 4 * ((base.shape[1] if n_base else 0) + (hist.shape[1] if n_hist else 0))
        for budget in (S * row_bytes * 2, S * row_bytes // 2):
            ctx.cache_configure(budget)
            for _ in range(3):
                got = ctx.score(cur, base, hist, strategy=strat, metric_type=mt, want_aux=True, row_ids=ids)
                assert got["rec"].tobytes() == ref["rec"].tobytes() and got["aux"].tobytes() == ref["aux"].tobytes()
            perm = rng.permutation(S)
            got = ctx.score(cur[perm], None if base is None else base[perm], None if hist is None else hist[perm],
                            strategy=strat, metric_type=mt[perm], want_aux=True, row_ids=ids[perm])
            assert got["rec"].tobytes() == ref["rec"][perm].tobytes()
