"""The brain mirror end to end on the real library: documents in, statuses out."""
import pytest

from foremast_1_b200 import brain as B
from foremast_1_b200.engine import Context
from test_brain_host import DEPLOYED_ENV, check_scored, make_docs

pytestmark = pytest.mark.gpu


def test_score_documents_on_gpu():
    docs, fetch = make_docs()
    with Context(0) as ctx:
        g, table, names = B.params_from_env(DEPLOYED_ENV)
        ctx.set_params(g, table)
        check_scored(B.score_documents(ctx, docs, fetch, names))


def test_worker_on_gpu():
    docs, fetch = make_docs()
    store = B.MemoryStore(docs)
    with Context(0) as ctx:
        w = B.BrainWorker(ctx, store, fetch, env=DEPLOYED_ENV, now=lambda: 100.0)
        res = w.poll_once()
        check_scored(res)
        assert store.docs["bad"]["status"] == B.ST_UNHEALTH and store.docs["ok"]["status"] == B.ST_HEALTH


def test_worker_keeps_models_resident_between_polls():
    """MAX_CACHE_SIZE (foremast-brain/README.md:30): a job that is re-judged on the next poll
    (healthy before its endTime, docs/guides/design.md:43) keeps its baseline and history rows in HBM;
    the second poll fetches the current windows only and judges the same."""
    docs, fetch = make_docs()
    for d in docs:
        d["endTime"] = B.format_go_time(10_000.0)
    calls = []

    def counting_fetch(url):
        calls.append(url)
        return fetch(url)

    store = B.MemoryStore(docs)
    clock = [100.0]
    with Context(0) as ctx:
        env = dict(DEPLOYED_ENV, MAX_CACHE_SIZE="1000", ML_BASELINE_POINTS="60", ML_HISTORICAL_POINTS="600")
        w = B.BrainWorker(ctx, store, counting_fetch, env=env, now=lambda: clock[0])
        res1 = w.poll_once()
        check_scored(res1)
        first = list(calls)
        assert any("/h_lat" in u for u in first) and any("/b_lat" in u for u in first)
        assert store.docs["ok"]["status"] == B.ST_INITIAL            # healthy before endTime: judged again
        del calls[:]
        clock[0] = 200.0
        res2 = w.poll_once()
        assert set(res2) == {"ok"} and res2["ok"]["status"] == B.ST_HEALTH
        assert sorted(calls) == ["http://p/c_err", "http://p/c_lat_ok"]      # no baseline / history fetch
        for alias in ("latency", "error5xx"):
            a, b = res1["ok"]["metrics"][alias], res2["ok"]["metrics"][alias]
            assert a == b                                             # same rows, same record, bit for bit
        assert ctx.cache_stats()["hits"] >= 2
