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
