"""Host logic of the brain mirror (no GPU): config strings, env knobs, Prometheus
decoding, status mapping, job loop.  Scoring is delegated to a stand-in context
that answers with the CPU oracle, so only the host code is under test here; the
same flow against the real library is tests/test_gpu_brain.py."""
import ctypes as C
import json
import re

import numpy as np
import pytest

import oracle_lib as O
from foremast_1_b200 import _abi as A
from foremast_1_b200 import brain as B
from foremast_1_b200 import gauges as gauges_mod
from foremast_1_b200.engine import _as_matrix


class OracleCtx:
    """engine.Context look-alike backed by the oracle (tests only)."""

    def __init__(self):
        self.g, self.table = A.default_params()

    def set_params(self, g, table):
        self.g, self.table = g, table

    def score(self, cur, base=None, hist=None, *, strategy, cur_len=None, base_len=None,
              hist_len=None, metric_type=None, **kw):
        cur, nc = _as_matrix(cur)
        base, nb = _as_matrix(base)
        hist, nh = _as_matrix(hist)
        b = A.Batch()
        b.struct_size = C.sizeof(A.Batch)
        b.strategy, b.n_series = strategy, cur.shape[0]
        b.cur, b.n_cur, b.cur_pitch = A.ptr(cur), nc, cur.shape[1]
        keep = [cur, base, hist, cur_len, base_len, hist_len, metric_type]
        if base is not None:
            b.base, b.n_base, b.base_pitch = A.ptr(base), nb, base.shape[1]
        if hist is not None:
            b.hist, b.n_hist, b.hist_pitch = A.ptr(hist), nh, hist.shape[1]
        b.cur_len, b.base_len, b.hist_len = A.ptr(cur_len), A.ptr(base_len), A.ptr(hist_len)
        b.metric_type = A.ptr(metric_type)
        out = O.score_batch(self.g, self.table, b, n_threads=2)
        out["_keep"] = keep
        return out


def prom(ts0, values, step=60, metric=None):
    return json.dumps({"status": "success", "data": {"resultType": "matrix", "result": [
        {"metric": metric or {}, "values": [[ts0 + i * step, str(v)] for i, v in enumerate(values)]}]}})


def test_config_string_round_trip():
    d = {"error5xx": "http://prom/api/v1/query_range?query=a%7Bb%3D%22c%22%7D&start=1&end=2&step=60",
         "latency": "http://prom/api/v1/query_range?query=l&start=1&end=2&step=60"}
    s = B.format_config(d)
    assert s.count(B.CONFIG_SEPARATOR) == 1 and "error5xx== http" in s      # main.go:28-31
    assert B.parse_config(s) == d
    assert B.parse_config("") == {}
    with pytest.raises(ValueError):
        B.parse_config("no separator here")


def test_external_status_mapping():
    # pkg/converter/converter.go:11-30
    m = {"initial": "new", "preprocess_inprogress": "inprogress", "postprocess_inprogress": "inprogress",
         "preprocess_completed": "inprogress", "completed_health": "success", "completed_unhealth": "anomaly",
         "completed_unknown": "abort", "preprocess_failed": "abort", "abort": "abort", "whatever": "inprogress"}
    for k, v in m.items():
        assert B.external_status(k) == v


# the env block of deploy/foremast/3_brain/foremast-brain.yaml:24-81
DEPLOYED_ENV = {"ML_ALGORITHM": "moving_average_all", "threshold": "2.0", "min_lower_bound": "0", "bound": "1",
           "metric_type_threshold_count": "5",
           "metric_type0": "error5xx", "threshold0": "2", "bound0": "1", "min_lower_bound0": "0",
           "metric_type1": "error4xx", "threshold1": "3", "bound1": "1", "min_lower_bound1": "0",
           "metric_type2": "latency", "threshold2": "10", "bound2": "3", "min_lower_bound2": "0",
           "metric_type3": "cpu", "threshold3": "5", "bound3": "1", "min_lower_bound3": "0",
           "metric_type4": "memory", "threshold4": "5", "bound4": "1", "min_lower_bound4": "0",
           "MIN_MANN_WHITE_DATA_POINTS": "20", "MIN_WILCOXON_DATA_POINTS": "20",
           "MIN_KRUSKAL_DATA_POINTS": "5", "MAX_STUCK_IN_SECONDS": "90"}


def test_params_from_deployed_env():
    g, table, names = B.params_from_env(DEPLOYED_ENV)
    dg, dt = A.default_params()
    assert bytes(g) == bytes(dg) and bytes(table) == bytes(dt)
    assert names == A.METRIC_TYPES
    assert B.metric_type_id("latency", names) == 3
    assert B.metric_type_id("namespace_app_per_pod:http_server_requests_error5xx", names) == 1
    assert B.metric_type_id("something_else", names) == 0
    with pytest.raises(ValueError):
        B.params_from_env({"ML_ALGORITHM": "prophet"})
    with pytest.raises(ValueError):
        B.params_from_env({"bound": "4"})
    g2, _, _ = B.params_from_env({"ML_PAIRWISE_ALGORITHM": "mann_white", "ML_PAIRWISE_THRESHOLD": "0.01"})
    assert g2.pairwise_mode == A.FM_PW_MANN_WHITE and g2.alpha == 0.01


def test_decode_query_range_drops_gaps_and_merges_pods():
    body = json.dumps({"status": "success", "data": {"result": [
        {"metric": {"pod": "a"}, "values": [[100, "1.5"], [160, "NaN"], [220, "2.5"]]},
        {"metric": {"pod": "b"}, "values": [[130, "7"], [190, "+Inf"]]}]}})
    res = B.decode_query_range(body)
    assert [r[2].tolist() for r in res] == [[1.5, 2.5], [7.0]]
    # several pods: one row per alias on the 60 s grid, mean over the pods that reported there
    # (samples at 100 and 130 both snap to grid point 120; 160 and 190 are gaps and stay gaps)
    ts, x = B.merge_series(res)
    assert ts.tolist() == [120, 240] and x.tolist() == [4.25, 2.5]
    # one pod, on-grid timestamps (what Prometheus returns for the reference's step-aligned
    # queries, metricsquery.go:43-46): untouched; a duplicate slot keeps its last sample
    one = [({}, np.array([60, 120, 120, 300], np.int64), np.array([1, 2, 3, 4], np.float32))]
    ts, x = B.merge_series(one)
    assert ts.tolist() == [60, 120, 300] and x.tolist() == [1.0, 3.0, 4.0]
    assert B.decode_query_range(json.dumps(body))[0][2].tolist() == [1.5, 2.5]   # double-encoded proxy reply
    with pytest.raises(ValueError):
        B.decode_query_range('{"status":"error","error":"boom"}')
    assert B.merge_series([])[1].size == 0


def make_docs():
    rng = np.random.default_rng(0)
    hist = rng.normal(100, 5, 600)
    store = {
        "http://p/h_lat": prom(0, hist), "http://p/b_lat": prom(39960, rng.normal(100, 5, 60)),
        "http://p/c_lat_ok": prom(49980, rng.normal(100, 5, 30)),
        "http://p/c_lat_bad": prom(49980, np.concatenate([rng.normal(100, 5, 28), [400.0, 420.0]])),
        "http://p/h_err": prom(0, rng.integers(0, 3, 600) / 60), "http://p/c_err": prom(49980, rng.integers(0, 3, 30) / 60),
        "http://p/empty": json.dumps({"status": "success", "data": {"result": []}}),
    }
    docs = [
        {"id": "ok", "strategy": "canary", "status": "initial", "modified_at": "2018-11-03T16:00:00Z",
         "currentConfig": B.format_config({"latency": "http://p/c_lat_ok", "error5xx": "http://p/c_err"}),
         "baselineConfig": B.format_config({"latency": "http://p/b_lat"}),
         "historicalConfig": B.format_config({"latency": "http://p/h_lat", "error5xx": "http://p/h_err"})},
        {"id": "bad", "strategy": "rollingUpdate", "status": "initial", "modified_at": "2018-11-03T16:00:00Z",
         "currentConfig": B.format_config({"latency": "http://p/c_lat_bad"}), "baselineConfig": "",
         "historicalConfig": B.format_config({"latency": "http://p/h_lat"})},
        {"id": "nocur", "strategy": "continuous", "status": "initial", "modified_at": "2018-11-03T16:00:00Z",
         "currentConfig": B.format_config({"latency": "http://p/empty"}), "baselineConfig": "",
         "historicalConfig": B.format_config({"latency": "http://p/h_lat"})},
        {"id": "broken", "strategy": "canary", "status": "initial", "modified_at": "2018-11-03T16:00:00Z",
         "currentConfig": B.format_config({"latency": "http://p/404"}), "baselineConfig": "", "historicalConfig": ""},
    ]
    return docs, (lambda url: store[url])


def check_scored(res):
    assert res["ok"]["status"] == B.ST_HEALTH and res["ok"]["anomalyInfo"] == ""
    assert set(res["ok"]["metrics"]) == {"latency", "error5xx"}
    assert res["bad"]["status"] == B.ST_UNHEALTH
    info = json.loads(res["bad"]["anomalyInfo"])
    vals = info["latency"]["values"]          # [t0, v0, t1, v1] (Barrelman.go:593-620)
    assert vals == [49980 + 28 * 60, 400.0, 49980 + 29 * 60, 420.0]
    assert "latency: 2 anomalous" in res["bad"]["reason"]
    assert res["nocur"]["status"] == B.ST_UNKNOWN and "no current metric" in res["nocur"]["reason"]
    assert res["broken"]["status"] == B.ST_PRE_FAILED and "KeyError" in res["broken"]["reason"]


def test_score_documents_statuses_and_anomaly_info():
    docs, fetch = make_docs()
    check_scored(B.score_documents(OracleCtx(), docs, fetch))


T0 = 1541260800.0          # 2018-11-03T16:00:00Z
RFC3339_UTC = re.compile(r"^\d{4}-\d\d-\d\dT\d\d:\d\d:\d\d(\.\d{1,9})?Z$")


def test_go_time_round_trip():
    """time.Time as json.Marshal writes it (models.go:86-89) and as common.StrToTime reads the
    request's startTime / endTime (timeutils.go:9-19)."""
    assert B.parse_go_time("2018-11-03T16:00:00Z") == T0
    assert B.parse_go_time("2018-11-03T09:00:00-07:00") == T0
    assert B.parse_go_time("2018-11-03T16:00:00.123456789Z") == pytest.approx(T0 + 0.123456789, abs=1e-6)
    assert B.parse_go_time("0001-01-01T00:00:00Z") is None          # Go's zero time: "not set"
    assert B.parse_go_time("") is None and B.parse_go_time(None) is None
    assert B.parse_go_time(12.5) == 12.5                             # stamps of older workers of this package
    with pytest.raises(ValueError):
        B.parse_go_time("yesterday")
    assert B.format_go_time(T0) == "2018-11-03T16:00:00Z"            # zero fraction is left out
    assert B.format_go_time(T0 + 0.25) == "2018-11-03T16:00:00.25Z"  # trailing zeros trimmed
    for t in (T0, T0 + 0.5, T0 + 86399.000001):
        assert B.parse_go_time(B.format_go_time(t)) == pytest.approx(t, abs=1e-6)
        assert RFC3339_UTC.match(B.format_go_time(t))


def test_worker_claim_takeover_and_fail_fast():
    docs, fetch = make_docs()
    for d in docs:
        d["endTime"] = B.format_go_time(T0 + 1000)
    docs.append(dict(docs[0], id="stuck", status=B.ST_PRE_INPROGRESS, modified_at=B.format_go_time(T0 + 5)))
    docs.append(dict(docs[0], id="busy", status=B.ST_POST_INPROGRESS, modified_at=B.format_go_time(T0 + 95)))
    store = B.MemoryStore(docs)
    clock = [T0 + 100.0]
    w = B.BrainWorker(OracleCtx(), store, fetch, env=DEPLOYED_ENV, now=lambda: clock[0])
    res = w.poll_once()
    # "stuck" (95 s old) is taken over, "busy" (5 s old) is left to its owner
    assert "stuck" in res and "busy" not in res and store.docs["busy"]["status"] == B.ST_POST_INPROGRESS
    # fail fast: unhealthy completes now; healthy before endTime goes back to the queue
    assert store.docs["bad"]["status"] == B.ST_UNHEALTH
    assert store.docs["ok"]["status"] == B.ST_INITIAL and res["ok"]["status"] == B.ST_HEALTH
    assert store.docs["nocur"]["status"] == B.ST_UNKNOWN
    assert store.docs["broken"]["status"] == B.ST_PRE_FAILED and "KeyError" in store.docs["broken"]["reason"]
    # no time field is ever written as a number
    for d in store.docs.values():
        assert isinstance(d["modified_at"], str) and RFC3339_UTC.match(d["modified_at"]), d
    clock[0] = T0 + 2000.0
    res = w.poll_once()
    assert store.docs["ok"]["status"] == B.ST_HEALTH and B.external_status(store.docs["ok"]["status"]) == "success"
    assert "bad" not in res                       # completed documents are not re-scored
    # an optimistic-lock loser does not claim
    assert store.update("ok", expect_modified_at="1999-01-01T00:00:00Z", status="x") is False
    assert w.errors == []


class FakeES:
    """The two Elasticsearch calls of ElasticsearchStore over a dict of _source documents; keeps the
    order of the statuses every document went through."""

    def __init__(self, docs):
        self.docs = docs
        self.calls, self.trail = [], {k: [] for k in docs}
        self.fail_ids = set()

    def __call__(self, method, url, body):
        self.calls.append((method, url, body))
        if url.endswith("/documents/_search"):
            want = set(body["query"]["terms"]["status"])
            hits = [{"_id": k, "_source": dict(v)} for k, v in self.docs.items() if v["status"] in want]
            return {"hits": {"hits": hits[: body["size"]]}}
        doc_id = url.split("/")[-2]
        if doc_id in self.fail_ids:
            raise OSError("connection reset")
        p = body["script"]["params"] if "script" in body else {"expect": None, "doc": body["doc"]}
        if p["expect"] is not None and self.docs[doc_id]["modified_at"] != p["expect"]:
            return {"result": "noop"}
        self.docs[doc_id].update(p["doc"])
        if "status" in p["doc"]:
            self.trail[doc_id].append(p["doc"]["status"])
        return {"result": "updated"}


def test_elasticsearch_store_speaks_the_documents_index():
    """ElasticsearchStore against a fake transport: search by status, optimistic update on the
    RFC3339 stamp."""
    es = FakeES({"a1": {"status": "initial", "modified_at": "2018-11-03T16:00:05Z", "appName": "x"},
                 "b2": {"status": "completed_health", "modified_at": "2018-11-03T16:00:06Z"}})
    st = B.ElasticsearchStore("http://es:9200/", transport=es)
    got = st.by_status({"initial", "preprocess_inprogress"})
    assert [d["id"] for d in got] == ["a1"] and got[0]["appName"] == "x"
    assert es.calls[0][1] == "http://es:9200/documents/_search"
    assert st.update("a1", expect_modified_at="2018-11-03T16:00:05Z", status="preprocess_inprogress",
                     modified_at="2018-11-03T16:00:09Z")
    assert es.docs["a1"]["status"] == "preprocess_inprogress"
    assert not st.update("a1", expect_modified_at="2018-11-03T16:00:05Z", status="initial",
                         modified_at="2018-11-03T16:00:10Z")   # lost the race
    assert es.calls[-1][1] == "http://es:9200/documents/document/a1/_update"
    assert st.update("b2", reason="ok")
    assert es.docs["b2"]["reason"] == "ok"


def test_worker_on_a_go_marshalled_document():
    """tests/golden/document_go_marshal.json is json.Marshal(models.Document) byte for byte as
    CreateNewDoc indexes it (elasticsearchstore.go:36-52; field order and names from the struct
    tags, models.go:83-101): time.Time fields are RFC3339 strings, startTime / endTime keep the
    offset of the request.  The worker must claim it, walk the states of the diagram, re-queue it
    while endTime lies ahead, take it over when its owner is stuck, and leave every time field a
    string that time.Time can unmarshal."""
    import os
    raw = open(os.path.join(os.path.dirname(__file__), "golden", "document_go_marshal.json"), "rb").read()
    src = json.loads(raw)
    assert list(src)[:6] == ["id", "appName", "created_at", "startTime", "endTime", "modified_at"]
    docs0, fetch = make_docs()
    src["currentConfig"], src["baselineConfig"], src["historicalConfig"] = (
        docs0[0]["currentConfig"], docs0[0]["baselineConfig"], docs0[0]["historicalConfig"])
    end = B.parse_go_time(src["endTime"])
    assert end == B.parse_go_time("2018-11-03T16:44:00Z")          # 09:44 at -07:00
    es = FakeES({src["id"]: dict(src)})
    store = B.ElasticsearchStore("http://es:9200", transport=es)
    clock = [B.parse_go_time(src["created_at"]) + 10.0]
    reg = gauges_mod.GaugeRegistry()
    w = B.BrainWorker(OracleCtx(), store, fetch, env=DEPLOYED_ENV, now=lambda: clock[0], gauges=reg)
    res = w.poll_once()
    did = src["id"]
    # healthy before endTime: back to the queue, having passed every state of the diagram
    assert es.trail[did] == [B.ST_PRE_INPROGRESS, B.ST_PRE_COMPLETED, B.ST_POST_INPROGRESS, B.ST_INITIAL]
    assert res[did]["status"] == B.ST_HEALTH and res[did]["written_status"] == B.ST_INITIAL
    d = es.docs[did]
    for f in ("created_at", "startTime", "endTime", "modified_at"):
        assert isinstance(d[f], str) and B.parse_go_time(d[f]) is not None
    assert RFC3339_UTC.match(d["modified_at"]) and d["created_at"] == src["created_at"]
    assert d["endTime"] == src["endTime"] and d["startTime"] == src["startTime"]    # not ours to rewrite
    # another worker dies holding the job: nothing happens for 90 s, then it is taken over
    stuck = B.format_go_time(clock[0] + 1.0)
    d["status"], d["modified_at"] = B.ST_POST_INPROGRESS, stuck
    clock[0] += 60.0
    assert w.poll_once() == {}
    clock[0] += 40.0
    res = w.poll_once()
    assert res[did]["written_status"] == B.ST_INITIAL and es.docs[did]["modified_at"] != stuck
    # endTime reached: the healthy verdict completes the job
    clock[0] = end + 1.0
    res = w.poll_once()
    assert es.docs[did]["status"] == B.ST_HEALTH and res[did]["written_status"] == B.ST_HEALTH
    assert B.external_status(es.docs[did]["status"]) == "success"          # converter.go:11-30
    # the bounds of both metrics were published for :8000/metrics
    text = reg.text()
    assert "foremastbrain:latency_upper" in text and "foremastbrain:error5xx_lower" in text
    # a job somebody else took over between scoring and write-back is not overwritten
    es.docs[did]["status"] = B.ST_INITIAL
    orig_update = store.update

    def racing_update(doc_id, expect_modified_at=None, **fields):
        if fields.get("status") in (B.ST_HEALTH, B.ST_UNHEALTH, B.ST_UNKNOWN):
            es.docs[doc_id]["modified_at"] = "2018-11-03T23:59:59Z"      # the other worker's stamp
        return orig_update(doc_id, expect_modified_at=expect_modified_at, **fields)

    store.update = racing_update
    res = w.poll_once()
    assert res[did]["written_status"] is None and es.docs[did]["status"] == B.ST_POST_INPROGRESS
    assert w.errors == []


def test_one_bad_write_back_does_not_strand_the_batch():
    docs, fetch = make_docs()
    es = FakeES({d["id"]: dict(d) for d in docs})
    es.fail_ids = {"bad"}
    w = B.BrainWorker(OracleCtx(), B.ElasticsearchStore("http://es:9200", transport=es), fetch, env=DEPLOYED_ENV,
                      now=lambda: T0 + 100.0)
    res = w.poll_once()
    assert es.docs["ok"]["status"] == B.ST_HEALTH and es.docs["nocur"]["status"] == B.ST_UNKNOWN
    assert "bad" not in res or res["bad"].get("written_status") is None
    assert w.errors and w.errors[0][0] == "bad"


def test_metrics_endpoint_serves_what_the_worker_scored():
    """GET /metrics (foremast-brain.yaml:87-122: port 8000, path /metrics) returns the gauges the UI
    plots (metrics.js:14-24): bounds of every judged metric and one _anomaly sample per anomalous
    point whose value is the point's unix timestamp (App.js:231-241)."""
    import urllib.request
    docs, fetch = make_docs()
    docs[1]["appName"] = "demo"
    reg = gauges_mod.GaugeRegistry()
    srv = gauges_mod.MetricsServer(reg, host="127.0.0.1", port=0).start()
    try:
        w = B.BrainWorker(OracleCtx(), B.MemoryStore(docs), fetch, env=DEPLOYED_ENV, now=lambda: T0, gauges=reg)
        w.poll_once()
        body = urllib.request.urlopen(f"http://127.0.0.1:{srv.port}/metrics", timeout=5).read().decode()
        assert "# TYPE foremastbrain:latency_upper gauge" in body
        anomalies = [ln for ln in body.splitlines() if ln.startswith("foremastbrain:latency_anomaly")]
        assert [int(ln.split()[-1]) for ln in anomalies] == [49980 + 28 * 60, 49980 + 29 * 60]
        assert 'foremastbrain:latency_anomaly{app="demo"}' in body
        with pytest.raises(urllib.error.HTTPError):
            urllib.request.urlopen(f"http://127.0.0.1:{srv.port}/other", timeout=5)
    finally:
        srv.stop()


def _payload(rng, n_series, n_points, style=0):
    import json
    res = []
    t0 = 1554076800
    for s in range(n_series):
        vals = []
        for i in range(n_points):
            r = rng.random()
            if r < 0.03:
                v = rng.choice(["NaN", "+Inf", "-Inf", "Inf"])
            elif r < 0.2:
                v = repr(float(rng.integers(0, 50)))
            elif r < 0.4:
                v = "%.6e" % rng.normal(0, 1e-5)
            elif r < 0.45:
                v = rng.choice(["1e-400", "-0", "0.1", "123456789012345678901234567890", "4.9e-324", "1e400"])
            else:
                v = repr(float(rng.normal(100, 30)))
            ts = t0 + 60 * i if style == 0 else t0 + 60 * i + 0.123
            vals.append([ts, v])
        res.append({"metric": {"__name__": 'na"me\\\\{x}', "pod": "p-%d" % s, "nested": "a]b}c"}, "values": vals})
    body = {"status": "success", "data": {"resultType": "matrix", "result": res}, "warnings": ["x", {"y": [1, 2]}]}
    return json.dumps(body, indent=None if style == 0 else 1).encode()


def test_decode_query_range_fast_equals_python():
    """C scanner (fm_decode_query_range) against the json.loads decoder: every sample bit for bit,
    gaps dropped, escapes / nested values in labels skipped, compact and indented bodies."""
    from foremast_1_b200 import brain
    rng = np.random.default_rng(3)
    for style in (0, 1):
        body = _payload(rng, 7, 300, style)
        ref = brain.decode_query_range(body)
        got = brain.decode_query_range_fast(body)
        assert len(got) == len(ref) == 7
        for (_, t0, x0), (_, t1, x1) in zip(ref, got):
            assert np.array_equal(t0, t1)
            assert x0.tobytes() == x1.tobytes()
    assert brain.decode_query_range_fast(b'{"status":"success","data":{"resultType":"matrix","result":[]}}') == []
    one = brain.decode_query_range_fast(b'{"data":{"result":[{"metric":{},"value":[12.5,"3"]}]}}')
    assert one[0][1].tolist() == [12] and one[0][2].tolist() == [3.0]


def test_decode_query_range_fast_errors():
    from foremast_1_b200 import brain
    for bad in (b'{"status":"error","errorType":"bad_data","error":"x"}', b'[1,2]', b'{"data":{"result":[{"values":[[1,"2"]',
                b'{"data":{"result":[{"values":[[1,"abc"]]}]}}', b''):
        with pytest.raises(ValueError):
            brain.decode_query_range_fast(bad)
