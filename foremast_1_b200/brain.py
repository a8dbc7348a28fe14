"""Host-side mirror of the brain's interface for the scoring path.

The reference decouples the brain through the Elasticsearch document
(foremast-service/pkg/models/models.go:83-101): the service writes
status "initial" plus three config strings, the brain fetches the series,
scores them and writes status / reason / anomalyInfo back
(docs/guides/design.md:35-49).  Everything here is host logic (string
parsing, JSON decoding, document bookkeeping); every number comes from
libforemast_cuda.so through engine.Context.

  parse_config          inverse of convertMetricQuerys (cmd/manager/main.go:28-74)
  params_from_env       the env block of deploy/foremast/3_brain/foremast-brain.yaml:21-81
  decode_query_range    Prometheus query_range JSON -> (timestamps, values) rows
                        (response shape: foremast-browser/src/App.js:278-292)
  merge_series          the samples of one query on the 60 s grid the reference queries on
                        (foremast-barrelman/pkg/client/metrics/metricsquery.go:43-78), one row per alias
  parse_go_time /       Document time fields as Go's time.Time marshals them (RFC3339, nanoseconds;
  format_go_time        foremast-service/pkg/models/models.go:86-89)
  score_documents       one poll of "initial" documents -> one coalesced batch ->
                        status / reason / anomalyInfo per document
  BrainWorker           poll / claim / takeover loop over a document store
                        (pkg/search/elasticsearchstore.go:124-149; design.md:37-43)
"""
import calendar
import ctypes as C
import json
import re
import time
import urllib.request

import numpy as np

from . import _abi as A

# foremast-service/cmd/manager/main.go:28,31
CONFIG_SEPARATOR = " ||"
KV_SEPARATOR = "== "

# internal document statuses (pkg/converter/converter.go:11-30)
ST_INITIAL = "initial"
ST_PRE_INPROGRESS = "preprocess_inprogress"
ST_PRE_COMPLETED = "preprocess_completed"
ST_PRE_FAILED = "preprocess_failed"
ST_POST_INPROGRESS = "postprocess_inprogress"
ST_HEALTH = "completed_health"
ST_UNHEALTH = "completed_unhealth"
ST_UNKNOWN = "completed_unknown"
ST_ABORT = "abort"

VERDICT_STATUS = {A.FM_HEALTHY: ST_HEALTH, A.FM_UNHEALTHY: ST_UNHEALTH, A.FM_UNKNOWN: ST_UNKNOWN}

# strategy spellings: this snapshot of the reference (metricsquery.go:17-19) and BASELINE.json
CANARY_STRATEGIES = {"canary"}
HISTORICAL_STRATEGIES = {"rollingupdate", "rollover", "continuous", "hpa"}


def external_status(status):
    """pkg/converter/converter.go:11-30 (what barrelman sees)."""
    if status == ST_INITIAL:
        return "new"
    if status == ST_HEALTH:
        return "success"
    if status == ST_UNHEALTH:
        return "anomaly"
    if status in (ST_UNKNOWN, ST_PRE_FAILED, ST_ABORT):
        return "abort"
    return "inprogress"


# Document.CreatedAt / StartTime / EndTime / ModifiedAt are time.Time (models.go:86-89): in the
# 'documents' index they are RFC3339 strings with up to nine fractional digits, "Z" for UTC
# (CreateNewDoc writes time.Now().UTC(), elasticsearchstore.go:42-45) or a numeric offset
# (StartTime / EndTime are parsed from the request, common/timeutils.go:9-19).
_GO_TIME = re.compile(r"^(\d{4})-(\d\d)-(\d\d)[Tt ](\d\d):(\d\d):(\d\d)(?:[.,](\d{1,9}))?([Zz]|[+-]\d\d:?\d\d)$")


def parse_go_time(v):
    """RFC3339 string -> unix seconds (float); None for empty values and Go's zero time
    (0001-01-01T00:00:00Z, what an unset time.Time marshals to).  Numbers pass through, so that
    documents written by older workers of this package are still understood."""
    if v is None or v == "":
        return None
    if isinstance(v, (int, float)):
        return float(v)
    m = _GO_TIME.match(str(v).strip())
    if not m:
        raise ValueError(f"not an RFC3339 time: {v!r}")
    y, mo, d, hh, mi, ss = (int(m.group(i)) for i in range(1, 7))
    if y <= 1:
        return None
    t = calendar.timegm((y, mo, d, hh, mi, ss, 0, 0, 0))
    if m.group(7):
        t += int(m.group(7).ljust(9, "0")) / 1e9
    z = m.group(8)
    if z not in "Zz":
        sign = -1 if z[0] == "-" else 1
        digits = z[1:].replace(":", "")
        t -= sign * (int(digits[:2]) * 3600 + int(digits[2:]) * 60)
    return float(t)


def format_go_time(t):
    """unix seconds -> what json.Marshal(time.Time in UTC) writes: RFC3339 with the fractional
    part trimmed of trailing zeros and left out when it is zero (Go's RFC3339Nano layout)."""
    t = float(t)
    sec = int(t // 1)
    frac = int(round((t - sec) * 1e6)) * 1000          # a float clock resolves microseconds at best
    if frac >= 1_000_000_000:
        sec, frac = sec + 1, 0
    st = time.gmtime(sec)
    out = "%04d-%02d-%02dT%02d:%02d:%02d" % (st.tm_year, st.tm_mon, st.tm_mday, st.tm_hour, st.tm_min, st.tm_sec)
    if frac:
        out += "." + ("%09d" % frac).rstrip("0")
    return out + "Z"


def parse_config(s):
    """'alias== url ||alias2== url2' -> {alias: url} (main.go:49-74)."""
    out = {}
    if not s:
        return out
    for part in s.split(CONFIG_SEPARATOR):
        if not part.strip():
            continue
        if KV_SEPARATOR not in part:
            raise ValueError(f"config entry without '{KV_SEPARATOR}': {part!r}")
        k, v = part.split(KV_SEPARATOR, 1)
        out[k.strip()] = v.strip()
    return out


def format_config(d):
    """The service-side formatting, for tests and fakes (main.go:49-74)."""
    return CONFIG_SEPARATOR.join(f"{k}{KV_SEPARATOR}{v}" for k, v in d.items())


_PAIRWISE = {"ALL": A.FM_PW_ALL, "ANY": A.FM_PW_ANY, "MANN_WHITE": A.FM_PW_MANN_WHITE,
             "WILCOXON": A.FM_PW_WILCOXON, "KRUSKAL": A.FM_PW_KRUSKAL, "NONE": A.FM_PW_NONE,
             "FRIEDMAN": A.FM_PW_FRIEDMAN}   # rollover (pod-group) jobs only
_ALGORITHMS = {"moving_average_all": A.FM_MA_ALL, "moving_average": A.FM_MA_WINDOW,
               "exponential_smoothing": A.FM_MODEL_EWMA,
               "double_exponential_smoothing": A.FM_MODEL_DOUBLE_EXP,
               "holt_winters": A.FM_MODEL_HOLT_WINTERS}


def params_from_env(env):
    """Env knobs of the brain deployment -> (GlobalParams, TypeParams table, {type name: id}).

    Names and the indexed per-metric-type override scheme
    (metric_type{i}/threshold{i}/bound{i}/min_lower_bound{i},
    metric_type_threshold_count) follow foremast-brain.yaml:24-79 and
    foremast-brain/README.md:22-38.  Index 0 of the table is the default row.
    """
    g, _ = A.default_params()
    algo = env.get("ML_ALGORITHM", "moving_average_all")
    if algo not in _ALGORITHMS:
        raise ValueError(f"unsupported ML_ALGORITHM {algo!r}")
    g.ma_mode = _ALGORITHMS[algo]
    g.ma_window = int(env.get("ML_MOVING_AVERAGE_WINDOW", "0"))
    g.season_len = int(env.get("ML_SEASON_LENGTH", "1440" if g.ma_mode == A.FM_MODEL_HOLT_WINTERS else "0"))
    for name, field in (("ML_SMOOTHING_ALPHA", "smooth_alpha"), ("ML_SMOOTHING_BETA", "smooth_beta"),
                        ("ML_SMOOTHING_GAMMA", "smooth_gamma")):
        if name in env:
            setattr(g, field, float(env[name]))
    g.min_hist_points = int(env.get("MIN_HISTORICAL_DATA_POINT_TO_MEASURE", "10"))
    pw = env.get("ML_PAIRWISE_ALGORITHM", "ANY").upper()
    if pw not in _PAIRWISE:
        raise ValueError(f"unsupported ML_PAIRWISE_ALGORITHM {pw!r}")
    g.pairwise_mode = _PAIRWISE[pw]
    g.alpha = float(env.get("ML_PAIRWISE_THRESHOLD", "0.05"))
    g.min_mw = int(env.get("MIN_MANN_WHITE_DATA_POINTS", "20"))
    g.min_wx = int(env.get("MIN_WILCOXON_DATA_POINTS", "20"))
    g.min_kw = int(env.get("MIN_KRUSKAL_DATA_POINTS", "5"))
    k0 = float(env.get("ML_THRESHOLD", env.get("threshold", "2.0")))
    b0 = int(env.get("ML_BOUND", env.get("bound", "1")))
    lo0 = float(env.get("min_lower_bound", "0"))
    n = int(env.get("metric_type_threshold_count", "0"))
    if n + 1 > A.FM_MAX_TYPES:
        raise ValueError("too many metric types")
    table = (A.TypeParams * (n + 1))()
    names = {"default": 0}

    def fill(row, k, b, lo):
        if b not in (1, 2, 3):
            raise ValueError(f"bound must be 1 (upper), 2 (lower) or 3 (both), got {b}")
        row.k, row.k_low, row.min_lower_bound, row.bound_mask = k, 0.5 * k, lo, b

    fill(table[0], k0, b0, lo0)
    for i in range(n):
        name = env.get(f"metric_type{i}")
        if not name:
            raise ValueError(f"metric_type{i} missing")
        names[name] = i + 1
        fill(table[i + 1], float(env.get(f"threshold{i}", k0)), int(env.get(f"bound{i}", b0)),
             float(env.get(f"min_lower_bound{i}", lo0)))
    return g, table, names


def metric_type_id(alias, names):
    """Metric alias -> row of the threshold table: exact name, else the longest
    type name contained in the alias (aliases like 'error5xx', 'latency',
    'namespace_app_per_pod:cpu_usage' ...)."""
    if alias in names:
        return names[alias]
    a = alias.lower()
    best, bl = 0, 0
    for n, i in names.items():
        if i and n.lower() in a and len(n) > bl:
            best, bl = i, len(n)
    return best


def decode_query_range(payload):
    """Prometheus query_range body -> list of (labels, ts int64[], values float32[]).

    Shape: {"status":"success","data":{"result":[{"metric":{..},"values":[[ts,"v"],..]}]}}
    (foremast-browser/src/App.js:278-292).  NaN / Inf samples (Prometheus gaps,
    divisions by zero in recording rules) are dropped: the ABI wants NaN-free rows.
    """
    if isinstance(payload, (bytes, bytearray)):
        payload = payload.decode("utf-8")
    if isinstance(payload, str):
        payload = json.loads(payload)
    if isinstance(payload, str):          # the service's proxy double-encodes (App.js:281)
        payload = json.loads(payload)
    if payload.get("status", "success") != "success":
        raise ValueError(f"query_range failed: {payload.get('error', payload.get('status'))}")
    out = []
    for r in payload.get("data", {}).get("result", []):
        vals = r.get("values", [])
        ts = np.fromiter((int(float(v[0])) for v in vals), dtype=np.int64, count=len(vals))
        x = np.fromiter((float(v[1]) for v in vals), dtype=np.float64, count=len(vals))
        keep = np.isfinite(x)
        out.append((r.get("metric", {}), ts[keep], x[keep].astype(np.float32)))
    return out


def decode_query_range_fast(body):
    """Same result as decode_query_range for a raw response body (bytes), through the C scanner
    fm_decode_query_range of libforemast_cuda.so: no Python objects per sample.  The labels are not
    decoded ({} in their place): merge_series and the scoring path do not use them."""
    from . import _ffi
    if isinstance(body, str):
        body = body.encode("utf-8")
    body = bytes(body)
    cap = len(body) // 8 + 1                      # a sample takes at least [0,"0"], = 8 bytes
    cap_series = body.count(b'"value') + 2
    ts = np.empty(cap, np.int64)
    x = np.empty(cap, np.float32)
    start = np.zeros(cap_series, np.uint32)
    n_series = C.c_uint32(0)
    n = _ffi.lib().fm_decode_query_range(body, len(body), ts.ctypes.data, x.ctypes.data, cap,
                                         start.ctypes.data, cap_series, C.byref(n_series))
    if n < 0:
        msg = _ffi.lib().fm_last_error(None)
        raise ValueError((msg or b"fm_decode_query_range failed").decode())
    return [({}, ts[start[i]:start[i + 1]].copy(), x[start[i]:start[i + 1]].copy())
            for i in range(n_series.value)]


def decode_response(payload):
    """What the job loop calls on a fetched body: the C scanner for raw JSON object bodies, the
    json.loads decoder for everything else (already parsed dicts, the service proxy's
    double-encoded strings, App.js:281)."""
    if isinstance(payload, (bytes, bytearray)) and bytes(payload[:64]).lstrip()[:1] == b"{":
        return decode_query_range_fast(payload)
    return decode_query_range(payload)


STEP_SECONDS = 60        # metricsquery.go:43: every query is made with step 60 on a step-aligned range


def align_to_grid(ts, x, step=STEP_SECONDS):
    """Samples of ONE result series -> (grid index int64[], value float32[]), ascending, one sample
    per grid slot.  The reference asks Prometheus for start/end that are multiples of the step
    (metricsquery.go:43-46, 52-78), so samples sit on the grid already; anything else (a proxy that
    shifts timestamps, a sub-step scrape) is snapped to the nearest slot and the last sample of a
    slot wins.  Gaps stay gaps: the rows handed to the library are compacted, never interpolated."""
    if ts.size == 0:
        return np.zeros(0, np.int64), np.zeros(0, np.float32)
    g = (ts.astype(np.int64) + step // 2) // step
    o = np.argsort(g, kind="stable")
    g, x = g[o], x[o]
    last = np.ones(g.size, bool)
    last[:-1] = g[1:] != g[:-1]
    return g[last], x[last]


def merge_series(results, step=STEP_SECONDS):
    """All result series of one query -> ONE row per metric alias on the step grid.

    Declared rule: a query that matches several pods
    (`namespace_pod:...{pod=~"a|b"}`, metricsquery.go:52-64) is judged as one sample set whose value
    at a grid point is the MEAN over the pods that reported there — the same aggregation the
    `namespace_app_per_pod:*` recording rules apply to the historical series the bounds come from
    (deploy/foremast/2_barrelman/metrics-rules-default.yaml:12-84), so current, baseline and
    history stay comparable.  Grid points where no pod reported are dropped (gap compaction).
    Returns (timestamps int64[] = grid index * step, values float32[])."""
    rows = [align_to_grid(r[1], r[2], step) for r in results]
    rows = [r for r in rows if r[0].size]
    if not rows:
        return np.zeros(0, np.int64), np.zeros(0, np.float32)
    if len(rows) == 1:
        return rows[0][0] * step, rows[0][1]
    g = np.concatenate([r[0] for r in rows])
    x = np.concatenate([r[1] for r in rows]).astype(np.float64)
    ug, inv = np.unique(g, return_inverse=True)
    tot = np.bincount(inv, weights=x, minlength=ug.size)
    cnt = np.bincount(inv, minlength=ug.size)
    return ug * step, (tot / cnt).astype(np.float32)


def http_fetch(url, timeout=30):
    with urllib.request.urlopen(url, timeout=timeout) as r:
        return r.read()


def _rows(ctx, role, mats, width):
    """Ragged rows -> (matrix [S x pitch] fp32, lens uint32).  With a real engine.Context the matrix
    is a view of PINNED host memory owned by the ctx (fm_alloc_pinned, reused from poll to poll), so
    fm_score_batch copies straight out of it; the stand-in contexts of the CPU tests get numpy."""
    shape = (len(mats), max(4, A.row_pitch(width)))
    alloc = getattr(ctx, "pinned_rows", None)
    m = alloc(role, shape) if alloc else np.zeros(shape, np.float32)
    lens = np.zeros(len(mats), np.uint32)
    for i, x in enumerate(mats):
        n = min(x.size, width)
        m[i, :n] = x[-n:] if n else x[:0]     # keep the most recent points when over-long
        lens[i] = n
    return m, lens


def _failed(reason):
    return {"status": ST_PRE_FAILED, "reason": reason, "anomalyInfo": "", "processingContent": "", "metrics": {},
            "gauges": []}


def row_id(doc_id, alias):
    """64-bit id of one metric of one job for the device-resident row cache (FNV-1a)."""
    h = 0xCBF29CE484222325
    for b in f"{doc_id}|{alias}".encode("utf-8"):
        h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return h


def fetch_documents(docs, fetch=http_fetch, resident=None):
    """Preprocessing (the state diagram's preprocess_* half): parse the three config strings of every
    document, fetch and decode the series.  Returns (series, failed): series = list of
    (doc index, alias, ts_cur, cur, base, hist); failed = {doc id: result} for documents whose configs
    cannot be parsed or fetched ("preprocess_failed" with the reason; the others go on).
    resident(ids) -> bool array (engine.Context.cache_query): the baseline / history of a metric
    whose rows are still in HBM from the last poll are not fetched again (empty placeholders)."""
    series, failed = [], {}
    empty = np.zeros(0, np.float32)
    for di, doc in enumerate(docs):
        try:
            cur_cfg = parse_config(doc.get("currentConfig", ""))
            if not cur_cfg:
                raise ValueError("currentConfig is empty")
            strat = (doc.get("strategy") or "").lower()
            base_cfg = parse_config(doc.get("baselineConfig", "")) if strat in CANARY_STRATEGIES else {}
            # Document says historicalConfig, DocumentResponse historicConfig (models.go:92,130)
            hist_cfg = parse_config(doc.get("historicalConfig") or doc.get("historicConfig") or "")
            mine = []
            here = None
            if resident is not None:
                here = resident(np.array([row_id(doc["id"], a) for a in cur_cfg], np.uint64))
            for k, (alias, url) in enumerate(cur_cfg.items()):
                ts, cur = merge_series(decode_response(fetch(url)))
                skip = here is not None and bool(here[k])
                base = (empty if skip else merge_series(decode_response(fetch(base_cfg[alias])))[1]) if alias in base_cfg else None
                hist = (empty if skip else merge_series(decode_response(fetch(hist_cfg[alias])))[1]) if alias in hist_cfg else None
                mine.append((di, alias, ts, cur, base, hist))
            series += mine
        except Exception as e:  # host I/O problem: the document fails, the batch goes on
            failed[doc["id"]] = _failed(f"{type(e).__name__}: {e}")
    return series, failed


def score_series(ctx, docs, series, failed=None, type_names=None, max_hist=10080, max_base=None, cached=False):
    """Postprocessing: every fetched metric of every document in ONE coalesced batch per strategy
    kind, then per document the A.9 roll-up over its metrics, reason and anomalyInfo.
    cached: score through the row cache (fm_score_batch_cached) — rows are then packed to the fixed
    widths max_base / max_hist so that the resident layout never changes between polls."""
    type_names = type_names or dict(A.METRIC_TYPES)
    results = dict(failed or {})
    out = {}
    # canary series (with a baseline) and historical-only series are two batches
    for canary in (True, False):
        idx = [i for i, s in enumerate(series) if (s[4] is not None) == canary]
        if not idx:
            continue
        tag = "c" if canary else "h"
        cur_w = max(1, max(series[i][3].size for i in idx))
        if cur_w > A.FM_MAX_CUR:
            cur_w = A.FM_MAX_CUR
        cur, cl = _rows(ctx, tag + "cur", [series[i][3] for i in idx], cur_w)
        hist = hl = base = bl = None
        if any(series[i][5] is not None for i in idx):
            hw = min(max_hist, max(1, max((series[i][5].size if series[i][5] is not None else 0) for i in idx)))
            if cached:
                hw = max_hist
            hist, hl = _rows(ctx, tag + "hist",
                             [series[i][5] if series[i][5] is not None else np.zeros(0, np.float32) for i in idx], hw)
        if canary:
            bw = max(1, max(series[i][4].size for i in idx))
            bw = min(bw, max_base or (A.FM_MAX_RANK_N - cur_w))
            if cached and max_base:
                bw = max_base
            base, bl = _rows(ctx, tag + "base", [series[i][4] for i in idx], bw)
        mt = np.array([metric_type_id(series[i][1], type_names) for i in idx], np.uint8)
        extra = {}
        if cached:
            extra["row_ids"] = np.array([row_id(docs[series[i][0]]["id"], series[i][1]) for i in idx], np.uint64)
        r = ctx.score(cur, base, hist, strategy=A.FM_STRAT_CANARY if canary else A.FM_STRAT_HISTORICAL,
                      cur_len=cl, base_len=bl, hist_len=hl, metric_type=mt, **extra)
        for j, i in enumerate(idx):
            out[i] = r["rec"][j].copy()
    # per document: A.9 roll-up over its metrics, reason, anomalyInfo (models.go:60-63;
    # values are [t0, v0, t1, v1, ...] as decoded by Barrelman.go:593-620)
    for di, doc in enumerate(docs):
        if doc["id"] in results:
            continue
        mine = [(i, s) for i, s in enumerate(series) if s[0] == di]
        verdicts = [int(out[i]["verdict"]) for i, _ in mine]
        if any(v == A.FM_UNHEALTHY for v in verdicts):
            verdict = A.FM_UNHEALTHY
        elif verdicts and all(v == A.FM_UNKNOWN for v in verdicts):
            verdict = A.FM_UNKNOWN
        elif not verdicts:
            verdict = A.FM_UNKNOWN
        else:
            verdict = A.FM_HEALTHY
        anomaly, reasons, metrics, gauge_rows = {}, [], {}, []
        for i, (_, alias, ts, cur, _b, _h) in mine:
            rec = out[i]
            metrics[alias] = {k: (float(rec[k]) if rec[k].dtype.kind == "f" else int(rec[k])) for k in rec.dtype.names}
            anom_ts = []
            if rec["verdict"] == A.FM_UNHEALTHY:
                n = min(cur.size, A.FM_MAX_CUR)
                off = cur.size - n
                x = cur[off:].astype(np.float64)
                tid = metric_type_id(alias, type_names)
                bm = int(ctx.table[tid if tid < len(ctx.table) else 0].bound_mask)
                flag = ((bm & 1) != 0) & (x > rec["upper"]) | ((bm & 2) != 0) & (x < rec["lower"])
                vals = []
                for k in np.nonzero(flag)[0]:
                    vals += [int(ts[off + k]), float(x[k])]
                    anom_ts.append(int(ts[off + k]))
                anomaly[alias] = {"tags": "", "values": vals}
                reasons.append(f"{alias}: {int(rec['n_anom'])} anomalous point(s) outside "
                               f"[{rec['lower']:.6g}, {rec['upper']:.6g}]")
            elif rec["verdict"] == A.FM_UNKNOWN:
                reasons.append(f"{alias}: " + ("no current metric" if rec["status"] & A.FM_ST_NO_CURRENT
                                               else "not enough historical data"))
            if not (rec["status"] & A.FM_ST_HIST_INSUFFICIENT):
                gauge_rows.append((alias, rec, {"app": doc.get("appName", "")} if doc.get("appName") else {}, anom_ts))
        results[doc["id"]] = {"status": VERDICT_STATUS[verdict], "reason": "; ".join(reasons),
                              "anomalyInfo": json.dumps(anomaly) if anomaly else "",
                              "processingContent": "", "metrics": metrics, "gauges": gauge_rows}
    return results


def score_documents(ctx, docs, fetch=http_fetch, type_names=None, max_hist=10080, max_base=None):
    """fetch_documents + score_series: one poll of documents -> {id: {"status", "reason",
    "anomalyInfo", "processingContent", "metrics", "gauges"}}."""
    series, failed = fetch_documents(docs, fetch)
    return score_series(ctx, docs, series, failed, type_names, max_hist, max_base)


class MemoryStore:
    """Stand-in for the Elasticsearch 'documents' index (elasticsearchstore.go:16-19)
    with the two operations the brain needs: search by status, update."""

    def __init__(self, docs=()):
        self.docs = {d["id"]: dict(d) for d in docs}

    def by_status(self, statuses, limit=1000):
        return [dict(d) for d in self.docs.values() if d.get("status") in statuses][:limit]

    def update(self, doc_id, expect_modified_at=None, **fields):
        d = self.docs[doc_id]
        if expect_modified_at is not None and d.get("modified_at") != expect_modified_at:
            return False                                  # somebody else claimed it
        d.update(fields)
        return True


class ElasticsearchStore:
    """The same two operations against the 'documents' index of the Elasticsearch instance
    foremast-service writes to (elasticsearchstore.go:16-19: index "documents", type "document";
    search by status as in ByStatus, elasticsearchstore.go:124-149).

    by_status   POST /documents/_search  {"query": {"terms": {"status": [...]}}, "size": limit}
    update      POST /documents/document/<id>/_update with a script that applies the fields only
                while modified_at still has the value the caller saw (the claim / takeover race
                between brain workers); Elasticsearch answers "noop" when somebody else won.
                modified_at travels as the RFC3339 string the service wrote / this worker writes,
                and is compared as that string.

    `transport(method, url, body_dict) -> dict` defaults to urllib; tests inject a fake."""

    def __init__(self, base_url, index="documents", doc_type="document", transport=None):
        self.base = base_url.rstrip("/")
        self.index, self.doc_type = index, doc_type
        self.transport = transport or self._urllib

    @staticmethod
    def _urllib(method, url, body):
        req = urllib.request.Request(url, data=json.dumps(body).encode(), method=method,
                                     headers={"Content-Type": "application/json"})
        with urllib.request.urlopen(req, timeout=30) as r:
            return json.loads(r.read().decode())

    def by_status(self, statuses, limit=1000):
        body = {"query": {"terms": {"status": sorted(statuses)}}, "size": int(limit)}
        res = self.transport("POST", f"{self.base}/{self.index}/_search", body)
        out = []
        for hit in res.get("hits", {}).get("hits", []):
            d = dict(hit.get("_source", {}))
            d.setdefault("id", hit.get("_id"))
            out.append(d)
        return out

    def update(self, doc_id, expect_modified_at=None, **fields):
        url = f"{self.base}/{self.index}/{self.doc_type}/{doc_id}/_update"
        if expect_modified_at is None and "modified_at" not in fields:
            res = self.transport("POST", url, {"doc": fields})
        else:
            script = ("if (params.expect == null || ctx._source.modified_at == params.expect) "
                      "{ ctx._source.putAll(params.doc) } else { ctx.op = 'noop' }")
            res = self.transport("POST", url, {"script": {"lang": "painless", "source": script,
                                                          "params": {"expect": expect_modified_at, "doc": fields}}})
        return res.get("result") != "noop"


class BrainWorker:
    """Poll / claim / score / write-back loop (design.md:37-43; state diagram
    .gitbook/assets/foremastrequeststatediagram.png; status names converter.go:13-26).

      initial --claim--> preprocess_inprogress --fetched--> preprocess_completed
              --> postprocess_inprogress --scored--> completed_health | completed_unhealth |
              completed_unknown        (fetch problems: preprocess_failed)

    - every transition is an optimistic update on `modified_at`: the stamp this worker wrote last
      must still be there, else another worker has taken the job over and this one drops it.
      Stamps are RFC3339 UTC strings, the way the service's time.Time fields are stored
      (models.go:86-89); times are never written as numbers;
    - a document that sits in an *_inprogress / preprocess_completed state for longer than
      MAX_STUCK_IN_SECONDS (foremast-brain.yaml:80-81) is taken over;
    - fail fast: unhealthy completes the job at once; a healthy verdict before endTime puts the job
      back to "initial" so it is re-judged on the next poll ("continuous to monitor ... until
      endTime is reached", design.md:43);
    - all claimed documents of one poll are scored as one batch; a document whose write-back fails
      does not stop the others;
    - `gauges` (a gauges.GaugeRegistry, optional) receives the bounds and anomaly points of every
      scored metric: what :8000/metrics serves (foremast-brain.yaml:87-122).
    """

    IN_FLIGHT = (ST_PRE_INPROGRESS, ST_PRE_COMPLETED, ST_POST_INPROGRESS)

    def __init__(self, ctx, store, fetch=http_fetch, env=None, now=time.time, gauges=None):
        self.ctx, self.store, self.fetch, self.now, self.gauges = ctx, store, fetch, now, gauges
        env = env or {}
        self.max_stuck = float(env.get("MAX_STUCK_IN_SECONDS", "90"))
        g, table, names = params_from_env(env)
        self.type_names = names
        ctx.set_params(g, table)
        self.errors = []          # (doc id, message) of write-backs that failed
        # MAX_CACHE_SIZE ("max cached model size", foremast-brain/README.md:30): that many metric rows
        # (baseline + history) stay in HBM between polls; rows are packed to fixed widths then
        self.base_width = int(env.get("ML_BASELINE_POINTS", "1440"))
        self.hist_width = int(env.get("ML_HISTORICAL_POINTS", "10080"))
        self.cache_rows = int(env.get("MAX_CACHE_SIZE", "0"))
        self.cached = self.cache_rows > 0 and hasattr(ctx, "cache_configure")
        if self.cached:
            ctx.cache_configure(self.cache_rows * 4 * (A.row_pitch(self.base_width) + A.row_pitch(self.hist_width)))

    def _stamp(self):
        return format_go_time(self.now())

    def _move(self, d, status, **fields):
        """One optimistic transition; d carries the stamp it was last seen / written with."""
        stamp = self._stamp()
        if stamp == d.get("modified_at"):        # two transitions inside one clock tick: stay distinguishable
            stamp = format_go_time(parse_go_time(stamp) + 1e-6)
        try:
            ok = self.store.update(d["id"], expect_modified_at=d.get("modified_at"), status=status,
                                   modified_at=stamp, **fields)
        except Exception as e:                    # store trouble on one document does not strand the batch
            self.errors.append((d["id"], f"{type(e).__name__}: {e}"))
            return False
        if ok:
            d["status"], d["modified_at"] = status, stamp
        return ok

    def claim(self):
        now = self.now()
        claimed = []
        for d in self.store.by_status({ST_INITIAL, *self.IN_FLIGHT}):
            if d["status"] != ST_INITIAL:
                try:
                    seen = parse_go_time(d.get("modified_at"))
                except ValueError:
                    seen = None
                if seen is not None and now - seen <= self.max_stuck:
                    continue                               # another worker is on it
            if self._move(d, ST_PRE_INPROGRESS):
                claimed.append(d)
        return claimed

    def poll_once(self):
        docs = self.claim()
        if not docs:
            return {}
        series, failed = fetch_documents(docs, self.fetch, self.ctx.cache_query if self.cached else None)
        live = []
        for d in docs:
            if d["id"] in failed:
                r = failed[d["id"]]
                self._move(d, ST_PRE_FAILED, reason=r["reason"])
                r["written_status"] = d["status"]
            elif self._move(d, ST_PRE_COMPLETED) and self._move(d, ST_POST_INPROGRESS):
                live.append(d)
        keep = {id(d) for d in live}
        index = {i: j for j, i in enumerate(i for i, d in enumerate(docs) if id(d) in keep)}
        series = [(index[s[0]],) + tuple(s[1:]) for s in series if s[0] in index]
        res = score_series(self.ctx, live, series, None, self.type_names, max_hist=self.hist_width,
                           max_base=self.base_width if self.cached else None, cached=self.cached)
        now = self.now()
        for d in live:
            r = res[d["id"]]
            status = r["status"]
            try:
                end = parse_go_time(d.get("endTime"))
            except ValueError:
                end = None
            if status == ST_HEALTH and end is not None and now < end:
                status = ST_INITIAL                       # keep watching until endTime
            # the claim stamp must still be ours: a worker that took the job over meanwhile wins
            if self._move(d, status, reason=r["reason"], anomalyInfo=r["anomalyInfo"]):
                r["written_status"] = status
                if self.gauges is not None:
                    self.gauges.update(d["id"], r["gauges"], final=status != ST_INITIAL)
            else:
                r["written_status"] = None
        res.update(failed)
        return res
