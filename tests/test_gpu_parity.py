"""GPU parity: libforemast_cuda.so (through the C ABI) against the CPU oracle on
identical inputs.  Sizes are chosen so the oracle finishes in seconds; the
full-size BASELINE shapes are covered by tests/test_gpu_fullsize.py.
"""
import copy
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
from foremast_1_b200 import _abi as A
from foremast_1_b200.engine import Context, synth_desc, synth_host
from parity import assert_records_match, bound_masks

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["auto", "auto-tiny", "auto-notie", "auto-alltie", "auto-1stage", "sort", "generic", "elect", "elect-1stage"])
def ctx(request):
    """auto: the counting canary kernel (fm_canary_count.cu) wherever it applies; auto-tiny: the same
    with an election table of 64 slots, so that every series goes through all the overflow rounds and
    the direct count behind them; auto-notie: without the class path for heavily tied series (they then
    pass the singleton filter with every key a suspect); auto-alltie: every series tries the class path
    first (continuous ones fail it and fall back to the search path); auto-1stage: its free-running launch shape (no barrier per series); sort: the register-sort canary kernel (fm_canary_sort.cu); elect:
    the CTA-per-series election kernel (fm_canary_elect.cu); generic: the shared-memory sort kernel
    for everything (the FOREMAST_CUDA_* switches are read when the ctx is created)."""
    import os
    name = request.param.split("-")
    os.environ["FOREMAST_CUDA_PATH"] = name[0]
    os.environ["FOREMAST_CUDA_STAGES"] = "1" if request.param.endswith("1stage") else "0"
    if request.param.endswith("tiny"):
        os.environ["FOREMAST_CUDA_SLOTS"] = "64"
        os.environ["FOREMAST_CUDA_STAGES"] = "0"
    if request.param.endswith("notie"):
        os.environ["FOREMAST_CUDA_TIE_RUNS"] = "0"
    if request.param.endswith("alltie"):
        os.environ["FOREMAST_CUDA_TIE_RUNS"] = "65"
    c = Context(0)
    for k in ("FOREMAST_CUDA_PATH", "FOREMAST_CUDA_STAGES", "FOREMAST_CUDA_SLOTS", "FOREMAST_CUDA_TIE_RUNS"):
        os.environ.pop(k, None)
    yield c
    c.close()


def params(**kw):
    g, table = A.default_params()
    for k, v in kw.items():
        setattr(g, k, v)
    return g, table


def both(ctx, g, table, cur, base=None, hist=None, **kw):
    ctx.set_params(g, table)
    out = ctx.score(cur, base, hist, want_aux=True, **kw)
    ora = O.score_batch(g, table, out["batch"], n_threads=8)
    b = out["batch"]
    S = int(b.n_series)
    curm = np.ctypeslib.as_array(C.cast(b.cur, C.POINTER(C.c_float)), shape=(S, b.cur_pitch)) if S else np.zeros((0, 4), np.float32)
    cl = kw.get("cur_len")
    listed = assert_records_match(out["rec"], ora["rec"], curm[:, : b.n_cur], cl,
                                  bound_masks(table, kw.get("metric_type"), S),
                                  out["aux"], ora["aux"], alpha=g.alpha)
    if out["api"] is not None:
        assert (out["api"] == ora["api"]).all()
    if out["svc"] is not None:
        assert (out["svc"] == ora["svc"]).all()
    return out, ora, listed


def test_config1_single_job(ctx):
    """BASELINE configs[0]: 1 service x 4 metrics, 60 vs 1440, MA + Mann-Whitney."""
    d = synth_desc(1, 4, 60, 1440, 0)
    cur, base, _, mt = synth_host(d)
    g, t = params(pairwise_mode=A.FM_PW_MANN_WHITE)
    out, ora, _ = both(ctx, g, t, cur, base, metric_type=mt,
                       api_id=np.zeros(4, np.uint32), api_service=np.zeros(1, np.uint32),
                       n_apis=1, n_services=1)
    assert (out["rec"]["status"] & A.FM_ST_MW_RUN).all()


@pytest.mark.parametrize("mode", [A.FM_PW_ANY, A.FM_PW_ALL, A.FM_PW_MANN_WHITE, A.FM_PW_WILCOXON,
                                  A.FM_PW_KRUSKAL, A.FM_PW_NONE])
def test_config3_shape_modes(ctx, mode):
    d = synth_desc(3, 3000, 60, 1440, 0)
    cur, base, _, mt = synth_host(d)
    g, t = params(pairwise_mode=mode)
    out, ora, _ = both(ctx, g, t, cur, base, metric_type=mt)
    # the synthetic mix must exercise both verdicts and the pairwise decision
    if mode == A.FM_PW_ANY:
        v = out["rec"]["verdict"]
        assert (v == A.FM_UNHEALTHY).any() and (v == A.FM_HEALTHY).any()
        assert (out["rec"]["status"] & A.FM_ST_DIFFER).any()


def test_config2_shape_separate_hist(ctx):
    d = synth_desc(2, 300, 60, 1440, 10080)
    cur, base, hist, mt = synth_host(d)
    g, t = params(pairwise_mode=A.FM_PW_MANN_WHITE)
    both(ctx, g, t, cur, base, hist, metric_type=mt)


def test_historical_strategy(ctx):
    d = synth_desc(5, 2000, 60, 0, 480)
    cur, _, hist, mt = synth_host(d)
    g, t = params()
    out, _, _ = both(ctx, g, t, cur, None, hist, strategy=A.FM_STRAT_HISTORICAL, metric_type=mt,
                     record_kind=A.FM_REC_COMPACT)
    out2, _, _ = both(ctx, g, t, cur, None, hist, strategy=A.FM_STRAT_HISTORICAL, metric_type=mt)
    assert (out["rec"]["verdict"] == out2["rec"]["verdict"]).all()


def test_ragged_and_empty_rows(ctx):
    rng = np.random.default_rng(7)
    S = 600
    d = synth_desc(3, S, 60, 500, 0)
    cur, base, _, mt = synth_host(d)
    cl = rng.integers(0, 61, S).astype(np.uint32)
    bl = rng.integers(0, 501, S).astype(np.uint32)
    cl[:8] = [0, 1, 4, 5, 19, 20, 21, 60]
    bl[:8] = [500, 0, 3, 9, 10, 19, 20, 500]
    g, t = params()
    out, _, _ = both(ctx, g, t, cur, base, cur_len=cl, base_len=bl, metric_type=mt)
    r = out["rec"]
    assert r["verdict"][0] == A.FM_UNKNOWN and (r["status"][0] & A.FM_ST_NO_CURRENT)
    assert r["verdict"][1] == A.FM_UNKNOWN and (r["status"][1] & A.FM_ST_HIST_INSUFFICIENT)


def test_heavy_ties_and_degenerate(ctx):
    rng = np.random.default_rng(11)
    S = 400
    cur = rng.integers(0, 4, (S, 60)).astype(np.float32)
    base = rng.integers(0, 4, (S, 1440)).astype(np.float32)
    cur[0] = 1.0; base[0] = 1.0            # everything tied: degenerate, p = 1
    cur[1] = 0.0; base[1] = 0.0
    cur[2, ::2] = -0.0; cur[2, 1::2] = 0.0; base[2] = 0.0; base[2, ::3] = -0.0   # -0 == +0
    cur[3] = base[3, :60]                  # all differences zero: Wilcoxon skipped
    g, t = params()
    out, _, _ = both(ctx, g, t, cur, base)
    r = out["rec"]
    assert r["p_u"][0] == 1.0 and (r["status"][0] & A.FM_ST_DEGENERATE)
    assert r["p_u"][2] == 1.0
    assert not (r["status"][3] & A.FM_ST_WX_RUN)


def test_many_ties_medium_cardinality(ctx):
    rng = np.random.default_rng(5)
    S = 300
    cur = (rng.integers(0, 300, (S, 64)) / 7.0).astype(np.float32)
    base = (rng.integers(0, 300, (S, 1000)) / 7.0).astype(np.float32)
    g, t = params()
    both(ctx, g, t, cur, base)


@pytest.mark.parametrize("nc,nb", [(1, 1), (5, 7), (20, 20), (33, 100), (64, 64), (65, 2047),
                                   (128, 2048), (300, 3000), (1000, 1000), (60, 10080)])
def test_shapes(ctx, nc, nb):
    rng = np.random.default_rng(nc * 7919 + nb)
    S = 64
    cur = rng.normal(5, 1, (S, nc)).astype(np.float32)
    base = rng.normal(5, 1, (S, nb)).astype(np.float32)
    base[: S // 2] = np.round(base[: S // 2] * 4) / 4   # half the series tie heavily
    cur[: S // 2] = np.round(cur[: S // 2] * 4) / 4
    g, t = params(min_hist_points=1)
    both(ctx, g, t, cur, base, metric_type=rng.integers(0, 6, S).astype(np.uint8))


@pytest.mark.parametrize("nb", [383, 384, 385, 512, 513, 768, 769, 1024, 1025, 1439, 1536, 1537, 2047, 2048])
def test_sort_kernel_tiers(ctx, nb):
    """Every (keys per lane, pruned registers) instantiation of the register-sort kernel and both
    sides of each boundary, with ragged lengths, long tie runs that cross lanes, values equal to
    the row maximum (next to the +inf padding) and current points tied with baseline points."""
    rng = np.random.default_rng(nb)
    S = 96
    cur = rng.normal(100, 3, (S, 60)).astype(np.float32)
    base = rng.normal(100, 3, (S, nb)).astype(np.float32)
    base[:32] = np.round(base[:32])                      # ~20 distinct values: runs of 50-200 keys
    cur[:32] = np.round(cur[:32])
    base[32:48, : nb // 2] = base[32:48, :1]             # one value fills half of the row
    cur[32:48, :30] = base[32:48, :1]
    base[48:56] = np.maximum(base[48:56], 101.0)         # a long run AT the minimum
    base[56:64] = np.minimum(base[56:64], 99.0)          # and at the maximum (beside the padding)
    cur[56:64, ::2] = 99.0
    bl = np.full(S, nb, np.uint32)
    bl[64:] = rng.integers(1, nb + 1, S - 64)
    bl[64:68] = [1, 2, min(nb, 33), nb - 1]
    g, t = params(min_hist_points=1)
    both(ctx, g, t, cur, base, base_len=bl, metric_type=rng.integers(0, 6, S).astype(np.uint8))


@pytest.mark.parametrize("nb", [2049, 2176, 4095, 4096, 8000, 10080, 16320])
def test_long_baselines(ctx, nb):
    """Baselines above 2048 points (BASELINE configs[1] read literally: 60 points against the 7-day
    window) go to the team form of the counting kernel: both sides of its row boundaries, heavy ties
    (runs of thousands of keys, p^2 sums beyond 32 bits), ragged lengths down to one point."""
    rng = np.random.default_rng(nb)
    S = 40
    cur = rng.normal(100, 3, (S, 60)).astype(np.float32)
    base = rng.normal(100, 3, (S, nb)).astype(np.float32)
    base[:10] = np.round(base[:10])
    cur[:10] = np.round(cur[:10])
    base[10:14] = 7.0                                       # one class of nb keys
    cur[10:12] = 7.0
    base[14:18, : nb // 2] = base[14:18, :1]
    bl = np.full(S, nb, np.uint32)
    bl[24:] = rng.integers(1, nb + 1, S - 24)
    bl[24:28] = [1, 20, 2049, nb - 1]
    g, t = params(min_hist_points=1)
    both(ctx, g, t, cur, base, base_len=bl, metric_type=rng.integers(0, 6, S).astype(np.uint8))


def test_param_variants(ctx):
    d = synth_desc(3, 800, 60, 1440, 0)
    cur, base, _, mt = synth_host(d)
    for kw in (dict(ddof=1), dict(pair_rule=A.FM_PAIR_TAIL), dict(cc_mw=0, cc_wx=1),
               dict(ma_mode=A.FM_MA_WINDOW, ma_window=120), dict(min_anomaly_points=3),
               dict(alpha=0.5), dict(min_mw=61), dict(ma_mode=A.FM_MODEL_EWMA),
               dict(ma_mode=A.FM_MODEL_DOUBLE_EXP)):
        g, t = params(**kw)
        both(ctx, g, t, cur, base, metric_type=mt)
    g, t = params()
    for i in range(len(t)):
        t[i].bound_mask = 1 + (i % 3)
        t[i].min_lower_bound = 0.25 * i
        t[i].k = 1.5 + 0.5 * i
        t[i].k_low = 1.0
    both(ctx, g, t, cur, base, metric_type=mt)


def test_memory_like_cancellation(ctx):
    """mean ~1e9, sigma/mean ~1e-6: the pivoted moments must keep 1e-6 on sigma."""
    rng = np.random.default_rng(3)
    S = 200
    base = (1.0e9 + rng.integers(-2000, 2000, (S, 1440)) * 64.0).astype(np.float32)
    cur = (1.0e9 + rng.integers(-2000, 2000, (S, 60)) * 64.0).astype(np.float32)
    base[0] = 1.0e9                       # zero variance
    g, t = params()
    both(ctx, g, t, cur, base, metric_type=np.full(S, 5, np.uint8))


def test_rollup(ctx):
    S = 4096
    d = synth_desc(5, S, 60, 0, 240)
    cur, _, hist, mt = synth_host(d)
    api = (np.arange(S) // 4).astype(np.uint32)          # 4 metrics per API
    svc = (np.arange(S // 4) // 8).astype(np.uint32)     # 8 APIs per service
    cl = np.full(S, 60, np.uint32)
    cl[:32] = 0                                          # one whole service unknown
    g, t = params()
    out, ora, _ = both(ctx, g, t, cur, None, hist, strategy=A.FM_STRAT_HISTORICAL, cur_len=cl,
                       metric_type=mt, api_id=api, api_service=svc, n_apis=S // 4, n_services=S // 32)
    assert out["svc"]["severity"][0] == A.FM_UNKNOWN
    assert (out["svc"]["n_children"] == 8).all()
    assert (out["svc"]["severity"][1:] != A.FM_UNKNOWN).all()
    assert set(np.unique(out["api"]["severity"])) == {A.FM_HEALTHY, A.FM_UNHEALTHY, A.FM_UNKNOWN}


def test_multi_chunk_pipeline(ctx):
    """More series than one host->device chunk: both staging slots get reused."""
    S = 40000
    d = synth_desc(3, S, 60, 1440, 0)
    cur, base, _, mt = synth_host(d)
    g, t = params(pairwise_mode=A.FM_PW_MANN_WHITE)
    ctx.set_params(g, t)
    out = ctx.score(cur, base, metric_type=mt, want_aux=True)
    # oracle on a strided subset (full 40k would take a while on few cores)
    idx = np.arange(0, S, 37)
    sub = ctx.score(cur[idx], base[idx], metric_type=mt[idx], want_aux=True)
    ora = O.score_batch(g, t, sub["batch"], n_threads=8)
    assert_records_match(sub["rec"], ora["rec"], cur[idx][:, :60], None, bound_masks(t, mt[idx], idx.size),
                         sub["aux"], ora["aux"])
    assert (out["rec"][idx] == sub["rec"]).all()
    assert (out["aux"][idx] == sub["aux"]).all()


def test_groups_kruskal(ctx):
    rng = np.random.default_rng(17)
    V = 500
    K = rng.integers(3, 9, V)
    off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
    G = int(off[-1])
    ids = np.concatenate([(v << 4) + np.arange(K[v]) for v in range(V)]).astype(np.uint64)
    d = synth_desc(4, G, 60, 0, 0, row_ids=ids.ctypes.data)
    cur, _, _, _ = synth_host(d)
    hids = (np.arange(V, dtype=np.uint64) << np.uint64(4))
    dh = synth_desc(4, V, 60, 0, 1440, row_ids=hids.ctypes.data)
    _, _, hist, mt = synth_host(dh)
    # make a few services differ: shift one group
    for v in range(0, V, 9):
        cur[off[v]] += np.float32(1.0 / 60.0) * 3
    cl = np.full(G, 60, np.uint32)
    cl[off[1]] = 3          # a group below MIN_KRUSKAL_DATA_POINTS: test skipped for that service
    cl[off[2]] = 0
    g, t = params(pairwise_mode=A.FM_PW_KRUSKAL)
    ctx.set_params(g, t)
    out = ctx.score_groups(cur, off, hist, cur_len=cl, metric_type=mt)
    ora = O.score_groups(g, t, out["batch"], n_threads=8)
    gmt = np.repeat(mt, K)
    assert_records_match(out["rec"], ora["rec"], cur[:, :60], cl, bound_masks(t, gmt, G))
    assert (out["svc"] == ora["svc"]).all()
    assert (out["rec"]["status"] & A.FM_ST_DIFFER).any()
    assert (out["rec"]["status"][off[1]] & A.FM_ST_PAIRWISE_SKIPPED)


@pytest.mark.parametrize("n_cur", [12, 60, 64, 100])
def test_groups_mixed_shapes(ctx, n_cur):
    """Services of 1..16 groups in one batch (more than 8 groups go through the deferred list to the
    general kernel), ragged group lengths incl. empty ones, continuous and heavily tied values,
    no history for the model (hist = None is legal: bounds are then skipped)."""
    rng = np.random.default_rng(1000 + n_cur)
    V = 300
    K = rng.integers(1, 17, V)
    K[:4] = [1, 2, 8, 9]
    off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
    G = int(off[-1])
    cur = rng.normal(50, 5, (G, n_cur)).astype(np.float32)
    tied = rng.random(V) < 0.5
    for v in np.nonzero(tied)[0]:
        cur[off[v]:off[v + 1]] = rng.poisson(2.0, (K[v], n_cur)).astype(np.float32)
    cur[off[5]:off[6]] = 7.0                           # one value everywhere: degenerate, p = 1
    hist = rng.normal(50, 5, (V, 500)).astype(np.float32)
    hist[tied] = rng.poisson(2.0, (int(tied.sum()), 500)).astype(np.float32)
    cl = rng.integers(0, n_cur + 1, G).astype(np.uint32)
    cl[rng.random(G) < 0.7] = n_cur
    hl = rng.integers(0, 501, V).astype(np.uint32)
    hl[rng.random(V) < 0.8] = 500
    mt = rng.integers(0, 6, V).astype(np.uint8)
    g, t = params(pairwise_mode=A.FM_PW_KRUSKAL)
    ctx.set_params(g, t)
    out = ctx.score_groups(cur, off, hist, cur_len=cl, hist_len=hl, metric_type=mt)
    ora = O.score_groups(g, t, out["batch"], n_threads=8)
    assert_records_match(out["rec"], ora["rec"], cur, cl, bound_masks(t, np.repeat(mt, K), G))
    assert (out["svc"] == ora["svc"]).all()
    assert (out["rec"]["status"] & A.FM_ST_KW_RUN).any()


def test_groups_multi_chunk_pipeline(monkeypatch):
    """fm_score_groups cuts a batch at service boundaries into chunks that are uploaded, scored and
    read back on three streams: with 1 MB chunks a 3000-service batch takes ~20 of them, both buffer
    slots are reused, and the records equal the one-chunk result byte for byte."""
    rng = np.random.default_rng(77)
    V = 3000
    K = rng.integers(1, 13, V)                      # > 8 groups: deferred list inside a chunk
    off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
    G = int(off[-1])
    cur = rng.poisson(3.0, (G, 60)).astype(np.float32) / np.float32(60)
    cur[rng.random(G) < 0.3] = rng.normal(5, 1, 60).astype(np.float32)
    hist = rng.normal(0.05, 0.03, (V, 1440)).astype(np.float32)
    cl = rng.integers(0, 61, G).astype(np.uint32)
    cl[rng.random(G) < 0.8] = 60
    hl = np.full(V, 1440, np.uint32); hl[::17] = 5
    mt = rng.integers(0, 6, V).astype(np.uint8)
    g, t = params(pairwise_mode=A.FM_PW_KRUSKAL)
    with Context(0) as one:
        one.set_params(g, t)
        ref = one.score_groups(cur, off, hist, cur_len=cl, hist_len=hl, metric_type=mt)
    monkeypatch.setenv("FOREMAST_CUDA_CHUNK_MB", "1")
    with Context(0) as many:
        many.set_params(g, t)
        out = many.score_groups(cur, off, hist, cur_len=cl, hist_len=hl, metric_type=mt)
        again = many.score_groups(cur, off, hist, cur_len=cl, hist_len=hl, metric_type=mt)
    assert out["timing"].launches > 3 * ref["timing"].launches
    assert out["rec"].tobytes() == ref["rec"].tobytes() and out["svc"].tobytes() == ref["svc"].tobytes()
    assert again["rec"].tobytes() == ref["rec"].tobytes() and again["svc"].tobytes() == ref["svc"].tobytes()
    ora = O.score_groups(g, t, ref["batch"], n_threads=8)
    assert_records_match(ref["rec"], ora["rec"], cur, cl, bound_masks(t, np.repeat(mt, K), G))
    assert (ref["svc"] == ora["svc"]).all()


@pytest.mark.parametrize("distinct", [1, 2, 7, 31, 32, 33, 40, 63, 64, 65, 100, 200])
def test_groups_distinct_values(ctx, distinct):
    """The rollover kernel ranks a pooled sample with at most 64 distinct values from a hash multiset and
    anything else with the sorting network: both sides of that limit, values that collide in the table
    (many services, arbitrary bit patterns), +-0, +-inf and denormals as values, ragged and empty groups."""
    rng = np.random.default_rng(4000 + distinct)
    V = 400
    K = rng.integers(2, 9, V)
    off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
    G = int(off[-1])
    cur = np.zeros((G, 64), np.float32)
    for v in range(V):
        kind = v % 4
        if kind == 0:      # counts per second
            vals = (np.arange(distinct) / 60.0).astype(np.float32)
        elif kind == 1:    # arbitrary finite bit patterns
            vals = rng.integers(0, 0x7F000000, distinct).astype(np.uint32).view(np.float32) * rng.choice([-1, 1], distinct).astype(np.float32)
        elif kind == 2:    # specials among ordinary values
            vals = rng.normal(0, 1, distinct).astype(np.float32)
            sp = np.array([0.0, -0.0, np.inf, -np.inf, 1e-45, -1e-45, 3e-39], np.float32)
            vals[:min(distinct, sp.size)] = sp[:min(distinct, sp.size)]
        else:              # exactly `distinct` values, every one present
            vals = rng.permutation(4096)[:distinct].astype(np.float32)
        rows = slice(off[v], off[v + 1])
        pick = rng.integers(0, distinct, (K[v], 64))
        if kind == 3 and K[v] * 64 >= distinct:
            pick.ravel()[:distinct] = np.arange(distinct)
        cur[rows] = vals[pick]
    cl = rng.integers(5, 65, G).astype(np.uint32)
    cl[rng.random(G) < 0.6] = 64
    cl[off[7]] = 0
    hist = rng.normal(0.5, 0.2, (V, 300)).astype(np.float32)
    mt = rng.integers(0, 6, V).astype(np.uint8)
    g, t = params(pairwise_mode=A.FM_PW_KRUSKAL)
    ctx.set_params(g, t)
    out = ctx.score_groups(cur, off, hist, cur_len=cl, metric_type=mt)
    ora = O.score_groups(g, t, out["batch"], n_threads=8)
    assert_records_match(out["rec"], ora["rec"], cur, cl, bound_masks(t, np.repeat(mt, K), G))
    assert (out["svc"] == ora["svc"]).all()
    assert (out["rec"]["status"] & A.FM_ST_KW_RUN).any()


@pytest.mark.parametrize("n_cur", [20, 60, 100])
def test_groups_friedman(ctx, n_cur):
    """Friedman chi-square (docs/guides/design.md:92) on both rollover kernels: 3..16 groups, tied and
    continuous blocks, services whose groups differ in length (not evaluated), a shifted group."""
    rng = np.random.default_rng(2000 + n_cur)
    V = 240
    K = rng.integers(1, 17, V)
    K[:5] = [1, 2, 3, 8, 9]
    off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
    G = int(off[-1])
    cur = rng.normal(50, 5, (G, n_cur)).astype(np.float32)
    for v in range(0, V, 2):
        cur[off[v]:off[v + 1]] = rng.poisson(2.0, (K[v], n_cur)).astype(np.float32)
    for v in range(5, V, 7):
        cur[off[v]] += 4.0                                  # one pod group off: the test fires
    cur[off[6]:off[7]] = 1.0                                # degenerate
    hist = rng.normal(50, 5, (V, 300)).astype(np.float32)
    cl = np.full(G, n_cur, np.uint32)
    for v in range(10, V, 11):
        cl[off[v]] = n_cur - 1                              # unequal lengths
    mt = rng.integers(0, 6, V).astype(np.uint8)
    g, t = params(pairwise_mode=A.FM_PW_FRIEDMAN)
    ctx.set_params(g, t)
    out = ctx.score_groups(cur, off, hist, cur_len=cl, metric_type=mt)
    ora = O.score_groups(g, t, out["batch"], n_threads=8)
    assert_records_match(out["rec"], ora["rec"], cur, cl, bound_masks(t, np.repeat(mt, K), G))
    assert (out["svc"] == ora["svc"]).all()
    st = out["rec"]["status"]
    assert (st & A.FM_ST_KW_RUN).any() and (st & A.FM_ST_DIFFER).any() and (st & A.FM_ST_PAIRWISE_SKIPPED).any()


def test_friedman_mode_on_series_batches_is_skipped(ctx):
    """The series (canary) path has no related samples: FRIEDMAN there evaluates no test."""
    d = synth_desc(3, 64, 60, 1440, 0)
    cur, base, _, mt = synth_host(d)
    g, t = params(pairwise_mode=A.FM_PW_FRIEDMAN)
    out, _, _ = both(ctx, g, t, cur, base, metric_type=mt)
    assert (out["rec"]["status"] & A.FM_ST_PAIRWISE_SKIPPED).all()


@pytest.mark.parametrize("mode", [A.FM_MODEL_EWMA, A.FM_MODEL_DOUBLE_EXP])
@pytest.mark.parametrize("nh", [1, 2, 3, 31, 32, 33, 64, 100, 1440, 10080])
def test_smoothing_models_scan(ctx, mode, nh):
    """Exponential / double exponential smoothing evaluated as a warp scan of affine maps against
    the oracle's sequential recurrence: every chunking edge (fewer steps than lanes, one step per
    lane, ragged last chunk), ragged history lengths, trending and level-shifted rows."""
    rng = np.random.default_rng(300 + nh)
    S = 48
    t = np.arange(nh, dtype=np.float64)
    hist = (100 + 0.01 * t[None, :] * rng.normal(1, 0.5, (S, 1)) + rng.normal(0, 2, (S, nh))).astype(np.float32)
    hist[: S // 4, nh // 2:] += 15.0
    cur = (hist[:, -1:] + rng.normal(0, 3, (S, 30))).astype(np.float32)
    hl = np.full(S, nh, np.uint32)
    hl[S // 2:] = rng.integers(0, nh + 1, S - S // 2)
    for al, be in ((0.3, 0.1), (0.9, 0.5), (0.05, 0.01)):
        g, tb = params(ma_mode=mode, smooth_alpha=al, smooth_beta=be, min_hist_points=1, pairwise_mode=A.FM_PW_NONE)
        both(ctx, g, tb, cur, None, hist, hist_len=hl, strategy=A.FM_STRAT_HISTORICAL,
             metric_type=rng.integers(0, 6, S).astype(np.uint8))


@pytest.mark.parametrize("m,nh", [(2, 4), (2, 5), (7, 14), (24, 100), (60, 1440), (1440, 10080)])
def test_holt_winters(ctx, m, nh):
    """Additive Holt-Winters (docs/guides/design.md:60-66; A.2c) against the oracle's sequential
    recurrence: centre of every current point, sigma, per-point flags; ragged histories (shorter
    than two seasons -> unknown), a trend, a level shift, as the model of a canary batch too."""
    rng = np.random.default_rng(900 + m + nh)
    S = 40
    t = np.arange(nh + 64, dtype=np.float64)
    wave = 50 + 0.002 * t[None, :] * rng.normal(1, 1, (S, 1)) + 8 * np.sin(2 * np.pi * t[None, :] / m + rng.uniform(0, 6, (S, 1)))
    x = (wave + rng.normal(0, 0.5, (S, nh + 64))).astype(np.float32)
    hist = x[:, :nh].copy()
    hist[: S // 4, nh // 2:] += 5.0
    cur = x[:, nh: nh + 40].copy()
    cur[::5, 7] += 12.0                                     # a spike the seasonal bounds must catch
    hl = np.full(S, nh, np.uint32)
    hl[S // 2:] = rng.integers(0, nh + 1, S - S // 2)
    for al, be, ga in ((0.2, 0.05, 0.3), (0.9, 0.5, 0.9)):
        g, tb = params(ma_mode=A.FM_MODEL_HOLT_WINTERS, season_len=m, smooth_alpha=al, smooth_beta=be, smooth_gamma=ga,
                       min_hist_points=1, pairwise_mode=A.FM_PW_NONE)
        out, _, _ = both(ctx, g, tb, cur, None, hist, hist_len=hl, strategy=A.FM_STRAT_HISTORICAL,
                         metric_type=rng.integers(0, 6, S).astype(np.uint8))
        assert (out["rec"]["verdict"] == A.FM_UNHEALTHY).any()
    if nh <= 2048:      # the baseline row as the model window of a canary batch, pairwise tests beside it
        g, tb = params(ma_mode=A.FM_MODEL_HOLT_WINTERS, season_len=m, min_hist_points=1)
        both(ctx, g, tb, cur, hist, metric_type=rng.integers(0, 6, S).astype(np.uint8))


@pytest.mark.parametrize("nh,nc", [(3, 1), (10, 5), (100, 60), (1440, 60), (10080, 64), (333, 300)])
def test_bivariate_pairs(ctx, nh, nc):
    """fm_score_pairs (bivariate normal, docs/guides/design.md:67-76) against the oracle: moments
    within 1e-6, flags exact; ragged lengths, correlated / independent / collinear / constant
    histories, memory-like magnitudes."""
    from test_oracle import pair_batch
    rng = np.random.default_rng(nh * 131 + nc)
    P = 64
    z = rng.normal(0, 1, (P, nh + nc, 2))
    rho = rng.uniform(-0.95, 0.95, (P, 1))
    x = 1.0e3 + 25 * z[..., 0]
    y = 7.0 + 2 * (rho * z[..., 0] + np.sqrt(1 - rho ** 2) * z[..., 1])
    x[:8] = 1.0e9 + 6.4e4 * np.round(z[:8, :, 0] * 20)           # memory-like: sigma / mean ~ 1e-3, quantised
    hx, hy = x[:, :nh].astype(np.float32), y[:, :nh].astype(np.float32)
    cx, cy = x[:, nh:].astype(np.float32), y[:, nh:].astype(np.float32)
    cx[::3, 0] += 200.0
    hy[8] = 3.0 * hx[8]                                            # collinear
    hx[9] = 5.0                                                    # constant
    hl = np.full(P, nh, np.uint32); hl[32:] = rng.integers(0, nh + 1, P - 32)
    cl = np.full(P, nc, np.uint32); cl[48:] = rng.integers(0, nc + 1, P - 48)
    mt = rng.integers(0, 6, P).astype(np.uint8)
    g, t = params(min_hist_points=3)
    ctx.set_params(g, t)
    out = ctx.score_pairs(cx, cy, hx, hy, cur_len=cl, hist_len=hl, metric_type=mt)
    b, keep = pair_batch(cx, cy, hx, hy, cl, hl, mt)
    ora = O.score_pairs(g, t, b)
    got = out["rec"]
    assert (got["status"] == ora["status"]).all()
    scale = np.maximum(np.abs(ora["mean_x"]) + np.abs(ora["mean_y"]), 1.0)
    for f in ("mean_x", "mean_y", "var_x", "var_y", "cov_xy"):
        assert np.isclose(got[f], ora[f], rtol=1e-6, atol=1e-12 * scale * scale).all(), f
    # knife edge: a pair whose distance is within 1e-6 of k^2 may flip; anything else is exact
    k2 = np.array([t[int(m)].k ** 2 for m in mt])
    edge = np.abs(ora["d2_max"] - k2) <= 1e-6 * k2
    same = (got["n_anom"] == ora["n_anom"]) & (got["anom_mask"] == ora["anom_mask"]) & (got["verdict"] == ora["verdict"]) & \
           (got["first_idx"] == ora["first_idx"])
    assert (same | edge).all()
    assert np.isclose(got["d2_max"], ora["d2_max"], rtol=1e-5, atol=1e-9).all()
    if nh >= 100:
        assert (got["verdict"] == A.FM_UNHEALTHY).any() and (got["verdict"] == A.FM_UNKNOWN).any()


def test_rank_dump(ctx):
    rng = np.random.default_rng(2)
    for nc, nb in [(5, 9), (60, 1440), (64, 64), (1, 0), (0, 3)]:
        cur = np.round(rng.normal(0, 2, nc)).astype(np.float32)
        base = np.round(rng.normal(0, 2, nb)).astype(np.float32)
        assert (ctx.rank_dump(cur, base) == O.rank_dump(cur, base)).all()


def test_ma_window_series(ctx):
    rng = np.random.default_rng(4)
    x = (1.0e6 + rng.normal(0, 30, 5000)).astype(np.float32)
    for w, ddof in [(1, 0), (10, 0), (60, 1), (1440, 0), (5000, 1)]:
        m, s = ctx.ma_window_series(x, w, ddof)
        mo, so = O.ma_window_series(x, w, ddof)
        assert np.array_equal(np.isnan(m), np.isnan(mo))
        ok = ~np.isnan(mo)
        assert np.allclose(m[ok], mo[ok], rtol=1e-6, atol=0)
        assert np.allclose(s[ok], so[ok], rtol=1e-6, atol=1e-9)


def test_synth_device_matches_host(ctx):
    for cfg, nb, nh in [(3, 1440, 0), (2, 100, 300), (5, 0, 77)]:
        S = 257
        d = synth_desc(cfg, S, 60, nb, nh, first_series=12345)
        cur, base, hist, mt = synth_host(d)
        dc = ctx.dev_alloc(S * d.cur_pitch * 4)
        db = ctx.dev_alloc(max(1, S * d.base_pitch * 4))
        dh = ctx.dev_alloc(max(1, S * d.hist_pitch * 4))
        dm = ctx.dev_alloc(S)
        ctx.synth_device(d, dc, db if nb else None, dh if nh else None, dm)
        ctx.sync()
        gc = np.zeros_like(cur); ctx.d2h(gc, dc)
        assert gc.tobytes() == cur.tobytes()
        if nb:
            gb = np.zeros_like(base); ctx.d2h(gb, db)
            assert gb.tobytes() == base.tobytes()
        if nh:
            gh = np.zeros_like(hist); ctx.d2h(gh, dh)
            assert gh.tobytes() == hist.tobytes()
        gm = np.zeros_like(mt); ctx.d2h(gm, dm)
        assert (gm == mt).all()
        for p in (dc, db, dh, dm):
            ctx.dev_free(p)


def test_errors(ctx):
    g, t = params()
    ctx.set_params(g, t)
    from foremast_1_b200._ffi import ForemastCudaError
    with pytest.raises(ForemastCudaError) as e:
        ctx.score(np.zeros((2, 5000), np.float32), np.zeros((2, 100), np.float32))
    assert e.value.code == A.FM_E_UNSUPPORTED
    g2 = copy.copy(g)
    g2.ddof = 7
    with pytest.raises(ForemastCudaError) as e:
        ctx.set_params(g2, t)
    assert e.value.code == A.FM_E_INVALID
    # layout contract: struct_size, 16-byte alignment, pitch a multiple of 4 and >= roundup4(n)
    import ctypes as C
    lib = ctx._lib
    cur = np.zeros((2, 8), np.float32)
    rec = np.zeros(2, A.SERIES_REC)
    b = A.Batch()
    b.struct_size = C.sizeof(A.Batch) - 8
    b.n_series, b.cur, b.n_cur, b.cur_pitch = 2, cur.ctypes.data, 8, 8
    assert lib.fm_score_batch(ctx._h, C.byref(b), rec.ctypes.data, None, None, None, None) == A.FM_E_INVALID
    assert b"struct_size" in lib.fm_last_error(ctx._h)
    b.struct_size = C.sizeof(A.Batch)
    b.cur_pitch = 6
    assert lib.fm_score_batch(ctx._h, C.byref(b), rec.ctypes.data, None, None, None, None) == A.FM_E_INVALID
    b.cur_pitch = 8
    b.cur = cur.ctypes.data + 4
    assert lib.fm_score_batch(ctx._h, C.byref(b), rec.ctypes.data, None, None, None, None) == A.FM_E_INVALID
    b.cur = cur.ctypes.data
    b.strategy = 7
    assert lib.fm_score_batch(ctx._h, C.byref(b), rec.ctypes.data, None, None, None, None) == A.FM_E_INVALID
    b.strategy = A.FM_STRAT_HISTORICAL
    assert lib.fm_score_batch(ctx._h, C.byref(b), None, None, None, None, None) == A.FM_E_INVALID
    # a batch without history is legal: every series is "unknown" (no model), not an error
    assert lib.fm_score_batch(ctx._h, C.byref(b), rec.ctypes.data, None, None, None, None) == A.FM_OK
    assert (rec["verdict"] == A.FM_UNKNOWN).all() and (rec["status"] & A.FM_ST_HIST_INSUFFICIENT).all()
    b.n_series = 0
    assert lib.fm_score_batch(ctx._h, C.byref(b), rec.ctypes.data, None, None, None, None) == A.FM_OK
    # the ctx stays usable after an error
    ctx.set_params(g, t)
    out = ctx.score(np.ones((1, 8), np.float32), np.ones((1, 16), np.float32))
    assert out["rec"]["verdict"][0] == A.FM_HEALTHY
