"""Pins the CPU oracle: hand-derived known answers, scipy (float64 inputs) as a
second opinion, and the frozen golden fixtures under tests/golden/.

The reference holds no implementation and no test vectors for this path
("parity unpinned", oracle/fm_oracle.h), so these are the anchors.
"""
import json
import os
from fractions import Fraction

import numpy as np
import pytest
from scipy import stats

import oracle_lib as O
from foremast_1_b200 import _abi as A

HERE = os.path.dirname(os.path.abspath(__file__))


def params(**kw):
    g, t = A.default_params()
    for k, v in kw.items():
        setattr(g, k, v)
    return g, t


# ---------------------------------------------------------------- A.1 / A.2 --
def test_moments_known_answer():
    m, s = O.moments([2, 4, 4, 4, 5, 5, 7, 9], 0)
    assert m == 5.0 and s == 2.0
    m, s = O.moments([2, 4, 4, 4, 5, 5, 7, 9], 1)
    assert m == 5.0 and s == pytest.approx(np.sqrt(32 / 7), rel=1e-15)
    assert O.moments([3.5], 0) == (3.5, 0.0)
    assert O.moments([3.5], 1) == (3.5, 0.0)


def test_window_series_matches_numpy():
    rng = np.random.default_rng(0)
    x = rng.normal(10, 2, 300).astype(np.float32)
    m, s = O.ma_window_series(x, 25, 0)
    assert np.isnan(m[:24]).all() and not np.isnan(m[24:]).any()
    for t in (24, 100, 299):
        win = x[t - 24: t + 1].astype(np.float64)
        assert m[t] == pytest.approx(win.mean(), rel=1e-14)
        assert s[t] == pytest.approx(win.std(), rel=1e-12)


# ---------------------------------------------------------------------- A.3 --
def test_midranks_known_answer():
    # sorted: 1 2 2 3 3 3 7 -> ranks 1, 2.5, 2.5, 5, 5, 5, 7
    r2, T = O.midranks([3, 1, 2, 3, 2, 7, 3])
    assert r2.tolist() == [10, 2, 5, 10, 5, 14, 10]
    assert T == (2 ** 3 - 2) + (3 ** 3 - 3)
    r2, T = O.midranks([0.0, -0.0, 0.0])      # -0 ties with +0
    assert r2.tolist() == [4, 4, 4] and T == 24


def test_midranks_vs_scipy_rankdata():
    rng = np.random.default_rng(1)
    for n in (1, 2, 17, 500):
        v = np.round(rng.normal(0, 3, n))
        r2, T = O.midranks(v)
        assert np.array_equal(r2, (2 * stats.rankdata(v)).astype(np.uint32))
        _, cnt = np.unique(v, return_counts=True)
        assert T == int((cnt.astype(np.int64) ** 3 - cnt).sum())


# ---------------------------------------------------------------------- A.4 --
def exact_mwu(x, y):
    """Mann-Whitney U1 and tie term by exact rational arithmetic (tiny inputs)."""
    u = Fraction(0)
    for a in x:
        for b in y:
            u += 1 if a > b else (Fraction(1, 2) if a == b else 0)
    allv = list(x) + list(y)
    T = sum(c ** 3 - c for c in (allv.count(v) for v in set(allv)))
    return u, T


def test_mann_whitney_hand_case():
    x = np.array([1.0, 2.0, 2.0, 5.0] * 5, np.float32)
    y = np.array([0.0, 2.0, 3.0, 3.0, 4.0] * 5, np.float32)
    g, t = params(pairwise_mode=A.FM_PW_MANN_WHITE, min_hist_points=1)
    rec, aux = O.score_series(g, t[0], A.FM_STRAT_CANARY, x, y)
    u, T = exact_mwu(x.tolist(), y.tolist())
    assert Fraction(rec["u1"]).limit_denominator(2) == u
    assert aux["t_mw"] == T
    n1, n2 = len(x), len(y)
    assert aux["r2_cur"] == int(2 * u + n1 * (n1 + 1))


@pytest.mark.parametrize("seed", range(6))
def test_tests_vs_scipy(seed):
    rng = np.random.default_rng(seed)
    nc, nb = int(rng.integers(20, 80)), int(rng.integers(20, 400))
    if seed % 2:
        cur = rng.integers(0, 6, nc).astype(np.float32)
        base = rng.integers(0, 6, nb).astype(np.float32)
    else:
        cur = rng.normal(0.3, 1, nc).astype(np.float32)
        base = rng.normal(0, 1, nb).astype(np.float32)
    g, t = params(pairwise_mode=A.FM_PW_ANY, min_hist_points=1)
    rec, aux = O.score_series(g, t[0], A.FM_STRAT_CANARY, cur, base)
    c64, b64 = cur.astype(np.float64), base.astype(np.float64)
    r = stats.mannwhitneyu(c64, b64, use_continuity=True, alternative="two-sided", method="asymptotic")
    assert rec["u1"] == r.statistic and rec["p_u"] == pytest.approx(r.pvalue, rel=1e-12)
    n = min(nc, nb)
    d = c64[:n] - b64[:n]
    if (d != 0).sum() >= 20:
        wx = stats.wilcoxon(c64[:n], b64[:n], zero_method="wilcox", correction=False, method="asymptotic")
        assert rec["w"] == wx.statistic and rec["p_w"] == pytest.approx(wx.pvalue, rel=1e-12)
    k = stats.kruskal(c64, b64)
    assert rec["h"] == pytest.approx(k.statistic, rel=1e-12) and rec["p_h"] == pytest.approx(k.pvalue, rel=1e-10)


def test_chi2_sf_vs_scipy():
    for dof in range(1, 16):
        for x in (1e-9, 0.3, 1.0, 5.5, 20.0, 80.0, 400.0):
            assert O.chi2_sf(x, dof) == pytest.approx(stats.chi2.sf(x, dof), rel=1e-12, abs=1e-300)
    assert O.chi2_sf(0.0, 3) == 1.0 and O.chi2_sf(-1.0, 2) == 1.0


def test_degenerate_and_eligibility():
    g, t = params(min_hist_points=1)
    one = np.ones(30, np.float32)
    rec, aux = O.score_series(g, t[0], A.FM_STRAT_CANARY, one, np.ones(100, np.float32))
    assert rec["p_u"] == 1.0 and rec["p_h"] == 1.0 and rec["status"] & A.FM_ST_DEGENERATE
    assert aux["n_wx"] == 0 and not (rec["status"] & A.FM_ST_WX_RUN)
    assert rec["status"] & A.FM_ST_PAIRWISE_SKIPPED and not (rec["status"] & A.FM_ST_DIFFER)
    # fewer points than MIN_MANN_WHITE / MIN_WILCOXON (20) but enough for Kruskal (5)
    rng = np.random.default_rng(0)
    rec, _ = O.score_series(g, t[0], A.FM_STRAT_CANARY, rng.normal(0, 1, 10), rng.normal(0, 1, 50))
    assert not (rec["status"] & (A.FM_ST_MW_RUN | A.FM_ST_WX_RUN)) and rec["status"] & A.FM_ST_KW_RUN


# ------------------------------------------------------------- A.6 .. A.8 ----
def test_threshold_lowering_and_verdict():
    rng = np.random.default_rng(3)
    base = rng.normal(10, 1, 1440).astype(np.float32)
    mu, sd = O.moments(base)
    cur_same = rng.normal(10, 1, 60).astype(np.float32)
    cur_same[5] = np.float32(mu + 1.5 * sd)       # inside k=2, outside k_low=1
    g, t = params(pairwise_mode=A.FM_PW_MANN_WHITE)
    rec, _ = O.score_series(g, t[1], A.FM_STRAT_CANARY, cur_same, base)
    if not (rec["status"] & A.FM_ST_DIFFER):
        assert rec["upper"] == pytest.approx(mu + 2 * sd, rel=1e-14)
    cur_shift = (rng.normal(10, 1, 60) + 0.8).astype(np.float32)   # distributions differ -> k_low
    rec, _ = O.score_series(g, t[1], A.FM_STRAT_CANARY, cur_shift, base)
    assert rec["status"] & A.FM_ST_DIFFER and rec["upper"] == pytest.approx(mu + 1 * sd, rel=1e-14)
    assert rec["verdict"] == A.FM_UNHEALTHY and rec["n_anom"] == int((cur_shift.astype(np.float64) > rec["upper"]).sum())
    # bound mask 3 (latency): lower side too, floored by min_lower_bound
    rec, _ = O.score_series(g, t[3], A.FM_STRAT_HISTORICAL, np.array([-5.0, 10.0], np.float32), None, base)
    assert rec["lower"] == pytest.approx(mu - 10 * sd, rel=1e-12)
    assert rec["n_anom"] == 1 and rec["first_idx"] == 0 and rec["anom_mask"] == 1
    rec, _ = O.score_series(g, t[3], A.FM_STRAT_HISTORICAL, np.array([-5.0, 1.0], np.float32), None, base - 9)
    assert rec["lower"] == 0.0 and rec["n_anom"] == 1          # floored by min_lower_bound = 0
    # unknown: no current points / too little history (design.md:25)
    rec, _ = O.score_series(g, t[0], A.FM_STRAT_HISTORICAL, np.zeros(0, np.float32), None, base)
    assert rec["verdict"] == A.FM_UNKNOWN and rec["status"] & A.FM_ST_NO_CURRENT
    rec, _ = O.score_series(g, t[0], A.FM_STRAT_HISTORICAL, cur_same, None, base[:5])
    assert rec["verdict"] == A.FM_UNKNOWN and rec["status"] & A.FM_ST_HIST_INSUFFICIENT


def test_demo_fixture_flags_the_two_spikes():
    """examples/spring-boot-demo data1.txt (normal) vs data2.txt (two 40.x spikes):
    with k = 2, upper bound only, exactly the two spikes are anomalous
    (docs/guides/installation.md:84,126-134 describes the Unhealthy outcome)."""
    with open(os.path.join(HERE, "golden", "demo_data.json")) as f:
        d = json.load(f)
    g, t = params(min_hist_points=10)
    rec, _ = O.score_series(g, t[1], A.FM_STRAT_HISTORICAL, d["data2"], None, d["data1"])
    spikes = [i for i, v in enumerate(d["data2"]) if v > 40]
    assert len(spikes) == 2
    assert rec["verdict"] == A.FM_UNHEALTHY and rec["n_anom"] == 2
    assert rec["first_idx"] == spikes[0]
    assert rec["anom_mask"] == sum(1 << i for i in spikes)


# ------------------------------------------------------------------ A.9 ------
def test_rollup_rule():
    import ctypes as C
    v = np.array([0, 0, 1, 2, 2, 2, 0, 2], np.uint8)
    parent = np.array([0, 0, 0, 1, 1, 2, 2, 3], np.uint32)
    out = np.zeros(5, A.ROLLUP_REC)
    O.lib().fmo_rollup(C.c_void_p(v.ctypes.data), C.c_void_p(parent.ctypes.data), C.c_uint64(v.size),
                       C.c_void_p(out.ctypes.data), C.c_uint32(5))
    assert out["severity"].tolist() == [A.FM_UNHEALTHY, A.FM_UNKNOWN, A.FM_HEALTHY, A.FM_UNKNOWN, A.FM_HEALTHY]
    assert out["n_children"].tolist() == [3, 2, 2, 1, 0]


# ------------------------------------------------------------- golden files --
def test_golden_records():
    """Frozen oracle outputs on seeded synthetic batches (tests/golden/make_golden.py)."""
    from make_golden import CASES, run_case
    z = np.load(os.path.join(HERE, "golden", "oracle_records.npz"))
    for name in CASES:
        out = run_case(name)
        for k, v in out.items():
            assert v.tobytes() == z[f"{name}__{k}"].tobytes(), (name, k)


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_friedman_vs_scipy(seed):
    """Friedman chi-square of the group path (docs/guides/design.md:92) against
    scipy.stats.friedmanchisquare on float64 copies: continuous and heavily tied blocks."""
    rng = np.random.default_rng(seed)
    V = 12
    K = rng.integers(3, 17, V)
    off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
    G = int(off[-1])
    n = 40
    cur = rng.normal(10, 2, (G, n)).astype(np.float32)
    for v in range(0, V, 2):
        cur[off[v]:off[v + 1]] = rng.poisson(1.5, (K[v], n)).astype(np.float32)
    cur[off[1]:off[2]] = 3.0                                   # every block fully tied: degenerate
    cl = np.full(G, n, np.uint32)
    cl[off[3]] = n - 1                                         # unequal lengths: not evaluated
    g, t = A.default_params()
    g.pairwise_mode = A.FM_PW_FRIEDMAN
    rec, svc = O.score_groups_arrays(g, t, cur, off, None, cur_len=cl)
    for v in range(V):
        r = rec[off[v]]
        if v == 3:
            assert not (r["status"] & A.FM_ST_KW_RUN) and (r["status"] & A.FM_ST_PAIRWISE_SKIPPED)
            continue
        assert r["status"] & A.FM_ST_KW_RUN
        if v == 1:
            assert (r["status"] & A.FM_ST_DEGENERATE) and r["p_h"] == 1.0
            continue
        q, p = stats.friedmanchisquare(*[cur[off[v] + k].astype(np.float64) for k in range(K[v])])
        assert r["h"] == pytest.approx(q, rel=1e-12, abs=1e-12)
        assert r["p_h"] == pytest.approx(p, rel=1e-10, abs=1e-300)
        assert bool(r["status"] & A.FM_ST_DIFFER) == (r["p_h"] < g.alpha)


# ------------------------------------------------------------------- A.2c ---
def holt_winters_py(x, m, a, b, c, ddof=0):
    """The textbook additive recurrence in plain Python (second opinion for the C restatement)."""
    x = [float(v) for v in x]
    n = len(x)
    level = sum(x[:m]) / m
    trend = (sum(x[m:2 * m]) / m - level) / m
    season = [x[i] - level for i in range(m)]
    sse = 0.0
    for t in range(m, n):
        s = season[t % m]
        e = x[t] - (level + trend + s)
        sse += e * e
        nl = a * (x[t] - s) + (1 - a) * (level + trend)
        trend = b * (nl - level) + (1 - b) * trend
        season[t % m] = c * (x[t] - nl) + (1 - c) * s
        level = nl
    cnt = n - m
    sigma = (sse / (cnt - ddof)) ** 0.5 if cnt > ddof else 0.0
    centres = [level + (i + 1) * trend + season[(n + i) % m] for i in range(64)]
    return centres, sigma


def test_holt_winters_known_answer_and_seasonal_flags():
    # hand case: m = 2, x = 0 2 0 2: level 1, trend 0, season (-1, +1); every forecast is exact
    g, t = params(ma_mode=A.FM_MODEL_HOLT_WINTERS, season_len=2, smooth_alpha=0.5, smooth_beta=0.5,
                  smooth_gamma=0.5, min_hist_points=1, pairwise_mode=A.FM_PW_NONE)
    rec, _ = O.score_series(g, t[3], A.FM_STRAT_HISTORICAL, np.array([0, 2, 0, 2.5], np.float32), None,
                            np.array([0, 2, 0, 2], np.float32))
    assert rec["mean"] == 0.0 and rec["sigma"] == 0.0           # centre of the first current point: 1 + 0 - 1
    assert rec["n_anom"] == 1 and rec["first_idx"] == 3        # 2.5 against its own centre 2, sigma 0
    # a daily pattern: the same value is normal at the peak and anomalous in the trough
    rng = np.random.default_rng(1)
    m, n = 48, 48 * 7
    tt = np.arange(n + 60)
    wave = 100 + 0.01 * tt + 30 * np.sin(2 * np.pi * tt / m)
    x = (wave + rng.normal(0, 1.0, tt.size)).astype(np.float32)
    hist, cur = x[:n], x[n:n + 60].copy()
    g, t = params(ma_mode=A.FM_MODEL_HOLT_WINTERS, season_len=m, smooth_alpha=0.2, smooth_beta=0.05,
                  smooth_gamma=0.3, pairwise_mode=A.FM_PW_NONE)
    rec, _ = O.score_series(g, t[4], A.FM_STRAT_HISTORICAL, cur, None, hist)      # cpu: k = 5, upper only
    centres, sigma = holt_winters_py(hist, m, 0.2, 0.05, 0.3)
    assert rec["mean"] == pytest.approx(centres[0], rel=1e-12) and rec["sigma"] == pytest.approx(sigma, rel=1e-12)
    assert rec["verdict"] == A.FM_HEALTHY and 0.5 < sigma < 3.0
    trough = int(np.argmin(wave[n:n + 48]))
    cur[trough] = np.float32(wave[n:n + 60].max())                # the peak's value in the trough
    rec2, _ = O.score_series(g, t[4], A.FM_STRAT_HISTORICAL, cur, None, hist)
    assert rec2["verdict"] == A.FM_UNHEALTHY and rec2["n_anom"] == 1 and rec2["first_idx"] == trough
    # the plain moving average cannot see it: the value lies inside mean + 5 sigma of the whole week
    g0, _ = params(pairwise_mode=A.FM_PW_NONE)
    rec3, _ = O.score_series(g0, t[4], A.FM_STRAT_HISTORICAL, cur, None, hist)
    assert rec3["verdict"] == A.FM_HEALTHY
    # fewer than two seasons of history: no model
    rec4, _ = O.score_series(g, t[4], A.FM_STRAT_HISTORICAL, cur, None, hist[: 2 * m - 1])
    assert rec4["verdict"] == A.FM_UNKNOWN and rec4["status"] & A.FM_ST_HIST_INSUFFICIENT


# ------------------------------------------------------------------- A.2d ---
def pair_batch(cx, cy, hx, hy, cl=None, hl=None, mt=None):
    import ctypes as C
    from foremast_1_b200.engine import _as_matrix
    cx, nc = _as_matrix(cx); cy, _ = _as_matrix(cy); hx, nh = _as_matrix(hx); hy, _ = _as_matrix(hy)
    b = A.PairBatch()
    b.struct_size = C.sizeof(A.PairBatch)
    b.n_pairs = cx.shape[0]
    b.cur_x, b.cur_y, b.n_cur, b.cur_pitch = cx.ctypes.data, cy.ctypes.data, nc, cx.shape[1]
    b.hist_x, b.hist_y, b.n_hist, b.hist_pitch = hx.ctypes.data, hy.ctypes.data, nh, hx.shape[1]
    keep = [cx, cy, hx, hy, cl, hl, mt]
    b.cur_len, b.hist_len, b.metric_type = A.ptr(cl), A.ptr(hl), A.ptr(mt)
    return b, keep


def test_bivariate_normal_vs_numpy():
    """Two metrics judged jointly (docs/guides/design.md:67-76): covariance against numpy, the
    Mahalanobis rule against a direct evaluation; a point inside both 1-D bounds but off the
    correlation axis is caught; singular / short histories are unknown."""
    rng = np.random.default_rng(8)
    P, nh, nc = 6, 500, 30
    z = rng.normal(0, 1, (P, nh + nc, 2))
    x = (100 + 10 * z[..., 0]).astype(np.float32)
    y = (50 + 4 * (0.9 * z[..., 0] + np.sqrt(1 - 0.81) * z[..., 1])).astype(np.float32)   # correlation 0.9
    hx, hy, cx, cy = x[:, :nh].copy(), y[:, :nh].copy(), x[:, nh:].copy(), y[:, nh:].copy()
    cx[0, 3], cy[0, 3] = 112.0, 44.0          # +1.2 sigma in x, -1.5 sigma in y: against the correlation
    hy[4] = 2.0 * hx[4] + 1.0                 # collinear history
    hl = np.full(P, nh, np.uint32); hl[5] = 2
    g, t = params()
    mt = np.full(P, 2, np.uint8)              # error4xx: k = 3
    b, keep = pair_batch(cx, cy, hx, hy, None, hl, mt)
    rec = O.score_pairs(g, t, b)
    for p in range(4):
        c = np.cov(np.stack([hx[p].astype(np.float64), hy[p].astype(np.float64)]), ddof=0)
        assert rec["var_x"][p] == pytest.approx(c[0, 0], rel=1e-12) and rec["cov_xy"][p] == pytest.approx(c[0, 1], rel=1e-12)
        d = np.stack([cx[p] - rec["mean_x"][p], cy[p] - rec["mean_y"][p]]).astype(np.float64)
        d2 = np.einsum("it,ij,jt->t", d, np.linalg.inv(c), d)
        assert rec["d2_max"][p] == pytest.approx(d2.max(), rel=1e-9)
        assert rec["n_anom"][p] == int((d2 > 9.0).sum())
    assert rec["verdict"][0] == A.FM_UNHEALTHY and (rec["anom_mask"][0] >> 3) & 1
    sx, sy = np.sqrt(rec["var_x"][0]), np.sqrt(rec["var_y"][0])
    assert abs(112.0 - rec["mean_x"][0]) < 3 * sx and abs(44.0 - rec["mean_y"][0]) < 3 * sy   # each metric alone passes
    assert rec["verdict"][4] == A.FM_UNKNOWN and rec["status"][4] & A.FM_ST_DEGENERATE
    assert rec["verdict"][5] == A.FM_UNKNOWN and rec["status"][5] & A.FM_ST_HIST_INSUFFICIENT
