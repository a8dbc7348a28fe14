"""Record comparison used by the parity tests (north_star tolerances).

Bit-exact: rank sums, tie terms, U, W, anomaly counts / masks / first index,
verdicts, status bits, roll-ups.  1e-6 relative: mean, sigma, bounds, H,
p-values.  Two knife edges are tolerated, and every use is LISTED in the report
that assert_records_match returns:

* a series whose anomaly flags differ is accepted only if every disagreeing
  point lies within 1e-6 relative of the bound it was compared to; only the
  fields derived from the flags (n_anom, first_idx, anom_mask, verdict) are
  excused for it — U, W and the status bits stay exact;
* a series whose DIFFER bit differs (the pairwise decision p < alpha) is accepted
  only if a p-value of an evaluated test lies within 1e-6 relative of alpha;
  its bounds then follow different thresholds (k against k_low), so bounds and
  flags are excused for it, nothing else.
"""
import numpy as np

from foremast_1_b200 import _abi as A

RTOL = 1e-6
EXACT_FIELDS = ("u1", "w", "n_anom", "first_idx", "verdict", "status", "anom_mask")
CLOSE_FIELDS = ("mean", "sigma", "upper", "lower", "p_u", "p_w", "h", "p_h")
AUX_FIELDS = ("r2_cur", "t_mw", "w2_plus", "w2_minus", "t_wx", "n_wx")


def _close(a, b, scale):
    return np.isclose(a, b, rtol=RTOL, atol=1e-12 * scale)


def flag_mismatch_report(gpu, ora, cur, cur_len, bound_mask):
    """For series whose anomaly outputs differ: is every differing point within
    tolerance of its bound?  Returns (ok_mask, report list)."""
    bad = (gpu["n_anom"] != ora["n_anom"]) | (gpu["anom_mask"] != ora["anom_mask"]) | \
          (gpu["first_idx"] != ora["first_idx"]) | (gpu["verdict"] != ora["verdict"])
    ok = np.ones(gpu.size, bool)
    report = []
    for s in np.nonzero(bad)[0]:
        n = int(cur_len[s]) if cur_len is not None else cur.shape[1]
        x = cur[s, :n].astype(np.float64)
        bm = int(bound_mask[s])
        fo = ((bm & 1) != 0) & (x > ora["upper"][s]) | ((bm & 2) != 0) & (x < ora["lower"][s])
        fg = ((bm & 1) != 0) & (x > gpu["upper"][s]) | ((bm & 2) != 0) & (x < gpu["lower"][s])
        fine = True
        for i in np.nonzero(fo != fg)[0]:
            gaps = []
            if bm & 1:
                gaps.append(abs(x[i] - ora["upper"][s]) / max(abs(ora["upper"][s]), 1e-300))
            if bm & 2:
                gaps.append(abs(x[i] - ora["lower"][s]) / max(abs(ora["lower"][s]), 1e-300))
            gap = min(gaps) if gaps else np.inf
            report.append({"series": int(s), "point": int(i), "value": float(x[i]),
                           "upper": float(ora["upper"][s]), "lower": float(ora["lower"][s]),
                           "rel_gap": float(gap)})
            fine &= gap <= RTOL
        # the flags recomputed from each side's own bounds must explain the difference
        fine &= int(fg.sum()) == int(gpu["n_anom"][s]) and int(fo.sum()) == int(ora["n_anom"][s])
        ok[s] = fine
    return bad, ok, report


def differ_flip_report(gpu, ora, alpha):
    """Series whose DIFFER bit differs: is a p-value of an evaluated test within tolerance of alpha?"""
    flip = ((gpu["status"] ^ ora["status"]) & A.FM_ST_DIFFER) != 0
    ok = np.ones(gpu.size, bool)
    report = []
    for s in np.nonzero(flip)[0]:
        gaps = [abs(float(ora[f][s]) - alpha) / alpha
                for f, bit in (("p_u", A.FM_ST_MW_RUN), ("p_w", A.FM_ST_WX_RUN), ("p_h", A.FM_ST_KW_RUN))
                if ora["status"][s] & bit]
        gap = min(gaps) if gaps else np.inf
        report.append({"series": int(s), "kind": "differ", "alpha": float(alpha), "rel_gap": float(gap),
                       "p": [float(ora[f][s]) for f in ("p_u", "p_w", "p_h")]})
        ok[s] = alpha is not None and gap <= RTOL
    return flip, ok, report


def assert_records_match(gpu, ora, cur=None, cur_len=None, bound_mask=None, aux_gpu=None, aux_ora=None, alpha=None):
    assert gpu.shape == ora.shape
    scale = np.maximum(np.abs(ora["mean"]) + np.abs(ora["sigma"]), 1.0) if "mean" in ora.dtype.names else 1.0
    listed = []
    skip_flags = np.zeros(gpu.size, bool)      # excused: the flag-derived fields only
    flip = np.zeros(gpu.size, bool)            # excused: bounds and flags (the threshold k differs)
    if "mean" in gpu.dtype.names:
        if alpha is not None:
            flip, ok, rep = differ_flip_report(gpu, ora, alpha)
            assert ok.all(), ("pairwise decision differs beyond tolerance", rep[:10])
            listed += rep
        for f in CLOSE_FIELDS:
            c = _close(gpu[f], ora[f], scale) | (flip if f in ("upper", "lower") else False)
            assert c.all(), (f, np.nonzero(~c)[0][:10], gpu[f][~c][:5], ora[f][~c][:5])
        if cur is not None and bound_mask is not None:
            bad, ok, rep = flag_mismatch_report(gpu, ora, cur, cur_len, bound_mask)
            assert (ok | flip).all(), ("anomaly flags differ beyond tolerance", rep[:10])
            listed += rep
            skip_flags = bad
    fields = [f for f in EXACT_FIELDS if f in gpu.dtype.names]
    for f in fields:
        if f in ("n_anom", "first_idx", "anom_mask", "verdict"):
            eq = (gpu[f] == ora[f]) | skip_flags | flip
        elif f == "status":
            eq = ((gpu[f] ^ ora[f]) & ~(np.where(flip, A.FM_ST_DIFFER, 0).astype(gpu[f].dtype))) == 0
        else:
            eq = gpu[f] == ora[f]
        assert eq.all(), (f, np.nonzero(~eq)[0][:10], gpu[f][~eq][:5], ora[f][~eq][:5])
    if aux_gpu is not None and aux_ora is not None:
        for f in AUX_FIELDS:
            eq = aux_gpu[f] == aux_ora[f]
            assert eq.all(), (f, np.nonzero(~eq)[0][:10], aux_gpu[f][~eq][:5], aux_ora[f][~eq][:5])
    return listed


def bound_masks(table, metric_type, n):
    bm = np.array([t.bound_mask for t in table], np.uint32)
    if metric_type is None:
        return np.full(n, bm[0], np.uint32)
    mt = np.asarray(metric_type, np.int64)
    mt = np.where(mt < len(table), mt, 0)
    return bm[mt]
