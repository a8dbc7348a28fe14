"""BASELINE.json's full sizes on one GPU, checked through size-independent properties:
a strided sample against the oracle, shard equivalence (two halves == the whole, byte for
byte), conservation laws of the statistics and of the roll-up."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
from foremast_1_b200 import _abi as A
from foremast_1_b200.engine import Context, synth_desc, synth_host
from parity import assert_records_match, bound_masks

pytestmark = pytest.mark.gpu


def device_batch(ctx, d, strat, api=False):
    S = int(d.n_series)
    p = {"cur": ctx.dev_alloc(S * d.cur_pitch * 4), "mt": ctx.dev_alloc(S),
         "base": ctx.dev_alloc(S * d.base_pitch * 4) if d.n_base else None,
         "hist": ctx.dev_alloc(S * d.hist_pitch * 4) if d.n_hist else None}
    ctx.synth_device(d, p["cur"], p["base"], p["hist"], p["mt"])
    b = A.Batch()
    b.struct_size = C.sizeof(A.Batch)
    b.strategy, b.n_series = strat, S
    b.cur, b.n_cur, b.cur_pitch = p["cur"], d.n_cur, d.cur_pitch
    if p["base"]:
        b.base, b.n_base, b.base_pitch = p["base"], d.n_base, d.base_pitch
    if p["hist"]:
        b.hist, b.n_hist, b.hist_pitch = p["hist"], d.n_hist, d.hist_pitch
    b.metric_type = p["mt"]
    return b, p


def free_all(ctx, p):
    for v in p.values():
        if v:
            ctx.dev_free(v)


def test_config3_one_million_series():
    S, nc, nb = 1_000_000, 60, 1440
    with Context(0) as ctx:
        g, t = A.default_params()
        ctx.set_params(g, t)
        d = synth_desc(3, S, nc, nb, 0)
        b, p = device_batch(ctx, d, A.FM_STRAT_CANARY)
        api = (np.arange(S) // 4).astype(np.uint32)
        svc = (np.arange(S // 4) // 8).astype(np.uint32)
        p["api"], p["svc"] = ctx.dev_alloc(api.nbytes), ctx.dev_alloc(svc.nbytes)
        ctx.h2d(p["api"], api); ctx.h2d(p["svc"], svc)
        b.api_id, b.api_service, b.n_apis, b.n_services = p["api"], p["svc"], S // 4, S // 32
        p["out"], p["aux"] = ctx.dev_alloc(S * 96), ctx.dev_alloc(S * 48)
        p["oapi"], p["osvc"] = ctx.dev_alloc(S // 4 * 16), ctx.dev_alloc(S // 32 * 16)
        tm = A.Timing()
        ctx.score_device(b, p["out"], p["aux"], p["oapi"], p["osvc"], tm)
        rec = np.zeros(S, A.SERIES_REC); ctx.d2h(rec, p["out"])
        aux = np.zeros(S, A.RANK_AUX); ctx.d2h(aux, p["aux"])
        rapi = np.zeros(S // 4, A.ROLLUP_REC); ctx.d2h(rapi, p["oapi"])
        rsvc = np.zeros(S // 32, A.ROLLUP_REC); ctx.d2h(rsvc, p["osvc"])

        # conservation laws
        assert ((rec["u1"] >= 0) & (rec["u1"] <= nc * nb)).all()
        assert (aux["r2_cur"] == (2 * rec["u1"] + nc * (nc + 1)).astype(np.int64)).all()
        ran = (rec["status"] & A.FM_ST_WX_RUN) != 0          # Wilcoxon needs >= 20 non-zero differences
        nwx = aux["n_wx"].astype(np.int64)
        assert ran.mean() > 0.7 and (ran == (nwx >= 20)).all()
        assert ((aux["w2_plus"] + aux["w2_minus"])[ran] == (nwx * (nwx + 1))[ran]).all()
        assert (aux["w2_plus"][~ran] == 0).all() and (aux["w2_minus"][~ran] == 0).all()
        for f in ("p_u", "p_w", "p_h"):
            assert ((rec[f] >= 0) & (rec[f] <= 1)).all()
        assert ((rec["verdict"] == A.FM_UNHEALTHY) == (rec["n_anom"] > 0)).all()
        pop = np.array([bin(int(m)).count("1") for m in rec["anom_mask"][:20000]])
        assert (pop == rec["n_anom"][:20000]).all()
        # roll-up: counts add up, severities follow rule A.9
        assert rapi["n_children"].sum() == S and (rapi["n_children"] == 4).all()
        unhealthy_api = np.add.reduceat((rec["verdict"] == A.FM_UNHEALTHY).astype(np.int64), np.arange(0, S, 4))
        assert (rapi["n_unhealthy"] == unhealthy_api).all()
        assert ((rapi["severity"] == A.FM_UNHEALTHY) == (unhealthy_api > 0)).all()
        assert (rsvc["n_children"] == 8).all()
        assert (rsvc["n_unhealthy"] == np.add.reduceat((rapi["severity"] == A.FM_UNHEALTHY).astype(np.int64), np.arange(0, S // 4, 8))).all()

        # shard equivalence: two halves scored separately == the whole
        half = S // 2
        for lo in (0, half):
            hb = A.Batch.from_buffer_copy(b)
            hb.n_series = half
            hb.cur = p["cur"] + lo * d.cur_pitch * 4
            hb.base = p["base"] + lo * d.base_pitch * 4
            hb.metric_type = p["mt"] + lo
            hb.api_id = None
            o2 = ctx.dev_alloc(half * 96)
            ctx.score_device(hb, o2)
            ctx.sync()
            r2 = np.zeros(half, A.SERIES_REC); ctx.d2h(r2, o2)
            ctx.dev_free(o2)
            assert r2.tobytes() == rec[lo: lo + half].tobytes()

        # a strided sample against the oracle (rows regenerated on the host from their ids)
        ids = np.arange(0, S, 3989, dtype=np.uint64)
        ds = synth_desc(3, ids.size, nc, nb, 0, row_ids=ids.ctypes.data)
        cur, base, _, mt = synth_host(ds)
        from make_golden_shim import host_batch
        ora = O.score_batch(g, t, host_batch(A.FM_STRAT_CANARY, cur, base, None, mt), n_threads=8)
        k = ids.astype(np.int64)
        assert_records_match(rec[k], ora["rec"], cur[:, :nc], None, bound_masks(t, mt, ids.size), aux[k], ora["aux"],
                             alpha=g.alpha)
        free_all(ctx, p)


def test_config5_ma_sweep_and_rollup():
    """Config 5 at size: MA-only bounds over three window lengths WITH the series -> API -> service
    roll-up (4 metrics per API, 8 APIs per service: 250 000 APIs, 31 250 services).  Conservation laws
    over the whole batch, and a strided sample of whole services against the oracle: series records,
    API records and service records."""
    S = 1_000_000
    api = (np.arange(S) // 4).astype(np.uint32)
    svc = (np.arange(S // 4) // 8).astype(np.uint32)
    with Context(0) as ctx:
        g, t = A.default_params()
        ctx.set_params(g, t)
        for nh in (60, 480, 2880):
            d = synth_desc(5, S, 60, 0, nh)
            b, p = device_batch(ctx, d, A.FM_STRAT_HISTORICAL)
            b.record_kind = A.FM_REC_COMPACT
            p["api"], p["svc"] = ctx.dev_alloc(api.nbytes), ctx.dev_alloc(svc.nbytes)
            ctx.h2d(p["api"], api); ctx.h2d(p["svc"], svc)
            b.api_id, b.api_service, b.n_apis, b.n_services = p["api"], p["svc"], S // 4, S // 32
            p["out"] = ctx.dev_alloc(S * 16)
            p["oapi"], p["osvc"] = ctx.dev_alloc(S // 4 * 16), ctx.dev_alloc(S // 32 * 16)
            ctx.score_device(b, p["out"], None, p["oapi"], p["osvc"], A.Timing())
            rec = np.zeros(S, A.COMPACT_REC); ctx.d2h(rec, p["out"])
            rapi = np.zeros(S // 4, A.ROLLUP_REC); ctx.d2h(rapi, p["oapi"])
            rsvc = np.zeros(S // 32, A.ROLLUP_REC); ctx.d2h(rsvc, p["osvc"])
            assert ((rec["verdict"] == A.FM_UNHEALTHY) == (rec["n_anom"] > 0)).all()
            assert (rec["status"] == 0).all()
            # roll-up conservation: every series counted once, severities follow rule A.9
            assert (rapi["n_children"] == 4).all() and (rsvc["n_children"] == 8).all()
            un_api = np.add.reduceat((rec["verdict"] == A.FM_UNHEALTHY).astype(np.int64), np.arange(0, S, 4))
            assert (rapi["n_unhealthy"] == un_api).all() and (rapi["n_unknown"] == 0).all()
            assert ((rapi["severity"] == A.FM_UNHEALTHY) == (un_api > 0)).all()
            un_svc = np.add.reduceat((rapi["severity"] == A.FM_UNHEALTHY).astype(np.int64), np.arange(0, S // 4, 8))
            assert (rsvc["n_unhealthy"] == un_svc).all()
            assert ((rsvc["severity"] == A.FM_UNHEALTHY) == (un_svc > 0)).all()
            assert (rsvc["severity"] == A.FM_UNHEALTHY).any() and (rsvc["severity"] == A.FM_HEALTHY).any()
            # whole services (32 consecutive series each) against the oracle, roll-up included
            pick = np.arange(0, S // 32, 251)
            ids = (pick[:, None] * 32 + np.arange(32)[None, :]).ravel().astype(np.uint64)
            ds = synth_desc(5, ids.size, 60, 0, nh, row_ids=ids.ctypes.data)
            cur, _, hist, mt = synth_host(ds)
            from make_golden_shim import host_batch
            hb = host_batch(A.FM_STRAT_HISTORICAL, cur, None, hist, mt)
            hb.record_kind = A.FM_REC_COMPACT
            lapi = (np.arange(ids.size) // 4).astype(np.uint32)
            lsvc = (np.arange(ids.size // 4) // 8).astype(np.uint32)
            hb.api_id, hb.api_service, hb.n_apis, hb.n_services = lapi.ctypes.data, lsvc.ctypes.data, ids.size // 4, pick.size
            ora = O.score_batch(g, t, hb, n_threads=8, want_aux=False)
            k = ids.astype(np.int64)
            got = rec[k]
            same = got == ora["rec"]
            # compact records carry no bounds: allow the documented 1e-6 flag tolerance by re-checking
            # disagreeing series with full records
            if not same.all():
                bad = np.nonzero(~same)[0]
                full = ctx.score(cur[bad], None, hist[bad], strategy=A.FM_STRAT_HISTORICAL, metric_type=mt[bad])
                fb = host_batch(A.FM_STRAT_HISTORICAL, cur[bad], None, hist[bad], mt[bad])
                fo = O.score_batch(g, t, fb, n_threads=2)
                assert_records_match(full["rec"], fo["rec"], cur[bad][:, :60], None, bound_masks(t, mt[bad], bad.size))
            # roll-up records of the sampled services: exact wherever their series agree exactly
            svc_ok = same.reshape(pick.size, 32).all(axis=1)
            api_ok = same.reshape(-1, 4).all(axis=1)
            gapi = rapi[(pick[:, None] * 8 + np.arange(8)[None, :]).ravel()]
            assert (gapi[api_ok] == ora["api"][api_ok]).all()
            assert (rsvc[pick][svc_ok] == ora["svc"][svc_ok]).all() and svc_ok.mean() > 0.99
            free_all(ctx, p)


def test_config4_hundred_thousand_services():
    V = 100_000
    rng = np.random.default_rng(44)
    K = rng.integers(3, 9, V)
    off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
    G = int(off[-1])
    svc_of = np.repeat(np.arange(V, dtype=np.uint64), K)
    ids = (svc_of << np.uint64(4)) + (np.arange(G, dtype=np.uint64) - off[:-1].astype(np.uint64)[svc_of.astype(np.int64)])
    hids = np.arange(V, dtype=np.uint64) << np.uint64(4)
    with Context(0) as ctx:
        g, t = A.default_params()
        g.pairwise_mode = A.FM_PW_KRUSKAL
        ctx.set_params(g, t)
        d_ids, d_hids = ctx.dev_alloc(ids.nbytes), ctx.dev_alloc(hids.nbytes)
        ctx.h2d(d_ids, ids); ctx.h2d(d_hids, hids)
        dg = synth_desc(4, G, 60, 0, 0, row_ids=d_ids)
        dh = synth_desc(4, V, 60, 0, 1440, row_ids=d_hids)
        d_cur, d_hist, d_mt = ctx.dev_alloc(G * dg.cur_pitch * 4), ctx.dev_alloc(V * dh.hist_pitch * 4), ctx.dev_alloc(V)
        ctx.synth_device(dg, d_cur, None, None, None)
        ctx.synth_device(dh, None, None, d_hist, d_mt)
        d_off = ctx.dev_alloc(off.nbytes); ctx.h2d(d_off, off)
        gb = A.GroupBatch()
        gb.struct_size = C.sizeof(A.GroupBatch)
        gb.n_services, gb.n_groups, gb.group_offsets = V, G, d_off
        gb.cur, gb.n_cur, gb.cur_pitch = d_cur, 60, dg.cur_pitch
        gb.hist, gb.n_hist, gb.hist_pitch = d_hist, 1440, dh.hist_pitch
        gb.metric_type = d_mt
        d_out, d_svc = ctx.dev_alloc(G * 96), ctx.dev_alloc(V * 16)
        ctx.score_groups_device(gb, d_out, d_svc, A.Timing())
        rec = np.zeros(G, A.SERIES_REC); ctx.d2h(rec, d_out)
        rsvc = np.zeros(V, A.ROLLUP_REC); ctx.d2h(rsvc, d_svc)
        assert (rsvc["n_children"] == K).all()
        assert (rec["status"] & A.FM_ST_KW_RUN).all()
        assert ((rec["p_h"] >= 0) & (rec["p_h"] <= 1)).all() and (rec["h"] >= -1e-9).all()   # H0 can round to -1e-13
        assert (rec["h"][off[:-1]] == rec["h"][off[1:] - 1]).all()          # one H per service
        # a sample of services against the oracle
        pick = np.arange(0, V, 997)
        rows = np.concatenate([np.arange(off[v], off[v + 1]) for v in pick])
        sub_ids = ids[rows].copy()
        sub_hids = hids[pick].copy()
        cur, _, _, _ = synth_host(synth_desc(4, rows.size, 60, 0, 0, row_ids=sub_ids.ctypes.data))
        _, _, hist, mt = synth_host(synth_desc(4, pick.size, 60, 0, 1440, row_ids=sub_hids.ctypes.data))
        soff = np.concatenate([[0], np.cumsum(K[pick])]).astype(np.uint32)
        hb = A.GroupBatch()
        hb.struct_size = C.sizeof(A.GroupBatch)
        hb.n_services, hb.n_groups, hb.group_offsets = pick.size, rows.size, soff.ctypes.data
        hb.cur, hb.n_cur, hb.cur_pitch = cur.ctypes.data, 60, cur.shape[1]
        hb.hist, hb.n_hist, hb.hist_pitch = hist.ctypes.data, 1440, hist.shape[1]
        hb.metric_type = mt.ctypes.data
        ora = O.score_groups(g, t, hb, n_threads=8)
        assert_records_match(rec[rows], ora["rec"], cur[:, :60], None, bound_masks(t, np.repeat(mt, K[pick]), rows.size))
        assert (rsvc[pick] == ora["svc"]).all()
        for q in (d_ids, d_hids, d_cur, d_hist, d_mt, d_off, d_out, d_svc):
            ctx.dev_free(q)
