#!/usr/bin/env python
"""bench.py — series scored / second on BASELINE.json's headline workload.

A step is one pass of the scoring path over one batch of synthetic series:
config 3 (1M series x 60-point canary vs 1440-point baseline; moving-average
bounds + Mann-Whitney + Wilcoxon (+ 2-group Kruskal), verdict) per GPU, sharded
by series id with no collective (weak scaling: every rank owns 1M series of an
N x 1M job).  `value` is measured with the rows resident in HBM (CUDA events on
the library's stream, max over ranks); `e2e` goes through fm_score_batch with
pinned HOST rows, host<->device copies inside the timed region.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

WORKLOADS = {
    # name: (config id, n_cur, n_base, n_hist, strategy, pairwise mode name, default series/GPU)
    "c3": dict(config=3, n_cur=60, n_base=1440, n_hist=0, canary=True, mode="ANY", series=1_000_000,
               desc="config 3: 1M series x (60-pt canary vs 1440-pt baseline), MA + Mann-Whitney + Wilcoxon"),
    "c2": dict(config=2, n_cur=60, n_base=1440, n_hist=10080, canary=True, mode="MANN_WHITE", series=10_000,
               desc="config 2: 10k series, 60-pt canary vs 1440 baseline, MA on 10080-pt history + Mann-Whitney"),
    "c5": dict(config=5, n_cur=60, n_base=0, n_hist=1440, canary=False, mode="NONE", series=1_000_000,
               desc="config 5 slice: MA-only bounds on 1440-pt history + roll-up"),
}


def algorithmic_bytes(w, rec_bytes):
    """4*(n_cur + n_base + n_hist') + R, each sample counted once."""
    return 4 * (w["n_cur"] + w["n_base"] + w["n_hist"]) + rec_bytes


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([time.perf_counter()] + [x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        return self.window()

    def window(self, t0=None, t1=None):
        """Clock record of the samples that arrived in [t0, t1] (all of them by default); a window
        shorter than the sampling period falls back to the samples around it."""
        rows = [r for r in self.rows if (t0 is None or r[0] >= t0 - 0.15) and (t1 is None or r[0] <= t1 + 0.15)]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            r = r[1:]
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(kernel, workload, n_series):
    """DRAM bytes of one launch from the committed ncu capture of that kernel ON THAT WORKLOAD
    (None if this combination was not captured)."""
    try:
        t = None
        for rnd in ("r02", "r01"):
            path = os.path.join(ROOT, "profiles", rnd, "traffic.json")
            if os.path.exists(path):
                with open(path) as f:
                    t = json.load(f).get(kernel, {}).get(workload)
            if t is not None:
                break
        if t is None:
            return None
        return int(t["bytes_per_series"]) * int(n_series)
    except Exception:
        return None


def params_for(w):
    from foremast_1_b200 import _abi as A
    g, table = A.default_params()
    g.pairwise_mode = getattr(A, "FM_PW_" + w["mode"])
    return g, table


def cpu_baseline(w, budget_s=12.0, threads=None):
    """The oracle (a port: the reference tree holds no implementation) on the host cores,
    on a bounded sample of the same synthetic workload."""
    import oracle_lib as O
    from foremast_1_b200 import _abi as A
    from foremast_1_b200.engine import synth_desc
    g, table = params_for(w)
    threads = threads or os.cpu_count() or 1
    strat = A.FM_STRAT_CANARY if w["canary"] else A.FM_STRAT_HISTORICAL
    hist_is_base = w["canary"] and w["n_hist"] == 0
    d = synth_desc(w["config"], 32 * threads, w["n_cur"], w["n_base"], w["n_hist"])
    secs, _ = O.bench_synth(g, table, d, strat, hist_is_base, threads)
    rate = d.n_series / max(secs, 1e-9)
    n = int(max(32 * threads, min(rate * budget_s, 4_000_000)))
    d = synth_desc(w["config"], n, w["n_cur"], w["n_base"], w["n_hist"])
    secs, chk = O.bench_synth(g, table, d, strat, hist_is_base, threads)
    return {"value": n / secs, "unit": "series/s", "cores": threads, "kind": "port",
            "sample": f"{n} series of the same synthetic workload (ids 0..{n - 1}), scoring time only, "
                      f"{secs:.2f} s on {threads} threads"}, n, secs


def run_reference(args, w, rank):
    """--impl reference: the reference's CPU path = the oracle port, all host threads."""
    if rank != 0:
        return
    import oracle_lib as O
    from foremast_1_b200 import _abi as A
    from foremast_1_b200.engine import synth_desc
    g, table = params_for(w)
    threads = os.cpu_count() or 1
    strat = A.FM_STRAT_CANARY if w["canary"] else A.FM_STRAT_HISTORICAL
    hist_is_base = w["canary"] and w["n_hist"] == 0
    d = synth_desc(w["config"], 32 * threads, w["n_cur"], w["n_base"], w["n_hist"])
    secs, _ = O.bench_synth(g, table, d, strat, hist_is_base, threads)
    rate = d.n_series / max(secs, 1e-9)
    total_steps = args.steps + args.warmup
    per_step_s = max(1.0, min(8.0, 90.0 / max(total_steps, 1)))
    n = int(max(32 * threads, rate * per_step_s))
    d = synth_desc(w["config"], n, w["n_cur"], w["n_base"], w["n_hist"])
    for _ in range(args.warmup):
        O.bench_synth(g, table, d, strat, hist_is_base, threads)
    tot = 0.0
    for _ in range(args.steps):
        s, _ = O.bench_synth(g, table, d, strat, hist_is_base, threads)
        tot += s
    v = n * args.steps / tot
    line = {"impl": "reference", "metric": "series scored/sec", "value": v, "unit": "series/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["desc"], "series_per_step": n, "n_cur": w["n_cur"],
                       "n_base": w["n_base"], "n_hist": w["n_hist"], "pairwise": w["mode"]},
            "cpu_baseline": {"value": v, "unit": "series/s", "cores": threads, "kind": "port",
                             "sample": f"{n} series per step, oracle port (reference holds no brain code)"},
            "e2e": {"value": v, "unit": "series/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def canary_kernel_name(n_cur, n_base):
    """Which kernel plan_score (csrc/fm_api.cu) picks for a canary batch of this shape."""
    if n_cur <= 63 and n_base <= 2048:
        return "k_canary_count"
    if n_cur <= 63 and n_base <= 16320:
        return "k_canary_team"
    if n_cur <= 64 and n_base <= 2048:
        r = 16 if n_base <= 512 else 32 if n_base <= 1024 else 64
        return "k_canary_sort<%d, %d>" % (r, 3 * r // 4 if n_base <= 24 * r else r)
    return "k_score_series"


def secondary(ctx, sampler, peak, steps=5, warmup=3):
    """The other BASELINE.json configs under the same clock sampler as the headline (device-resident
    rows, CUDA events on the library stream, inputs larger than L2): config 2 (10k series, 10080-point
    history; and read literally, Mann-Whitney against a 10080-point baseline), config 4 (100k services
    through fm_score_groups_device) and config 5 (MA-only, 1M series, four window lengths, WITH the
    series -> API -> service roll-up: 4 metrics x 8 APIs per service)."""
    import ctypes as C
    import numpy as np
    from foremast_1_b200 import _abi as A
    from foremast_1_b200.engine import synth_desc
    out = {}
    g0, table = A.default_params()

    def timed(run, t):
        for _ in range(warmup):
            run()
        ctx.sync()
        # at least `steps` runs and at least ~0.25 s of them, so that the clock sampler sees the case
        # under load (a 0.1 ms kernel timed five times would be over before its first sample)
        n = max(steps, min(2000, int(250.0 / max(t.main_kernel_ms, 1e-3))))
        t0 = time.perf_counter()
        ms = []
        for _ in range(n):
            run()
            ms.append(t.main_kernel_ms)
        ctx.sync()
        return statistics.median(ms), dict(sampler.window(t0, time.perf_counter()), runs=n)

    def series_case(name, config, S, n_cur, n_base, n_hist, mode, canary, compact, rollup):
        g, _ = A.default_params()
        g.pairwise_mode = getattr(A, "FM_PW_" + mode)
        ctx.set_params(g, table)
        d = synth_desc(config, S, n_cur, n_base, n_hist)
        ptrs = [ctx.dev_alloc(S * d.cur_pitch * 4), ctx.dev_alloc(S * d.base_pitch * 4) if n_base else None,
                ctx.dev_alloc(S * d.hist_pitch * 4) if n_hist else None, ctx.dev_alloc(S)]
        rdt = A.COMPACT_REC if compact else A.SERIES_REC
        d_out = ctx.dev_alloc(S * rdt.itemsize)
        ctx.synth_device(d, *ptrs)
        b = A.Batch()
        b.struct_size = C.sizeof(A.Batch)
        b.strategy = A.FM_STRAT_CANARY if canary else A.FM_STRAT_HISTORICAL
        b.n_series = S
        b.cur, b.n_cur, b.cur_pitch = ptrs[0], n_cur, d.cur_pitch
        if n_base:
            b.base, b.n_base, b.base_pitch = ptrs[1], n_base, d.base_pitch
        if n_hist:
            b.hist, b.n_hist, b.hist_pitch = ptrs[2], n_hist, d.hist_pitch
        b.metric_type = ptrs[3]
        b.record_kind = A.FM_REC_COMPACT if compact else A.FM_REC_FULL
        extra = []
        d_api = d_svc = None
        if rollup:
            api = (np.arange(S) // 4).astype(np.uint32)
            svc = (np.arange(S // 4) // 8).astype(np.uint32)
            p_api, p_svc = ctx.dev_alloc(api.nbytes), ctx.dev_alloc(svc.nbytes)
            ctx.h2d(p_api, api); ctx.h2d(p_svc, svc)
            b.api_id, b.api_service, b.n_apis, b.n_services = p_api, p_svc, S // 4, S // 32
            d_api, d_svc = ctx.dev_alloc(S // 4 * 16), ctx.dev_alloc(S // 32 * 16)
            extra = [p_api, p_svc, d_api, d_svc]
        t = A.Timing()
        ms, clk = timed(lambda: ctx.score_device(b, d_out, None, d_api, d_svc, timing=t), t)
        bps = 4 * (n_cur + n_base + n_hist) + rdt.itemsize
        res = {"series": S, "n_cur": n_cur, "n_base": n_base, "n_hist": n_hist, "pairwise": mode,
               "kernel": canary_kernel_name(n_cur, n_base) if canary else
               "k_ma_only<%d>" % (16 if n_hist >= 1200 else 8 if n_hist >= 400 else 4),
               "ms": ms, "series_per_s": S / (ms / 1e3), "bytes_per_series": bps,
               "achieved_GBps": S * bps / (ms / 1e3) / 1e9, "frac": S * bps / (ms / 1e3) / 1e9 / peak,
               "launches_per_step": int(t.launches), "clocks": clk}
        if rollup:
            sv = np.zeros(S // 32, A.ROLLUP_REC)
            ctx.d2h(sv, d_svc)
            res["rollup"] = {"apis": S // 4, "services": S // 32,
                             "service_severity_histogram": np.bincount(sv["severity"], minlength=3).tolist()}
        for q in ptrs + [d_out] + extra:
            ctx.dev_free(q)
        out[name] = res

    series_case("c2", 2, 10_000, 60, 1440, 10080, "MANN_WHITE", True, False, False)
    series_case("c2_base10080", 2, 10_000, 60, 10080, 0, "MANN_WHITE", True, False, False)
    for nh in (60, 240, 1440, 10080):
        series_case("c5_%d" % nh, 5, 1_000_000, 60, 0, nh, "NONE", False, True, True)

    # config 4: 100k services x 3..8 pod groups x 60 integer counts, Kruskal-Wallis + per-group bounds
    V = 100_000
    rng = np.random.default_rng(44)
    K = rng.integers(3, 9, V)
    off = np.concatenate([[0], np.cumsum(K)]).astype(np.uint32)
    G = int(off[-1])
    svc_of = np.repeat(np.arange(V, dtype=np.uint64), K)
    ids = (svc_of << np.uint64(4)) + (np.arange(G, dtype=np.uint64) - off[:-1].astype(np.uint64)[svc_of.astype(np.int64)])
    hids = np.arange(V, dtype=np.uint64) << np.uint64(4)
    g, _ = A.default_params()
    g.pairwise_mode = A.FM_PW_KRUSKAL
    ctx.set_params(g, table)
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
    t = A.Timing()
    ms, clk = timed(lambda: ctx.score_groups_device(gb, d_out, d_svc, t), t)
    bytes_alg = 4 * (G * 60 + V * 1440) + G * 96 + V * 16
    out["c4"] = {"services": V, "groups": G, "kernel": "k_groups_sort + k_score_groups (deferred list)", "ms": ms,
                 "services_per_s": V / (ms / 1e3), "bytes_per_service": bytes_alg / V,
                 "achieved_GBps": bytes_alg / (ms / 1e3) / 1e9, "frac": bytes_alg / (ms / 1e3) / 1e9 / peak,
                 "launches_per_step": int(t.launches), "clocks": clk}
    # config 4 end to end: pinned host rows through fm_score_groups
    try:
        from foremast_1_b200.engine import Pinned
        from foremast_1_b200 import _ffi
        p_cur = Pinned((G, dg.cur_pitch), np.float32)
        p_hist = Pinned((V, dh.hist_pitch), np.float32)
        p_mt = Pinned((V,), np.uint8)
        p_off = Pinned((V + 1,), np.uint32)
        p_rec = Pinned((G,), A.SERIES_REC)
        p_svc = Pinned((V,), A.ROLLUP_REC)
        p_off.array[:] = off
        hg = synth_desc(4, G, 60, 0, 0, row_ids=ids.ctypes.data)
        hh = synth_desc(4, V, 60, 0, 1440, row_ids=hids.ctypes.data)
        lib = _ffi.lib()
        assert lib.fm_synth_fill_host(C.byref(hg), p_cur.ptr, None, None, None, 0) == 0
        assert lib.fm_synth_fill_host(C.byref(hh), None, None, p_hist.ptr, p_mt.ptr, 0) == 0
        hb = A.GroupBatch()
        hb.struct_size = C.sizeof(A.GroupBatch)
        hb.n_services, hb.n_groups, hb.group_offsets = V, G, p_off.ptr
        hb.cur, hb.n_cur, hb.cur_pitch = p_cur.ptr, 60, dg.cur_pitch
        hb.hist, hb.n_hist, hb.hist_pitch = p_hist.ptr, 1440, dh.hist_pitch
        hb.metric_type = p_mt.ptr
        te = A.Timing()

        def e2e_run():
            _ffi.check(lib.fm_score_groups(ctx._h, C.byref(hb), C.c_void_p(p_rec.ptr), C.c_void_p(p_svc.ptr), C.byref(te)), ctx._h)
        e2e_run()
        w0 = time.perf_counter()
        for _ in range(3):
            e2e_run()
        e_s = (time.perf_counter() - w0) / 3
        out["c4"]["e2e"] = {"services_per_s": V / e_s, "ms_per_step": 1e3 * e_s, "h2d_bytes_per_step": int(te.h2d_bytes),
                            "d2h_bytes_per_step": int(te.d2h_bytes), "launches_per_step": int(te.launches)}
        for pb in (p_cur, p_hist, p_mt, p_off, p_rec, p_svc):
            pb.free()
    except Exception as e:  # the device-resident numbers stand on their own
        out["c4"]["e2e"] = {"error": f"{type(e).__name__}: {e}"}
    for q in (d_ids, d_hids, d_cur, d_hist, d_mt, d_off, d_out, d_svc):
        ctx.dev_free(q)
    ctx.set_params(g0, table)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--series", type=int, default=0, help="series per GPU (default: the workload's)")
    ap.add_argument("--e2e-series", type=int, default=0, help="series per fm_score_batch call (default: --series)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the config 2 / 4 / 5 block")
    ap.add_argument("--mode", default="", help="override the workload's pairwise mode (diagnostics only)")
    ap.add_argument("--n-hist", type=int, default=-1, help="override the history length (config 5 sweep)")
    ap.add_argument("--compact", action="store_true", help="16-byte records (config 5 sweep)")
    args = ap.parse_args()
    w = dict(WORKLOADS[args.workload])
    if args.mode:
        w["mode"] = args.mode
    if args.n_hist >= 0:
        w["n_hist"] = args.n_hist
        w["desc"] += f" [n_hist={args.n_hist}]"
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, w, rank)
        return
    if args.warmup < 3:
        args.warmup = 3

    import ctypes as C
    import numpy as np
    import torch
    from foremast_1_b200 import _abi as A
    from foremast_1_b200.engine import Context, Pinned, synth_desc

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    numa_cpus = None
    if world > 1:   # every rank next to its own GPU's PCIe root before any pinned buffer exists
        from foremast_1_b200.shard import bind_host_to_gpu
        pr = torch.cuda.get_device_properties(local)
        if hasattr(pr, "pci_bus_id"):
            numa_cpus = bind_host_to_gpu("%04x:%02x:%02x.0" % (getattr(pr, "pci_domain_id", 0), pr.pci_bus_id,
                                                               getattr(pr, "pci_device_id", 0)))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def rank_max(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    S = args.series or w["series"]
    ctx = Context(local)
    g, table = params_for(w)
    ctx.set_params(g, table)
    strat = A.FM_STRAT_CANARY if w["canary"] else A.FM_STRAT_HISTORICAL
    d = synth_desc(w["config"], S, w["n_cur"], w["n_base"], w["n_hist"], first_series=rank * S)
    d_cur = ctx.dev_alloc(S * d.cur_pitch * 4)
    d_base = ctx.dev_alloc(S * d.base_pitch * 4) if w["n_base"] else None
    d_hist = ctx.dev_alloc(S * d.hist_pitch * 4) if w["n_hist"] else None
    d_mt = ctx.dev_alloc(S)
    rec_dtype = A.COMPACT_REC if args.compact else A.SERIES_REC
    d_out = ctx.dev_alloc(S * rec_dtype.itemsize)
    ctx.synth_device(d, d_cur, d_base, d_hist, d_mt)
    ctx.sync()

    b = A.Batch()
    b.struct_size = C.sizeof(A.Batch)
    b.strategy = strat
    b.n_series = S
    b.cur, b.n_cur, b.cur_pitch = d_cur, w["n_cur"], d.cur_pitch
    if d_base:
        b.base, b.n_base, b.base_pitch = d_base, w["n_base"], d.base_pitch
    if d_hist:
        b.hist, b.n_hist, b.hist_pitch = d_hist, w["n_hist"], d.hist_pitch
    b.metric_type = d_mt
    b.record_kind = A.FM_REC_COMPACT if args.compact else A.FM_REC_FULL

    t = A.Timing()
    for _ in range(args.warmup):
        ctx.score_device(b, d_out, timing=t)
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    wall0 = time.perf_counter()
    dev_ms, main_ms, launches = 0.0, 0.0, 0
    for _ in range(args.steps):
        ctx.score_device(b, d_out, timing=t)
        dev_ms += t.kernel_ms
        main_ms += t.main_kernel_ms
        launches += t.launches
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - wall0)
    clocks = sampler.stop()
    dev_ms_max = rank_max(dev_ms)
    main_ms_max = rank_max(main_ms)
    value = world * S * args.steps / (dev_ms_max / 1e3)

    # sanity: verdict histogram of the last step (also proves the kernel wrote results)
    rec = np.zeros(min(S, 65536), rec_dtype)
    ctx.d2h(rec, d_out)
    hist_v = np.bincount(rec["verdict"], minlength=3).tolist()

    # ---- end to end: pinned host rows through fm_score_batch ----------------
    e2e = None
    if not args.no_e2e:
        Se = args.e2e_series or S
        try:
            import psutil
            need = Se * 4 * (d.cur_pitch + d.base_pitch + d.hist_pitch) * max(1, world)
            if need * 3 > psutil.virtual_memory().available:
                Se = max(4096, int(Se * psutil.virtual_memory().available / (need * 3)))
        except Exception:
            pass
        de = synth_desc(w["config"], Se, w["n_cur"], w["n_base"], w["n_hist"], first_series=rank * S)
        p_cur = Pinned((Se, de.cur_pitch), np.float32)
        p_base = Pinned((Se, de.base_pitch), np.float32) if w["n_base"] else None
        p_hist = Pinned((Se, de.hist_pitch), np.float32) if w["n_hist"] else None
        p_mt = Pinned((Se,), np.uint8)
        p_out = Pinned((Se,), rec_dtype)
        from foremast_1_b200 import _ffi
        rc = _ffi.lib().fm_synth_fill_host(C.byref(de), p_cur.ptr, p_base.ptr if p_base else None,
                                           p_hist.ptr if p_hist else None, p_mt.ptr, 0)
        assert rc == 0
        hb = A.Batch()
        hb.struct_size = C.sizeof(A.Batch)
        hb.strategy = strat
        hb.n_series = Se
        hb.cur, hb.n_cur, hb.cur_pitch = p_cur.ptr, w["n_cur"], de.cur_pitch
        if p_base:
            hb.base, hb.n_base, hb.base_pitch = p_base.ptr, w["n_base"], de.base_pitch
        if p_hist:
            hb.hist, hb.n_hist, hb.hist_pitch = p_hist.ptr, w["n_hist"], de.hist_pitch
        hb.metric_type = p_mt.ptr
        hb.record_kind = b.record_kind
        te = A.Timing()
        ctx.score_host(hb, p_out.ptr, timing=te)   # warm-up (arena growth, first touch)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            ctx.score_host(hb, p_out.ptr, timing=te)
            launches_e2e = te.launches
        barrier()
        e_s = rank_max(time.perf_counter() - t0)
        # the first records of the e2e pass must equal the device-resident pass
        same = bool((p_out.array[: rec.size] == rec[: min(rec.size, Se)]).all()) if Se >= rec.size else None
        e2e = {"value": world * Se * args.e2e_steps / e_s, "unit": "series/s",
               "h2d_bytes_per_step": int(te.h2d_bytes), "d2h_bytes_per_step": int(te.d2h_bytes),
               "series_per_step_per_gpu": Se, "steps": args.e2e_steps, "ms_per_step": 1e3 * e_s / args.e2e_steps,
               "launches_per_step": int(launches_e2e), "matches_device_pass": same}
        # second poll of the same jobs: baselines resident in HBM (fm_score_batch_cached), only the
        # current rows cross PCIe
        try:
            p_ids = Pinned((Se,), np.uint64)
            p_ids.array[:] = (np.arange(Se, dtype=np.uint64) + np.uint64(rank * S + 1)) * np.uint64(0x9E3779B97F4A7C15)
            ctx.cache_configure(Se * 4 * (de.base_pitch + de.hist_pitch) + (64 << 20))
            tw = A.Timing()
            ctx.score_host(hb, p_out.ptr, timing=tw, row_ids=p_ids.ptr)       # cold: fills the cache
            cold_same = bool((p_out.array[: rec.size] == rec[: min(rec.size, Se)]).all()) if Se >= rec.size else None
            ctx.score_host(hb, p_out.ptr, timing=tw, row_ids=p_ids.ptr)       # warm-up of the warm path
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.e2e_steps):
                ctx.score_host(hb, p_out.ptr, timing=tw, row_ids=p_ids.ptr)
            barrier()
            w_s = rank_max(time.perf_counter() - t0)
            st = ctx.cache_stats()
            warm_same = bool((p_out.array[: rec.size] == rec[: min(rec.size, Se)]).all()) if Se >= rec.size else None
            e2e["warm"] = {"value": world * Se * args.e2e_steps / w_s, "unit": "series/s",
                           "what": "re-poll of the same row ids through fm_score_batch_cached: baselines resident in HBM",
                           "h2d_bytes_per_step": int(tw.h2d_bytes), "d2h_bytes_per_step": int(tw.d2h_bytes),
                           "ms_per_step": 1e3 * w_s / args.e2e_steps, "launches_per_step": int(tw.launches),
                           "kernel_ms_per_step": float(tw.kernel_ms), "resident_rows": st["resident_rows"],
                           "matches_device_pass": warm_same and cold_same}
            ctx.cache_configure(0)
            p_ids.free()
        except Exception as e:  # the cold number stands on its own
            e2e["warm"] = {"error": f"{type(e).__name__}: {e}"}
        for pbuf in (p_cur, p_base, p_hist, p_mt, p_out):
            if pbuf:
                pbuf.free()

    sec = None
    if world == 1 and not args.no_secondary and args.workload == "c3" and args.n_hist < 0 and not args.mode:
        for q in (d_cur, d_base, d_hist, d_mt, d_out):
            if q:
                ctx.dev_free(q)
        s2 = ClockSampler(local)
        s2.start()
        time.sleep(0.3)
        try:
            sec = secondary(ctx, s2, measured_peak()[0])
        except Exception as e:  # the headline stands on its own
            sec = {"error": f"{type(e).__name__}: {e}"}
        s2.stop()

    if rank == 0:
        peak, peak_src = measured_peak()
        rec_bytes = rec_dtype.itemsize
        bps = algorithmic_bytes(w, rec_bytes)
        achieved = S * bps / (main_ms_max / args.steps / 1e3) / 1e9
        kernel = (canary_kernel_name(w["n_cur"], w["n_base"]) if w["canary"] else
                  ("k_ma_only<%d>" % (16 if w["n_hist"] >= 1200 else 8 if w["n_hist"] >= 400 else 4)))
        traffic = measured_traffic(kernel, args.workload, S) if (args.n_hist < 0 and not args.mode) else None
        line = {
            "metric": "series scored/sec", "value": value, "unit": "series/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": w["desc"], "series_per_gpu": S, "n_cur": w["n_cur"], "n_base": w["n_base"],
                       "n_hist": w["n_hist"], "pairwise": w["mode"], "sharding": "by series id, no collective",
                       "l2": "inputs (%.1f GB/GPU) larger than L2" % (S * (bps - rec_bytes) / 1e9),
                       "timing": "CUDA events on the library stream, per step, max over ranks",
                       "host_numa_binding": (len(numa_cpus) if numa_cpus else 0)},
            "wall_ms_per_step": wall_ms / args.steps,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "bytes_per_series": bps,
                         "kernel": kernel,
                         "kernel_ms": main_ms_max / args.steps},
            "verdict_histogram_first_64k": hist_v,
        }
        if e2e is not None:
            line["e2e"] = e2e
        if sec is not None:
            line["secondary"] = sec
        if world == 1 and not args.no_cpu:
            cb, _, _ = cpu_baseline(w)
            line["cpu_baseline"] = cb
        print(json.dumps(line), flush=True)
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
