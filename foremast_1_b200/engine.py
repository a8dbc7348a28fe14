"""Scoring context over the C ABI: numpy rows in, verdict records out.

This is the thin host layer the brain's job loop (foremast_1_b200/brain.py) and
the tests call; it adds nothing to the arithmetic — every number comes from
libforemast_cuda.so.
"""
import ctypes as C

import numpy as np

from . import _abi as A
from . import _ffi


def pack_rows(rows, n=None):
    """Ragged list of 1-D float arrays -> (matrix [S x pitch] fp32, lens uint32).

    NaNs (Prometheus gaps) are compacted out, as the ABI requires NaN-free rows.
    """
    rows = [np.asarray(r, dtype=np.float32).ravel() for r in rows]
    rows = [r[~np.isnan(r)] for r in rows]
    lens = np.array([r.size for r in rows], dtype=np.uint32)
    width = int(lens.max()) if len(rows) and n is None else int(n or 0)
    pitch = max(4, A.row_pitch(width))
    m = np.zeros((len(rows), pitch), dtype=np.float32)
    for i, r in enumerate(rows):
        m[i, : min(r.size, width)] = r[:width]
    return m, np.minimum(lens, width).astype(np.uint32), width


def _as_matrix(x):
    """2-D array -> C-contiguous fp32 matrix whose pitch is a multiple of 4, 16-B aligned."""
    if x is None:
        return None, 0
    x = np.asarray(x, dtype=np.float32)
    if x.ndim == 1:
        x = x[None, :]
    n = x.shape[1]
    pitch = max(4, A.row_pitch(n))
    if pitch != n or not x.flags.c_contiguous or (x.ctypes.data & 15):
        m = np.zeros((x.shape[0], pitch), dtype=np.float32)
        m[:, :n] = x
        x = m
    return x, n


class Context:
    """One GPU, one stream, grow-only device arenas (fm_ctx)."""

    def __init__(self, device=0):
        self._lib = _ffi.lib()
        h = C.c_void_p()
        rc = self._lib.fm_ctx_create(device, C.byref(h))
        _ffi.check(rc, None)
        self._h = h
        self.device = device
        self.g, self.table = A.default_params()
        self._pinned = {}          # role -> Pinned, grow-only (pinned_rows)

    def close(self):
        if getattr(self, "_h", None):
            for p in self._pinned.values():
                p.free()
            self._pinned = {}
            self._lib.fm_ctx_destroy(self._h)
            self._h = None

    def pinned_rows(self, role, shape):
        """A zeroed fp32 matrix of `shape` in pinned host memory (fm_alloc_pinned), reused between
        calls with the same role: what the job loop packs fetched rows into, so that fm_score_batch
        copies them to the device without a staging copy."""
        need = int(shape[0]) * int(shape[1])
        p = self._pinned.get(role)
        if p is None or p.array.size < need:
            if p is not None:
                p.free()
            p = Pinned((max(need, 1024) * 3 // 2,), np.float32)
            self._pinned[role] = p
        m = p.array[:need].reshape(shape)
        m[...] = 0
        return m

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        _ffi.check(rc, self._h)

    # ---- parameters -----------------------------------------------------
    def set_params(self, g, table):
        self._ck(self._lib.fm_set_params(self._h, C.byref(g), table, len(table)))
        self.g, self.table = g, table

    # ---- host-row scoring -------------------------------------------------
    def score(self, cur, base=None, hist=None, *, strategy=A.FM_STRAT_CANARY, cur_len=None,
              base_len=None, hist_len=None, metric_type=None, api_id=None, api_service=None,
              n_apis=0, n_services=0, record_kind=A.FM_REC_FULL, want_aux=False, row_ids=None):
        """row_ids (uint64 per series): score through the device-resident row cache
        (fm_score_batch_cached; see cache_configure)."""
        cur, n_cur = _as_matrix(cur)
        base, n_base = _as_matrix(base)
        hist, n_hist = _as_matrix(hist)
        S = cur.shape[0]
        keep = [cur, base, hist]
        b = A.Batch()
        b.struct_size = C.sizeof(A.Batch)
        b.strategy = strategy
        b.n_series = S
        b.cur, b.n_cur, b.cur_pitch = A.ptr(cur), n_cur, cur.shape[1]
        if base is not None:
            b.base, b.n_base, b.base_pitch = A.ptr(base), n_base, base.shape[1]
        if hist is not None:
            b.hist, b.n_hist, b.hist_pitch = A.ptr(hist), n_hist, hist.shape[1]

        def u32(x):
            if x is None:
                return None
            x = np.ascontiguousarray(x, dtype=np.uint32)
            keep.append(x)
            return x

        cl, bl, hl = u32(cur_len), u32(base_len), u32(hist_len)
        b.cur_len, b.base_len, b.hist_len = A.ptr(cl), A.ptr(bl), A.ptr(hl)
        mt = None if metric_type is None else np.ascontiguousarray(metric_type, dtype=np.uint8)
        b.metric_type = A.ptr(mt)
        keep.append(mt)
        ai, asv = u32(api_id), u32(api_service)
        b.api_id, b.api_service = A.ptr(ai), A.ptr(asv)
        b.n_apis, b.n_services = int(n_apis), int(n_services)
        b.record_kind = record_kind
        rec = np.zeros(S, A.COMPACT_REC if record_kind == A.FM_REC_COMPACT else A.SERIES_REC)
        aux = np.zeros(S, A.RANK_AUX) if want_aux else None
        api = np.zeros(int(n_apis), A.ROLLUP_REC) if ai is not None else None
        svc = np.zeros(int(n_services), A.ROLLUP_REC) if (ai is not None and asv is not None) else None
        t = A.Timing()
        if row_ids is None:
            self._ck(self._lib.fm_score_batch(self._h, C.byref(b), A.ptr(rec), A.ptr(aux), A.ptr(api),
                                              A.ptr(svc), C.byref(t)))
        else:
            ids = np.ascontiguousarray(row_ids, dtype=np.uint64)
            assert ids.size == S
            keep.append(ids)
            self._ck(self._lib.fm_score_batch_cached(self._h, C.byref(b), A.ptr(ids), A.ptr(rec), A.ptr(aux),
                                                     A.ptr(api), A.ptr(svc), C.byref(t)))
        return {"rec": rec, "aux": aux, "api": api, "svc": svc, "timing": t, "batch": b, "_keep": keep}

    # ---- device-resident row cache (MAX_CACHE_SIZE) ---------------------------
    def cache_configure(self, max_bytes):
        self._ck(self._lib.fm_cache_configure(self._h, int(max_bytes)))

    def cache_stats(self):
        st = A.CacheStats()
        self._ck(self._lib.fm_cache_get_stats(self._h, C.byref(st)))
        return {n: int(getattr(st, n)) for n, _ in A.CacheStats._fields_}

    def cache_query(self, row_ids):
        ids = np.ascontiguousarray(row_ids, dtype=np.uint64)
        res = np.zeros(ids.size, np.uint8)
        self._ck(self._lib.fm_cache_query(self._h, A.ptr(ids), ids.size, A.ptr(res)))
        return res.astype(bool)

    def score_groups(self, cur, group_offsets, hist, *, cur_len=None, hist_len=None, metric_type=None):
        cur, n_cur = _as_matrix(cur)
        hist, n_hist = _as_matrix(hist)
        off = np.ascontiguousarray(group_offsets, dtype=np.uint32)
        V, G = off.size - 1, cur.shape[0]
        b = A.GroupBatch()
        b.struct_size = C.sizeof(A.GroupBatch)
        b.n_services, b.n_groups = V, G
        b.group_offsets = A.ptr(off)
        b.cur, b.n_cur, b.cur_pitch = A.ptr(cur), n_cur, cur.shape[1]
        cl = None if cur_len is None else np.ascontiguousarray(cur_len, dtype=np.uint32)
        hl = None if hist_len is None else np.ascontiguousarray(hist_len, dtype=np.uint32)
        mt = None if metric_type is None else np.ascontiguousarray(metric_type, dtype=np.uint8)
        b.cur_len, b.hist_len, b.metric_type = A.ptr(cl), A.ptr(hl), A.ptr(mt)
        if hist is not None:
            b.hist, b.n_hist, b.hist_pitch = A.ptr(hist), n_hist, hist.shape[1]
        rec = np.zeros(G, A.SERIES_REC)
        svc = np.zeros(V, A.ROLLUP_REC)
        t = A.Timing()
        self._ck(self._lib.fm_score_groups(self._h, C.byref(b), A.ptr(rec), A.ptr(svc), C.byref(t)))
        return {"rec": rec, "svc": svc, "timing": t, "batch": b, "_keep": [cur, hist, off, cl, hl, mt]}

    def score_pairs(self, cur_x, cur_y, hist_x, hist_y, *, cur_len=None, hist_len=None, metric_type=None):
        """Two metrics judged jointly (bivariate normal, fm_score_pairs): rows of pair p are
        cur_x[p], cur_y[p], hist_x[p], hist_y[p]."""
        cx, n_cur = _as_matrix(cur_x)
        cy, _ = _as_matrix(cur_y)
        hx, n_hist = _as_matrix(hist_x)
        hy, _ = _as_matrix(hist_y)
        P = cx.shape[0]
        b = A.PairBatch()
        b.struct_size = C.sizeof(A.PairBatch)
        b.n_pairs = P
        b.cur_x, b.cur_y, b.n_cur, b.cur_pitch = A.ptr(cx), A.ptr(cy), n_cur, cx.shape[1]
        b.hist_x, b.hist_y, b.n_hist, b.hist_pitch = A.ptr(hx), A.ptr(hy), n_hist, hx.shape[1]
        cl = None if cur_len is None else np.ascontiguousarray(cur_len, dtype=np.uint32)
        hl = None if hist_len is None else np.ascontiguousarray(hist_len, dtype=np.uint32)
        mt = None if metric_type is None else np.ascontiguousarray(metric_type, dtype=np.uint8)
        b.cur_len, b.hist_len, b.metric_type = A.ptr(cl), A.ptr(hl), A.ptr(mt)
        rec = np.zeros(P, A.PAIR_REC)
        t = A.Timing()
        self._ck(self._lib.fm_score_pairs(self._h, C.byref(b), A.ptr(rec), C.byref(t)))
        return {"rec": rec, "timing": t, "batch": b, "_keep": [cx, cy, hx, hy, cl, hl, mt]}

    def rank_dump(self, cur, base):
        cur = np.ascontiguousarray(cur, dtype=np.float32)
        base = np.ascontiguousarray(base, dtype=np.float32)
        r2 = np.zeros(cur.size + base.size, np.uint32)
        self._ck(self._lib.fm_rank_dump(self._h, A.ptr(cur), cur.size, A.ptr(base), base.size, A.ptr(r2)))
        return r2

    def ma_window_series(self, x, w, ddof=0):
        x = np.ascontiguousarray(x, dtype=np.float32)
        m = np.empty(x.size, np.float64)
        s = np.empty(x.size, np.float64)
        self._ck(self._lib.fm_ma_window_series(self._h, A.ptr(x), x.size, w, ddof, A.ptr(m), A.ptr(s)))
        return m, s

    # ---- device-resident helpers (bench, large parity runs) ----------------
    def dev_alloc(self, nbytes):
        p = C.c_void_p()
        self._ck(self._lib.fm_device_alloc(self._h, int(nbytes), C.byref(p)))
        return p.value

    def dev_free(self, p):
        if p:
            self._ck(self._lib.fm_device_free(self._h, C.c_void_p(p)))

    def h2d(self, dptr, arr):
        arr = np.ascontiguousarray(arr)
        self._ck(self._lib.fm_memcpy_h2d(self._h, C.c_void_p(dptr), A.ptr(arr), arr.nbytes))

    def d2h(self, arr, dptr):
        assert arr.flags.c_contiguous
        self._ck(self._lib.fm_memcpy_d2h(self._h, A.ptr(arr), C.c_void_p(dptr), arr.nbytes))

    def sync(self):
        self._ck(self._lib.fm_ctx_sync(self._h))

    def synth_device(self, desc, d_cur, d_base, d_hist, d_mt):
        self._ck(self._lib.fm_synth_fill_device(self._h, C.byref(desc), C.c_void_p(d_cur), C.c_void_p(d_base),
                                                C.c_void_p(d_hist), C.c_void_p(d_mt)))

    def score_device(self, batch, d_out, d_aux=None, d_api=None, d_svc=None, timing=None):
        self._ck(self._lib.fm_score_batch_device(self._h, C.byref(batch), C.c_void_p(d_out), C.c_void_p(d_aux),
                                                 C.c_void_p(d_api), C.c_void_p(d_svc),
                                                 C.byref(timing) if timing is not None else None))

    def score_host(self, batch, out, aux=None, api=None, svc=None, timing=None, row_ids=None):
        """fm_score_batch (fm_score_batch_cached with row_ids, a host pointer to uint64 ids) on raw
        host pointers (pinned buffers from alloc_pinned)."""
        tp = C.byref(timing) if timing is not None else None
        if row_ids is None:
            self._ck(self._lib.fm_score_batch(self._h, C.byref(batch), C.c_void_p(out), C.c_void_p(aux),
                                              C.c_void_p(api), C.c_void_p(svc), tp))
        else:
            self._ck(self._lib.fm_score_batch_cached(self._h, C.byref(batch), C.c_void_p(row_ids), C.c_void_p(out),
                                                     C.c_void_p(aux), C.c_void_p(api), C.c_void_p(svc), tp))

    def score_groups_device(self, gbatch, d_out, d_svc=None, timing=None):
        self._ck(self._lib.fm_score_groups_device(self._h, C.byref(gbatch), C.c_void_p(d_out), C.c_void_p(d_svc),
                                                  C.byref(timing) if timing is not None else None))


class Pool:
    """Several GPUs driven from this one process (fm_pool): one ctx and one host thread per device,
    a batch cut into contiguous ranges of series, records written into disjoint slices of one
    output array.  devices=None: every visible GPU."""

    def __init__(self, devices=None):
        self._lib = _ffi.lib()
        h = C.c_void_p()
        if devices is None:
            rc = self._lib.fm_pool_create(None, 0, C.byref(h))
        else:
            arr = (C.c_int32 * len(devices))(*devices)
            rc = self._lib.fm_pool_create(arr, len(devices), C.byref(h))
        _ffi.check(rc, None)
        self._h = h
        self.size = self._lib.fm_pool_size(h)
        self.g, self.table = A.default_params()

    def close(self):
        if getattr(self, "_h", None):
            self._lib.fm_pool_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != A.FM_OK:
            raise _ffi.ForemastCudaError(rc, (self._lib.fm_pool_last_error(self._h) or b"").decode("utf-8", "replace"))

    def set_params(self, g, table):
        self._ck(self._lib.fm_pool_set_params(self._h, C.byref(g), table, len(table)))
        self.g, self.table = g, table

    def cache_configure(self, max_bytes_per_device):
        self._ck(self._lib.fm_pool_cache_configure(self._h, int(max_bytes_per_device)))

    def score_host(self, batch, out, aux=None, api=None, svc=None, row_ids=None):
        """fm_pool_score_batch on raw host pointers; returns the per-device fm_timing array."""
        t = (A.Timing * self.size)()
        self._ck(self._lib.fm_pool_score_batch(self._h, C.byref(batch), C.c_void_p(row_ids), C.c_void_p(out),
                                               C.c_void_p(aux), C.c_void_p(api), C.c_void_p(svc), t))
        return t


def synth_desc(config, n_series, n_cur, n_base, n_hist, first_series=0, seed=None, row_ids=None):
    d = A.SynthDesc()
    d.struct_size = C.sizeof(A.SynthDesc)
    d.config = config
    d.seed = (A.FM_SYNTH_SEED ^ config) if seed is None else seed
    d.first_series = first_series
    d.n_series = n_series
    d.n_cur, d.n_base, d.n_hist = n_cur, n_base, n_hist
    d.cur_pitch, d.base_pitch, d.hist_pitch = (A.row_pitch(n_cur), A.row_pitch(n_base), A.row_pitch(n_hist))
    d.row_ids = row_ids
    return d


def synth_host(desc, want_hist=True, n_threads=0):
    """Synthetic rows on the host via fm_synth_fill_host (no GPU needed)."""
    lib = _ffi.lib()
    S = int(desc.n_series)
    cur = np.zeros((S, desc.cur_pitch), np.float32) if desc.n_cur else None
    base = np.zeros((S, desc.base_pitch), np.float32) if desc.n_base else None
    hist = np.zeros((S, desc.hist_pitch), np.float32) if (desc.n_hist and want_hist) else None
    mt = np.zeros(S, np.uint8)
    rc = lib.fm_synth_fill_host(C.byref(desc), A.ptr(cur), A.ptr(base), A.ptr(hist), A.ptr(mt), n_threads)
    if rc != 0:
        raise _ffi.ForemastCudaError(rc, "fm_synth_fill_host")
    return cur, base, hist, mt


class Pinned:
    """Pinned host buffer from fm_alloc_pinned, viewed as a numpy array."""

    def __init__(self, shape, dtype):
        self._lib = _ffi.lib()
        self.dtype = np.dtype(dtype)
        self.shape = tuple(int(s) for s in np.atleast_1d(shape))
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        p = C.c_void_p()
        rc = self._lib.fm_alloc_pinned(max(nbytes, 1), C.byref(p))
        if rc != 0:
            raise _ffi.ForemastCudaError(rc, "fm_alloc_pinned")
        self.ptr = p.value
        buf = (C.c_char * max(nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self.ptr:
            self.array = None
            self._lib.fm_free_pinned(C.c_void_p(self.ptr))
            self.ptr = None
