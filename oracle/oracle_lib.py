"""ctypes loader for the CPU oracle (oracle/fm_oracle.c) — test infrastructure.

Builds oracle/_build/libfm_oracle.so on first use (gcc, a second or two).
"""
import ctypes as C
import os
import subprocess

import numpy as np

from foremast_1_b200 import _abi as A

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))  # repo root
_SO = os.path.join(_ROOT, "oracle", "_build", "libfm_oracle.so")
_lib = None


def build():
    src = os.path.join(_ROOT, "oracle", "fm_oracle.c")
    deps = [src, os.path.join(_ROOT, "oracle", "fm_oracle.h"),
            os.path.join(_ROOT, "include", "foremast_cuda.h"),
            os.path.join(_ROOT, "include", "fm_synth.h")]
    if os.path.exists(_SO) and all(os.path.getmtime(_SO) >= os.path.getmtime(d) for d in deps):
        return _SO
    subprocess.run(["make", "-C", os.path.join(_ROOT, "oracle")], check=True,
                   stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.fmo_chi2_sf.restype = C.c_double
        _lib.fmo_chi2_sf.argtypes = [C.c_double, C.c_uint32]
        for name in ("fmo_score_batch", "fmo_score_groups", "fmo_bench_synth"):
            getattr(_lib, name).restype = C.c_int32
    return _lib


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def moments(x, ddof=0):
    x = _f32(x)
    m, s = C.c_double(), C.c_double()
    lib().fmo_moments(C.c_void_p(x.ctypes.data), C.c_uint32(x.size), C.c_uint32(ddof),
                      C.byref(m), C.byref(s))
    return m.value, s.value


def ma_window_series(x, w, ddof=0):
    x = _f32(x)
    m = np.empty(x.size, np.float64)
    s = np.empty(x.size, np.float64)
    lib().fmo_ma_window_series(C.c_void_p(x.ctypes.data), C.c_uint32(x.size), C.c_uint32(w),
                               C.c_uint32(ddof), C.c_void_p(m.ctypes.data), C.c_void_p(s.ctypes.data))
    return m, s


def midranks(v):
    v = np.ascontiguousarray(v, dtype=np.float64)
    r2 = np.zeros(v.size, np.uint32)
    T = C.c_int64()
    lib().fmo_midranks(C.c_void_p(v.ctypes.data), C.c_uint32(v.size), C.c_void_p(r2.ctypes.data),
                       C.byref(T))
    return r2, T.value


def rank_dump(cur, base):
    cur, base = _f32(cur), _f32(base)
    r2 = np.zeros(cur.size + base.size, np.uint32)
    lib().fmo_rank_dump(C.c_void_p(cur.ctypes.data), C.c_uint32(cur.size),
                        C.c_void_p(base.ctypes.data), C.c_uint32(base.size),
                        C.c_void_p(r2.ctypes.data))
    return r2


def chi2_sf(x, dof):
    return lib().fmo_chi2_sf(x, dof)


def score_series(g, tp, strategy, cur, base=None, hist=None):
    """One series -> (record as numpy scalar, aux as numpy scalar)."""
    cur, base, hist = _f32(cur), _f32(base), _f32(hist)
    rec = np.zeros(1, A.SERIES_REC)
    aux = np.zeros(1, A.RANK_AUX)
    lib().fmo_score_series(C.byref(g), C.byref(tp), C.c_uint32(strategy),
                           C.c_void_p(cur.ctypes.data if cur is not None and cur.size else None),
                           C.c_uint32(0 if cur is None else cur.size),
                           C.c_void_p(None if base is None else base.ctypes.data),
                           C.c_uint32(0 if base is None else base.size),
                           C.c_void_p(None if hist is None else hist.ctypes.data),
                           C.c_uint32(0 if hist is None else hist.size),
                           C.c_void_p(rec.ctypes.data), C.c_void_p(aux.ctypes.data))
    return rec[0], aux[0]


def score_batch(g, table, batch, n_threads=4, want_aux=True):
    """batch: a _abi.Batch with HOST pointers.  Returns dict of numpy arrays."""
    S = int(batch.n_series)
    rec = np.zeros(S, A.COMPACT_REC if batch.record_kind == A.FM_REC_COMPACT else A.SERIES_REC)
    aux = np.zeros(S, A.RANK_AUX) if want_aux else None
    api = np.zeros(batch.n_apis, A.ROLLUP_REC) if batch.api_id else None
    svc = np.zeros(batch.n_services, A.ROLLUP_REC) if (batch.api_id and batch.api_service) else None
    rc = lib().fmo_score_batch(C.byref(g), table, C.c_uint32(len(table)), C.byref(batch),
                               C.c_void_p(rec.ctypes.data), C.c_void_p(A.ptr(aux)),
                               C.c_void_p(A.ptr(api)), C.c_void_p(A.ptr(svc)), C.c_uint32(n_threads))
    assert rc == 0, rc
    return {"rec": rec, "aux": aux, "api": api, "svc": svc}


def score_pairs(g, table, pbatch):
    """pbatch: a _abi.PairBatch with HOST pointers -> PAIR_REC array."""
    rec = np.zeros(int(pbatch.n_pairs), A.PAIR_REC)
    rc = lib().fmo_score_pairs(C.byref(g), table, C.c_uint32(len(table)), C.byref(pbatch), C.c_void_p(rec.ctypes.data))
    assert rc == 0, rc
    return rec


def score_groups(g, table, gbatch, n_threads=4):
    rec = np.zeros(int(gbatch.n_groups), A.SERIES_REC)
    svc = np.zeros(int(gbatch.n_services), A.ROLLUP_REC)
    rc = lib().fmo_score_groups(C.byref(g), table, C.c_uint32(len(table)), C.byref(gbatch),
                                C.c_void_p(rec.ctypes.data), C.c_void_p(svc.ctypes.data),
                                C.c_uint32(n_threads))
    assert rc == 0, rc
    return {"rec": rec, "svc": svc}


def score_groups_arrays(g, table, cur, group_offsets, hist, cur_len=None, hist_len=None, metric_type=None,
                        n_threads=4):
    """score_groups on plain numpy arrays (rows packed with pitch = padded width); returns (rec, svc)."""
    off = np.ascontiguousarray(group_offsets, dtype=np.uint32)
    V, G = len(off) - 1, int(off[-1])

    def pack(m):
        m = np.asarray(m, np.float32)
        pitch = (m.shape[1] + 3) & ~3
        out = np.zeros((m.shape[0], pitch), np.float32)
        out[:, : m.shape[1]] = m
        return out, m.shape[1], pitch
    b = A.GroupBatch()
    b.struct_size = C.sizeof(A.GroupBatch)
    b.n_services, b.n_groups, b.group_offsets = V, G, off.ctypes.data
    cm, b.n_cur, b.cur_pitch = pack(cur)
    b.cur = cm.ctypes.data
    keep = [off, cm]
    if hist is not None:
        hm, b.n_hist, b.hist_pitch = pack(hist)
        b.hist = hm.ctypes.data
        keep.append(hm)
    for name, arr, dt in (("cur_len", cur_len, np.uint32), ("hist_len", hist_len, np.uint32),
                          ("metric_type", metric_type, np.uint8)):
        if arr is not None:
            a = np.ascontiguousarray(arr, dtype=dt)
            keep.append(a)
            setattr(b, name, a.ctypes.data)
    r = score_groups(g, table, b, n_threads)
    return r["rec"], r["svc"]


def bench_synth(g, table, desc, strategy, hist_is_base, n_threads):
    secs = C.c_double()
    chk = C.c_uint64()
    rc = lib().fmo_bench_synth(C.byref(g), table, C.c_uint32(len(table)), C.byref(desc),
                               C.c_uint32(strategy), C.c_int(1 if hist_is_base else 0),
                               C.c_uint32(n_threads), C.byref(secs), C.byref(chk))
    assert rc == 0, rc
    return secs.value, chk.value
