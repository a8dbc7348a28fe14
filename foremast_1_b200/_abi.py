"""ctypes mirror of include/foremast_cuda.h (POD structs and constants).

Field order and widths must match the header exactly; tests/test_abi.py checks
the sizes against the compiled library.
"""
import ctypes as C

import numpy as np

FM_OK = 0
FM_E_INVALID, FM_E_CUDA, FM_E_NOMEM, FM_E_UNSUPPORTED, FM_E_NODEVICE = -1, -2, -3, -4, -5

FM_HEALTHY, FM_UNHEALTHY, FM_UNKNOWN = 0, 1, 2

FM_ST_HIST_INSUFFICIENT = 0x01
FM_ST_NO_CURRENT = 0x02
FM_ST_DIFFER = 0x04
FM_ST_PAIRWISE_SKIPPED = 0x08
FM_ST_MW_RUN = 0x10
FM_ST_WX_RUN = 0x20
FM_ST_KW_RUN = 0x40
FM_ST_DEGENERATE = 0x80

FM_STRAT_CANARY, FM_STRAT_HISTORICAL = 0, 1
FM_MA_ALL, FM_MA_WINDOW, FM_MODEL_EWMA, FM_MODEL_DOUBLE_EXP, FM_MODEL_HOLT_WINTERS = 0, 1, 2, 3, 4
FM_PW_NONE, FM_PW_MANN_WHITE, FM_PW_WILCOXON, FM_PW_KRUSKAL, FM_PW_ALL, FM_PW_ANY, FM_PW_FRIEDMAN = range(7)
FM_PAIR_HEAD, FM_PAIR_TAIL = 0, 1
FM_REC_FULL, FM_REC_COMPACT = 0, 1

FM_MAX_TYPES = 64
FM_MAX_CUR = 4096
FM_MAX_RANK_N = 16384
FM_MAX_GROUPS = 16

FM_SYNTH_SEED = 0xF04E3A57


class TypeParams(C.Structure):
    _fields_ = [("k", C.c_double), ("k_low", C.c_double), ("min_lower_bound", C.c_double),
                ("bound_mask", C.c_uint32), ("reserved", C.c_uint32)]


class GlobalParams(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("ddof", C.c_uint32), ("ma_mode", C.c_uint32),
                ("ma_window", C.c_uint32), ("min_hist_points", C.c_uint32),
                ("pairwise_mode", C.c_uint32), ("min_mw", C.c_uint32), ("min_wx", C.c_uint32),
                ("min_kw", C.c_uint32), ("cc_mw", C.c_uint32), ("cc_wx", C.c_uint32),
                ("pair_rule", C.c_uint32), ("min_anomaly_points", C.c_uint32),
                ("season_len", C.c_uint32), ("alpha", C.c_double), ("smooth_alpha", C.c_double),
                ("smooth_beta", C.c_double), ("smooth_gamma", C.c_double)]


class Batch(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("strategy", C.c_uint32), ("n_series", C.c_uint64),
                ("cur", C.c_void_p), ("cur_len", C.c_void_p), ("n_cur", C.c_uint32), ("cur_pitch", C.c_uint32),
                ("base", C.c_void_p), ("base_len", C.c_void_p), ("n_base", C.c_uint32), ("base_pitch", C.c_uint32),
                ("hist", C.c_void_p), ("hist_len", C.c_void_p), ("n_hist", C.c_uint32), ("hist_pitch", C.c_uint32),
                ("metric_type", C.c_void_p), ("api_id", C.c_void_p), ("api_service", C.c_void_p),
                ("n_apis", C.c_uint32), ("n_services", C.c_uint32), ("record_kind", C.c_uint32),
                ("reserved", C.c_uint32)]


class GroupBatch(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("reserved", C.c_uint32), ("n_services", C.c_uint64),
                ("n_groups", C.c_uint64), ("group_offsets", C.c_void_p),
                ("cur", C.c_void_p), ("cur_len", C.c_void_p), ("n_cur", C.c_uint32), ("cur_pitch", C.c_uint32),
                ("hist", C.c_void_p), ("hist_len", C.c_void_p), ("n_hist", C.c_uint32), ("hist_pitch", C.c_uint32),
                ("metric_type", C.c_void_p)]


class Timing(C.Structure):
    _fields_ = [("kernel_ms", C.c_float), ("launches", C.c_uint32), ("h2d_ms", C.c_float),
                ("d2h_ms", C.c_float), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("main_kernel_ms", C.c_float), ("reserved", C.c_uint32)]


class PairBatch(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("reserved", C.c_uint32), ("n_pairs", C.c_uint64),
                ("cur_x", C.c_void_p), ("cur_y", C.c_void_p), ("cur_len", C.c_void_p),
                ("n_cur", C.c_uint32), ("cur_pitch", C.c_uint32),
                ("hist_x", C.c_void_p), ("hist_y", C.c_void_p), ("hist_len", C.c_void_p),
                ("n_hist", C.c_uint32), ("hist_pitch", C.c_uint32), ("metric_type", C.c_void_p)]


class CacheStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("max_bytes", "row_bytes", "capacity_rows", "resident_rows",
                                          "hits", "misses", "evictions", "uncached")]


class SynthDesc(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("config", C.c_uint32), ("seed", C.c_uint64),
                ("first_series", C.c_uint64), ("n_series", C.c_uint64),
                ("n_cur", C.c_uint32), ("n_base", C.c_uint32), ("n_hist", C.c_uint32),
                ("cur_pitch", C.c_uint32), ("base_pitch", C.c_uint32), ("hist_pitch", C.c_uint32),
                ("row_ids", C.c_void_p)]


# numpy views of the result records (same layout as the C structs)
SERIES_REC = np.dtype([("mean", "<f8"), ("sigma", "<f8"), ("upper", "<f8"), ("lower", "<f8"),
                       ("u1", "<f8"), ("p_u", "<f8"), ("w", "<f8"), ("p_w", "<f8"),
                       ("h", "<f8"), ("p_h", "<f8"), ("n_anom", "<u4"), ("first_idx", "<u2"),
                       ("verdict", "u1"), ("status", "u1"), ("anom_mask", "<u8")])
COMPACT_REC = np.dtype([("anom_mask", "<u8"), ("n_anom", "<u4"), ("verdict", "u1"),
                        ("status", "u1"), ("first_idx", "<u2")])
RANK_AUX = np.dtype([("r2_cur", "<i8"), ("t_mw", "<i8"), ("w2_plus", "<i8"), ("w2_minus", "<i8"),
                     ("t_wx", "<i8"), ("n_wx", "<u4"), ("reserved", "<u4")])
ROLLUP_REC = np.dtype([("n_children", "<u4"), ("n_unhealthy", "<u4"), ("n_unknown", "<u4"),
                       ("severity", "u1"), ("reserved", "u1", (3,))])

PAIR_REC = np.dtype([("mean_x", "<f8"), ("mean_y", "<f8"), ("var_x", "<f8"), ("var_y", "<f8"), ("cov_xy", "<f8"),
                     ("d2_max", "<f8"), ("n_anom", "<u4"), ("first_idx", "<u2"), ("verdict", "u1"),
                     ("status", "u1"), ("anom_mask", "<u8")])

assert SERIES_REC.itemsize == 96 and COMPACT_REC.itemsize == 16 and PAIR_REC.itemsize == 64
assert RANK_AUX.itemsize == 48 and ROLLUP_REC.itemsize == 16

# metric-type ids of the default table (foremast-brain.yaml:32-73)
METRIC_TYPES = {"default": 0, "error5xx": 1, "error4xx": 2, "latency": 3, "cpu": 4, "memory": 5}


def row_pitch(n: int) -> int:
    """Row pitch in elements: rows must start 16-byte aligned."""
    return (int(n) + 3) & ~3


def ptr(a):
    """Raw address of a numpy array (or None)."""
    if a is None:
        return None
    return a.ctypes.data


def default_params():
    """Deployed defaults of deploy/foremast/3_brain/foremast-brain.yaml:24-79.

    Kept in Python too so that host logic and CPU-only tests do not need the
    CUDA library; tests check it against fm_default_params().
    """
    g = GlobalParams()
    g.struct_size = C.sizeof(GlobalParams)
    g.ddof = 0
    g.ma_mode = FM_MA_ALL
    g.ma_window = 0
    g.min_hist_points = 10
    g.pairwise_mode = FM_PW_ANY
    g.min_mw, g.min_wx, g.min_kw = 20, 20, 5
    g.cc_mw, g.cc_wx = 1, 0
    g.pair_rule = FM_PAIR_HEAD
    g.min_anomaly_points = 1
    g.alpha = 0.05
    g.smooth_alpha = 0.3
    g.smooth_beta = 0.1
    g.smooth_gamma = 0.1
    g.season_len = 0
    rows = [(2.0, 1), (2.0, 1), (3.0, 1), (10.0, 3), (5.0, 1), (5.0, 1)]
    table = (TypeParams * len(rows))()
    for i, (k, b) in enumerate(rows):
        table[i].k = k
        table[i].k_low = 0.5 * k
        table[i].min_lower_bound = 0.0
        table[i].bound_mask = b
    return g, table
