"""The C-ABI library loads without a GPU and exports every symbol declared in
include/foremast_cuda.h; the ctypes mirror has the compiled struct sizes."""
import ctypes as C
import os
import re
import pytest

import numpy as np

from foremast_1_b200 import _abi as A
from foremast_1_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "foremast_cuda.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fm_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    lib = _ffi.lib()
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in foremast_cuda.h but not exported"
        assert n in _ffi.SYMBOLS, f"{n} has no ctypes signature in _ffi.SYMBOLS"
    assert sorted(_ffi.SYMBOLS) == names


def test_struct_sizes_match_compiled_library():
    lib = _ffi.lib()
    mirror = [C.sizeof(A.GlobalParams), C.sizeof(A.TypeParams), C.sizeof(A.Batch), C.sizeof(A.GroupBatch),
              A.SERIES_REC.itemsize, A.COMPACT_REC.itemsize, A.RANK_AUX.itemsize, A.ROLLUP_REC.itemsize,
              C.sizeof(A.Timing), C.sizeof(A.SynthDesc), C.sizeof(A.CacheStats), C.sizeof(A.PairBatch),
              A.PAIR_REC.itemsize]
    assert [lib.fm_sizeof(i) for i in range(13)] == mirror
    assert lib.fm_sizeof(99) == 0
    assert lib.fm_version() == (0 << 16) | 1


def test_default_params_match_python_mirror():
    lib = _ffi.lib()
    g = A.GlobalParams()
    t = (A.TypeParams * 8)()
    n = C.c_uint32(8)
    assert lib.fm_default_params(C.byref(g), t, C.byref(n)) == 0 and n.value == 6
    pg, pt = A.default_params()
    assert bytes(g) == bytes(pg)
    assert bytes(t)[: 6 * C.sizeof(A.TypeParams)] == bytes(pt)
    # deployed values of deploy/foremast/3_brain/foremast-brain.yaml:26-73
    assert [t[i].k for i in range(6)] == [2.0, 2.0, 3.0, 10.0, 5.0, 5.0]
    assert [t[i].bound_mask for i in range(6)] == [1, 1, 1, 3, 1, 1]
    n = C.c_uint32(3)
    assert lib.fm_default_params(C.byref(g), t, C.byref(n)) == A.FM_E_INVALID


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        return
    lib = _ffi.lib()
    n = C.c_int32(-1)
    assert lib.fm_device_count(C.byref(n)) == 0 and n.value == 0
    h = C.c_void_p()
    assert lib.fm_ctx_create(0, C.byref(h)) == A.FM_E_NODEVICE
    assert b"no CPU fallback" in lib.fm_last_error(None)


def test_synth_host_is_deterministic_and_typed():
    from foremast_1_b200.engine import synth_desc, synth_host
    d = synth_desc(3, 64, 60, 1440, 0, first_series=1000)
    a = synth_host(d, n_threads=1)
    b = synth_host(d, n_threads=3)
    for x, y in zip(a, b):
        assert (x is None and y is None) or x.tobytes() == y.tobytes()
    cur, base, _, mt = a
    assert mt.tolist() == [[1, 3, 4, 5][i & 3] for i in range(1000, 1064)]
    err = base[mt == 1]
    assert np.allclose(err * 60, np.round(err * 60)) and (err == 0).any()      # integer counts / 60
    mem = base[mt == 5].astype(np.float64)
    assert 2e8 < mem.mean() < 2e9 and 1e-4 < (mem.std(axis=1) / mem.mean(axis=1)).mean() < 1e-2
    assert (base[mt == 3] > 0).all() and (base[mt == 4] > 0).all() and (base[mt == 4] < 4.5).all()


def test_headline_kernel_does_not_spill():
    """k_canary_sort<64, 48> sits at the 128-register limit; a build that spills more than a few
    words costs 10 % and more, so a change that tips it over should be noticed."""
    import re
    import shutil
    import subprocess
    from foremast_1_b200 import _ffi
    tool = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(tool):
        pytest.skip("cuobjdump not available")
    _ffi.lib()
    out = subprocess.run([tool, "-res-usage", _ffi.lib_path()], capture_output=True, text=True).stdout
    m = re.search(r"k_canary_sortILi64ELi48E.*?\n\s*REG:(\d+) STACK:(\d+)", out, re.S)
    assert m, "kernel not found in the library"
    assert int(m.group(1)) <= 128 and int(m.group(2)) <= 64, m.groups()
