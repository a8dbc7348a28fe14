"""Multi-rank path on CPU: world size 2 over gloo.  Each rank owns a range of series
(cut at service boundaries), scores it, and the records are gathered; the result must
equal the single-rank result byte for byte.  The scorer here is the oracle (this is a
test of the sharding / gather logic, which is host code); the same flow on GPUs is
bench.py --gpus N and tests/test_gpu_shard.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _score_range(lo, hi):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests", "golden")]
    import oracle_lib as O
    from foremast_1_b200 import _abi as A
    from foremast_1_b200.engine import synth_desc, synth_host
    from make_golden import host_batch
    g, t = A.default_params()
    d = synth_desc(3, hi - lo, 60, 300, 0, first_series=lo)
    cur, base, _, mt = synth_host(d, n_threads=1)
    return O.score_batch(g, t, host_batch(A.FM_STRAT_CANARY, cur, base, None, mt), n_threads=1)["rec"]


def _worker(rank, world, port, n, q):
    import torch.distributed as dist
    sys.path[:0] = [ROOT]
    from foremast_1_b200 import shard
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    svc = np.arange(n) // 7                      # 7 series per service
    out = shard.score_sharded(_score_range, rank, world, n, svc, shard.torch_gather())
    if rank == 0:
        q.put(out.tobytes())
    dist.barrier()
    dist.destroy_process_group()


def test_shard_bounds_respect_services():
    sys.path[:0] = [ROOT]
    from foremast_1_b200.shard import shard_bounds
    svc = np.arange(100) // 7
    for world in (1, 2, 3, 8):
        b = shard_bounds(100, world, svc)
        assert b[0][0] == 0 and b[-1][1] == 100
        assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
        for lo, hi in b[1:]:
            assert lo == 100 or lo == 0 or svc[lo] != svc[lo - 1]
    assert shard_bounds(10, 4) == [(0, 2), (2, 5), (5, 7), (7, 10)]
    assert shard_bounds(0, 2) == [(0, 0), (0, 0)]


@pytest.mark.timeout(180)
def test_two_ranks_equal_one_rank():
    n, world = 200, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=150)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = _score_range(0, n)
    assert got == want.tobytes()


def test_bind_host_to_gpu_reads_sysfs(tmp_path):
    """NUMA binding helper: parses local_cpulist, ignores boxes that do not say (numa_node = -1)."""
    import os
    from foremast_1_b200 import shard
    assert shard._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    dev = tmp_path / "0000:1b:00.0"
    dev.mkdir()
    (dev / "numa_node").write_text("-1\n")
    (dev / "local_cpulist").write_text("0-1\n")
    before = os.sched_getaffinity(0)
    assert shard.bind_host_to_gpu("0000:1B:00.0", sysfs=str(tmp_path)) is None
    assert shard.bind_host_to_gpu("0000:ff:00.0", sysfs=str(tmp_path)) is None      # unknown device
    (dev / "numa_node").write_text("0\n")
    (dev / "local_cpulist").write_text(",".join(str(c) for c in sorted(before)) + "\n")
    assert shard.bind_host_to_gpu("0000:1b:00.0", sysfs=str(tmp_path)) is None      # whole allowed set: no change
    if len(before) > 1:
        one = min(before)
        (dev / "local_cpulist").write_text("%d\n" % one)
        try:
            assert shard.bind_host_to_gpu("0000:1b:00.0", sysfs=str(tmp_path)) == {one}
            assert os.sched_getaffinity(0) == {one}
        finally:
            os.sched_setaffinity(0, before)
