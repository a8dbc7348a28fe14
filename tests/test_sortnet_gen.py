"""The generated 1536-key network (fm_sort48_net.cuh): its simulation sorts, its shared-memory
swizzle is conflict free, and the committed header is what the generator writes."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN = os.path.join(ROOT, "foremast_1_b200", "csrc", "gen", "gen_sort48.py")


def _gen():
    spec = importlib.util.spec_from_file_location("gen_sort48", GEN)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_network_simulation_sorts_and_swizzle_is_conflict_free():
    g = _gen()
    g.check_banks()
    g.check_network()
    assert len(g.local_merge64_pruned()) == 128
    assert [len(g.striped_phase(K)[1]) for K in (256, 512, 1024, 2048)] == [56, 80, 104, 128]


def test_committed_header_is_current(tmp_path):
    g = _gen()
    out = tmp_path / "net.cuh"
    g.emit(str(out))
    with open(os.path.join(ROOT, "foremast_1_b200", "csrc", "fm_sort48_net.cuh")) as f:
        assert f.read() == out.read_text()
