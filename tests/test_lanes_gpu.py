"""The lanes engine (one source per THREAD, sms_b200/csrc/ocw_search_lanes.cuh) on the GPU through the C ABI: with
SMSGPU_LANES=1 it runs in front of the pairs engine for every source list; walks, labels and counters must equal the
oracle's exactly like every other engine's.  (Its sequential core is also checked on the CPU: tests/test_lanes_core.py.)"""
import numpy as np
import pytest

from sms_b200 import instances as I
from test_gpu_parity import check_raw, gpu  # noqa: F401  (fixture)

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mirror", ["1", "0"])
def test_lanes_engine_equals_oracle(gpu, orc, monkeypatch, mirror):
    """tie-free x* (twins mirrored: one search per vertex), tie-heavy x* (ties break the mirror, buckets overflow: the
    pairs / runs engines take those sources), a degree-40 graph (several groups of four arcs per row), a lattice"""
    monkeypatch.setenv("SMSGPU_LANES", "1")
    monkeypatch.setenv("SMSGPU_MIRROR", mirror)
    insts = (I.random_graph(3000, 12000, seed=11), I.random_graph(200, 4000, seed=12), I.torus3d(7))
    for inst in (insts if mirror == "1" else insts[:1]):
        for xf in (I.xstar_uniform, I.xstar_lp_like, I.xstar_half_integral):
            _, _, st, tiers = check_raw(inst, xf(inst.m, seed=6), orc, start_tier=1)
            assert st["tier_sources"][0] == 0
            if xf is I.xstar_uniform and mirror == "1" and inst.n >= 3000:
                assert st["mirrored_sources"] > 0.4 * st["sources_run"]


def test_lanes_engine_batches_and_sets(gpu, orc, monkeypatch):
    """more work items than tiles: several batches alternating between two tile sets (stamps tell the searches of a tile
    apart), sources on a 3-stride (no twins: every second entry goes to the pairs engine)"""
    monkeypatch.setenv("SMSGPU_LANES", "1")
    monkeypatch.setenv("SMSGPU_LANES_TILES", "1024")
    inst = I.random_graph(3000, 12000, seed=11)
    x = I.xstar_uniform(inst.m, seed=9)
    check_raw(inst, x, orc, start_tier=1)
    check_raw(inst, x, orc, start_tier=1, sources=np.arange(0, inst.n, 3))
    check_raw(inst, x, orc, start_tier=0)


def test_lanes_engine_rounds_with_history(gpu, orc, monkeypatch):
    """rounds on one context: issue order from the previous round's work, heavy sources beside tier 1, tables never cleared"""
    from sms_b200.separator import GpuOcwSeparator
    monkeypatch.setenv("SMSGPU_LANES", "1")
    monkeypatch.setenv("SMSGPU_LANES_TILES", "1024")
    inst = I.random_graph(3000, 12000, seed=11)
    xs = [I.xstar_uniform(inst.m, seed=7), I.xstar_lp_like(inst.m, seed=6), I.xstar_uniform(inst.m, seed=8), I.xstar_uniform(inst.m, seed=7)]
    bl = inst.root_blacklist()
    src = inst.handler_sources()
    sep = GpuOcwSeparator(inst.n, inst.rowptr, inst.col, inst.eid)
    try:
        for r, x in enumerate(xs):
            exp = orc.sweep(inst, x, blacklist=bl, sources=src, mode="I", canonical=True)
            sep.setRoundInput(x, bl)
            got = sep.separateRound(sources=src, dedup=False, start_tier=1, export_src_stats=1)
            assert got.walks() == exp.walks(), f"round {r}: walks differ"
            arcs, settled, tier = sep.ctx.src_stats()
            assert np.array_equal(arcs, exp.stats[:, 0]) and np.array_equal(settled, exp.stats[:, 1])
            st = sep.ctx.stats()
            assert st["sources_run"] == int((tier >= 0).sum()) == sum(st["tier_sources"])
    finally:
        sep.close()


def test_lanes_and_pairs_engines_agree_at_scale(gpu, monkeypatch):
    """Two engines that share no search code (two sources per warp in lock step / one source per thread) on a whole round
    of a 100 000-node graph, all sources, dedup on: the same cut list byte for byte, the same arcs scanned and nodes settled
    per source.  (The full 1M-source round of config 5 agrees as well: scripts/exp_lanes.py, profiles/r02_lanes_experiments.md.)"""
    from sms_b200.separator import GpuOcwSeparator
    inst = I.random_graph(100000, 400000, seed=21)
    x = I.xstar_uniform(inst.m, seed=22)
    bl = inst.root_blacklist()
    src = inst.handler_sources()
    out = []
    for lanes in ("0", "1"):
        monkeypatch.setenv("SMSGPU_LANES", lanes)
        monkeypatch.setenv("SMSGPU_LANES_TILES", "16384")
        sep = GpuOcwSeparator(inst.n, inst.rowptr, inst.col, inst.eid)
        try:
            sep.setRoundInput(x, bl)
            got = sep.separateRound(sources=src, dedup=True, start_tier=1, export_src_stats=1)
            arcs, settled, tier = sep.ctx.src_stats()
            st = sep.ctx.stats()
            out.append((got.walks(), arcs.copy(), settled.copy(), st["unique_cuts"], st["mirrored_sources"]))
        finally:
            sep.close()
    a, b = out
    assert a[3] == b[3] > 1000 and a[0] == b[0]
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    assert b[4] > 0.4 * len(src)
