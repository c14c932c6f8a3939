"""Parity tests proper (-m gpu): the CUDA path through the C ABI against the CPU oracle on the same seeded inputs.
Bar: bit-exact -- identical walks in identical order per source, identical (dist, hops) labels, identical
arcs-scanned / settled counters -- against the restatement in canonical tie mode, and against the UNMODIFIED
reference (oracle/_ref) on tie-free x*."""
import numpy as np
import pytest

from helpers import canonical, canonical_set, first_finder, walk_length
from sms_b200 import instances as I

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sms_b200 import build
    build.build()
    return True


def make_sep(inst):
    from sms_b200.separator import GpuOcwSeparator
    return GpuOcwSeparator(inst.n, inst.rowptr, inst.col, inst.eid)


def check_raw(inst, x, oracle, bl=None, sources=None, canonical_mode=True, **params):
    bl = inst.root_blacklist() if bl is None else np.asarray(bl, np.uint8)
    src = inst.handler_sources() if sources is None else np.asarray(sources, np.int32)
    kw = dict(canonical=True) if canonical_mode else {}
    exp = oracle.sweep(inst, x, blacklist=bl, sources=src, mode="I", want_labels=True, **kw)
    sep = make_sep(inst)
    try:
        sep.setRoundInput(x, bl)
        got = sep.separateRound(sources=src, dedup=False, export_labels=1, export_src_stats=1, **params)
        gw, ew = got.walks(), exp.walks()
        if gw != ew:
            bad = [k for k in range(min(len(gw), len(ew))) if gw[k] != ew[k]][:5]
            raise AssertionError(f"{inst.name}: walks differ: gpu {len(gw)} vs oracle {len(ew)}; first diffs at {bad}: "
                                 + "; ".join(f"gpu {gw[k]} / orc {ew[k]}" for k in bad))
        arcs, settled, tier = sep.ctx.src_stats()
        if exp.stats[:, 0].sum() > 0:  # the unmodified reference exports no counters
            assert np.array_equal(arcs, exp.stats[:, 0]), "arcs scanned per source differ"
            assert np.array_equal(settled, exp.stats[:, 1]), "settled per source differ"
        st = sep.ctx.stats()
        assert st["arcs_scanned"] == int(arcs.sum()) and st["settled"] == int(settled.sum())
        # labels: every label the GPU kept equals the oracle's, bit for bit
        lo, ln, ld, lh = sep.ctx.labels()
        for k in range(len(src)):
            ea, eb = int(exp.label_off[k]), int(exp.label_off[k + 1])
            omap = {int(n): (d, int(h)) for n, d, h in zip(exp.label_node[ea:eb], exp.label_dist[ea:eb], exp.label_hops[ea:eb])}
            for n, d, h in zip(ln[lo[k]:lo[k + 1]], ld[lo[k]:lo[k + 1]], lh[lo[k]:lo[k + 1]]):
                assert int(n) in omap, f"source {k}: GPU label for node {n} unknown to the oracle"
                od, oh = omap[int(n)]
                assert d == od and int(h) == oh, f"source {k} node {n}: gpu ({d},{h}) oracle ({od},{oh})"
                assert abs(d - od) <= 1e-9 * max(abs(od), 1e-300)  # north_star tolerance (implied by equality)
        # violation column
        for k, (_, nodes, stay) in enumerate(gw[:200]):
            assert abs((1.0 - walk_length(inst, x, nodes, stay)) - got.violation[k]) < 1e-12
        return got, exp, st, tier
    finally:
        sep.close()


SMALL = [lambda: I.torus2d(10), lambda: I.torus2d(16, 9), lambda: I.torus3d(6), lambda: I.random_graph(400, 1600, seed=3),
         lambda: I.random_graph(150, 1500, seed=5)]
XS = [("uniform", I.xstar_uniform), ("lp_like", I.xstar_lp_like), ("half", I.xstar_half_integral)]


@pytest.mark.parametrize("mk", SMALL)
@pytest.mark.parametrize("xname,xf", XS)
def test_raw_walks_equal_canonical_oracle(gpu, orc, mk, xname, xf):
    inst = mk()
    check_raw(inst, xf(inst.m, seed=7), orc)


@pytest.mark.parametrize("mk", SMALL)
def test_raw_walks_equal_unmodified_reference_on_tie_free_input(gpu, ref, mk):
    inst = mk()
    check_raw(inst, I.xstar_uniform(inst.m, seed=21), ref, canonical_mode=False)


def test_near_cut_points(gpu, orc):
    inst = I.torus2d(20)
    check_raw(inst, I.xstar_near_cut(inst, seed=3), orc)


@pytest.mark.parametrize("tier", [1, 2, 3])
def test_tiers_agree(gpu, orc, tier):
    inst = I.random_graph(300, 1300, seed=8)
    for xf in (I.xstar_uniform, I.xstar_lp_like):
        _, _, st, tiers = check_raw(inst, xf(inst.m, seed=4), orc, start_tier=tier)
        assert st["tier_sources"][0] == 0 and (tiers[tiers >= 0] >= tier).all()


@pytest.mark.parametrize("engine", ["pairs", "one_per_warp"])
def test_tier1_engines_agree(gpu, orc, engine, monkeypatch):
    """Tier 1 has two engines: the pairs engine (two sources per warp in lock step, ocw_search_pairs.cuh; the
    default) and the one-source-per-warp engine (SMSGPU_PAIRS=0).  Both must reproduce the oracle event for event,
    on tie-free and on tie-heavy x*, on a degree-8 graph (one pass per row) and a degree-40 one (several passes)."""
    monkeypatch.setenv("SMSGPU_PAIRS", "1" if engine == "pairs" else "0")
    for inst in (I.random_graph(3000, 12000, seed=11), I.random_graph(200, 4000, seed=12), I.torus3d(7)):
        for xf in (I.xstar_uniform, I.xstar_lp_like, I.xstar_half_integral):
            _, _, st, tiers = check_raw(inst, xf(inst.m, seed=6), orc, start_tier=1, sources=np.arange(0, inst.n, 3))
            assert st["tier_sources"][0] == 0


@pytest.mark.parametrize("start_tier", [0, 1])
def test_rounds_with_history_equal_the_oracle(gpu, orc, start_tier):
    """From the second round on a context issues its sources in the order of the previous round's work, skips the
    shared-memory tier when it was useless, and starts the pairs that needed tier 2 there at once, on a second
    stream beside tier 1 (smsgpu.cu "heavy").  None of it may change a result: three rounds with alternating x* on
    one context, each equal to the oracle walk for walk, with the overlapped tier-2 launch actually exercised
    (start_tier 1).  From start_tier 0 the first round also goes through the pair-wise rebuild of tier 0's overflow
    list (pairCompactKernel) with both sides of every vertex in the source list, i.e. with mirrored twins behind it."""
    from sms_b200.separator import GpuOcwSeparator
    inst = I.random_graph(3000, 12000, seed=11)
    xs = [I.xstar_lp_like(inst.m, seed=6), I.xstar_uniform(inst.m, seed=7), I.xstar_lp_like(inst.m, seed=6),
          I.xstar_lp_like(inst.m, seed=8)]
    bl = inst.root_blacklist()
    src = inst.handler_sources()
    sep = GpuOcwSeparator(inst.n, inst.rowptr, inst.col, inst.eid)
    heavy_rounds = 0
    try:
        for r, x in enumerate(xs):
            exp = orc.sweep(inst, x, blacklist=bl, sources=src, mode="I", canonical=True)
            sep.setRoundInput(x, bl)
            got = sep.separateRound(sources=src, dedup=False, start_tier=start_tier, export_src_stats=1)
            assert got.walks() == exp.walks(), f"round {r}: walks differ"
            arcs, settled, tier = sep.ctx.src_stats()
            assert np.array_equal(arcs, exp.stats[:, 0]) and np.array_equal(settled, exp.stats[:, 1])
            st = sep.ctx.stats()
            assert st["sources_run"] == int((tier >= 0).sum()) == sum(st["tier_sources"])
            assert [int((tier == t).sum()) for t in range(4)] == st["tier_sources"]
            if r > 0 and prev_heavy > 0:
                heavy_rounds += 1
            prev_heavy = int((tier >= 2).sum())
    finally:
        sep.close()
    if start_tier == 1:
        assert heavy_rounds >= 2, "no round started with sources known to need tier 2: the test instance is too easy"


def test_overflow_escalates_to_next_tier(gpu, orc):
    """LP-like x* on a degree-8 graph makes balls that do not fit shared memory: sources must escalate and
    still match."""
    inst = I.random_graph(2000, 8000, seed=7)
    _, _, st, tiers = check_raw(inst, I.xstar_lp_like(inst.m, seed=1), orc, sources=np.arange(0, 2000, 7))
    assert sum(st["tier_sources"][1:]) > 0


def test_blacklist_fixed_vertices_and_source_lists(gpu, orc):
    inst = I.torus2d(14)
    rng = np.random.default_rng(5)
    bl = inst.root_blacklist()
    bl[rng.choice(inst.n, 25, replace=False)] = 1
    x = I.xstar_uniform(inst.m, seed=9)
    check_raw(inst, x, orc, bl=bl)
    check_raw(inst, x, orc, bl=bl, sources=np.arange(inst.n)[::-1].copy())  # reversed order (mode I is order free)
    check_raw(inst, x, orc, bl=bl, sources=rng.permutation(inst.n)[:50])


def test_single_source_interface_and_reference_kats(gpu, orc):
    """ocw_separator_gtest.cpp:5-53 through the mirrored interface"""
    from sms_b200.separator import GpuOcwSeparator
    tri = I.from_edges("tri", 3, [(0, 1), (1, 2), (2, 0)])
    sep = GpuOcwSeparator(tri.n, tri.rowptr, tri.col, tri.eid)
    for (u, v) in [(0, 1), (1, 2), (2, 0)]:
        sep.updateWeights(u, v, 0.3)
    assert sep.getViolatedOcws(0) == []  # SimpleNoPath
    for (u, v) in [(0, 1), (1, 2), (2, 0)]:
        sep.updateWeights(u, v, 0.9)
    res = sep.getViolatedOcws(0)  # SimplePath
    assert len(res) >= 1 and res[0].isValid() and res[0].size() == 3
    sep.updateBlacklist(0, True)
    assert sep.getViolatedOcws(0) == []
    sep.close()


def test_dedup_set_order_and_representatives(gpu, orc):
    for inst, x in [(I.torus2d(24), I.xstar_uniform(24 * 24 * 2, seed=2)), (I.torus3d(7), I.xstar_lp_like(7 ** 3 * 3, seed=2)),
                    (I.random_graph(500, 2000, seed=2), I.xstar_uniform(2000, seed=3))]:
        exp = orc.sweep(inst, x, mode="I", canonical=True)
        sep = make_sep(inst)
        sep.setRoundInput(x, inst.root_blacklist())
        got = sep.separateRound()
        st = sep.ctx.stats()
        sep.close()
        ff = first_finder(exp.walks())
        gw = got.walks()
        assert st["raw_walks"] == exp.ncuts and st["unique_cuts"] == len(ff) == got.ncuts
        forms = [tuple(zip(n, s)) for _, n, s in gw]
        assert set(forms) == set(ff.keys()), "deduplicated cycle set differs from the oracle's"
        assert forms == [canonical(n, s) for _, n, s in gw], "output is not in canonical form"
        ranks = [ff[f] for f in forms]
        assert ranks == sorted(ranks), "output not ordered by (first source, emission index)"
        assert [r[0] for r in ranks] == got.src.tolist()
        assert np.array_equal(got.n_cross, np.array([len(s) - sum(s) for _, _, s in gw], np.int32))
        assert (got.violation > 1e-6 - 1e-12).all()  # every cut is violated by more than the reference's epsilon... of its walk
        for k in range(0, got.ncuts, max(1, got.ncuts // 100)):
            _, n, s = gw[k]
            assert abs(1.0 - walk_length(inst, x, n, s) - got.violation[k]) < 1e-12


def test_source_sharding_union_equals_full_round(gpu, orc):
    inst = I.torus2d(20)
    x = I.xstar_uniform(inst.m, seed=6)
    sep = make_sep(inst)
    sep.setRoundInput(x, inst.root_blacklist())
    full = sep.separateRound()
    parts = [sep.separateRound(rank=r, nranks=3) for r in range(3)]
    sep.close()
    best = {}
    for p in parts:
        for (src, n, s) in p.walks():
            f = tuple(zip(n, s))
            best[f] = min(best.get(f, src), src)
    fw = full.walks()
    assert {tuple(zip(n, s)) for _, n, s in fw} == set(best)
    assert all(best[tuple(zip(n, s))] == src for src, n, s in fw)


def test_small_cycle_scan_equals_oracle(gpu, orc):
    from sms_b200.separator import GpuOcwSeparator, GpuSmallCycleSeparator
    g = I.random_graph(80, 700, seed=4)
    tri, holes = orc.small_ccs(g)
    holes = holes[:3000]
    sep = GpuOcwSeparator(g.n, g.rowptr, g.col, g.eid)
    ts, hs = GpuSmallCycleSeparator(sep, 3), GpuSmallCycleSeparator(sep, 4)
    ts.setCycles(tri)
    hs.setCycles(holes)
    for xf in (I.xstar_uniform, I.xstar_lp_like, I.xstar_half_integral):
        x = xf(g.m, seed=3)
        x[::17] = 1.7   # clamped to 1
        x[5::19] = -0.2  # clamped to 0
        sep.setRoundInput(x, np.zeros(g.n, np.uint8))
        for s, cyc in ((ts, tri), (hs, holes)):
            exp = orc.small_cycle_scan(g, x, cyc)
            got = s.getViolated()
            assert [(n, st) for _, n, st in got.walks()] == [(n, st) for _, n, st in exp]
            # src = index of the cycle in the registered list
            assert all(tuple(cyc[i]) == n for (i, n, _) in got.walks())
    # reference KATs (triangles_gtest.cpp): (1,1,1) -> 1 ; (0,0,1) -> 1 ; nothing registered -> 0
    sep.close()
    t3 = I.from_edges("tri", 3, [(0, 1), (1, 2), (2, 0)])
    sep = GpuOcwSeparator(t3.n, t3.rowptr, t3.col, t3.eid)
    ts = GpuSmallCycleSeparator(sep, 3)
    sep.setRoundInput(np.ones(3), np.zeros(3, np.uint8))
    assert ts.getViolated().ncuts == 0
    ts.addTriangle(0, 1, 2)
    assert ts.getViolated().ncuts == 1
    sep.close()


# ---- BASELINE.json configs at full size ---------------------------------------------------------------------------
def test_config2_torus_100x100_full_round(gpu, orc):
    inst = I.torus2d(100)
    check_raw(inst, I.xstar_uniform(inst.m, seed=1), orc)


def test_config2_torus_100x100_lp_like_dedup(gpu, orc):
    inst = I.torus2d(100)
    x = I.xstar_lp_like(inst.m, seed=1)
    exp = orc.sweep(inst, x, mode="I", canonical=True)
    sep = make_sep(inst)
    sep.setRoundInput(x, inst.root_blacklist())
    got = sep.separateRound()
    sep.close()
    assert {tuple(zip(n, s)) for _, n, s in got.walks()} == canonical_set(exp.walks())


def test_config3_torus3d_20(gpu, orc):
    inst = I.torus3d(20)
    check_raw(inst, I.xstar_uniform(inst.m, seed=1), orc)


def test_config4_random_20k_full_round(gpu, orc):
    """config 4 in full: all 20 000 sources, walks / labels / counters against the oracle"""
    inst = I.random_graph(20000, 80000, seed=7)
    check_raw(inst, I.xstar_uniform(inst.m, seed=1), orc)


def test_config5_sources_that_need_the_large_tier(gpu, orc):
    """The 0.2 % of config 5's sources that overflow a bucket of the pairs engine finish in tier 2; a random sample
    almost never contains one.  A first full round reports the tier of every source; the ones with tier >= 2 and
    their twins (other side of the same vertex) are then checked walk for walk, label for label against the oracle."""
    inst = I.random_graph(1_000_000, 4_000_000, seed=7)
    x = I.xstar_uniform(inst.m, seed=1)
    sep = make_sep(inst)
    try:
        sep.setRoundInput(x, inst.root_blacklist())
        sep.separateRound(export_src_stats=1)
        _, _, tier = sep.ctx.src_stats()
        full = sep.ctx.stats()
    finally:
        sep.close()
    heavy = np.nonzero(tier >= 2)[0]
    assert full["tier_sources"][2] + full["tier_sources"][3] == len(heavy) > 100, "config 5 no longer has tier-2 sources: pick them another way"
    pick = heavy[:: max(1, len(heavy) // 700)][:700]
    src = np.unique(np.concatenate([pick, pick ^ 1])).astype(np.int32)
    _, _, st, tiers = check_raw(inst, x, orc, sources=src)
    assert (tiers[np.isin(src, pick)] >= 2).mean() > 0.9, "the picked sources did not go through the large tier again"


def test_config5_random_1m_sampled_sources_and_properties(gpu, orc):
    """1M nodes / 4M edges: exact parity on a seeded source sample; size-independent properties on the rest."""
    inst = I.random_graph(1_000_000, 4_000_000, seed=7)
    x = I.xstar_uniform(inst.m, seed=1)
    src = np.sort(np.random.default_rng(2).choice(1_000_000, 400, replace=False)).astype(np.int32)
    check_raw(inst, x, orc, sources=src)
    # properties: a bigger sample through dedup; every cut is a closed walk in the graph, odd #cross, length < 1,
    # canonical, unique; running it twice gives the identical list (determinism)
    src2 = np.arange(0, 1_000_000, 50, dtype=np.int32)
    sep = make_sep(inst)
    sep.setRoundInput(x, inst.root_blacklist())
    a = sep.separateRound(sources=src2)
    b = sep.separateRound(sources=src2)
    sep.close()
    assert a.ncuts > 1000
    for f in ("off", "node", "eid", "is_stay", "n_cross", "src", "violation"):
        assert np.array_equal(getattr(a, f), getattr(b, f)), f"round not deterministic in {f}"
    aw = a.walks()
    assert len({tuple(zip(n, s)) for _, n, s in aw}) == a.ncuts
    for k in range(0, a.ncuts, max(1, a.ncuts // 300)):
        _, n, s = aw[k]
        assert (len(s) - sum(s)) % 2 == 1 and tuple(zip(n, s)) == canonical(n, s)
        assert walk_length(inst, x, n, s) < 1.0  # chord-repaired cycles need < 1 only (ocw_seperator.cpp:286-303: no epsilon)
        assert abs(1.0 - walk_length(inst, x, n, s) - a.violation[k]) < 1e-12 and a.violation[k] > 0


def test_min_efficacy_prefilter(gpu, orc):
    """smsgpu_params.min_efficacy drops cuts with violation / sqrt(len) <= threshold on the device, before dedup:
    exactly the cuts SCIPisCutEfficacious (conshdlr_oscw.cpp:221,232) would refuse.  The surviving list equals the
    unfiltered list with those cuts removed (same order, same representatives)."""
    inst = I.random_graph(3000, 12000, seed=11)
    x = I.xstar_uniform(inst.m, seed=3)
    sep = make_sep(inst)
    try:
        sep.setRoundInput(x, inst.root_blacklist())
        full = sep.separateRound()
        for thr in (1e-6, 0.02, 0.1):
            got = sep.separateRound(min_efficacy=thr)
            lens = np.diff(full.off)
            keep = full.violation / np.sqrt(lens) > thr
            assert 0 < keep.sum() and (thr < 0.01 or keep.sum() < full.ncuts)
            fw, gw = [w for w, k in zip(full.walks(), keep) if k], got.walks()
            assert gw == fw
            assert np.array_equal(got.violation, full.violation[keep])
    finally:
        sep.close()
