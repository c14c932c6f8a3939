"""Mode V: the cut list of the reference's own SEQUENTIAL sweep (blacklist side effects carried from source to source:
ocw_seperator.cpp:113-114 + 363-382 blacklistIntReachable, :162-169 stem blacklisting, read at :98-99 and :334-335).

The device computes it as the fixed point of independent sweeps over a time-stamped blacklist (include/smsgpu.h
SMSGPU_MODE_VERBATIM).  Chain of evidence:
  CPU  the restatement's fixed-point sweep (oracle/ocw_oracle.hpp stamp mode) == its sequential sweep == the UNMODIFIED
       reference's sequential sweep (golden V_* arrays, produced by tests/golden/make_golden.py);
  GPU  smsgpu_separate(mode = 1), raw walk list == those golden V_* arrays (tie free) and == the restatement's
       sequential sweep with canonical ties (tie-heavy x*)."""
import numpy as np
import pytest

from sms_b200 import instances as I
from test_golden import INTROSORT_DEPENDENT, TIE_FREE, _id, load, walks_of

CASES = [(lambda: I.torus2d(10), I.xstar_uniform), (lambda: I.torus2d(10), I.xstar_lp_like), (lambda: I.torus2d(30), I.xstar_lp_like),
         (lambda: I.torus3d(6), I.xstar_half_integral), (lambda: I.random_graph(400, 1600, seed=3), I.xstar_lp_like),
         (lambda: I.random_graph(400, 1600, seed=3), I.xstar_uniform), (lambda: I.random_graph(150, 1500, seed=5), I.xstar_half_integral)]


@pytest.mark.parametrize("mk,xf", CASES)
def test_fixed_point_of_stamped_sweeps_is_the_sequential_sweep(orc, mk, xf):
    inst = mk()
    x = xf(inst.m, seed=7)
    for canonical in (False, True):
        seq = orc.sweep(inst, x, mode="V", canonical=canonical)
        fix, info = orc.sweep_fixpoint(inst, x, canonical=canonical)
        assert fix.walks() == seq.walks()
        assert 1 <= info["sweeps"] <= inst.n


@pytest.mark.parametrize("path", TIE_FREE, ids=_id)
def test_fixed_point_reproduces_the_reference_sequential_sweep(orc, path):
    inst, g = load(path)
    gold = walks_of(g["V_cut_off"], g["V_cut_src"], g["V_node"], g["V_stay"])
    fix, info = orc.sweep_fixpoint(inst, g["x"], blacklist=g["blacklist"], sources=g["sources"], canonical=False)
    assert fix.walks() == gold


def test_fixed_point_with_fixed_vertices_and_a_permuted_source_list(orc):
    inst = I.torus2d(14)
    rng = np.random.default_rng(5)
    bl = inst.root_blacklist()
    bl[rng.choice(inst.n, 25, replace=False)] = 1
    src = rng.permutation(inst.n).astype(np.int32)
    x = I.xstar_lp_like(inst.m, seed=9)
    seq = orc.sweep(inst, x, blacklist=bl, sources=src, mode="V", canonical=True)
    fix, _ = orc.sweep_fixpoint(inst, x, blacklist=bl, sources=src, canonical=True)
    assert fix.walks() == seq.walks()


# ---- GPU ---------------------------------------------------------------------------------------------------------------
def _gpu_sep(inst):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sms_b200 import build
    from sms_b200.separator import GpuOcwSeparator
    build.build()
    return GpuOcwSeparator(inst.n, inst.rowptr, inst.col, inst.eid)


@pytest.mark.gpu
@pytest.mark.parametrize("path", TIE_FREE, ids=_id)
def test_gpu_mode_v_reproduces_the_reference_sequential_sweep(path):
    inst, g = load(path)
    gold = walks_of(g["V_cut_off"], g["V_cut_src"], g["V_node"], g["V_stay"])
    sep = _gpu_sep(inst)
    try:
        sep.setRoundInput(g["x"], g["blacklist"])
        got = sep.separateRound(sources=g["sources"], dedup=False, mode=1)
        st = sep.ctx.stats()
        if _id(path) in INTROSORT_DEPENDENT:
            assert {w[0] for w in got.walks()} == {w[0] for w in gold}
        else:
            assert got.walks() == gold
        assert st["verbatim_sweeps"] >= 2
        # and mode I on the same context afterwards is untouched by it
        back = sep.separateRound(sources=g["sources"], dedup=False, mode=0)
        assert back.walks() == walks_of(g["I_cut_off"], g["I_cut_src"], g["I_node"], g["I_stay"]) or _id(path) in INTROSORT_DEPENDENT
    finally:
        sep.close()


@pytest.mark.gpu
@pytest.mark.parametrize("mk,xf", CASES + [(lambda: I.random_graph(3000, 12000, seed=11), I.xstar_uniform),
                                           (lambda: I.random_graph(3000, 12000, seed=11), I.xstar_lp_like),
                                           (lambda: I.torus2d(100), I.xstar_uniform), (lambda: I.torus3d(20), I.xstar_uniform)])
@pytest.mark.parametrize("start_tier", [0, 1])
def test_gpu_mode_v_equals_the_canonical_sequential_sweep(orc, mk, xf, start_tier):
    inst = mk()
    if start_tier == 1 and inst.n > 5000:
        pytest.skip("large instances once")
    x = xf(inst.m, seed=7)
    bl = inst.root_blacklist()
    exp = orc.sweep(inst, x, blacklist=bl, mode="V", canonical=True)
    sep = _gpu_sep(inst)
    try:
        sep.setRoundInput(x, bl)
        got = sep.separateRound(dedup=False, mode=1, start_tier=start_tier)
        assert got.walks() == exp.walks()
        # the shortcut (sources that find nothing against the initial blacklist are not searched again) is confirmed by
        # a full sweep inside the library; with it switched off the result is the same
        dd = sep.separateRound(dedup=True, mode=1, start_tier=start_tier)
        from helpers import canonical_set
        assert {tuple(zip(n, s)) for _, n, s in dd.walks()} == canonical_set(exp.walks())
    finally:
        sep.close()


@pytest.mark.gpu
def test_gpu_mode_v_with_fixed_vertices_and_a_permuted_source_list(orc):
    inst = I.torus2d(14)
    rng = np.random.default_rng(5)
    bl = inst.root_blacklist()
    bl[rng.choice(inst.n, 25, replace=False)] = 1
    src = rng.permutation(inst.n).astype(np.int32)
    x = I.xstar_lp_like(inst.m, seed=9)
    exp = orc.sweep(inst, x, blacklist=bl, sources=src, mode="V", canonical=True)
    sep = _gpu_sep(inst)
    try:
        sep.setRoundInput(x, bl)
        got = sep.separateRound(sources=src, dedup=False, mode=1)
        assert got.walks() == exp.walks()
    finally:
        sep.close()


@pytest.mark.gpu
@pytest.mark.parametrize("window", ["0", "37", "256"])
def test_gpu_mode_v_windows(orc, window, monkeypatch):
    """The fixed point is taken window by window of the source list (smsgpu.cu, mode V): any window length gives the
    sequential sweep -- 0 = one window (the whole list), 37 = odd and tiny (many windows, twins split), 256."""
    monkeypatch.setenv("SMSGPU_VERBATIM_WINDOW", window)
    for mk, xf in ((lambda: I.torus2d(30), I.xstar_lp_like), (lambda: I.random_graph(3000, 12000, seed=11), I.xstar_uniform),
                   (lambda: I.torus3d(6), I.xstar_half_integral)):
        inst = mk()
        x = xf(inst.m, seed=7)
        bl = inst.root_blacklist()
        exp = orc.sweep(inst, x, blacklist=bl, mode="V", canonical=True)
        sep = _gpu_sep(inst)
        try:
            sep.setRoundInput(x, bl)
            got = sep.separateRound(dedup=False, mode=1)
            assert got.walks() == exp.walks(), f"{inst.name} window {window}"
            st = sep.ctx.stats()
            assert st["verbatim_sweeps"] >= 1
            again = sep.separateRound(dedup=False, mode=1)  # a second round on the same context
            assert again.walks() == exp.walks()
        finally:
            sep.close()
