"""Golden vectors produced by the UNMODIFIED reference (tests/golden/make_golden.py, run where /root/reference
exists; the .npz files are committed).  CPU part: the restatement (oracle/) reproduces them -- this pins the oracle
on machines that have neither /root/reference nor oracle/_ref.  GPU part (-m gpu): the CUDA path through the C ABI
reproduces them directly, without any oracle in between."""
import glob
import os

import numpy as np
import pytest

from helpers import walk_length
from sms_b200 import instances as I

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz")))
TIE_FREE = [p for p in GOLD if "uniform" in p]
TIES = [p for p in GOLD if "lp_like" in p]


# std::sort of the chord list by INTEGER cycle length (ocw_seperator.cpp:309) is unstable in libstdc++ beyond 16
# elements; on the dense fixture cycles collect more good chords than that, so the reference's emission order there is
# an artefact of introsort even on tie-free x*.  The canonical contract (stable sort: length, then scan order) is
# compared on the other fixtures; on the dense one walks are compared for the verbatim tie mode only, labels always.
INTROSORT_DEPENDENT = ("dense_150_uniform",)


def _id(p):
    return os.path.basename(p)[:-4]


def load(path):
    g = np.load(path)
    inst = I.from_edges(_id(path), int(g["n"]), np.stack([g["eu"], g["ev"]], axis=1))
    return inst, g


def walks_of(off, src, node, stay):
    return [(int(src[k]), tuple(node[off[k]:off[k + 1]].tolist()), tuple(stay[off[k]:off[k + 1]].tolist()))
            for k in range(len(src))]


def test_fixtures_present():
    assert len(TIE_FREE) >= 6 and len(TIES) >= 2


@pytest.mark.parametrize("path", TIE_FREE, ids=_id)
@pytest.mark.parametrize("mode", ["I", "V"])
def test_oracle_reproduces_reference_walks(orc, path, mode):
    inst, g = load(path)
    for canonical in (False, True):  # tie-free input: both tie modes of the restatement must equal the reference
        s = orc.sweep(inst, g["x"], blacklist=g["blacklist"], sources=g["sources"], mode=mode, want_labels=True,
                      canonical=canonical)
        gold = walks_of(g[mode + "_cut_off"], g[mode + "_cut_src"], g[mode + "_node"], g[mode + "_stay"])
        if canonical and _id(path) in INTROSORT_DEPENDENT:
            assert {w[0] for w in s.walks()} == {w[0] for w in gold}
        else:
            assert s.walks() == gold
        if mode == "I" and "label_off" in g:
            assert np.array_equal(s.label_off, g["label_off"])
            for k in range(len(g["sources"])):
                a, b = int(s.label_off[k]), int(s.label_off[k + 1])
                o = np.argsort(s.label_node[a:b], kind="stable")
                r = np.argsort(g["label_node"][a:b], kind="stable")
                assert np.array_equal(s.label_node[a:b][o], g["label_node"][a:b][r])
                assert np.array_equal(s.label_dist[a:b][o], g["label_dist"][a:b][r])  # bit-equal fp64
                assert np.array_equal(s.label_hops[a:b][o], g["label_hops"][a:b][r])


@pytest.mark.parametrize("path", TIES, ids=_id)
def test_oracle_matches_tie_independent_properties(orc, path):
    inst, g = load(path)
    for canonical in (False, True):
        s = orc.sweep(inst, g["x"], blacklist=g["blacklist"], sources=g["sources"], mode="I", canonical=canonical)
        has = np.bincount(s.cut_src, minlength=len(g["sources"])) > 0
        assert np.array_equal(has, g["I_has_walk"])
        for _, nodes, stay in s.walks():
            assert walk_length(inst, g["x"], nodes, stay) < 1.0


@pytest.mark.gpu
@pytest.mark.parametrize("path", TIE_FREE, ids=_id)
def test_gpu_reproduces_reference_walks_and_labels(path):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sms_b200.separator import GpuOcwSeparator
    inst, g = load(path)
    sep = GpuOcwSeparator(inst.n, inst.rowptr, inst.col, inst.eid)
    try:
        sep.setRoundInput(g["x"], g["blacklist"])
        got = sep.separateRound(sources=g["sources"], dedup=False, export_labels=1)
        gold = walks_of(g["I_cut_off"], g["I_cut_src"], g["I_node"], g["I_stay"])
        if _id(path) in INTROSORT_DEPENDENT:
            assert {w[0] for w in got.walks()} == {w[0] for w in gold}
        else:
            assert got.walks() == gold
        if "label_off" in g:
            lo, ln, ld, lh = sep.ctx.labels()
            for k in range(len(g["sources"])):
                a, b = int(g["label_off"][k]), int(g["label_off"][k + 1])
                ref = {int(n): (d, int(h)) for n, d, h in zip(g["label_node"][a:b], g["label_dist"][a:b], g["label_hops"][a:b])}
                for n, d, h in zip(ln[lo[k]:lo[k + 1]], ld[lo[k]:lo[k + 1]], lh[lo[k]:lo[k + 1]]):
                    assert ref[int(n)] == (d, int(h))  # every label the GPU keeps is the reference's, bit for bit
    finally:
        sep.close()


@pytest.mark.gpu
@pytest.mark.parametrize("path", TIES, ids=_id)
def test_gpu_matches_tie_independent_properties(path):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sms_b200.separator import GpuOcwSeparator
    inst, g = load(path)
    sep = GpuOcwSeparator(inst.n, inst.rowptr, inst.col, inst.eid)
    try:
        sep.setRoundInput(g["x"], g["blacklist"])
        got = sep.separateRound(sources=g["sources"], dedup=False)
        has = np.zeros(len(g["sources"]), bool)
        for src, nodes, stay in got.walks():
            has[src] = True
            assert walk_length(inst, g["x"], nodes, stay) < 1.0
        assert np.array_equal(has, g["I_has_walk"])
    finally:
        sep.close()
