"""The NCCL path of libsmsgpu.so (smsgpu.cu mergeAcrossRanks: header all-gather, packed all-gather of the per-GPU cut
lists, unpack, global dedup) against the single-GPU round -- byte for byte, on every rank, with dedup on and off.

One context per GPU inside ONE process, one thread per context (the threading model include/smsgpu.h names for a SCIP
process); NCCL refuses two ranks on one device, so the test needs >= 2 GPUs (`gpurun --gpus 2 -- pytest -m gpu
tests/test_multi_gpu.py`) and skips on a 1-GPU box.  bench.py checks the same thing at every N through the hash of the
final arrays (config.cut_hash)."""
import threading

import numpy as np
import pytest

from sms_b200 import instances as I

pytestmark = pytest.mark.gpu

FIELDS = ("off", "node", "eid", "is_stay", "n_cross", "src", "violation")


def _ngpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _run_ranks(inst, x, bl, nranks, rounds, sources=None):
    """rounds: list of params dicts; returns per rank the list of Cuts, one per round"""
    from sms_b200 import capi
    ctxs = [capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=r) for r in range(nranks)]
    uid = ctxs[0].nccl_unique_id()
    out = [None] * nranks
    err = [None] * nranks

    def work(r):
        try:
            c = ctxs[r]
            c.comm_init(uid, r, nranks)
            c.set_sources(sources, r, nranks)
            res = []
            for kw in rounds:
                c.set_round_input(x, bl)
                res.append(c.separate(**kw))
            out[r] = res
        except Exception as e:  # noqa: BLE001
            err[r] = e

    th = [threading.Thread(target=work, args=(r,)) for r in range(nranks)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=600)
    for c in ctxs:
        c.close()
    for e in err:
        if e is not None:
            raise e
    assert all(o is not None for o in out), "a rank did not finish"
    return out


@pytest.mark.parametrize("mk,xf", [(lambda: I.torus2d(40), I.xstar_uniform), (lambda: I.random_graph(3000, 12000, seed=11), I.xstar_lp_like),
                                   (lambda: I.random_graph(20000, 80000, seed=7), I.xstar_uniform)])
def test_merged_list_equals_single_gpu_round(mk, xf):
    n = _ngpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (NCCL refuses two ranks on one device)")
    from sms_b200 import build, capi
    build.build()
    inst = mk()
    x = xf(inst.m, seed=5)
    bl = inst.root_blacklist()
    rounds = [dict(dedup=1), dict(dedup=0), dict(dedup=1)]  # the third round runs with the history of the first two
    one = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=0)
    one.set_sources(None, 0, 1)
    ref = []
    for kw in rounds:
        one.set_round_input(x, bl)
        ref.append(one.separate(**kw))
    one.close()
    for nranks in sorted({2, min(n, 4), n}):
        got = _run_ranks(inst, x, bl, nranks, rounds)
        for r in range(nranks):
            for k, kw in enumerate(rounds):
                assert got[r][k].ncuts == ref[k].ncuts > 0
                for f in FIELDS:
                    assert np.array_equal(getattr(got[r][k], f), getattr(ref[k], f)), f"{nranks} ranks, rank {r}, round {k} ({kw}): {f} differs"


def test_explicit_source_list_and_odd_count():
    n = _ngpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    from sms_b200 import capi
    inst = I.torus3d(8)
    x = I.xstar_uniform(inst.m, seed=3)
    bl = inst.root_blacklist()
    src = np.random.default_rng(4).permutation(inst.n)[:301].astype(np.int32)  # odd count, arbitrary order
    one = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=0)
    one.set_sources(src, 0, 1)
    one.set_round_input(x, bl)
    ref = one.separate(dedup=1)
    one.close()
    got = _run_ranks(inst, x, bl, 2, [dict(dedup=1)], sources=src)
    for r in range(2):
        for f in FIELDS:
            assert np.array_equal(getattr(got[r][0], f), getattr(ref, f)), f"rank {r}: {f} differs"
