"""N > 1 path on CPU: two processes over gloo (127.0.0.1) run the sharded round's host logic -- the cyclic deal of
the sources (sms_b200/sharding.py, the mirror of smsgpu_set_sources), an all-gather of the per-rank walk lists, the
merge + global dedup -- with the CPU oracle standing in for the per-rank search kernels (checker used as the test
double, never shipped).  The merged list must equal the single-rank round's deduplicated list on every rank.
The same property is tested on real GPUs with NCCL by tests/test_gpu_parity.py::test_source_sharding_*."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from oracle import oracle_py
    from sms_b200 import instances as I
    from sms_b200 import sharding
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        orc = oracle_py.orc()
        res = {}
        for name, inst, x in (("torus", I.torus2d(14), I.xstar_uniform(2 * 14 * 14, seed=3)),
                              ("random", I.random_graph(300, 1200, seed=9), I.xstar_lp_like(1200, seed=4))):
            bl = inst.root_blacklist()
            src = inst.handler_sources()
            mine = sharding.local_sources(src, rank, world)
            s = orc.sweep(inst, x, blacklist=bl, sources=mine, mode="I", canonical=True)
            gathered = [None] * world
            dist.all_gather_object(gathered, s.walks())
            merged = sharding.merge_walk_lists(gathered, world)
            # single-rank truth, computed redundantly on every rank
            full = orc.sweep(inst, x, blacklist=bl, sources=src, mode="I", canonical=True)
            truth = sharding.merge_walk_lists([full.walks()], 1)
            # the source ids of the merged list are positions in the round's source list
            res[name] = (merged == truth, len(merged), len(full.walks()), int(sum(len(g) for g in gathered)))
            cnt = [None] * world
            dist.all_gather_object(cnt, len(merged))
            assert len(set(cnt)) == 1
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def test_pair_cyclic_deal_partitions_the_sources():
    sys.path.insert(0, ROOT)
    from sms_b200 import sharding
    src = np.arange(1001, dtype=np.int32) * 3
    for world in (1, 2, 3, 8):
        parts = [sharding.local_sources(src, r, world) for r in range(world)]
        assert sum(len(p) for p in parts) == len(src)
        back = np.empty_like(src)
        for r, p in enumerate(parts):
            back[sharding.global_index(np.arange(len(p)), r, world)] = p
        assert np.array_equal(back, src)
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 2  # pairs are the unit of the deal
    with pytest.raises(ValueError):
        sharding.local_sources(src, 2, 2)


def test_two_rank_gloo_round_equals_single_rank_round(orc):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, res in out:
        for name, (same, nmerged, nraw, ngathered) in res.items():
            assert same, f"rank {rank} {name}: merged list differs from the single-rank round"
            assert ngathered == nraw and 0 < nmerged <= nraw
