"""Profiling driver (run plain first, then under ncu): a few separation rounds of one bench workload through the
C ABI on cuda:0.  --sources N limits the round to the first N source ids so that an ncu replay stays short."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from bench import WORKLOADS
from sms_b200 import capi
from sms_b200 import instances as I

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="random1m")
ap.add_argument("--sources", type=int, default=0)
ap.add_argument("--rounds", type=int, default=2)
ap.add_argument("--start-tier", type=int, default=0)
ap.add_argument("--rank", type=int, default=0, help="with --nranks: run this rank's shard of the sources only (no communicator)")
ap.add_argument("--nranks", type=int, default=1)
ap.add_argument("--xstar", default="uniform", choices=["uniform", "lp_like"])
a = ap.parse_args()
inst = WORKLOADS[a.workload][0]()
x = I.xstar_uniform(inst.m, seed=1) if a.xstar == "uniform" else I.xstar_lp_like(inst.m, seed=1)
ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=0)
if a.sources or a.nranks > 1:
    ctx.set_sources(np.arange(a.sources or inst.n, dtype=np.int32), rank=a.rank, nranks=a.nranks)
for r in range(a.rounds):
    ctx.set_round_input(x, inst.root_blacklist())
    ctx.separate_async(dedup=1, start_tier=a.start_tier)
    ctx.fetch()
    st = ctx.stats()
    print({k: st[k] for k in ("sources_run", "arcs_scanned", "settled", "raw_walks", "unique_cuts", "tier_sources", "ms_weights",
                              "ms_search", "ms_tier", "ms_dedup", "ms_total_device", "total_launches", "mirrored_sources", "mirrored_arcs")})
ctx.close()
