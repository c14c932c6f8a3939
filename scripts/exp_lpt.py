"""Experiment: does longest-first source order (by the previous round's per-source settled counts) shorten the round?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench import WORKLOADS
from sms_b200 import capi, instances as I
inst = WORKLOADS["random1m"][0]()
x = I.xstar_uniform(inst.m, seed=1)
ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=0)
def run(tag, **kw):
    ctx.set_round_input(x, inst.root_blacklist())
    ctx.separate_async(dedup=1, **kw)
    ctx.fetch()
    st = ctx.stats()
    print(tag, "search %.1f ms" % st["ms_search"], [round(v, 1) for v in st["ms_tier"]], st["tier_sources"], st["unique_cuts"], flush=True)
run("natural order  ")
run("natural order  ", export_src_stats=1)
arcs, settled, tier = ctx.src_stats()
order = np.argsort(-settled, kind="stable").astype(np.int32)
ctx.set_sources(order)
run("longest first  ")
run("longest first  ")
rng = np.random.default_rng(1)
key = (settled >> 4).astype(np.int64) * (1 << 21) + rng.integers(0, 1 << 20, len(settled))
order3 = np.argsort(-key, kind="stable").astype(np.int32)
ctx.set_sources(order3)
run("bins of 16, random inside, tier 1 start", start_tier=1)
ctx.set_sources(order)
run("exact, tier 1 start", start_tier=1)
# coarse: 16 bins
bins = np.minimum(settled // 512, 15)
order2 = np.argsort(-bins, kind="stable").astype(np.int32)
ctx.set_sources(order2)
run("16 bins, heavy first")
ctx.set_sources(order)
run("longest first, tier 1 start", start_tier=1)
ctx.close()
