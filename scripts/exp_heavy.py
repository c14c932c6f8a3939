"""Experiment: how heavy are the sources that leave the pairs tier?  Prints the distribution of settled nodes per
source for tier 1 and tier 2 of one config-5 round (or a shard of it)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from bench import WORKLOADS
from sms_b200 import capi
from sms_b200 import instances as I

inst = WORKLOADS["random1m"][0]()
x = I.xstar_uniform(inst.m, seed=1)
ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=0)
ctx.set_sources(np.arange(inst.n, dtype=np.int32), rank=0, nranks=4)
ctx.set_round_input(x, inst.root_blacklist())
ctx.separate_async(dedup=1, export_src_stats=1)
ctx.fetch()
arcs, settled, tier = ctx.src_stats()
for t in (1, 2, 3):
    s = settled[tier == t]
    if len(s) == 0:
        continue
    print("tier", t, "sources", len(s), "settled sum", int(s.sum()), "mean", float(s.mean()), "quantiles 50/90/99/99.9/max",
          [int(v) for v in np.quantile(s, [0.5, 0.9, 0.99, 0.999, 1.0])])
s2 = np.sort(settled[tier == 2])[::-1]
print("tier 2 top 20:", s2[:20].tolist())
print("tier 2 histogram (settled in k):", np.histogram(s2, bins=[0, 2000, 4000, 6000, 8000, 10000, 15000, 20000, 30000, 50000, 10**9])[0].tolist())
s1 = settled[tier == 1]
print("tier 1 above 6000:", int((s1 > 6000).sum()), "above 8000:", int((s1 > 8000).sum()), "max", int(s1.max()))
ctx.close()
