"""Debug helper: one separation round on a small instance through the C ABI (run under compute-sanitizer)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sms_b200 import instances as I
from sms_b200.separator import GpuOcwSeparator
n, m = int(sys.argv[1]), int(sys.argv[2])
kw = {}
if len(sys.argv) > 3: kw["start_tier"] = int(sys.argv[3])
if len(sys.argv) > 4: kw["tile_width"] = int(sys.argv[4])
if len(sys.argv) > 5: kw["debug_flags"] = int(sys.argv[5])
inst = I.random_graph(n, m, seed=3)
x = I.xstar_uniform(inst.m, seed=7)
sep = GpuOcwSeparator(inst.n, inst.rowptr, inst.col, inst.eid)
sep.setRoundInput(x, inst.root_blacklist())
got = sep.separateRound(dedup=False, **kw)
print("ok", got.ncuts, sep.ctx.stats())
