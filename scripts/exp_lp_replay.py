"""Per-round tier breakdown of the LP replay (the logged root-LP iterates of the HiGHS cutting loop) on one context."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from sms_b200 import capi
from sms_b200 import instances as I

name = sys.argv[1] if len(sys.argv) > 1 else "torus2d_100"
d = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", f"lpseq_{name}.npz"))
inst = I.torus2d(100) if name == "torus2d_100" else I.torus3d(20)
assert int(d["n"]) == inst.n and np.array_equal(d["eu"], inst.eu) and np.array_equal(d["ev"], inst.ev)
ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=0)
bl = np.ascontiguousarray(d["blacklist"], np.uint8)
for rep in range(2):
    for r, x in enumerate(d["xs"]):
        ctx.set_round_input(np.ascontiguousarray(x), bl)
        ctx.separate_async(dedup=1)
        ctx.fetch()
        st = ctx.stats()
        if rep == 1:
            print(f"round {r:2d}: search {st['ms_search']:7.2f} ms tiers {[round(v, 2) for v in st['ms_tier']]} sources {st['tier_sources']} "
                  f"arcs {st['arcs_scanned']} settled {st['settled']} cuts {st['unique_cuts']} launches {st['total_launches']} total {st['ms_total_device']:.2f}")
ctx.close()
