"""Experiment driver for the lanes engine (one source per thread): config 5 (random 1M / 4M, iid U(0,1) x*), a few rounds
per setting of SMSGPU_LANES / SMSGPU_LANES_TILES / SMSGPU_LANES_SETS on one context each; prints per-round search times
and the cut count (which must not depend on the setting)."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from bench import WORKLOADS
from sms_b200 import capi
from sms_b200 import instances as I

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="random1m")
ap.add_argument("--rounds", type=int, default=3)
ap.add_argument("--sources", type=int, default=0)
ap.add_argument("--settings", default="0:0:0,1:40960:2")
a = ap.parse_args()
t0 = time.time()
inst = WORKLOADS[a.workload][0]()
x = I.xstar_uniform(inst.m, seed=1)
bl = inst.root_blacklist()
print(f"instance {inst.n}/{inst.m} in {time.time() - t0:.1f} s", flush=True)
for setting in a.settings.split(","):
    f = setting.split(":") + ["0"] * 7
    lanes, tiles, sets, nb, tablg, park, l2f = f[:7]
    if l2f != "0":
        os.environ["SMSGPU_L2_FETCH"] = l2f
    else:
        os.environ.pop("SMSGPU_L2_FETCH", None)
    os.environ["SMSGPU_LANES_NB"] = nb if nb != "0" else "2048"
    os.environ["SMSGPU_LANES_TABLG"] = tablg
    os.environ["SMSGPU_LANES_QCAP"] = park if park != "0" else "8192"
    os.environ["SMSGPU_LANES"] = lanes
    os.environ["SMSGPU_LANES_TILES"] = tiles if tiles != "0" else "40960"
    os.environ["SMSGPU_LANES_SETS"] = sets if sets != "0" else "2"
    ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=0)
    if a.sources:
        ctx.set_sources(np.arange(a.sources, dtype=np.int32))
    for r in range(a.rounds):
        ctx.set_round_input(x, bl)
        t1 = time.time()
        ctx.separate_async(dedup=1, start_tier=1)
        ctx.fetch()
        wall = time.time() - t1
        st = ctx.stats()
        print(f"lanes={lanes} tiles={tiles} sets={sets} nb={nb} tablg={tablg} qcap={park} l2fetch={l2f} round {r}: search {st['ms_search']:.1f} ms tiers {[round(v, 1) for v in st['ms_tier']]} "
              f"wall {wall * 1e3:.0f} ms tier_sources {st['tier_sources']} cuts {st['unique_cuts']} arcs {st['arcs_scanned']} "
              f"settled {st['settled']} mirrored {st['mirrored_sources']}", flush=True)
    ctx.close()
