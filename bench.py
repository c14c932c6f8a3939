#!/usr/bin/env python
"""bench.py -- odd-cycle separation rounds/s (+ GTEPS) on synthetic instances of the BASELINE.json shapes.

    python bench.py --gpus N --steps K --warmup W [--workload random1m|torus100|torus3d20|random20k] [--impl reference]

A "step" is one root-node separation round: the round's x* and blacklist -> K1 arc records -> K2-K4 bounded
shortest paths + walk extraction + chord repair for every source 0..n-1 -> K5 canonical dedup
(-> NCCL all-gather + global dedup when N > 1).  N > 1 shards the SOURCES of the same round cyclically, in pairs, over
the ranks (each GPU holds a graph replica), so total work is fixed: "scaling": "strong".

  value    rounds/s with x*/blacklist already resident in HBM, result left on the device
  e2e      rounds/s through the C ABI with HOST buffers: H2D of x* + blacklist and D2H of the cut list inside
           the timed region
  roofline of the search kernel: algorithmic bytes = 12 B per arc scan + 8 B per settled node
  cpu_baseline / --impl reference: the UNMODIFIED reference separator (oracle/_ref, built from /root/reference)
           on the box's host cores over a bounded, seeded sample of the same round's sources, extrapolated
           linearly to the full round (sources are independent in mode I).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from sms_b200 import instances as I  # noqa: E402

WORKLOADS = {
    # name: (builder, description)
    "random1m": (lambda: I.random_graph(1_000_000, 4_000_000, seed=7),
                 "config 5: sparse random graph 1M nodes / 4M edges, iid U(0,1) x*, sources 0..n-1"),
    "torus100": (lambda: I.torus2d(100), "config 2: 100x100 toroidal grid (10k nodes, 20k edges), iid U(0,1) x*"),
    "torus3d20": (lambda: I.torus3d(20), "config 3: 20x20x20 toroidal grid (8k nodes, 24k edges), iid U(0,1) x*"),
    "random20k": (lambda: I.random_graph(20_000, 80_000, seed=7), "config 4: random graph 20k nodes / 80k edges, iid U(0,1) x*"),
}
DENSE_N = 500  # complete graph K_500: 124 750 edges, 20 708 500 triangles (the K6 workload)
WORKLOADS["dense"] = (lambda: I.from_edges(f"K{DENSE_N}", DENSE_N, np.stack(np.triu_indices(DENSE_N, 1), axis=1)),
                      f"dense instance: complete graph K_{DENSE_N}, all C({DENSE_N},3) triangles, iid U(0,1) x*: one triangle-inequality scan per step "
                      "(TriangleSeparator::getViolatedTriangles, triangles.hpp:34-53)")
WORKLOADS_N = {"random1m": 1_000_000, "torus100": 10_000, "torus3d20": 8_000, "random20k": 20_000}
# sources the reference arm times per step (bounded, seeded sample; the reference costs O(n) per source: at 1M nodes
# ~2.5 ms each, so 2048 sources are ~5 s of CPU work per step; SURVEY 7.3-9 asks for >= 1024).  The small configs run
# their whole round: no extrapolation.
REF_SAMPLE = {"random1m": 2048, "torus100": 10_000, "torus3d20": 8_000, "random20k": 20_000}


# dram bytes of the search of one full round, from the committed ncu capture (per launch; see roofline.traffic_note)
TRAFFIC_NCU = {"random1m": round(150.746169e9 * 1254134233 / 329005693)}  # 150.75 GB for the capture's 329.0 M (reference-equivalent) settled nodes
# the same capture's dram__bytes_read alone (120.31 GB), and what the microbenchmark of isolated accesses measured on this part
# (profiles/r02_sector_flavors.md: every isolated load brings a 128-byte line from DRAM; 33.9 G of them per second at most)
TRAFFIC_READ_NCU = {"random1m": round(120.307265e9 * 1254134233 / 329005693)}
ISOLATED_LINE_BYTES, ISOLATED_READS_PER_S = 128, 33.9e9


XSTAR_NOTE = {"fixed": "", "fresh": "; a NEW iid U(0,1) x* every step", "replay": "; x* sequence from the cutting loop (HiGHS root LP iterates, one per step)"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.idx = device_index
        self.proc = None
        self.lines = []

    def start(self):
        if os.environ.get("BENCH_NO_SMI"):  # experiment only: what the sampling itself costs
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", os.environ.get("BENCH_SMI_MS", "250"),
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def wait_first(self, timeout: float = 30.0):
        """Blocks until nvidia-smi has delivered its first sample: on a fresh box its start-up (NVML initialisation) takes
        seconds and, falling into the timed region, cost single steps tens to hundreds of ms."""
        t0 = time.perf_counter()
        while self.proc and not self.lines and self.proc.poll() is None and time.perf_counter() - t0 < timeout:
            time.sleep(0.05)

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the unmodified reference on host cores
# ------------------------------------------------------------------------------------------------------------------
LPSEQ = {"torus100": "lpseq_torus2d_100.npz", "torus3d20": "lpseq_torus3d_20.npz", "random20k": "lpseq_random20k.npz"}


def xstar_sequence(name: str, inst, mode: str, count: int):
    """The x* the timed steps feed, in order (step k uses entry k mod len):
      fixed   one iid U(0,1) point (seed 1), the same every step -- the mandated config-5 input
      fresh   a new iid U(0,1) point every step (seeds 1, 2, ...): nothing the previous round learnt about ball sizes
              carries over except what the graph itself determines
      replay  the logged x* sequence of the HiGHS root cutting loop on this instance (tests/golden/lpseq_*.npz,
              generated by tests/golden/make_lp_golden.py): x* changes every step the way a cutting loop changes it"""
    if mode == "fixed":
        return [I.xstar_uniform(inst.m, seed=1)]
    if mode == "fresh":
        return [I.xstar_uniform(inst.m, seed=1 + k) for k in range(max(1, min(count, 8)))]
    path = os.path.join(ROOT, "tests", "golden", LPSEQ.get(name, "-"))
    if not os.path.exists(path):
        raise SystemExit(f"bench.py: no logged LP sequence for workload {name} ({path})")
    g = np.load(path)
    if int(g["n"]) != inst.n or not np.array_equal(g["eu"], inst.eu) or not np.array_equal(g["ev"], inst.ev):
        raise SystemExit(f"bench.py: {path} was logged on a different instance")
    return [np.ascontiguousarray(x) for x in g["xs"]]


def _ref_worker(args):
    """One process = one separator (graph built once per call, outside the timed part) over its slice of the sample.
    Returns (set-up s, sum of per-source s, sum of squares, max per-source s, sources timed, walks)."""
    name, srcs, xmode, xidx = args
    from oracle import oracle_py  # checker library, executed here as the timed CPU baseline only
    inst = WORKLOADS[name][0]()
    seq = xstar_sequence(name, inst, xmode, xidx + 1)
    x = seq[xidx % len(seq)]
    eng = oracle_py.ref() if oracle_py.have_ref() else oracle_py.orc()
    # mode "V": the verbatim loop of ConshdlrOscw::sepaOddCycles (conshdlr_oscw.cpp:188-195) -- what the reference
    # actually executes per round; no per-source blacklist re-initialisation is charged to it
    s = eng.sweep(inst, x, sources=np.asarray(srcs, np.int32), mode="V", want_cuts=False)
    return (*s.timing, s.n_walks)


def reference_round(name: str, sample: int, procs: int, pool=None, xmode: str = "fixed", xidx: int = 0):
    """Times the reference on `sample` seeded sources of the workload's round and extrapolates to the full round:

        round seconds = set-up (m x updateEdgeWeight + blacklist reset: ONCE per round, not multiplied)
                        + (sum of the per-source times / sources timed) x sources of the round / procs

    i.e. from the MEAN cost of a source, not from the slowest of `procs` processes.  `procs` processes each own a separator
    (the reference is single-threaded; sources are dealt to processes as contiguous slices of its own sweep order).
    Returns a dict with the estimate, the sample's coefficient of variation and what was timed."""
    from oracle import oracle_py
    kind = "reference" if oracle_py.have_ref() else "port"
    n = WORKLOADS_N[name]
    sample = min(sample, n)
    rng = np.random.default_rng(12345)
    srcs = np.sort(rng.choice(n, sample, replace=False)) if sample < n else np.arange(n)
    chunks = [c for c in np.array_split(srcs, procs) if len(c)]
    jobs = [(name, c, xmode, xidx) for c in chunks]
    res = [_ref_worker(j) for j in jobs] if (pool is None or procs == 1) else pool.map(_ref_worker, jobs)
    setup = statistics.mean(r[0] for r in res)
    tsum = sum(r[1] for r in res)
    tsq = sum(r[2] for r in res)
    cnt = int(sum(r[4] for r in res))
    mean = tsum / max(cnt, 1)
    var = max(tsq / max(cnt, 1) - mean * mean, 0.0)
    est = setup + mean * n / procs  # (skipped sources -- the root -- cost nothing and are a 1e-6 share)
    return {"seconds": est, "setup_s": setup, "mean_source_s": mean, "cv_source": (var ** 0.5) / mean if mean > 0 else 0.0,
            "max_source_s": max(r[3] for r in res), "sources_timed": cnt, "n": n, "procs": procs, "kind": kind,
            "walks": int(sum(r[5] for r in res))}


def _ref_dense_worker(args):
    lo, hi = args
    from oracle import oracle_py
    inst = WORKLOADS["dense"][0]()
    eng = oracle_py.ref() if oracle_py.have_ref() else oracle_py.orc()
    # the reference's list order on a complete graph (small_ccs.cpp:18-43): u < x < y ascending, triangle {x, u, y}
    tri = dense_triangles(DENSE_N, lo, hi)
    nv, setup, scan = eng.small_cycle_scan_timed(inst, I.xstar_uniform(inst.m, seed=1), tri)
    return setup, scan, hi - lo, nv


def dense_triangles(c, lo, hi):
    """triangles lo..hi-1 of computeSmallCCs' list on the complete graph K_c with adjacency in ascending order"""
    out = np.empty((hi - lo, 3), np.int32)
    k = 0
    idx = 0
    for u in range(c):
        cnt_u = (c - 1 - u) * (c - 2 - u) // 2
        if idx + cnt_u <= lo:
            idx += cnt_u
            continue
        for x in range(u + 1, c):
            cnt_x = c - 1 - x
            if idx + cnt_x <= lo:
                idx += cnt_x
                continue
            ys = np.arange(x + 1, c)
            a, b = max(lo - idx, 0), min(hi - idx, cnt_x)
            if b > a:
                out[k:k + b - a, 0] = x
                out[k:k + b - a, 1] = u
                out[k:k + b - a, 2] = ys[a:b]
                k += b - a
            idx += cnt_x
            if idx >= hi:
                return out
    return out


def reference_dense_round(sample: int, procs: int, pool=None):
    """the reference's TriangleSeparator on `sample` triangles of the dense workload, `procs` processes with a slice each;
    round = the m updateWeights (once) + mean triangle cost x all triangles / procs"""
    from oracle import oracle_py
    T = DENSE_N * (DENSE_N - 1) * (DENSE_N - 2) // 6
    sample = min(sample, T)
    per = sample // procs
    start = np.random.default_rng(7).integers(0, T - per, size=procs)
    jobs = [(int(a), int(a) + per) for a in start]
    res = [_ref_dense_worker(j) for j in jobs] if (pool is None or procs == 1) else pool.map(_ref_dense_worker, jobs)
    setup = statistics.mean(r[0] for r in res)
    mean = sum(r[1] for r in res) / sum(r[2] for r in res)
    return {"seconds": setup + mean * T / procs, "setup_s": setup, "mean_triangle_s": mean, "triangles_timed": per * procs, "T": T,
            "procs": procs, "kind": "reference" if oracle_py.have_ref() else "port"}


def run_reference_arm_dense(args):
    import multiprocessing as mp
    procs = os.cpu_count() or 1
    runs = []
    with mp.get_context("spawn").Pool(procs) as pool:
        for it in range(args.warmup + args.steps):
            r = reference_dense_round(1_600_000, procs, pool)
            if it >= args.warmup:
                runs.append(r)
    sec = statistics.mean(r["seconds"] for r in runs)
    r0 = runs[-1]
    line = {"impl": "reference", "metric": "triangle-inequality scans/sec (K6 rounds/s)", "value": 1.0 / sec, "unit": "rounds/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": {"workload": WORKLOADS["dense"][1], "name": "dense"},
            "cpu_baseline": {"value": 1.0 / sec, "unit": "rounds/s", "cores": procs, "kind": r0["kind"],
                             "sample": f"{r0['triangles_timed']} of {r0['T']} triangles per step over {procs} processes; round = {r0['setup_s'] * 1e3:.1f} ms "
                                       f"(m updateWeights, once) + {r0['mean_triangle_s'] * 1e9:.0f} ns per triangle x {r0['T']} / {procs} cores"},
            "e2e": {"value": 1.0 / sec, "unit": "rounds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if args.workload == "dense":
        return run_reference_arm_dense(args)
    import multiprocessing as mp
    procs = os.cpu_count() or 1
    sample = REF_SAMPLE[args.workload]
    ctx = mp.get_context("spawn")
    runs = []
    with ctx.Pool(procs) as pool:
        pool.map(_ref_worker, [(args.workload, np.arange(2), "fixed", 0)] * procs)  # page in the libraries once per worker
        for it in range(args.warmup + args.steps):
            r = reference_round(args.workload, sample, procs, pool, args.xstar, max(it - args.warmup, 0))
            if it >= args.warmup:
                runs.append(r)
    sec = statistics.mean(r["seconds"] for r in runs)
    spread = (max(r["seconds"] for r in runs) - min(r["seconds"] for r in runs)) / sec if len(runs) > 1 else 0.0
    r0 = runs[-1]
    val = 1.0 / sec
    extrap = r0["n"] / max(r0["sources_timed"], 1)
    line = {
        "impl": "reference", "metric": "odd-cycle separation rounds/sec", "value": val, "unit": "rounds/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][1] + XSTAR_NOTE[args.xstar], "name": args.workload, "xstar": args.xstar,
                   "mode": "V (the reference's own sequential sweep, conshdlr_oscw.cpp:188-195) within each process slice; "
                           "the GPU arm runs mode I (independent sources), which is MORE work (no source is skipped "
                           "because an earlier one blacklisted it)"},
        "cpu_baseline": {"value": val, "unit": "rounds/s", "cores": procs, "kind": r0["kind"],
                         "sample": f"{r0['sources_timed']} seeded sources of {r0['n']} per step over {procs} processes (one separator each); "
                                   f"round = set-up {r0['setup_s'] * 1e3:.1f} ms (once, not multiplied) + mean source "
                                   f"{r0['mean_source_s'] * 1e3:.4f} ms x {r0['n']} / {procs} cores"
                                   + (f" (extrapolated x{extrap:.1f})" if extrap > 1.001 else " (the whole round, no extrapolation)"),
                         "cv_per_source": round(r0["cv_source"], 3), "spread_between_steps": round(spread, 4),
                         "mean_source_ms": r0["mean_source_s"] * 1e3, "setup_ms": r0["setup_s"] * 1e3},
        "e2e": {"value": val, "unit": "rounds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------------------------
# own arm
# ------------------------------------------------------------------------------------------------------------------
def time_rounds(ctx, torch, dist, world, steps, warmup, step_fn, flush):
    for _ in range(warmup):
        step_fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    per = []
    for _ in range(steps):
        flush()
        t1 = time.perf_counter()
        step_fn()
        if os.environ.get("BENCH_DEBUG"):
            torch.cuda.synchronize()
            ctx.fetch()
            st = ctx.stats()
            print(f"[bench] step wall {1e3 * (time.perf_counter() - t1):.1f} ms, device search {st['ms_search']:.1f} ms, tiers {st['tier_sources']}, "
                  f"ms_tier {[round(v, 1) for v in st['ms_tier']]}", file=sys.stderr, flush=True)
        per.append(None)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return dt, per


_INST = {}  # instances built once per process


def measure_workload(name, args, torch, dist, rank, world, device, full=True):
    from sms_b200 import capi
    if name not in _INST:
        _INST[name] = WORKLOADS[name][0]()
    inst = _INST[name]
    xs = xstar_sequence(name, inst, getattr(args, "xstar", "fixed"), args.steps + args.warmup)
    x = xs[0]
    bl = inst.root_blacklist()
    os.environ.pop("SMSGPU_NO_HISTORY", None)
    if getattr(args, "no_history", False):
        os.environ["SMSGPU_NO_HISTORY"] = "1"  # read when the context is created
    ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=device)
    if world > 1:
        uid = [ctx.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], rank, world)
    ctx.set_sources(None, rank, world)
    p = ctx.params(dedup=1, mode=getattr(args, "mode", 0))
    import ctypes as C
    lib = ctx.lib
    # L2 flush buffer (> 126 MB), written between timed iterations
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def flush():
        flush_buf.zero_()

    # ---- value leg: inputs resident in HBM, result left on the device ----
    dxs = [torch.from_numpy(v).cuda() for v in xs]
    dbl = torch.from_numpy(bl).cuda()
    torch.cuda.synchronize()
    kstep = [0, 0]
    agg = {"ms_search": 0.0, "arcs": 0, "settled": 0, "launches": 0, "search_launches": 0, "n": 0, "ms_dev": 0.0,
           "ms_weights": 0.0, "ms_dedup": 0.0}

    def step_dev():
        dx = dxs[kstep[0] % len(dxs)]
        kstep[0] += 1
        ctx.set_round_input_device(dx.data_ptr(), dbl.data_ptr())
        ctx._check(lib.smsgpu_separate_async(ctx.h, C.byref(p)))
        return None

    sampler = ClockSampler(device)
    sampler.start()  # before the warm-up: starting nvidia-smi (NVML init) inside the timed region once cost a step 400 ms
    # warm-up: allocations, tier scratch, and the history the engine keeps between rounds (issue order, tier-0 decision)
    for _ in range(max(args.warmup, 3)):
        flush()  # also warms the flush itself: its first call loads torch's fill kernel (measured: +7..75 ms on one step)
        step_dev()
    torch.cuda.synchronize()
    sampler.wait_first()
    kstep[0] = 0  # the timed steps walk through the sequence from its start
    if len(xs) > 1:  # ... behind one round on its last entry, so that step 0 also meets the history of ANOTHER x*
        kstep[0] = len(xs) - 1
        step_dev()
        kstep[0] = 0
    dt, _ = time_rounds(ctx, torch, dist, world, args.steps, 0, step_dev, flush)
    clocks = sampler.stop()
    # per-round device stats of the last round (events on the library's own stream)
    ctx.fetch()
    st = ctx.stats()
    ncuts, nnodes = ctx.cut_counts()
    res = {"dt": dt, "stats": st, "ncuts": ncuts, "nnodes": nnodes, "clocks": clocks, "n": inst.n, "m": inst.m}
    # aggregate algorithmic work over ranks (each rank searched its share of the sources)
    work = torch.tensor([st["arcs_scanned"], st["settled"], st["sources_run"], st["raw_walks"], st["mirrored_arcs"], st["mirrored_sources"]],
                        dtype=torch.float64, device="cuda")
    ms = torch.tensor([st["ms_search"], st["ms_weights"], st["ms_dedup"], st["ms_total_device"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    res["arcs"], res["settled"], res["sources_run"], res["raw_walks"], res["mirrored_arcs"], res["mirrored_sources"] = (int(v) for v in work.tolist())
    res["ms_search"], res["ms_weights"], res["ms_dedup"], res["ms_total_device"] = ms.tolist()

    # ---- e2e leg: host buffers through the C ABI, H2D and D2H inside the timed region ----
    # A handler gathers x* edge by edge (one SCIPgetSolVal per edge, conshdlr_oscw.cpp:175-180) into page-locked memory (the
    # context's own staging buffers, smsgpu_input_buffers, or its own); timed here: H2D of x* + blacklist -> round -> D2H of
    # the cut list.  Rank 0 is the consumer (SCIP's thread): the other ranks upload their replica of x* and run their share,
    # but do not read the merged list back.
    if full:
        # the round's inputs lie in page-locked host memory (the bench contract: "host->device copy of that step's inputs
        # from pinned host memory"): smsgpu_set_round_input DMAs them from where they lie
        pxs = [torch.from_numpy(v).pin_memory() for v in xs]
        pbl = torch.from_numpy(bl).pin_memory()
        hxs = [t.numpy() for t in pxs]
        hbl = pbl.numpy()

        def step_host():
            hx = hxs[kstep[1] % len(hxs)]
            kstep[1] += 1
            ctx.set_round_input(hx, hbl)
            if rank == 0:
                ctx.separate_nocopy(p)
            else:
                ctx._check(lib.smsgpu_separate_async(ctx.h, C.byref(p)))
            return None
        dt2, _ = time_rounds(ctx, torch, dist, world, args.steps, max(1, args.warmup // 2), step_host, flush)
        res["e2e_dt"] = dt2
        res["h2d"] = inst.m * 8 + inst.n
        if rank == 0:
            ncuts2, nnodes2 = ctx.cut_counts()
            res["d2h"] = (ncuts2 + 1) * 8 + nnodes2 * 9 + ncuts2 * 16
    # ---- the final arrays of every rank, hashed: all ranks must hold the identical list, and the same list at every N ----
    ctx.fetch()
    cuts = ctx.get_cuts()
    import hashlib
    h = hashlib.blake2b(digest_size=8)
    for f in ("off", "node", "eid", "is_stay", "src"):
        h.update(np.ascontiguousarray(getattr(cuts, f)).tobytes())
    res["cut_hash"] = h.hexdigest()
    if world > 1:
        hashes = [None] * world
        dist.all_gather_object(hashes, res["cut_hash"])
        res["cut_hashes"] = hashes
    else:
        res["cut_hashes"] = [res["cut_hash"]]
    ctx.close()
    return res


def measure_dense(args, torch, device):
    """K6: one triangle-inequality scan over all triangles of the dense workload per step"""
    from sms_b200 import build, capi
    inst = WORKLOADS["dense"][0]()
    x = I.xstar_uniform(inst.m, seed=1)
    ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid, device=device)
    t0 = time.perf_counter()
    T, _ = ctx.small_cycles_enumerate(False)  # scip_initsol: once per solve, not part of a round
    t_enum = time.perf_counter() - t0
    assert T == DENSE_N * (DENSE_N - 1) * (DENSE_N - 2) // 6
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    dx = torch.from_numpy(x).cuda()
    dbl = torch.zeros(inst.n, dtype=torch.uint8, device="cuda")
    hx, hbl = ctx.input_buffers()
    import ctypes as C
    lib = ctx.lib
    sampler = ClockSampler(device)
    sampler.start()

    def step_dev():  # x* resident, walks left on the device
        ctx.set_round_input_device(dx.data_ptr(), dbl.data_ptr())
        ctx._check(lib.smsgpu_small_cycles_scan_device(ctx.h, 3))

    np.copyto(hx, x)  # x* lies in the context's page-locked staging buffer

    def step_host():  # x* from the host (H2D), walks back on the host (D2H)
        ctx.set_round_input(hx, hbl)
        return ctx.small_cycles_scan_nocopy(3)

    for _ in range(max(args.warmup, 3)):
        flush_buf.zero_()
        step_dev()
    torch.cuda.synchronize()
    sampler.wait_first()
    ms_scan, ms_gather = [], []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        flush_buf.zero_()
        step_dev()
        tm = ctx.small_cycles_timing()  # CUDA events on the library's stream (the scan call waits for the counts)
        ms_scan.append(tm["ms_scan"])
        ms_gather.append(tm["ms_gather"])
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    clocks = sampler.stop()
    for _ in range(2):
        flush_buf.zero_()
        step_host()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        flush_buf.zero_()
        U, N = step_host()
    torch.cuda.synchronize()
    dt2 = time.perf_counter() - t0
    ctx.close()
    peak, peak_src = peaks()
    sec = dt / args.steps
    k_scan = statistics.mean(ms_scan) * 1e-3
    alg_scan = 12.0 * T + 8.0 * inst.m + 1.0 * T  # 3 edge ids per triangle + x* once + the code byte
    out_bytes = U * (12 + 12 + 3 + 4 + 4 + 8 + 8)
    line = {"metric": "triangle-inequality scans/sec (K6 rounds/s)", "value": 1.0 / sec, "unit": "rounds/s", "n_gpus": 1, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": WORKLOADS["dense"][1], "name": "dense", "n": inst.n, "m": inst.m, "triangles": T, "violated": U,
                       "l2": "256 MiB buffer rewritten between timed iterations; the list itself is 12 B x T = %.0f MB" % (12.0 * T / 1e6),
                       "enumeration_ms_once": t_enum * 1e3, "lib_source_hash": build.built_hash()},
            "gtriangles_per_s": T / sec / 1e9,
            "kernel_ms": {"scan": statistics.mean(ms_scan), "rank_and_gather": statistics.mean(ms_gather)},
            "roofline": {"bound": "hbm", "kernel": "smallCycleScanKernel<3>", "achieved": alg_scan / k_scan / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg_scan / k_scan / 1e9 / peak, "traffic": None, "peak_source": peak_src, "algorithmic_bytes": alg_scan,
                         "whole_scan": {"algorithmic_bytes": alg_scan + out_bytes, "note": "scan + rank + gather of the violated walks (51 B each out, 24 B each re-read)",
                                        "achieved": (alg_scan + out_bytes) / ((statistics.mean(ms_scan) + statistics.mean(ms_gather)) * 1e-3) / 1e9}},
            "e2e": {"value": args.steps / dt2, "unit": "rounds/s", "h2d_bytes_per_step": inst.m * 8 + inst.n, "d2h_bytes_per_step": int(out_bytes + 8)},
            "gpu_launches": 6 * args.steps, "clocks": clocks}
    if not args.no_cpu_baseline:
        r = reference_dense_round(400_000, 1)
        line["cpu_baseline"] = {"value": 1.0 / r["seconds"], "unit": "rounds/s", "cores": 1, "kind": r["kind"],
                                "sample": f"{r['triangles_timed']} of {r['T']} triangles, 1 thread; round = {r['setup_s'] * 1e3:.1f} ms (m updateWeights, once) + "
                                          f"{r['mean_triangle_s'] * 1e9:.0f} ns per triangle x {r['T']}"}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="random1m", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--xstar", default="fixed", choices=["fixed", "fresh", "replay"], help="x* of the timed steps (see xstar_sequence)")
    ap.add_argument("--no-history", action="store_true", help="the engine forgets the previous round (issue order, tier decisions): every round is a first round")
    ap.add_argument("--mode", type=int, default=0, choices=[0, 1], help="0: independent sources (default); 1: the reference's sequential sweep (SMSGPU_MODE_VERBATIM, 1 GPU)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary torus100 measurement")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the separation engine has no CPU fallback")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE {world}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    args.warmup = max(args.warmup, 3)

    from sms_b200 import build
    build.build()
    if args.workload == "dense":
        if world > 1:
            raise SystemExit("the dense (K6) workload is a single-GPU scan")
        return measure_dense(args, torch, local)
    res = measure_workload(args.workload, args, torch, dist, rank, world, local, full=True)
    extra = {}
    if world == 1 and not args.no_extra and args.workload == "random1m" and args.xstar == "fixed" and args.mode == 0 and not args.no_history:
        # the same engine away from its best case (the headline feeds the SAME x* every step, so the previous round predicts
        # the issue order and the tier of every source exactly): a new x* every step, and no memory of the previous round
        ns = argparse.Namespace
        r3 = measure_workload("random1m", ns(steps=3, warmup=3, xstar="fresh", no_history=False, mode=0), torch, dist, rank, world, local, full=False)
        extra["random1m_new_xstar_every_step"] = {"rounds_per_s_device": 3 / r3["dt"], "ms_search_last": r3["ms_search"], "tier_sources_last": r3["stats"]["tier_sources"]}
        r4 = measure_workload("random1m", ns(steps=3, warmup=3, xstar="fixed", no_history=True, mode=0), torch, dist, rank, world, local, full=False)
        extra["random1m_no_history"] = {"rounds_per_s_device": 3 / r4["dt"], "ms_search_last": r4["ms_search"], "tier_sources_last": r4["stats"]["tier_sources"]}
        r2 = measure_workload("torus100", ns(steps=50, warmup=5, xstar="fixed", no_history=False, mode=0), torch, dist, rank, world, local, full=True)
        extra["torus100"] = {"rounds_per_s_device": 50 / r2["dt"], "rounds_per_s_e2e": 50 / r2["e2e_dt"],
                             "ms_search": r2["ms_search"], "unique_cuts": r2["ncuts"], "arcs_scanned": r2["arcs"]}
        if os.path.exists(os.path.join(ROOT, "tests", "golden", LPSEQ["torus100"])):
            r5 = measure_workload("torus100", ns(steps=30, warmup=3, xstar="replay", no_history=False, mode=0), torch, dist, rank, world, local, full=True)
            extra["torus100_lp_replay"] = {"workload": "100x100 torus" + XSTAR_NOTE["replay"] + ", 30 LP solves", "rounds_per_s_device": 30 / r5["dt"],
                                           "rounds_per_s_e2e": 30 / r5["e2e_dt"], "ms_search_last": r5["ms_search"], "tier_sources_last": r5["stats"]["tier_sources"]}
    if len(set(res["cut_hashes"])) != 1:
        if rank == 0:
            print(f"bench.py: the ranks returned DIFFERENT cut lists: {res['cut_hashes']}", file=sys.stderr)
        if world > 1:
            dist.destroy_process_group()
        return 3
    if rank == 0:
        peak, peak_src = peaks()
        sec = res["dt"] / args.steps
        # roofline on the work actually done: the scans / settled nodes of mirrored twin sources are results obtained by
        # symmetry, not traffic (settled nodes of mirrored sources are not counted separately: same share as their arcs)
        done_share = 1.0 - res["mirrored_arcs"] / max(res["arcs"], 1)
        alg_bytes = (12.0 * res["arcs"] + 8.0 * res["settled"]) * done_share
        # the search kernels of all ranks run concurrently: aggregate bytes / slowest rank's kernel time / N GPUs
        ach = alg_bytes / (res["ms_search"] * 1e-3) / 1e9 / world if res["ms_search"] > 0 else 0.0
        line = {
            "metric": "odd-cycle separation rounds/sec", "value": 1.0 / sec, "unit": "rounds/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload][1] + XSTAR_NOTE[args.xstar], "name": args.workload, "xstar": args.xstar,
                       "history": "off (every round is a first round)" if args.no_history else "on (issue order / tier decisions from the previous round)",
                       "n": res["n"], "m": res["m"],
                       "sources": res["n"],
                       "mode": ("V (the reference's sequential sweep as a fixed point of stamped sweeps), canonical ties, dedup on" if args.mode == 1 else
                                "I (independent sources), canonical ties, dedup on; twin sources (v,0)/(v,1) searched once when tie-free"),
                       "l2": "256 MiB buffer rewritten between timed iterations; arc array alone is 16 B x 2m",
                       "parallelism": f"source pairs cyclic over {world} GPU(s), graph replicated",
                       # blake2b-64 of the final (off, node, eid, is_stay, src) arrays; identical on every rank (checked:
                       # the run fails otherwise) and, the inputs being seeded, identical at every N
                       "cut_hash": res["cut_hash"], "unique_cuts": res["ncuts"], "ranks_agree": True,
                       "lib_source_hash": build.built_hash()},
            "gteps": res["arcs"] / sec / 1e9,  # arc scans of the reference's sweep (its numberOfEdgesSeen) per second
            "gteps_scanned": (res["arcs"] - res["mirrored_arcs"]) / sec / 1e9,  # ... of which actually scanned by the GPU
            "mirrored_sources": res["mirrored_sources"], "mirrored_arcs": res["mirrored_arcs"],
            "arcs_scanned_per_round": res["arcs"], "settled_per_round": res["settled"], "raw_walks": res["raw_walks"],
            "unique_cuts": res["ncuts"],
            "kernel_ms": {"weights": res["ms_weights"], "search": res["ms_search"], "dedup_and_merge": res["ms_dedup"],
                          "device_total": res["ms_total_device"]},
            "tier_sources": res["stats"]["tier_sources"], "verbatim_sweeps": int(res["stats"].get("verbatim_sweeps", 0)),
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": TRAFFIC_NCU.get(args.workload), "peak_source": peak_src,
                         "kernel": "ocwSearchPairsKernel (tier 1; tiers 0/2: ocwSearchKernel)",
                         "traffic_note": "dram__bytes_read+write of the tier-1 launch from profiles/r02_pairs_tier1_random1m_256k_final.md "
                                         "(ncu --set full, 262144 sources), scaled to the round by settled nodes; every isolated 32-byte "
                                         "access brings a 128-byte line from DRAM (profiles/r02_sector_flavors.md); the path is bound by the "
                                         "latency of dependent accesses, not by bytes (DESIGN.md section 5)",
                         "algorithmic_bytes": alg_bytes, "per_gpu": True,
                         # second view (not the contract's `frac`): line fetches per second against the measured rate of isolated reads
                         "access_rate": ({"unit": "128-byte DRAM line fetches per second per GPU",
                                          "achieved": TRAFFIC_READ_NCU[args.workload] / ISOLATED_LINE_BYTES / (res["ms_search"] * 1e-3) / world,
                                          "ceiling": ISOLATED_READS_PER_S,
                                          "frac": TRAFFIC_READ_NCU[args.workload] / ISOLATED_LINE_BYTES / (res["ms_search"] * 1e-3) / world / ISOLATED_READS_PER_S,
                                          "source": "dram__bytes_read of profiles/r02_pairs_tier1_random1m_256k_final.md scaled to the round; "
                                                    "ceiling: scripts/sector_flavors.cu, profiles/r02_sector_flavors.md"}
                                         if args.workload in TRAFFIC_READ_NCU and res["ms_search"] > 0 and args.mode == 0 else None)},
            "e2e": {"value": args.steps / res["e2e_dt"], "unit": "rounds/s", "h2d_bytes_per_step": res["h2d"],
                    "d2h_bytes_per_step": res["d2h"]},
            "gpu_launches": int(res["stats"]["total_launches"]) * args.steps,
            "clocks": res["clocks"],
            "extra": extra,
        }
        if extra:  # the driver keeps `config`: the measurements away from the headline's best case travel with it
            line["config"]["other_inputs"] = extra
        if world == 1 and not args.no_cpu_baseline:
            r = reference_round(args.workload, min(REF_SAMPLE[args.workload], 1024), 1, None, args.xstar, 0)
            line["cpu_baseline"] = {"value": 1.0 / r["seconds"], "unit": "rounds/s", "cores": 1, "kind": r["kind"],
                                    "sample": f"{r['sources_timed']} seeded sources of {r['n']}, 1 thread (the reference is single-threaded), mode V; "
                                              f"round = set-up {r['setup_s'] * 1e3:.1f} ms (once) + mean source {r['mean_source_s'] * 1e3:.4f} ms x {r['n']}",
                                    "cv_per_source": round(r["cv_source"], 3)}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
