#!/usr/bin/env python
"""Markdown table of the bench lines under profiles/ (the table of DESIGN.md section 5 is this script's output):

    python scripts/summarize_bench.py [r02]

One row per workload / x* / mode: GPU rounds/s device-resident (`value`) and through the C ABI with host buffers (`e2e`),
the reference arm on the same box's host cores, and the ratio e2e / reference."""
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"


def load(path):
    try:
        return json.loads(open(path).read().strip().splitlines()[-1])
    except Exception:
        return None


def main():
    P = os.path.join(ROOT, "profiles")
    rows = [("random1m", "", "config 5: random 1M / 4M, iid U(0,1) x*"), ("torus100", "", "config 2: torus 100x100, iid U(0,1) x*"),
            ("torus100", "_lp_replay", "config 2: torus 100x100, root-LP iterates (30 LP solves, x* changes every step)"),
            ("torus3d20", "", "config 3: torus 20^3, iid U(0,1) x*"), ("torus3d20", "_lp_replay", "config 3: torus 20^3, root-LP iterates"),
            ("random20k", "", "config 4: random 20k / 80k, iid U(0,1) x*"), ("dense", "", "K6: triangle scan, K_500 (20.7 M triangles)")]
    print("| workload | mode | GPU device-resident | GPU through the C ABI (host buffers) | sweeps | unique cuts | reference, 16 host cores (its own sequential sweep) | C ABI / reference |")
    print("|---|---|---|---|---|---|---|---|")
    for w, x, title in rows:
        ref = load(os.path.join(P, f"{tag}_bench_reference_arm_{w}{x}.json"))
        for mode, suffix in (("I", ""), ("V", "_modeV")):
            d = load(os.path.join(P, f"{tag}_bench_{w}{x}{suffix}.json"))
            if d is None:
                continue
            rv = ref["value"] if ref else None
            unit = "rounds/s"
            cuts = d.get("unique_cuts", d.get("config", {}).get("violated", ""))
            sw = d.get("verbatim_sweeps") or ""
            print(f"| {title} | {mode if w != 'dense' else '-'} | {d['value']:.4g} {unit} ({d['ms_per_step']:.4g} ms) | {d['e2e']['value']:.4g} | {sw} | {cuts} | "
                  + (f"{rv:.4g} ({1e3 / rv:.4g} ms)" if rv else "-") + " | " + (f"x{d['e2e']['value'] / rv:.3g}" if rv else "-") + " |")
    d = load(os.path.join(P, f"{tag}_bench_random1m.json"))
    if d:
        r = d["roofline"]
        print()
        print(f"search kernel roofline (config 5): algorithmic {r['algorithmic_bytes'] / 1e9:.1f} GB per round / {d['kernel_ms']['search']:.1f} ms = "
              f"{r['achieved']:.1f} GB/s = {100 * r['frac']:.2f} % of the measured {r['peak']:.0f} GB/s; DRAM traffic {r['traffic'] / 1e9:.0f} GB per round")
        oi = d.get("config", {}).get("other_inputs", {})
        for k, v in oi.items():
            print(f"- {k}: " + ", ".join(f"{a} = {b:.4g}" if isinstance(b, float) else f"{a} = {b}" for a, b in v.items() if a != "workload"))
        cb = d.get("cpu_baseline")
        if cb:
            print(f"- cpu_baseline (1 thread): {cb['value']:.4g} rounds/s; {cb['sample']}")
    for n in (2, 4, 8):
        g = load(os.path.join(P, f"{tag}_bench_random1m_{n}gpu.json"))
        if g:
            print(f"- {n} GPUs: {g['value']:.4g} rounds/s device-resident ({g['ms_per_step']:.4g} ms), {g['e2e']['value']:.4g} through the C ABI, cut hash {g['config'].get('cut_hash')}")


if __name__ == "__main__":
    main()
