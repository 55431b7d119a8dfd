"""Summarise bench.py JSON lines (file arguments or stdin): one row per workload."""
import json
import sys


def row(name, d):
    if "error" in d:
        print(f"{name:5s} ERROR {d['error'][:200]}")
        return
    r = d["roofline"]
    pc = d.get("parity_check") or {}
    extra = ""
    if "sharded_output" in d:
        extra = " | sharded-C %.1f us allgather %.1f us balance %.3f ceiling %.2fx" % (
            d["sharded_output"]["ms_per_step"] * 1e3, d["reassembly_allgather_ms"] * 1e3, d["slab_balance"], d["model_scaling_ceiling"])
    if "e2e_pageable" in d:
        extra += " | pageable %.1f us (registered %.1f)" % (d["e2e_pageable"]["ms_per_step"] * 1e3,
                                                             d.get("e2e_pageable_registered", {}).get("ms_per_step", float("nan")) * 1e3)
    if "cpu_baseline" in d:
        extra += " | cpu %.2f GF/s (%s)" % (d["cpu_baseline"]["value"], d["cpu_baseline"]["kind"])
    print("%-5s %9.0f GF/s step %8.1f us kern %8.1f us (%s) warm %8.1f us e2e %8.1f us | %s frac %.3f hbm %.3f fp32 %.3f | parity %s %.1e upload %.0f ms%s" % (
        name, d["value"], d["ms_per_step"] * 1e3, r["kernel_us"], r.get("kernel", "?"), d["warm_l2"]["ms_per_step"] * 1e3,
        d["e2e"]["ms_per_step"] * 1e3, r["bound"], r["frac"], r.get("hbm_frac", r["frac"]), r["fp32_frac"],
        pc.get("ok"), pc.get("backward", float("nan")), d.get("upload_ms", float("nan")), extra))


def main():
    files = sys.argv[1:]
    lines = []
    for f in files:
        lines += open(f).read().splitlines()
    if not files:
        lines = sys.stdin.read().splitlines()
    for l in lines:
        if not l.strip():
            continue
        try:
            d = json.loads(l)
        except Exception:
            print(l.strip()[:300])
            continue
        if d.get("impl") == "reference":
            print("reference arm: %.2f GF/s %.1f ms/step (%s)" % (d["value"], d["ms_per_step"], d["config"]["workload"]))
            continue
        print("N = %d, scaling %s" % (d["n_gpus"], d["scaling"]))
        row(d["config"]["workload"], d)
        for name, sub in (d.get("per_workload") or {}).items():
            row(name, sub)


if __name__ == "__main__":
    main()
