#!/usr/bin/env python
"""Summarise an ncu launch list (--csv --log-file ... with gpu__time_duration.sum and, optionally,
dram__bytes_read.sum / dram__bytes_write.sum): launches, total time, share of the run and DRAM
traffic per kernel, plus the per-step figures when --steps says how many steps the run had.

    python tools/ncu_launches.py gpurun_out/r02_launches_c4.csv --steps 4 > profiles/r02_launches_c4.txt
"""
import argparse
import collections
import csv
import re


def short(name):
    name = re.sub(r"\(anonymous namespace\)::|<?unnamed>::", "", name)
    name = re.sub(r"^void\s+", "", name)
    m = re.match(r"([\w:]+)(<[^(]*>)?", name)
    base = m.group(1) if m else name
    targ = (m.group(2) or "") if m else ""
    targ = re.sub(r"\(int\)|\(bool\)", "", targ)
    targ = re.sub(r"b2g::hfft::StaticPlan<(\d+),[^>]*>", r"Plan\1", targ)
    return (base + targ)[:70]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--steps", type=int, default=1)
    a = ap.parse_args()
    with open(a.csv) as fh:
        lines = [ln for ln in fh if not ln.startswith("==")]
    rows = list(csv.reader(lines))
    head = rows[0]
    agg = collections.defaultdict(lambda: collections.defaultdict(float))
    seen = collections.defaultdict(set)
    for r in rows[1:]:
        if len(r) != len(head):
            continue
        d = dict(zip(head, r))
        k = short(d["Kernel Name"])
        v = float(d["Metric Value"].replace(",", ""))
        unit, metric = d["Metric Unit"], d["Metric Name"]
        if metric == "gpu__time_duration.sum":
            v *= {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0}.get(unit, 1.0)
        elif metric.startswith("dram__bytes"):
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1.0)
        agg[k][metric] += v
        seen[k].add(d["ID"])
    total = sum(m["gpu__time_duration.sum"] for m in agg.values())
    has_dram = any("dram__bytes_read.sum" in m for m in agg.values())
    print(f"# {a.csv}: {sum(len(s) for s in seen.values())} launches, {total:.1f} ms of kernel time "
          f"(serialised, cold cache), {a.steps} step(s) incl. warm-up")
    hdr = f"{'kernel':70s} {'launches':>8s} {'ms':>10s} {'share':>7s} {'ms/step':>9s}"
    if has_dram:
        hdr += f" {'GB read/step':>13s} {'GB written/step':>16s} {'TB/s':>6s}"
    print(hdr)
    for k, m in sorted(agg.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"]):
        t = m["gpu__time_duration.sum"]
        if t < 0.002 * total:
            continue
        line = f"{k:70s} {len(seen[k]):8d} {t:10.2f} {t / total:7.1%} {t / a.steps:9.2f}"
        if has_dram:
            rd, wr = m.get("dram__bytes_read.sum", 0.0), m.get("dram__bytes_write.sum", 0.0)
            line += f" {rd / a.steps / 1e9:13.2f} {wr / a.steps / 1e9:16.2f} {(rd + wr) / (t * 1e-3) / 1e12:6.2f}"
        print(line)


if __name__ == "__main__":
    main()
