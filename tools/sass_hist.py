#!/usr/bin/env python
"""SASS opcode histogram per kernel of an object file / shared library (cuobjdump -sass), so that
the instruction mix of the hot kernels (TMA: UBLKCP / UTMALDG / UTMASTG, mbarrier: SYNCS, packed
FMAs: FFMA2, tensor memory: UTCMMA / LDTM) can be checked without a GPU.

    python tools/sass_hist.py build/fft_planes.o [--match hfft_] [--top 25] > profiles/r02_sass_fft.txt
"""
import argparse
import collections
import re
import subprocess


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("obj")
    ap.add_argument("--match", default="")
    ap.add_argument("--top", type=int, default=25)
    a = ap.parse_args()
    out = subprocess.run(["cuobjdump", "-sass", a.obj], capture_output=True, text=True, check=True).stdout
    name, hist = None, {}
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            hist[name] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            hist[name][m.group(1).split(".")[0]] += 1
    for name, h in hist.items():
        if a.match and a.match not in name:
            continue
        total = sum(h.values())
        short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", ""))
        print(f"== {short}  ({total} instructions)")
        special = {k: v for k, v in h.items() if re.match(r"UBLKCP|UTMA|SYNCS|UTC|LDTM|STTM|FFMA2|FMUL2|FADD2|REDG|RED|LDGSTS|BAR|UCGABAR", k)}
        if special:
            print("   blackwell / async: " + ", ".join(f"{k} {v}" for k, v in sorted(special.items())))
        print("   " + ", ".join(f"{k} {v}" for k, v in h.most_common(a.top)))


if __name__ == "__main__":
    main()
