#!/usr/bin/env python3
"""Group the SASS of one kernel of an .ncu-rep into regions of (nearly) equal execution count and print, per region,
its share of the executed warp instructions, of the issue slots (FFMA2/FADD2/FMUL2 charged two, as measured by
tools/issue_slots_microbench.cu) and of the stall samples, plus the opcode mix.

    ncu -i x.ncu-rep --page source --csv --kernel-name regex:trace_secondary > k.csv
    python tools/ncu_regions.py k.csv [--tol 0.1] [--min 0.5]
"""
import argparse
import collections
import csv
import sys


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--tol", type=float, default=0.10, help="relative change of the execution count that opens a new region")
    ap.add_argument("--min", type=float, default=0.3, help="hide regions below this %% of the issue slots")
    ap.add_argument("--dump", type=int, default=-1, help="print the SASS of region k")
    a = ap.parse_args()
    rows = list(csv.reader(open(a.csv)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    col = {n: i for i, n in enumerate(hdr)}
    body = []
    for r in rows[hdr_i + 1:]:  # first launch of the kernel only
        if r and r[0] in ("Kernel Name", "Address"):
            break
        if len(r) >= len(hdr) - 2:
            body.append(r)
    ins = []
    for r in body:
        sass = r[col["Source"]].strip()
        toks = sass.split()
        op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
        ex = int(r[col["Instructions Executed"]] or 0)
        th = int(r[col["Thread Instructions Executed"]] or 0)
        smp = int(r[col["# Samples"]] or 0)
        ins.append((sass, op, ex, th, smp))
    regions, cur = [], []
    for it in ins:
        if cur:
            ref = cur[0][2]
            if abs(it[2] - ref) > a.tol * max(ref, it[2], 1):
                regions.append(cur)
                cur = []
        cur.append(it)
    if cur:
        regions.append(cur)

    def slots(it):
        base = it[1].split(".")[0]
        return it[2] * (2 if base in ("FFMA2", "FADD2", "FMUL2") else 1)

    tot_ex = sum(i[2] for i in ins) or 1
    tot_sl = sum(slots(i) for i in ins) or 1
    tot_sm = sum(i[4] for i in ins) or 1
    print(f"{len(ins)} SASS instructions, {tot_ex:.4g} executed, {tot_sl:.4g} issue slots, {tot_sm} samples")
    print(" reg  first   len   exec/instr  %instr %slots %samples  thr/instr  mix")
    pos = 0
    for k, reg in enumerate(regions):
        ex = sum(i[2] for i in reg)
        sl = sum(slots(i) for i in reg)
        sm = sum(i[4] for i in reg)
        th = sum(i[3] for i in reg)
        if 100.0 * sl / tot_sl >= a.min:
            mix = collections.Counter(i[1].split(".")[0] for i in reg).most_common(6)
            print(f"{k:4d} {pos:6d} {len(reg):5d} {reg[0][2]:12d} {100.0*ex/tot_ex:6.2f} {100.0*sl/tot_sl:6.2f} {100.0*sm/tot_sm:7.2f} "
                  f"{th/max(ex,1):9.1f}  " + " ".join(f"{o}:{c}" for o, c in mix))
        if k == a.dump:
            for i in reg:
                print("      ", i[2], i[0])
        pos += len(reg)


if __name__ == "__main__":
    sys.exit(main())
