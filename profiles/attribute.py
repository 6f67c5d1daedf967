#!/usr/bin/env python
"""Attributes an ncu SASS-level source page to source lines through inlining.

  nvdisasm -gi -c <cubin> > dis.txt          (cubin: cuobjdump -xelf all lib.so)
  ncu -i prof.ncu-rep --page source --csv --print-source sass > sass.csv
  python profiles/attribute.py dis.txt sass.csv <kernel substring> [file:depth ...]

Prints executed warp instructions and stall samples grouped by the source line
of `phys_step.cuh` (the phase of the step) and by the innermost line.
"""
import csv
import re
import sys
from collections import defaultdict


def parse_dis(path, kernel):
    chains, cur, on, last = {}, [], False, []
    for line in open(path, errors="replace"):
        if ".section" in line and ".text." in line:
            on = kernel in line
            continue
        if not on:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
        if m:
            cur.append((m.group(1).split("/")[-1], int(m.group(2))))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S+)", line)
        if m:
            if cur:                 # an instruction without markers of its own belongs to the line of the one before
                last = list(cur)
                cur = []
            chains[int(m.group(1), 16)] = (list(last), m.group(2))
    return chains


def main():
    dis, sass, kernel = sys.argv[1:4]
    chains = parse_dis(dis, kernel)
    rows = list(csv.reader(open(sass)))
    hdr = rows[1]
    ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    ithr = hdr.index("Thread Instructions Executed")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    by_phase = defaultdict(lambda: [0, 0, 0])
    by_inner = defaultdict(lambda: [0, 0, 0])
    by_mid = defaultdict(lambda: [0, 0, 0])
    stalls = defaultdict(int)
    last_chain = []
    tot = [0, 0, 0]
    base = None
    for r in rows[2:]:
        try:
            addr = int(r[ia], 16)
        except ValueError:
            continue
        if base is None:
            base = addr        # the page lists the kernel from its first instruction
        addr -= base
        n, s, t = int(r[ii] or 0), int(r[isamp] or 0), int(r[ithr] or 0)
        chain, op = chains.get(addr, ([], "?"))
        if chain:
            last_chain = chain
        else:
            chain = last_chain
        phase = next(((f, l) for f, l in reversed(chain) if f == "phys_step.cuh"), ("?", 0))
        inner = chain[0] if chain else ("?", 0)
        mid = next(((f, l) for f, l in reversed(chain) if f in ("phys_solver.cuh", "phys_narrow.cuh", "phys_world.cuh")), ("?", 0))
        for d, k in ((by_phase, phase), (by_inner, inner), (by_mid, mid)):
            d[k][0] += n; d[k][1] += s; d[k][2] += t
        tot[0] += n; tot[1] += s; tot[2] += t
        for i in stall_cols:
            if r[i]:
                stalls[hdr[i]] += int(r[i])
    print(f"total warp-inst {tot[0]:,}  samples {tot[1]:,}  avg threads/inst {tot[2] / max(tot[0], 1):.1f}")
    for name, d, top in (("phase (phys_step.cuh line)", by_phase, 40), ("outer function line", by_mid, 45),
                         ("innermost line", by_inner, 30)):
        print(f"\n== by {name}")
        for k, (n, s, t) in sorted(d.items(), key=lambda kv: -kv[1][1])[:top]:
            print(f"  {k[0]}:{k[1]:<5} inst {100 * n / tot[0]:5.1f}%  samples {100 * s / max(tot[1], 1):5.1f}%  thr/inst {t / max(n, 1):4.1f}")
    print("\n== stall reasons (all samples)")
    st = sum(stalls.values())
    for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:10]:
        print(f"  {k:<28}{100 * v / max(st, 1):5.1f}%")


if __name__ == "__main__":
    main()
