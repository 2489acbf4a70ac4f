#!/usr/bin/env python
"""Per-source-line instruction and stall-sample totals of one kernel in an .ncu-rep captured with --import-source on:
joins ncu's SASS page (counts per instruction) with nvdisasm's line table of the same kernel from the built library.
python tools/ncu_lines.py file.ncu-rep <cubin> <mangled kernel name substring> [top N]"""
import csv
import io
import re
import subprocess
import sys


OUTER = True


def main(rep, cubin, kern, top=40):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    inst = []
    for r in rows[2:]:
        try:
            inst.append((int(r[ia], 16), int(r[ii]), int(r[isamp]), r[hdr.index("Source")].strip()))
        except Exception:
            pass
    base = inst[0][0]
    dis = subprocess.run(["nvdisasm", "-gi", "-c", cubin], capture_output=True, text=True).stdout
    # find the function section, then track "//## File "...", line N" markers
    lines = dis.split("\n")
    # an instruction is preceded by its inline chain, innermost first; with outer=True the line of the kernel body
    # (the last marker, without "inlined at") is used, else the innermost one
    cur, infun, off2line, chain = None, False, {}, []
    for ln in lines:
        if ln.startswith(".text."):
            infun = kern in ln
        if not infun:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            chain.append((m.group(1).split("/")[-1], int(m.group(2))))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", ln)
        if m:
            if chain:
                cur = chain[-1] if OUTER else chain[0]
                chain = []
            if cur:
                off2line[int(m.group(1), 16)] = cur
    tot_i = sum(x[1] for x in inst)
    tot_s = sum(x[2] for x in inst) or 1
    agg = {}
    for a, n, s, src in inst:
        key = off2line.get(a - base, ("?", 0))
        v = agg.setdefault(key, [0, 0])
        v[0] += n
        v[1] += s
    print(f"total warp instructions {tot_i}, samples {tot_s}")
    for key, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{100 * n / tot_i:6.2f}% inst {100 * s / tot_s:6.2f}% samples  {key[0]}:{key[1]}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 40)
