#!/usr/bin/env python
"""Join an `ncu --page source --csv` dump (SASS rows with stall samples) with `nvdisasm -g -c` line info so the
samples can be read per CUDA source line.

  cuobjdump -xelf all trustworthy_dl/_C/libtdl_kernels.so ; nvdisasm -g -c attention.sm_100a.cubin > attention.dis
  ncu -i prof.ncu-rep --page source --csv > src.csv
  python bench/ncu_lines.py src.csv attention.dis attn_bwd_ws_kernelILi0E [source.cu]
"""
import csv
import re
import sys
from collections import defaultdict


def main():
    src_csv, dis, key = sys.argv[1:4]
    source = sys.argv[4] if len(sys.argv) > 4 else None
    nth = int(sys.argv[5]) if len(sys.argv) > 5 else 0        # which matching launch in the report
    # --- line number per SASS instruction (in order) from nvdisasm
    lines = []
    cur = None
    infunc = False
    for ln in open(dis, errors="replace"):
        if ln.startswith(".text.") and ln.rstrip().endswith(":"):
            infunc = key in ln
            continue
        if not infunc:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", ln):
            lines.append(cur)
    # --- samples per SASS instruction from ncu
    rows = list(csv.reader(open(src_csv)))
    kern, k = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            k = {"name": r[1], "rows": [], "hdr": None}; kern.append(k); continue
        if k is None:
            continue
        if k["hdr"] is None:
            k["hdr"] = r; continue
        k["rows"].append(r)
    # `key` is a substring of the mangled name in the disassembly; `match` (optional 7th argument, default: the
    # function name part of key) is a substring of the demangled name in the ncu report
    match = sys.argv[6] if len(sys.argv) > 6 else re.split(r"ILi|I[a-z]", key)[0]
    cands = [x for x in kern if match in x["name"]]
    k = cands[min(nth, len(cands) - 1)]
    h = k["hdr"]; iS = h.index("# Samples"); iI = h.index("Instructions Executed")
    R = k["rows"]
    print(f"{k['name'][:70]}: {len(R)} SASS rows, {len(lines)} disassembled", file=sys.stderr)
    per = defaultdict(lambda: [0, 0])
    n = min(len(R), len(lines))
    for i in range(n):
        per[lines[i]][0] += int(R[i][iS] or 0)
        per[lines[i]][1] += int(R[i][iI] or 0)
    tot = sum(v[0] for v in per.values()) or 1
    text = {}
    if source:
        for i, ln in enumerate(open(source), 1):
            text[i] = ln.rstrip()
    for (f, l), (s, ins) in sorted(per.items(), key=lambda kv: -kv[1][0])[:45]:
        t = text.get(l, "") if f and source and source.endswith(f) else ""
        print(f"{s:7d} {100 * s / tot:5.1f}%  inst {ins:10d}  {f}:{l:<5d} {t.strip()[:100]}")


if __name__ == "__main__":
    main()
