#!/bin/bash
# One SASS listing per kernel function of the built library -> profiles/sass/<kernel>.sass (+ index with the tcgen05 / TMA
# / multimem mnemonic counts per kernel).  Runs anywhere nvcc's cuobjdump is available (no GPU needed).
set -e
LIB=${1:-trustworthy_dl/_C/libtdl_kernels.so}
OUT=profiles/sass
mkdir -p $OUT
rm -f $OUT/*.sass
TMP=$(mktemp)
cuobjdump -sass "$LIB" > $TMP
python3 - "$OUT" "$TMP" <<'PY'
import re, subprocess, sys, collections
out = sys.argv[1]
cur, buf, funcs = None, [], {}
for line in open(sys.argv[2]):
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        if cur: funcs[cur] = buf
        cur, buf = m.group(1), []
    elif cur is not None:
        buf.append(line)
if cur: funcs[cur] = buf
names = list(funcs)
dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
keys = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAREDG", "LDTM", "STTM", "UTCCP", "SYNCS", "MUFU.EX2", "HMMA", "RED.E", "REDG", "ATOM", "LDG", "STG", "LD.E", "ST.E"]
index = []
used = collections.Counter()
for mangled, d in zip(names, dem):
    short = re.sub(r"\(anonymous namespace\)::", "", d)
    base = re.match(r"(?:void )?([\w:]+(?:<[^>(]*>)?)", short).group(1)
    fname = re.sub(r"[^\w]+", "_", base).strip("_")
    used[fname] += 1
    if used[fname] > 1: fname += f"_{used[fname]}"
    # keep address + instruction, drop the 128-bit encodings (two hex comments per instruction): 3x smaller, same content
    body = [re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "\n", l) for l in funcs[mangled]]
    body = [l for l in body if l.strip()]
    with open(f"{out}/{fname}.sass", "w") as f:
        f.write(f"// {d}\n// {mangled}\n"); f.writelines(body)
    text = "".join(body)
    counts = {k: len(re.findall(r"\b" + re.escape(k), text)) for k in keys}
    counts["LDGMC (multimem.ld_reduce)"] = len(re.findall(r"\bLDGMC", text))
    n_instr = len(re.findall(r"^\s+/\*[0-9a-f]{4}\*/", text, re.M))
    index.append((fname, n_instr, counts, d))
with open(f"{out}/INDEX.md", "w") as f:
    f.write("# SASS listings per kernel (cuobjdump -sass of libtdl_kernels.so, sm_100a)\n\n")
    f.write("| file | instructions | tensor-core / TMA / TMEM / NVLS evidence | kernel |\n|---|---|---|---|\n")
    for fname, n, c, d in sorted(index):
        ev = ", ".join(f"{k} x{v}" for k, v in c.items() if v and k in ("UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAREDG", "LDTM", "STTM", "LDGMC (multimem.ld_reduce)", "MUFU.EX2", "HMMA"))
        f.write(f"| {fname}.sass | {n} | {ev or '-'} | `{d[:140]}` |\n")
print(len(index), "kernels")
PY
rm -f $TMP
du -sh $OUT
