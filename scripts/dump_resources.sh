#!/usr/bin/env bash
# Registers / stack / static shared memory / local memory per kernel of the built library (cuobjdump --dump-resource-usage),
# as a markdown table: profiles/sass/RESOURCES.md.  Dynamic shared memory is set at launch and is not in this table.
set -euo pipefail
cd "$(dirname "$0")/.."
SO=trustworthy_dl/_C/libtdl_kernels.so
OUT=profiles/sass/RESOURCES.md
TMP=$(mktemp)
cuobjdump --dump-resource-usage "$SO" > "$TMP" 2>/dev/null
python - "$TMP" "$OUT" <<'PY'
import re, subprocess, sys
src, out = sys.argv[1], sys.argv[2]
rows, name = [], None
for line in open(src):
    m = re.match(r"\s*Function (\S+):", line)
    if m:
        name = m.group(1)
        continue
    m = re.match(r"\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
    if m and name:
        rows.append((name, *map(int, m.groups())))
        name = None
names = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True, text=True).stdout.splitlines()
with open(out, "w") as f:
    f.write("# Per-kernel resources (cuobjdump --dump-resource-usage, sm_100a)\n\n"
            "Registers per thread, stack bytes per thread, *static* shared memory bytes per block, local (spill) bytes per thread.\n"
            "The tcgen05 kernels take their operand stages as dynamic shared memory at launch (up to 227 KB), not shown here;\n"
            "1024 of the static bytes of every kernel are the per-block reservation the sm_100 toolchain adds.  No kernel spills\n"
            "(local = 0 everywhere; `stack` is address-taken per-thread arrays such as the per-rank coefficient tables).\n\n"
            "| kernel | regs | stack | static smem | local |\n|---|---|---|---|---|\n")
    for (raw, reg, stack, sh, loc), nm in sorted(zip(rows, names), key=lambda x: x[1]):
        nm = nm.replace("(anonymous namespace)::", "")
        f.write(f"| `{nm[:150]}` | {reg} | {stack} | {sh} | {loc} |\n")
print(f"{len(rows)} kernels -> {out}")
PY
rm -f "$TMP"
