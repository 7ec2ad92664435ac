#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel shares and, for the
GEMM launches of bench/profile_step.py (fixed launch order), TFLOP/s per shape."""
import collections
import csv
import re
import sys

path = sys.argv[1]
layers = int(sys.argv[2]) if len(sys.argv) > 2 else 4
M = int(sys.argv[3]) if len(sys.argv) > 3 else 32768
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
gemm = []
tot = 0.0
for row in csv.DictReader(lines):
    try:
        v = float(row["Metric Value"].replace(",", ""))
    except Exception:
        continue
    unit = row["Metric Unit"]
    v = v / 1e3 if unit == "ns" else (v * 1e3 if unit == "ms" else v)
    name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("<unnamed>::", "")[:48]
    agg[name][0] += 1
    agg[name][1] += v
    tot += v
    if "gemm_bf16" in name:
        gemm.append(v)
print(f"total {tot:.1f} us over {sum(n for n, _ in agg.values())} launches ({layers} layers, M={M})")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:16]:
    print(f"{t:10.1f} us {100 * t / tot:5.1f}%  n={n:4d}  avg {t / n:8.1f}  {k}")
C, V = 1024, 50304
fw = [("f_qkv", M, 3 * C, C), ("f_o", M, C, C), ("f_fc", M, 4 * C, C), ("f_proj", M, C, 4 * C)]
bw = [("w_proj", C, 4 * C, M), ("d_proj", M, 4 * C, C), ("w_fc", 4 * C, C, M), ("d_fc", M, C, 4 * C), ("w_o", C, C, M),
      ("d_o", M, C, C), ("w_qkv", 3 * C, C, M), ("d_qkv", M, C, 3 * C)]
names = fw * layers + [("f_head", M, V, C), ("d_head", M, C, V), ("w_head", V, C, M)] + bw * layers
if len(gemm) == len(names):
    per = collections.OrderedDict()
    for (n, m, nn, k), t in zip(names, gemm):
        per.setdefault(n, [m, nn, k, []])[3].append(t)
    for n, (m, nn, k, ts) in per.items():
        t = sum(ts) / len(ts)
        print(f"{n:8s} M={m:6d} N={nn:6d} K={k:6d}  {t:8.1f} us  {2 * m * nn * k / t / 1e6:7.1f} TFLOP/s")
