#!/usr/bin/env python
"""Measure the fp32 FMA ceiling of this GPU with tools/micro/ffma2_rate.cu and write profiles/ffma2_rate.json
(bench.py reads the roofline's fp32 denominator from that file, not from a literal).

    python tools/micro/measure_fp32_peak.py [out.json]        # on a GPU box (nvcc is in the image)"""
import json
import os
import re
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "ffma2_rate.json")
exe = os.path.join(HERE, "ffma2_rate.bin")
src = os.path.join(HERE, "ffma2_rate.cu")
if not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(src):
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-o", exe, src])
txt = subprocess.run([exe], capture_output=True, text=True, check=True).stdout
rows = []
for line in txt.splitlines():
    m = re.match(r"(.+?)\s+thr=\s*(\d+):\s+([\d.]+) ms\s+([\d.]+) TFLOP/s\s+cycles=(\d+)\s+=>\s+([\d.]+) scalar-FMA lanes/clk/SMSP\s+\(clk (\d+) MHz\)", line)
    if m:
        rows.append(dict(variant=m.group(1).strip(), threads=int(m.group(2)), ms=float(m.group(3)), tflops=float(m.group(4)),
                         lanes_per_clk_smsp=float(m.group(6)), sm_mhz=float(m.group(7))))
best = lambda key: max((r for r in rows if key in r["variant"]), key=lambda r: r["tflops"])  # noqa: E731
smi = subprocess.run(["nvidia-smi", "--query-gpu=name,clocks.max.sm", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
res = dict(when=time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()), gpu=smi, source="tools/micro/ffma2_rate.cu (148 CTAs x 128/256/512 threads; best of FFMA / FFMA2, register and LDS-broadcast operands)",
           ffma2_tflops=best("FFMA2 packed regs")["tflops"], ffma_tflops=best("FFMA regs")["tflops"],
           ffma2_bcast_lds_tflops=best("FFMA2 bcast")["tflops"], ffma_bcast_lds_tflops=best("FFMA c=LDS")["tflops"], rows=rows)
res["fp32_tflops_peak"] = max(r["tflops"] for r in rows)
with open(out, "w") as f:
    json.dump(res, f, indent=1)
print(json.dumps({k: v for k, v in res.items() if k != "rows"}))
