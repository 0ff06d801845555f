#!/bin/bash
# Round-2 evidence run on ONE B200 (under gpurun): tests, bench lines, ncu launch list and full captures of the top kernels.
# Everything lands in gpurun_out/ (copied into profiles/ afterwards).
set -u
mkdir -p gpurun_out
R=gpurun_out/r02
timeout 900 python -m pytest tests -m gpu -q > ${R}_gpu_tests.log 2>&1; tail -3 ${R}_gpu_tests.log
timeout 100 python tools/train_prof.py 512 > ${R}_trainprof.log 2>&1; head -3 ${R}_trainprof.log
timeout 400 python bench.py --steps 20 --warmup 5 > ${R}_bench.json 2> ${R}_bench.err; tail -c 300 ${R}_bench.err
timeout 200 python bench.py --config 4 --steps 20 --warmup 5 > ${R}_bench_config4.json 2> ${R}_bench_config4.err
timeout 300 python bench.py --config 5 --steps 10 --warmup 3 > ${R}_bench_config5_n1.json 2> ${R}_bench_config5.err
CMD="python bench.py --steps 2 --warmup 1 --no-train --no-cpu-baseline --no-small-batch"
$CMD > ${R}_ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file ${R}_ncu_launches.csv $CMD > ${R}_ncu_launches.log 2>&1
$CMD > ${R}_ncu_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_tp_scatter|k_sym_contract|k_tc_linear|k_radial_mlp" -s 8 -c 8 -o ${R}_fwd_top $CMD > ${R}_ncu_full.log 2>&1
CMD2="python tools/train_prof.py 128"
$CMD2 > ${R}_ncu_plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_tp_bwd2|k_sc_bwd_x|k_sc_wgrad|k_linear_wgrad" -s 12 -c 6 -o ${R}_bwd_top $CMD2 > ${R}_ncu_full_bwd.log 2>&1
ls -la gpurun_out/*.ncu-rep 2>/dev/null
python - <<'PY'
import json
for f in ("gpurun_out/r02_bench.json", "gpurun_out/r02_bench_config4.json", "gpurun_out/r02_bench_config5_n1.json"):
    try:
        d = json.load(open(f))
        print(f, round(d["value"]), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]), "roofline frac", round(d["roofline"]["frac"], 3),
              "train", d.get("train_step", {}).get("ms_per_step"))
    except Exception as e:
        print(f, "ERR", e)
PY
