# A/B timing of the TP forward launch order / L2 hints on one B200 (EEG_TP_MODE bit 0 = wave-interleaved output-l groups,
# bit 1 = L2 eviction hints; EEG_TP_WAVE = CTAs per wave).  Results of round 1: eeg_tp.cu launch_tp comment.
for cfg in "0 0" "2 0" "3 444" "3 296" "3 888"; do
  set -- $cfg
  EEG_TP_MODE=$1 EEG_TP_WAVE=$2 python bench.py --no-train --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/ab_tp_$1_$2.json 2>gpurun_out/ab_tp.err
  python - <<PY
import json
d=json.load(open("gpurun_out/ab_tp_$1_$2.json"))
print("mode $1 wave $2", d["ms_per_step"], {k:v["ms"] for k,v in d["kernels"].items() if "tp_scatter" in k})
PY
done
