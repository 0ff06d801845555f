# round-1 session-13 GPU pass: full gpu tests, bench line, ncu full capture of the TP forward (default launch order)
python -m pytest tests -m gpu -x -q > gpurun_out/r01_gpu_tests_s13b.log 2>&1; tail -3 gpurun_out/r01_gpu_tests_s13b.log
python bench.py > gpurun_out/r01_bench_s13b.json 2> gpurun_out/bench.err; cut -c1-200 gpurun_out/r01_bench_s13b.json; tail -c 400 gpurun_out/bench.err
ncu --set full --clock-control none --import-source on -k regex:k_tp_scatter -c 2 -o gpurun_out/r01_tp_fwd_s13c -f python bench.py --no-train --no-cpu-baseline --steps 1 --warmup 0 > gpurun_out/ncu_full.log 2>&1; tail -2 gpurun_out/ncu_full.log
