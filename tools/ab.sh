set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
PCC_EXACT_F2=0 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --dump-layers gpurun_out/layers_exact_f2_0.csv > gpurun_out/bench_exact_f2_0.log 2>&1; tail -1 gpurun_out/bench_exact_f2_0.log | cut -c1-250
PCC_EXACT_F2=1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --dump-layers gpurun_out/layers_exact_f2_1.csv > gpurun_out/bench_exact_f2_1.log 2>&1; tail -1 gpurun_out/bench_exact_f2_1.log | cut -c1-250
