set -x
timeout -s KILL 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
PCC_EXACT_F2=1 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --dump-layers gpurun_out/layers_f2_1.csv 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['achieved'], d['step_host_ms'])"
PCC_EXACT_F2=2 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --dump-layers gpurun_out/layers_f2_2.csv 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['achieved'], d['step_host_ms'])"
