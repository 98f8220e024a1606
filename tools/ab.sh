set -x
timeout -s KILL 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 2 --warmup 1 --frames-per-step 1 --radius 630 --grid 4096 --no-cpu-baseline --no-tf32-pass 2>&1 | tail -1 > gpurun_out/bench_5M.json; python -c "
import json; d=json.load(open('gpurun_out/bench_5M.json')); print(d['config']['workload'][:140]); print(round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['single_frame_latency_ms'], d['roofline']['achieved'], d['roofline']['frac']); import torch"
nvidia-smi --query-gpu=memory.used --format=csv
