set -x
timeout -s KILL 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for cfg in "1 1" "4 3" "6 3" "8 4"; do set -- $cfg
python bench.py --mode exact --steps 3 --warmup 2 --frames-per-step $1 --depth $2 --no-cpu-baseline --no-roofline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('exact B=$1 D=$2', round(d['value']), round(d['e2e']['value']), round(d['ms_per_step']/$1,1), d['step_host_ms'])"
python bench.py --mode tf32 --steps 3 --warmup 2 --frames-per-step $1 --depth $2 --no-cpu-baseline --no-roofline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('tf32 B=$1 D=$2', round(d['value']), round(d['e2e']['value']), round(d['ms_per_step']/$1,1), d['step_host_ms'])"
done
