set -x
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-roofline > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:conv_exact_kernel -s 100 -c 4 -o gpurun_out/prof_exact_f2 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-roofline > gpurun_out/ncu2.log 2>&1
python bench.py --mode tf32 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:conv_tf32_kernel -s 40 -c 4 -o gpurun_out/prof_tf32_b python bench.py --mode tf32 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline > gpurun_out/ncu3.log 2>&1
ls -la gpurun_out | tail -5
