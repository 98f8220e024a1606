set -x
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-roofline > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:conv_exact_kernel<.int.128, .bool.1" -s 60 -c 3 -o gpurun_out/prof_exact_f2 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-roofline > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
python bench.py --mode tf32 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:conv_tf32_kernel<.int.128" -s 56 -c 3 -o gpurun_out/prof_tf32_b python bench.py --mode tf32 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline > gpurun_out/ncu3.log 2>&1
tail -3 gpurun_out/ncu3.log
ls -la gpurun_out | tail -5
