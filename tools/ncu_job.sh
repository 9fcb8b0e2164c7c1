set -x
export PHB_BENCH_KERNEL_ONLY=1
python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/plain1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 3 -c 1 -o gpurun_out/r2q_sweep python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu1.log 2>&1
unset PHB_BENCH_KERNEL_ONLY
python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/plain2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2q_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu2.log 2>&1
python tools/configs_bench.py --cases cfg4_small --reps 1 > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:asr_down_sweep -s 1 -c 1 -o gpurun_out/r2q_down python tools/configs_bench.py --cases cfg4_small --reps 1 > gpurun_out/ncu3.log 2>&1
python tools/configs_bench.py --cases cfg5 --reps 1 > gpurun_out/plain4.log 2>&1 && ncu --set full --clock-control none -k regex:nni_edge_kernel -s 1 -c 1 -o gpurun_out/r2q_nni_edge python tools/configs_bench.py --cases cfg5 --reps 1 > gpurun_out/ncu4.log 2>&1
for f in gpurun_out/ncu1.log gpurun_out/ncu3.log gpurun_out/ncu4.log; do tail -n 2 $f; done
