#!/bin/bash
# The round's ncu captures (run under gpurun on ONE GPU).  Every capture follows a plain run of the same command that
# exited 0; numbers printed under ncu are never bench values.  Reports land in gpurun_out/; summarise them into
# profiles/ with tools/traffic_from_ncu.py, tools/launch_summary.py and `ncu -i ... --page raw --csv`.
set -x
T=${1:-r2f}
export PHB_BENCH_KERNEL_ONLY=1
# (an evaluation = two sweep launches, the full passes and the narrower rest: launches 2 and 3 are the second evaluation)
python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/plain1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 2 -c 2 -o gpurun_out/${T}_sweep python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu1.log 2>&1
unset PHB_BENCH_KERNEL_ONLY
python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/plain2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-configs > gpurun_out/ncu2.log 2>&1
python tools/configs_bench.py --cases cfg4_small --reps 1 > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:asr_down_sweep -s 1 -c 1 -o gpurun_out/${T}_down python tools/configs_bench.py --cases cfg4_small --reps 1 > gpurun_out/ncu3.log 2>&1
python tools/configs_bench.py --cases cfg5 --reps 1 > gpurun_out/plain4.log 2>&1 && ncu --set full --clock-control none -k regex:nni_edge_kernel -s 1 -c 1 -o gpurun_out/${T}_nni_edge python tools/configs_bench.py --cases cfg5 --reps 1 > gpurun_out/ncu4.log 2>&1
# 20-state CLV kernel: the launches of one evaluation (a launch per tree level), time only, then the longest in full
python tools/aa_bench.py --taxa 1000 --patterns 20000 --reps 1 > gpurun_out/plain6.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:aa_ -c 200 --csv --log-file gpurun_out/${T}_aa_launches.csv python tools/aa_bench.py --taxa 1000 --patterns 20000 --reps 1 > gpurun_out/ncu6.log 2>&1
for f in gpurun_out/ncu1.log gpurun_out/ncu3.log gpurun_out/ncu4.log gpurun_out/ncu6.log; do tail -n 2 $f | cut -c1-300; done
