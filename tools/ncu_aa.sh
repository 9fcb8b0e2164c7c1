#!/bin/bash
# ncu capture of the 20-state CLV and down kernels (run under gpurun; the plain command first)
set -x
python tools/aa_bench.py --taxa 1000 --patterns 20000 --reps 1 > gpurun_out/aa_plain.log 2>&1 || exit 1
tail -n 1 gpurun_out/aa_plain.log
ncu --set full --clock-control none --import-source on -k regex:aa_clv_kernel -s 40 -c 1 -o gpurun_out/r2_aa_clv python tools/aa_bench.py --taxa 1000 --patterns 20000 --reps 1 > gpurun_out/aa_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:aa_down_kernel -s 40 -c 1 -o gpurun_out/r2_aa_down python tools/aa_bench.py --taxa 1000 --patterns 20000 --reps 1 > gpurun_out/aa_ncu2.log 2>&1
tail -n 2 gpurun_out/aa_ncu1.log gpurun_out/aa_ncu2.log
