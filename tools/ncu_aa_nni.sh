#!/bin/bash
# launch list + one full capture of the upper-partial kernel of the 20-state NNI scoring (run under gpurun; plain command first)
set -x
python tools/aa_nni_probe.py > gpurun_out/aa_nni_plain.log 2>&1 || { tail -5 gpurun_out/aa_nni_plain.log; exit 1; }
tail -n 1 gpurun_out/aa_nni_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:aa_ -c 600 --csv --log-file gpurun_out/r2_aa_nni_launches.csv python tools/aa_nni_probe.py > gpurun_out/aa_nni_ncu0.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:aa_upper_kernel -s 30 -c 1 -o gpurun_out/r2_aa_upper python tools/aa_nni_probe.py > gpurun_out/aa_nni_ncu1.log 2>&1
tail -n 2 gpurun_out/aa_nni_ncu0.log gpurun_out/aa_nni_ncu1.log
