#!/bin/bash
# GPU box, 1 GPU: default bench line, then the launch list and a full ncu capture of K2 after the round-2 K2 changes
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_n1_default.json 2> gpurun_out/bench_n1_default.err || exit 1
CMD="python bench.py --steps 3 --warmup 3 --skip-train --skip-e2e --skip-cpu-baseline --skip-configs --skip-baselines --no-clock-probe --no-phase-timing"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 24 -c 40 --csv --log-file gpurun_out/r02b_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:k_orth_tall" -s 3 -c 1 -o gpurun_out/r02b_orth $CMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/*.ncu-rep
