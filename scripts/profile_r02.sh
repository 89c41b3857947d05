#!/bin/bash
# GPU box, 1 GPU: launch list and full ncu capture of one step's kernels (DistilBERT r=4), per the profiling recipe
CMD="python bench.py --steps 3 --warmup 3 --skip-train --skip-e2e --skip-cpu-baseline --skip-configs --skip-baselines --no-clock-probe --no-phase-timing"
mkdir -p gpurun_out
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 24 -c 40 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:k_p_from_m_stream|k_q_from_mt|k_orth_tall|k_finalize_stream" -s 12 -c 4 -o gpurun_out/r02_top4 $CMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/*.ncu-rep
