#!/bin/bash
# GPU box, 1 GPU: full ncu capture of one step's kernels on a small, L2-resident set (ResNet-18 r=2)
CMD="python bench.py --workload resnet18 --rank 2 --steps 3 --warmup 3 --skip-train --skip-e2e --skip-cpu-baseline --skip-configs --skip-baselines --no-clock-probe --no-phase-timing"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_small.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:k_p_from_m_stream|k_q_from_mt|k_orth|k_finalize_stream" -s 12 -c 4 -o gpurun_out/r02_small $CMD > gpurun_out/ncu_small.log 2>&1
ls -la gpurun_out/r02_small.ncu-rep
