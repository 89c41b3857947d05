CMD="python bench.py --rank 16 --steps 2 --warmup 3 --skip-train --skip-e2e --skip-cpu-baseline --skip-configs --skip-baselines --no-clock-probe --no-phase-timing"
$CMD > gpurun_out/plain16.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 30 --csv --log-file gpurun_out/r02_launches_r16.csv $CMD > gpurun_out/ncu16.log 2>&1
python - <<'PY'
import csv
rows=list(csv.DictReader(l for l in open('gpurun_out/r02_launches_r16.csv') if not l.startswith('==')))
for r in rows[-8:]:
    print(r['Kernel Name'][:70], r['Grid Size'], r['Block Size'], r['Metric Value'])
PY
