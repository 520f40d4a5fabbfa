#!/bin/bash
# per-kernel durations of the summary part of the step (ncu, duration + instruction count only)
python tools/run_summary_once.py && ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"bs_|group_fold|hist_kernel|collect|series_partial|tail_kernel|pick_kernel|finish|init_kernel" -s 10 -c ${1:-6} --csv --log-file gpurun_out/sum_launches.csv python tools/run_summary_once.py > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/sum_launches.csv")) if len(r)>10]
for r in rows[1:]: print(r[4][:50], r[-3], r[-1])
PY
