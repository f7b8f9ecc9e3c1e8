#!/bin/bash
cd "$(dirname "$0")/.."
for fam in "--steady-after 300" "--family walk" ""; do
  tag=$(echo "x$fam" | tr -d ' -')
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/ai_$tag.csv \
    python tools/small_n.py --mode summary --shapes 4096x100000 --chunks auto $fam > /dev/null 2>&1
  echo "== summary $fam"
  python - <<P
import csv
rows=[r for r in csv.reader(open("gpurun_out/ai_$tag.csv")) if len(r)>5 and r[0].isdigit()]
for r in rows[-10:]:
    print("  ", r[4][:90], r[-1])
P
done
