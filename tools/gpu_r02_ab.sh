#!/bin/bash
# Pass ab: where the time goes in full mode on steady / walk / quiet depths (ncu launch lists).
cd "$(dirname "$0")/.."
for fam in "--steady-after 300" "--family walk" "--quiet-after 300"; do
  tag=$(echo $fam | tr -d ' -')
  python tools/small_n.py --mode full --shapes 4096x100000 --chunks auto $fam
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/ab_$tag.csv \
    python tools/small_n.py --mode full --shapes 4096x100000 --chunks auto $fam > /dev/null 2>&1
  python - <<P
import csv
rows=[r for r in csv.reader(open("gpurun_out/ab_$tag.csv")) if len(r)>5 and r[0].isdigit()]
# the last call's launches: take the last 12 rows
for r in rows[-12:]:
    print("  ", r[4][:70], r[-1], r[-2])
P
done
