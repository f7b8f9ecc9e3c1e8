#!/bin/bash
cd "$(dirname "$0")/.."
for fam in "--steady-after 300" "--family walk"; do
  tag=$(echo $fam | tr -d ' -')
  python tools/small_n.py --mode full --shapes 4096x100000 --chunks auto $fam
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/ab_$tag.csv \
    python tools/small_n.py --mode full --shapes 4096x100000 --chunks auto $fam > /dev/null 2>&1
  python - <<P
import csv
rows=[r for r in csv.reader(open("gpurun_out/ab_$tag.csv")) if len(r)>5 and r[0].isdigit()]
for r in rows[-13:]:
    if "tp_" in r[4]: print("  ", r[4][:70], r[-1], r[-2])
P
done
