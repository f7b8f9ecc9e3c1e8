#!/bin/bash
cd "$(dirname "$0")/.."
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/an_bands.csv \
  python tools/small_n.py --mode full --shapes 4096x100000 --chunks auto > /dev/null 2>&1
python - <<P
import csv
rows=[r for r in csv.reader(open("gpurun_out/an_bands.csv")) if len(r)>5 and r[0].isdigit()]
for r in rows[-12:]:
    print("  ", r[4][:90], r[-1])
P
