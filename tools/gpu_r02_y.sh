SH=4096x100000,16384x40000,32768x20000,65536x10000
for mode in full_faults summary_faults; do
  python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto
  python tools/small_n.py --mode $mode --family walk --shapes $SH --chunks 0,auto
  python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto --quiet-after 300
  python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto --steady-after 300
done
