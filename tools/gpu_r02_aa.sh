#!/bin/bash
# Pass aa: the path fuzzer, the new test, the host topology, and bench with / without NUMA binding.
cd "$(dirname "$0")/.."
timeout 900 python tools/fuzz_paths.py --cases 400 --seed 1 2>&1 | cut -c1-700 | tail -30
timeout 300 python -m pytest tests/test_gpu_timepar.py -q -k "wider" 2>&1 | tail -3
nvidia-smi topo -m 2>&1 | head -30
lscpu | grep -i -E "numa|socket|^CPU\(s\)|model name"
for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/class 2>/dev/null)" = "0x030200" ]; then echo "$d numa $(cat $d/numa_node)"; fi; done
python -c "import os; print('affinity', len(os.sched_getaffinity(0)), sorted(os.sched_getaffinity(0))[:4], '...')"
python bench.py --steps 3 --warmup 3 --no-configs --no-cpu-baseline | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bind   ', d['config']['host_binding'], d['value'], d['e2e']['value'], d['e2e'].get('link_ceiling'))"
python bench.py --steps 3 --warmup 3 --no-configs --no-cpu-baseline --no-bind | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('no bind', d['config']['host_binding'], d['value'], d['e2e']['value'], d['e2e'].get('link_ceiling'))"
