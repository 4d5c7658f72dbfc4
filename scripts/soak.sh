#!/usr/bin/env bash
# Repeats the parity suite and a few bench legs to shake out rare races in the ring / look-back protocols.
#   gpurun -- 'bash scripts/soak.sh [rounds]'
R=${1:-5}
for i in $(seq 1 $R); do
  timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py tests/test_mapping_loss.py -x -q -m gpu 2>&1 | grep -E "passed|failed|Error|Timeout" | tail -1
done
for s in 1 3 10 16; do
  timeout 300 python bench.py --no-cpu-baseline --no-extras --streams $s --steps 30 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('streams $s', round(d['value'],1), round(d['e2e']['value'],1))"
done
