#!/usr/bin/env bash
# value / e2e of bench.py for priority lanes on and 4..8 view streams
for pl in 1; do for s in 4 5 6 8; do
  python bench.py --no-cpu-baseline --steps 10 --warmup 3 --streams $s --priority-lanes $pl 2>/dev/null > /tmp/sweep.json
  python -c "import json; d=json.loads(open('/tmp/sweep.json').read().strip().splitlines()[-1]); print('lanes $pl streams $s', d['value'], d['e2e']['value'], d['ms_per_step'])"
done; done
