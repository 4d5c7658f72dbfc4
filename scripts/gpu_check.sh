#!/usr/bin/env bash
# Parity suite, then the bench legs with one and five view streams (value / e2e and the per-stage times).
#   gpurun -- 'bash scripts/gpu_check.sh'
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | grep -E "passed|failed|Error" | tail -3
for s in 1 5; do
  python bench.py --no-cpu-baseline --no-extras --streams $s > gpurun_out/check_s$s.json 2> gpurun_out/check_s$s.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/check_s$s.json").read().strip().splitlines()[-1])
print("streams $s value %.1f e2e %.1f |" % (d["value"], d["e2e"]["value"]), " ".join("%s %.4f" % (k[:12], v) for k, v in d["stage_ms"].items()))
PY
done
