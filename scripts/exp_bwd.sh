#!/usr/bin/env bash
# A/B of the render-backward chunk sizes (SPLAT_BWD_CHUNK = 8 | 16): parity suite, then bench value / e2e and the
# per-stage times.
mkdir -p gpurun_out
for v in ${VARIANTS:-8 16}; do
  echo "=== variant $v: parity"
  SPLAT_BWD_CHUNK=$v python -m pytest tests -x -q -m gpu 2>&1 | tail -2
  for s in 1 5; do
    SPLAT_BWD_CHUNK=$v python bench.py --no-cpu-baseline --steps 10 --warmup 3 --streams $s > gpurun_out/exp_bwd_v${v}_s${s}.json 2> gpurun_out/exp_bwd_v${v}_s${s}.err
    python - <<PY
import json
d = json.loads(open("gpurun_out/exp_bwd_v${v}_s${s}.json").read().strip().splitlines()[-1])
st = d.get("stage_ms", {})
print("variant $v streams $s value %.1f e2e %.1f | " % (d["value"], d["e2e"]["value"]) + " ".join("%s %.4f" % (k[:14], x) for k, x in st.items()))
PY
  done
done
