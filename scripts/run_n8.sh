#!/usr/bin/env bash
# The 8-GPU measurements of the round: weak + strong window (replica), config 5 (scannet), each with the library's own
# exchange kernel and with NCCL.   gpurun --gpus 8 -- 'bash scripts/run_n8.sh'
N=${N:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err; tail -c 300 gpurun_out/$2.err | tr '\n' ' '; echo; }
mkdir -p gpurun_out
run 29601 r02_n${N}_replica --verify
run 29602 r02_n${N}_replica_nccl --exchange nccl --no-extras
run 29603 r02_n${N}_scannet --workload scannet --steps 10 --warmup 3 --verify
run 29604 r02_n${N}_scannet_nccl --workload scannet --steps 10 --warmup 3 --exchange nccl --no-extras
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/r02_n${N}_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "NO JSON", e); continue
    print(f, "value %.1f e2e %.1f ms/step %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), d.get("exchange"), d.get("strong", {}).get("value"), d.get("verify"))
PY
