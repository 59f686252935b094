#!/bin/bash
# Under gpurun --gpus N: end-to-end (host-buffer) throughput of the headline bench at N ranks for several host-worker counts.
# usage: bash tools/host_workers_sweep.sh N "0 2 4 8"
N=$1; OUT=gpurun_out; mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
for w in $2; do
  PVV_HOST_WORKERS=$w timeout 300 $RUN bench.py --gpus $N --steps 50 --warmup 5 --no-cpu > $OUT/hw_n${N}_w$w.json 2> $OUT/hw_n${N}_w$w.err
  python - <<PY
import json
try:
    d = json.load(open("$OUT/hw_n${N}_w$w.json"))
    print("N=$N workers=$w value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "h2d MB/step", round(d["e2e"]["h2d_bytes_per_step"] / 1e6, 1))
except Exception as e:
    print("N=$N workers=$w failed", e)
PY
done
