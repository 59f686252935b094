#!/bin/bash
# Under gpurun --gpus N: the headline bench at N ranks (value + e2e), JSON into gpurun_out/<tag>_bench_n<N>.json
N=$1; TAG=${2:-r02}; OUT=gpurun_out; mkdir -p $OUT
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 200 --warmup 10 --no-cpu > $OUT/${TAG}_bench_n$N.json 2> $OUT/${TAG}_bench_n$N.err
echo "exit $?"; nproc
python - <<PY
import json
d = json.load(open("$OUT/${TAG}_bench_n$N.json"))
print("N", d["n_gpus"], "value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "h2d MB", round(d["e2e"]["h2d_bytes_per_step"] / 1e6, 1))
PY
