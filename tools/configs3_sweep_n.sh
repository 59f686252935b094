#!/bin/bash
# Under gpurun --gpus N: the configs[3] mode of the bench and the quick configs[2..4] sweep at N ranks.
N=$1; TAG=${2:-r02}; OUT=gpurun_out; mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29566"
timeout 300 $RUN bench.py --gpus $N --steps 20 --warmup 3 --workload configs3 > $OUT/${TAG}_configs3_n$N.json 2> $OUT/${TAG}_configs3_n$N.err; echo "configs3 exit $?"
timeout 400 $RUN tools/sweep.py --quick > $OUT/${TAG}_sweep_n$N.jsonl 2> $OUT/${TAG}_sweep_n$N.err; echo "sweep exit $?"
python - <<PY
import json
d = json.load(open("$OUT/${TAG}_configs3_n$N.json"))
print("configs3 N", d["n_gpus"], "value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4))
PY
wc -l $OUT/${TAG}_sweep_n$N.jsonl
