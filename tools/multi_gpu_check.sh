#!/bin/bash
# One gpurun --gpus N call: headline bench, configs[3] mode, the configs[4] spot sweep and the concurrent H2D probe at N ranks.
# usage (under gpurun --gpus N):  bash tools/multi_gpu_check.sh N tag
N=$1; TAG=$2; OUT=gpurun_out; mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
[ "$N" = "1" ] && RUN="python"
nvidia-smi topo -m > $OUT/${TAG}_topo_n$N.txt 2>&1
timeout 400 $RUN bench.py --gpus $N --steps 200 --warmup 10 --no-cpu > $OUT/${TAG}_bench_n$N.json 2> $OUT/${TAG}_bench_n$N.err; echo "bench n$N exit $?"
timeout 400 $RUN bench.py --gpus $N --steps 20 --warmup 3 --workload configs3 > $OUT/${TAG}_configs3_n$N.json 2> $OUT/${TAG}_configs3_n$N.err; echo "configs3 n$N exit $?"
timeout 600 $RUN tools/sweep.py --quick > $OUT/${TAG}_sweep_n$N.jsonl 2> $OUT/${TAG}_sweep_n$N.err; echo "sweep n$N exit $?"
timeout 200 $RUN tools/h2d_probe.py > $OUT/${TAG}_h2d_n$N.json 2> $OUT/${TAG}_h2d_n$N.err; echo "h2d n$N exit $?"
python - <<PY
import json
for f in ("bench", "configs3"):
    try:
        d = json.load(open("$OUT/${TAG}_%s_n$N.json" % f))
        print(f, "n", d["n_gpus"], "value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), "e2e", (d.get("e2e") or {}).get("value"))
    except Exception as e:
        print(f, "unreadable", e)
print(open("$OUT/${TAG}_h2d_n$N.json").read())
PY
