#!/bin/bash
# 1 -> 8 scaling of the headline on one box (run under `gpurun --gpus 8`): the bench line at N = 1, 2, 4, 8 as the driver launches it,
# the 8-GPU line with the gather into every rank (--gather all; rank 0 only is the default), config 5 at 8 GPUs (WITH_C5=1) and the
# reference arm (WITH_REFERENCE=1).  Output: gpurun_out/${TAG}_*.json
TAG=${1:-r2s}
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
set -x
NS=${2:-"2 4 8"}
for n in $NS; do
  timeout 300 $T --nproc-per-node $n --master-port $((29500+n)) bench.py --gpus $n --steps 50 --warmup 5 --no-c5 > gpurun_out/${TAG}_bench_n$n.json 2> gpurun_out/${TAG}_bench_n$n.err
  echo "rc=$?"; tail -c 300 gpurun_out/${TAG}_bench_n$n.err
done
timeout 300 $T --nproc-per-node 8 --master-port 29520 bench.py --gpus 8 --steps 50 --warmup 5 --no-c5 --gather all > gpurun_out/${TAG}_bench_n8_all.json 2> gpurun_out/${TAG}_bench_n8_all.err; echo "rc=$?"
if [ -n "${WITH_C5:-}" ]; then
  timeout 900 $T --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --config c5 --steps 3 --warmup 1 > gpurun_out/${TAG}_c5_n8.json 2> gpurun_out/${TAG}_c5_n8.err; echo "rc=$?"; tail -c 400 gpurun_out/${TAG}_c5_n8.err
fi
if [ -n "${WITH_REFERENCE:-}" ]; then
  timeout 600 $T --nproc-per-node 8 --master-port 29522 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/${TAG}_reference_n8.json 2> gpurun_out/${TAG}_reference_n8.err; echo "rc=$?"
fi
python - <<PY
import json
for n in ("n1","n2","n4","n8","n8_all"):
    try:
        d=json.loads(open("gpurun_out/${TAG}_bench_%s.json"%n).read().strip().splitlines()[-1])
        print(n, "ms", round(d["ms_per_step"],4), "e2e ms", round(d["e2e"]["ms_per_step"],4), "skip", {k:round(v["ms_per_step"],4) for k,v in d["measured"]["with_skipping"].items()})
    except Exception as e: print(n, "ERR", e)
PY
