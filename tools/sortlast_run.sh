set -x
for n in 2 8; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) tools/bench_sortlast.py --steps 10 > gpurun_out/r1c_sortlast_n$n.jsonl 2> gpurun_out/r1c_sortlast_n$n.err
  echo "rc=$?"; cut -c1-300 gpurun_out/r1c_sortlast_n$n.jsonl; tail -3 gpurun_out/r1c_sortlast_n$n.err
done
