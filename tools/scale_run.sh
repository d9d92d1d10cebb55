set -x
for n in 2 4 8; do
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 50 --warmup 5 > gpurun_out/r1e_bench_n$n.json 2> gpurun_out/r1e_bench_n$n.err
  echo "rc=$?"; tail -c 1500 gpurun_out/r1e_bench_n$n.json
done
