set -x
python bench.py > gpurun_out/r1d_bench.json 2> gpurun_out/r1d_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1d_bench_ref.json 2>> gpurun_out/r1d_bench.err
python tools/sat_bench.py > gpurun_out/r1d_sat.jsonl 2>> gpurun_out/r1d_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1d_launches_bench_noskip.csv python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:rc1pass_kernel -s 10 -c 2 -o gpurun_out/r1d_rc1pass python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu2.log 2>&1
ncu -i gpurun_out/r1d_rc1pass.ncu-rep --page raw --csv > gpurun_out/r1d_rc1pass_raw.csv
ncu --set full --clock-control none --import-source on -k regex:sat -c 10 -o gpurun_out/r1d_sat python tools/sat_bench.py --once > gpurun_out/ncu3.log 2>&1
ncu -i gpurun_out/r1d_sat.ncu-rep --page raw --csv > gpurun_out/r1d_sat_raw.csv
ls -la gpurun_out
