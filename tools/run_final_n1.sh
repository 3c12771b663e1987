set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 20 --warmup 5 > gpurun_out/r2b_bench_n1.json 2> gpurun_out/r2b_bench_n1.err; echo bench rc $?
tail -c 600 gpurun_out/r2b_bench_n1.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2b_launches.csv python bench.py --steps 2 --warmup 3 --c5-small > gpurun_out/r2b_ncu_list.log 2>&1; echo ncu rc $?
