# Final single-GPU evidence of the round (run under gpurun, one GPU). Outputs go to gpurun_out/.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 > gpurun_out/r2b_bench_n1.json 2> gpurun_out/r2b_bench_n1.err; echo bench rc $?
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2b_bench_ref_n1.json 2> gpurun_out/r2b_bench_ref_n1.err; echo ref rc $?
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2b_launches.csv \
    python bench.py --steps 2 --warmup 3 --c5-small > gpurun_out/r2b_ncu_list.log 2>&1; echo ncu list rc $?
ncu --set full --import-source on --clock-control none -k regex:CostKernelPersistent --launch-skip 1 --launch-count 1 \
    -o gpurun_out/r2b_cost_kernel -f python tools/ab_cost.py --once "HG_TILE_OUT=1024" > gpurun_out/r2b_ncu_cost.log 2>&1; echo ncu cost rc $?
HG_TRACE=1 python tools/trace_c4.py --device-only > gpurun_out/r2b_c4_trace.log 2>&1
python tools/ab_cost.py "HG_NO_STENCILS=0" "HG_NO_STENCILS=1" 2>&1 | tail -2
