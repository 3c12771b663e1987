# Round-2 ncu evidence (run under gpurun, one GPU). Outputs go to gpurun_out/.
set -x
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --c5-small > gpurun_out/r2_ncu_list.log 2>&1
HG_CONSTRAINT_THREADS=1 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_c5_launches.csv \
    python tools/trace_c5.py 2 100 > gpurun_out/r2_c5_ncu.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:CostKernelPersistent --launch-skip 1 --launch-count 1 \
    -o gpurun_out/r2_cost_kernel -f python tools/ab_cost.py --once "HG_TILE_OUT=1024" > gpurun_out/r2_ncu_cost.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:ScoreLatticePackedKernel --launch-skip 1 --launch-count 1 \
    -o gpurun_out/r2_lattice_kernel -f python tools/trace_c4.py > gpurun_out/r2_ncu_lattice.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:"DiveKernel|ExpandKernel" --launch-skip 7 --launch-count 4 \
    -o gpurun_out/r2_search_kernels -f python tools/trace_c4.py > gpurun_out/r2_ncu_search.log 2>&1
ls -la gpurun_out/*.ncu-rep
