# Round-2 ncu evidence (run under gpurun, one GPU). Outputs go to gpurun_out/.
set -x
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --c5-small > gpurun_out/r2_plain.json 2> gpurun_out/r2_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --c5-small > gpurun_out/r2_ncu_list.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:CostKernel --launch-skip 3 --launch-count 1 \
    -o gpurun_out/r2_cost_kernel -f python tools/ab_cost.py --once "HG_TILE_OUT=1024" > gpurun_out/r2_ncu_cost.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:ScoreLatticePackedKernel --launch-skip 1 --launch-count 1 \
    -o gpurun_out/r2_lattice_kernel -f python tools/trace_c4.py > gpurun_out/r2_ncu_lattice.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:"DiveLevelKernel|ExpandKernel" --launch-skip 8 --launch-count 4 \
    -o gpurun_out/r2_search_kernels -f python tools/trace_c4.py > gpurun_out/r2_ncu_search.log 2>&1
ls -la gpurun_out/*.ncu-rep
