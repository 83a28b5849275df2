#!/usr/bin/env bash
# Round-2 evidence run (one GPU): bench line, then the ncu launch list of the same command, then --set full captures of the chain
# kernels on the 12-picture > L2 batch and of the search kernels.  Outputs under gpurun_out/$TAG_*.
set -u
TAG=${1:-r02}
mkdir -p gpurun_out
python bench.py --no-stripes > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${TAG}_launches_bench_4k10.csv \
    python bench.py --steps 1 --warmup 1 --passes 2 --no-stripes > gpurun_out/${TAG}_ncu_list.log 2>&1
python tools/prof_batch.py > /dev/null || exit 1
# the last pass of the 12-picture batch: launches 2*12*2 .. (deblock, cdef3 alternate); capture 6 of the last pass
ncu --set full --clock-control none --import-source on -k regex:'deblock_kernel|cdef3_kernel' --launch-skip 60 --launch-count 6 \
    -o gpurun_out/${TAG}_chain_batch -f python tools/prof_batch.py > gpurun_out/${TAG}_ncu_chain.log 2>&1
PROF_PICTURES=3 PROF_PASSES=1 ncu --set full --clock-control none --import-source on -k regex:'cdef_search3_kernel|search_fold|pick_' --launch-skip 0 --launch-count 12 \
    -o gpurun_out/${TAG}_search -f python tools/prof_batch.py search > gpurun_out/${TAG}_ncu_search.log 2>&1
ls -la gpurun_out | tail -8
