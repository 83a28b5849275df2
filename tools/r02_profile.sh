#!/usr/bin/env bash
# Round-2 evidence run (one GPU): the ncu launch list of the bench command, --set full captures of the chain kernels on the 12-picture
# > L2 batch and of the search / pick kernels, and the steady-state DRAM traffic per launch (no cache control, one pass).
# Outputs under gpurun_out/$TAG_*.  The bench line itself is taken by a separate run without ncu.
set -u
TAG=${1:-r02}
mkdir -p gpurun_out
python tools/prof_batch.py > /dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${TAG}_launches_bench_4k10.csv \
    python bench.py --steps 1 --warmup 1 --passes 2 --no-stripes --no-cpu > gpurun_out/${TAG}_ncu_list.log 2>&1
# the last pass of the 12-picture batch (deblock, cdef5 alternate: 24 launches per pass, 3 passes)
ncu --set full --clock-control none --import-source on -k regex:'deblock_kernel|cdef5_kernel' --launch-skip 60 --launch-count 6 \
    -o gpurun_out/${TAG}_chain_batch -f python tools/prof_batch.py > gpurun_out/${TAG}_ncu_chain.log 2>&1
ncu --set full --clock-control none -k regex:'dbk_edge_plane' --launch-skip 6 --launch-count 2 \
    -o gpurun_out/${TAG}_edge_plane -f python tools/prof_batch.py > gpurun_out/${TAG}_ncu_edge.log 2>&1
[ -n "${SKIP_SEARCH:-}" ] || PROF_PICTURES=3 PROF_PASSES=1 ncu --set full --clock-control none --import-source on -k regex:'cdef_search3_kernel|search_fold|pick_' --launch-skip 0 --launch-count 8 \
    -o gpurun_out/${TAG}_search -f python tools/prof_batch.py search > gpurun_out/${TAG}_ncu_search.log 2>&1
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --cache-control none --clock-control none \
    -k regex:'deblock_kernel|cdef5_kernel' --launch-skip 48 --launch-count 24 --csv --log-file gpurun_out/${TAG}_traffic_batch.csv \
    python tools/prof_batch.py > gpurun_out/${TAG}_traffic.log 2>&1
[ -n "${SKIP_SEARCH:-}" ] || PROF_PICTURES=3 PROF_PASSES=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --cache-control none --clock-control none \
    -k regex:'cdef_search3_kernel' --csv --log-file gpurun_out/${TAG}_traffic_search.csv python tools/prof_batch.py search >> gpurun_out/${TAG}_traffic.log 2>&1
ls -la gpurun_out | grep ${TAG}
