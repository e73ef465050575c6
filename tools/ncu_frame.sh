#!/bin/bash
# ncu captures of ONE frame of the C3 workload (view 0) on the GPU box (run via gpurun, after the plain command exited 0).
#   tools/ncu_frame.sh <tag>   ->  gpurun_out/<tag>_plain.json, <tag>_launches.csv, <tag>_full.ncu-rep, <tag>_full_raw.csv
# then here:  python tools/ncu_summarise.py gpurun_out/<tag>_full_raw.csv profiles/<tag> "<what>" gpurun_out/<tag>_plain.json
set -u
TAG=${1:-r2}
CMD="python bench.py --steps 2 --warmup 3 --views-per-rank 1 --streams 1 --no-e2e --no-cpu-baseline --no-extras"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
# one frame = 8 launches of the library (preprocess, tile scan, scatter, tile sort, large-tile sort, render fwd, render bwd,
# preprocess bwd); only those kernels are counted (-k regex on the base name).  Launches before the window: the workload
# counters' forward (6) + 3 warm-up steps (24) + the unsampled region of 2 steps (16) = 46, so the window is the first
# frame of the next timed region: steady state, view 0.
ncu --set full --clock-control none --import-source on -k regex:"^(preprocess|tile_|render_)" --launch-skip 46 -c 8 -f -o gpurun_out/${TAG}_full $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2>/dev/null
tail -3 gpurun_out/${TAG}_ncu2.log
