#!/bin/bash
# ncu launch list (gpu__time_duration per launch) of the default bench command, after the same command ran clean.
# Usage (GPU box):  bash profiles/capture_launches.sh <tag>  -> gpurun_out/<tag>_bench.json, gpurun_out/<tag>_launches.csv
set -e
tag=${1:-r01}
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mardyn-gpu --e2e-steps 1 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mardyn-gpu --e2e-steps 1 > gpurun_out/${tag}_launches_ncu.log 2>&1
