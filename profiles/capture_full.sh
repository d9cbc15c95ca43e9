#!/bin/bash
# One `ncu --set full` capture of the dominant kernel (nl_force_kernel) of the default bench workload (10.06M molecules).
# Usage (GPU box):  bash profiles/capture_full.sh <tag> [kernel-regex] [launch-skip]   -> gpurun_out/<tag>_plain.json, gpurun_out/<tag>.ncu-rep
# The plain run comes first and must exit 0; numbers printed under ncu are never bench values.
set -e
tag=${1:-r01_nl_force_10M}
kern=${2:-nl_force_kernel}
skip=${3:-4}
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mardyn-gpu --e2e-steps 1 > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err
# launch-skip 4: past the warm-up launches, a steady-state force pass in the frozen list layout
ncu --set full --clock-control none --import-source on -k regex:$kern --launch-skip $skip --launch-count 1 \
    -f -o gpurun_out/${tag} python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-mardyn-gpu --e2e-steps 1 > gpurun_out/${tag}_ncu.log 2>&1
