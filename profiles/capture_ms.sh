#!/bin/bash
# One `ncu --set full` capture of the multi-site list kernel on examples/CO2 replicated to 1 048 576 molecules.
# Usage (GPU box):  bash profiles/capture_ms.sh <tag>   -> gpurun_out/<tag>_plain.json, gpurun_out/<tag>.ncu-rep
set -e
tag=${1:-r02m_nl_force_ms_co2}
mkdir -p gpurun_out
python bench.py --workload co2 --steps 6 --warmup 2 --no-cpu-baseline --no-mardyn-gpu > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err
ncu --set full --clock-control none --import-source on -k regex:nl_force_ms_kernel --launch-skip 3 --launch-count 1 \
    -f -o gpurun_out/${tag} python bench.py --workload co2 --steps 6 --warmup 2 --no-cpu-baseline --no-mardyn-gpu > gpurun_out/${tag}_ncu.log 2>&1
