#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
T=${TAG:-r2g}
timeout 600 python tools/gpu_probe_latency.py > gpurun_out/${T}_latency.json 2> gpurun_out/${T}_latency.err; echo "latency rc=$?"
cat gpurun_out/${T}_latency.json; tail -3 gpurun_out/${T}_latency.err
NCU="ncu --set full --clock-control none --import-source on"
timeout 900 $NCU -k regex:msm_accumulate -s 3 -c 2 -o gpurun_out/${T}_accumulate_g2 -f python tools/ncu_targets.py --what g2msm --reps 2 > gpurun_out/${T}_ncu_g2.log 2>&1; echo "ncu g2 rc=$?"
ncu -i gpurun_out/${T}_accumulate_g2.ncu-rep --page raw --csv > gpurun_out/${T}_accumulate_g2.raw.csv 2>/dev/null; rm -f gpurun_out/${T}_accumulate_g2.ncu-rep
