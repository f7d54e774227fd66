set -u
T=r2m
NCU="ncu --set full --clock-control none"
for w in msm16 msm17 msm18; do
timeout 600 $NCU -k regex:msm_accumulate -s 2 -c 1 -o gpurun_out/${T}_acc_$w -f python tools/ncu_targets.py --what $w --reps 2 > gpurun_out/${T}_ncu_$w.log 2>&1; echo "ncu $w rc=$?"
ncu -i gpurun_out/${T}_acc_$w.ncu-rep --page raw --csv > gpurun_out/${T}_acc_$w.raw.csv 2>/dev/null; rm -f gpurun_out/${T}_acc_$w.ncu-rep
done
