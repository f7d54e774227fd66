set -u
T=${TAG:-r2q}
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${T}_smoke.log
timeout 900 python tools/gpu_probe_large_degree.py > gpurun_out/${T}_large_degree.json 2> gpurun_out/${T}_large_degree.err; echo "large degree rc=$?"
tail -5 gpurun_out/${T}_large_degree.err
NCU="ncu --set full --clock-control none --import-source on"
timeout 900 $NCU -k regex:msm_accumulate -c 8 -o gpurun_out/${T}_accumulate_g2 -f python tools/ncu_targets.py --what g2msm --reps 1 > gpurun_out/${T}_ncu_g2.log 2>&1; echo "ncu g2 rc=$?"
ncu -i gpurun_out/${T}_accumulate_g2.ncu-rep --page raw --csv > gpurun_out/${T}_accumulate_g2.raw.csv 2>/dev/null; rm -f gpurun_out/${T}_accumulate_g2.ncu-rep
python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/${T}_accumulate_g2.raw.csv')))
h=rows[0]
for r in rows[2:]:
    d=dict(zip(h,r)); print(d['Kernel Name'][:60], d['launch__grid_size'], d['gpu__time_duration.sum'], d['sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed'], d['launch__registers_per_thread'])
PY
