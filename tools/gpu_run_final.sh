#!/usr/bin/env bash
# round-end validation: the whole GPU suite, smoke(), the bench (both arms) and the ncu launch list of the bench command
set -u
mkdir -p gpurun_out
T=${TAG:-r2final}
timeout 2700 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/${T}_pytest_gpu.log
cp gpurun_out/benchmark-bls12381-gpu.txt gpurun_out/${T}_reference_benchmark_bls12381_on_gpu.txt 2>/dev/null
cp gpurun_out/benchmark-bn254-gpu.txt gpurun_out/${T}_reference_benchmark_bn254_on_gpu.txt 2>/dev/null
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"
tail -c 600 gpurun_out/${T}_bench_n1.err
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_bench_n1.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['phases_ms'], d['launches_per_msm'], d['roofline']['frac'])
print(d['strong']['msm_2p22']['ms'], d['commit']['value'], d['commit']['e2e'], d['commit']['proofs_per_s_e2e'])
for k,v in d['secondary'].items(): print(k, round(v['ms'],3), round(v.get('frac',0),3))
PY
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench_reference_arm.err; echo "ref arm rc=$?"
head -c 1200 gpurun_out/${T}_bench_reference_arm.json; echo
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches_bench.csv python bench.py --steps 2 --warmup 1 > gpurun_out/${T}_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
python tools/launch_summary.py gpurun_out/${T}_launches_bench.csv 2>/dev/null | head -30
ls -la gpurun_out | tail -15
