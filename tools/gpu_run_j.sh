#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
T=${TAG:-r2j}
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/${T}_pytest_gpu.log
cp gpurun_out/benchmark-bls12381-gpu.txt gpurun_out/${T}_reference_benchmark_bls12381_on_gpu.txt 2>/dev/null
cp gpurun_out/benchmark-bn254-gpu.txt gpurun_out/${T}_reference_benchmark_bn254_on_gpu.txt 2>/dev/null
timeout 900 python bench.py > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"
tail -c 600 gpurun_out/${T}_bench_n1.err
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_bench_n1.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['phases_ms'], d['launches_per_msm'], d['roofline']['traffic'])
print(d['strong']['msm_2p22']['ms'], d['commit']['value'], d['commit']['e2e'], d['commit']['proofs_per_s_e2e'])
for k,v in d['secondary'].items(): print(k, round(v['ms'],3), round(v.get('frac',0),3))
PY
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench_reference_arm.err; echo "ref arm rc=$?"
cat gpurun_out/${T}_bench_reference_arm.json | head -c 1500
