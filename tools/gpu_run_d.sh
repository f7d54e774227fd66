#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
T=${TAG:-r2d}
(lscpu | head -20; nproc; free -g | head -2) > gpurun_out/${T}_host.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/${T}_pytest_gpu.log
cp gpurun_out/benchmark-bls12381-gpu.txt gpurun_out/${T}_reference_benchmark_bls12381_on_gpu.txt 2>/dev/null
cp gpurun_out/benchmark-bn254-gpu.txt gpurun_out/${T}_reference_benchmark_bn254_on_gpu.txt 2>/dev/null
cat gpurun_out/${T}_reference_benchmark_bls12381_on_gpu.txt
timeout 900 python bench.py > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"
tail -c 600 gpurun_out/${T}_bench_n1.err
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_bench_n1.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['phases_ms'], d['launches_per_msm'])
print(d['strong']['msm_2p22']['ms'], d['commit']['value'], d['commit']['e2e'], d['commit']['proofs_per_s_e2e'])
for k,v in d['secondary'].items(): print(k, round(v['ms'],3), round(v.get('frac',0),3))
print(d.get('cpu_baseline'))
PY
# config 1: the reference's own CPU benchmark, unmodified, on this box's host cores
(time timeout 900 ./oracle/_ref/benchmark-bn254) > gpurun_out/${T}_reference_benchmark_bn254_cpu.txt 2>&1; echo "cpu bench rc=$?"
tail -25 gpurun_out/${T}_reference_benchmark_bn254_cpu.txt
