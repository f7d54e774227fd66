set -u
T=${TAG:-r2p}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tail_and_group.py tests/test_gpu_kat_large.py tests/test_zz_no_alloc.py -x -q --timeout 900 > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${T}_pytest.log
timeout 600 python tools/gpu_probe_latency.py > gpurun_out/${T}_latency.json 2> gpurun_out/${T}_latency.err; echo "latency rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${T}_latency.json')); print(d['us_per_step']); print({k:v['ms'] for k,v in d['msm_ms'].items() if 'sleep64' in k})"
timeout 900 python tools/gpu_probe.py --sizes 12,13,14,15,16,17,18,19,20 --precompute 1 --out gpurun_out/${T}_probe_sizes.json > gpurun_out/${T}_probe_sizes.log 2>&1; echo "probe rc=$?"
grep -o '"log2n": [0-9]*\|"ms": [0-9.]*\|"phases": {[^}]*}' gpurun_out/${T}_probe_sizes.log | paste - - - | head -20
timeout 900 python bench.py > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"
tail -c 400 gpurun_out/${T}_bench_n1.err
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_bench_n1.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['phases_ms'], d['launches_per_msm'])
print(d['strong']['msm_2p22']['ms'], d['commit']['value'], d['commit']['e2e'], d['commit']['proofs_per_s_e2e'])
for k,v in d['secondary'].items(): print(k, round(v['ms'],3), round(v.get('frac',0),3))
PY
