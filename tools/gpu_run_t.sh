set -u
T=${TAG:-r2y}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tail_and_group.py tests/test_gpu_kat_large.py -x -q --timeout 900 > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/${T}_pytest.log
timeout 600 python tools/gpu_probe_g2.py > gpurun_out/${T}_probe_g2.json 2> gpurun_out/${T}_probe_g2.err; echo "g2 probe rc=$?"; tail -3 gpurun_out/${T}_probe_g2.err
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_probe_g2.json'))
for k,v in d.items():
    if 'latency' in k: print(k, v)
    elif isinstance(v,dict): print(k, v['ms'], {a:round(b,3) for a,b in (v['phases'] or {}).items()})
    elif not isinstance(v,dict): print(k,[round(x,2) for x in v])
PY
