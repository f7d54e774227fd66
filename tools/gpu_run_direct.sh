set -u
T=${TAG:-r2dir}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tail_and_group.py -x -q --timeout 600 -k "direct" > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${T}_pytest.log
timeout 900 python tools/gpu_probe_direct.py ${BUDGETS:-40} > gpurun_out/${T}_probe_direct.json 2> gpurun_out/${T}_probe_direct.err; echo "probe rc=$?"; tail -3 gpurun_out/${T}_probe_direct.err
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_probe_direct.json'))
for k,v in d.items():
    print(k, v if not isinstance(v,dict) else {a:b for a,b in v.items() if a!='phases'}, (v.get('phases') if isinstance(v,dict) else ''))
PY
