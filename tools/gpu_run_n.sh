set -u
T=${TAG:-r2n}
timeout 1200 python -m pytest tests/test_gpu_tail_and_group.py tests/test_gpu_parity.py -q --timeout 900 -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${T}_pytest.log
timeout 900 python tools/gpu_probe.py --sizes 12,13,14,15,16,17,18,19,20,21,22 --precompute 1 --out gpurun_out/${T}_probe_sizes.json > gpurun_out/${T}_probe_sizes.log 2>&1; echo "probe rc=$?"
grep -o '"log2n": [0-9]*\|"ms": [0-9.]*\|"phases": {[^}]*}' gpurun_out/${T}_probe_sizes.log | paste - - - | head -20
timeout 600 python bench.py --no-strong --no-secondary --no-cpu --blobs 4096 > gpurun_out/${T}_bench_quick.json 2> gpurun_out/${T}_bench_quick.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_quick.json')); print(d['ms_per_step'], d['phases_ms'], d['commit']['value'], d['commit']['phases_ms'])"
