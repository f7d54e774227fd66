#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
T=${TAG:-r2b}
timeout 900 python -m pytest tests/test_gpu_tail_and_group.py -q --timeout 600 > gpurun_out/${T}_pytest_new.log 2>&1; echo "new tests rc=$?"
tail -15 gpurun_out/${T}_pytest_new.log
timeout 900 python tools/gpu_probe.py --sizes 12,13,14,15,16,17,18,19,20 --precompute 1 --out gpurun_out/${T}_probe_sizes.json > gpurun_out/${T}_probe_sizes.log 2>&1; echo "probe rc=$?"
grep -o '"log2n": [0-9]*\|"ms": [0-9.]*\|"phases": {[^}]*}' gpurun_out/${T}_probe_sizes.log | paste - - - | head -20
timeout 900 python tools/gpu_probe.py --sizes 12,14,16,17 --precompute 1 --windows 8,9,10,11,12,13,14,15,16 --out gpurun_out/${T}_probe_windows.json > gpurun_out/${T}_probe_windows.log 2>&1; echo "probe windows rc=$?"
grep -o '"log2n": [0-9]*\|"window_bits": [0-9]*\|"ms": [0-9.]*\|"phases": {[^}]*}' gpurun_out/${T}_probe_windows.log | paste - - - - | head -40
timeout 900 python tools/gpu_probe.py --sizes 12,16 --precompute 1 --chunks 1,2,4,8,16 --out gpurun_out/${T}_probe_chunks.json > gpurun_out/${T}_probe_chunks.log 2>&1; echo "probe chunks rc=$?"
grep -o '"log2n": [0-9]*\|"task_chunk": [0-9]*\|"ms": [0-9.]*\|"phases": {[^}]*}' gpurun_out/${T}_probe_chunks.log | paste - - - - | head -40
timeout 600 python bench.py --no-strong --no-secondary --no-cpu --blobs 512 > gpurun_out/${T}_bench_quick.json 2> gpurun_out/${T}_bench_quick.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_quick.json')); print(d['ms_per_step'], d['phases_ms'], d['launches_per_msm'], d['commit']['value'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches_bench.csv python bench.py --steps 2 --warmup 3 --blobs 0 --no-cpu --no-strong --no-secondary > gpurun_out/${T}_ncu1.log 2>&1; echo "ncu rc=$?"
