#!/usr/bin/env bash
# one gpurun call: new tests first, then the full GPU suite, bench, size sweep, launch list
set -u
mkdir -p gpurun_out
T=${TAG:-r2a}
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/${T}_gpu.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_tail_and_group.py -q --timeout 600 > gpurun_out/${T}_pytest_new.log 2>&1; echo "new tests rc=$?"
tail -30 gpurun_out/${T}_pytest_new.log
timeout 900 python -m pytest tests/test_gpu_kat_large.py -q --timeout 600 > gpurun_out/${T}_pytest_kat.log 2>&1; echo "kat tests rc=$?"
tail -30 gpurun_out/${T}_pytest_kat.log
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 --deselect tests/test_gpu_tail_and_group.py --deselect tests/test_gpu_kat_large.py > gpurun_out/${T}_pytest_rest.log 2>&1; echo "rest rc=$?"
tail -15 gpurun_out/${T}_pytest_rest.log
timeout 600 python bench.py > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/${T}_bench_n1.err
cat gpurun_out/${T}_bench_n1.json | head -c 6000
timeout 600 python tools/gpu_probe.py --sizes 12,13,14,15,16,17,18,19,20 --precompute 1 --out gpurun_out/${T}_probe_sizes.json > gpurun_out/${T}_probe_sizes.log 2>&1; echo "probe rc=$?"
grep -o '"log2n": [0-9]*\|"ms": [0-9.]*\|"phases": {[^}]*}' gpurun_out/${T}_probe_sizes.log | paste - - - | head -20
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches_bench.csv python bench.py --steps 2 --warmup 3 --blobs 0 --no-cpu --no-strong --no-secondary > gpurun_out/${T}_ncu1.log 2>&1; echo "ncu rc=$?"
