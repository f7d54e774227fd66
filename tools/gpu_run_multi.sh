#!/usr/bin/env bash
# multi-GPU call: group tests on real devices, the NCCL two-rank test, then bench.py at N GPUs
set -u
N=${N:-2}
T=${TAG:-r2c}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${T}_topo.txt 2>&1
timeout 1200 python -m pytest tests/test_gpu_tail_and_group.py -q --timeout 900 -k "group or nccl" > gpurun_out/${T}_pytest_group.log 2>&1; echo "group tests rc=$?"
tail -15 gpurun_out/${T}_pytest_group.log
timeout 900 python bench.py --single-process --gpus $N --steps 5 > gpurun_out/${T}_bench_single_process_n${N}.json 2> gpurun_out/${T}_bench_single_process_n${N}.err; echo "single-process rc=$?"
cat gpurun_out/${T}_bench_single_process_n${N}.json
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu > gpurun_out/${T}_bench_n${N}.json 2> gpurun_out/${T}_bench_n${N}.err; echo "bench rc=$?"
tail -c 800 gpurun_out/${T}_bench_n${N}.err
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_bench_n${N}.json'))
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e'])
print(d.get('strong')); print({k:v for k,v in d['commit'].items() if k in ('value','e2e','proofs_per_s_e2e','ms_resident','ms_e2e')})
PY
