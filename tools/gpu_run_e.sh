#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
T=${TAG:-r2e}
timeout 1500 python -m pytest tests/test_gpu_dropin.py tests/test_gpu_tail_and_group.py tests/test_gpu_parity.py -q --timeout 900 > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/${T}_pytest.log
cp gpurun_out/benchmark-bls12381-gpu.txt gpurun_out/${T}_reference_benchmark_bls12381_on_gpu.txt 2>/dev/null
cp gpurun_out/benchmark-bn254-gpu.txt gpurun_out/${T}_reference_benchmark_bn254_on_gpu.txt 2>/dev/null
cat gpurun_out/${T}_reference_benchmark_bls12381_on_gpu.txt
# gather layout experiment: 96-byte table entries vs one 128-byte line per entry
for pad in 0 1; do
  timeout 600 python tools/gpu_probe.py --sizes 20 --precompute 1 --table-pad $pad --reps 5 --out gpurun_out/${T}_probe_pad$pad.json > gpurun_out/${T}_probe_pad$pad.log 2>&1; echo "probe pad$pad rc=$?"
  grep -o '"ms": [0-9.]*\|"phases": {[^}]*}' gpurun_out/${T}_probe_pad$pad.log | paste - - | tail -1
done
NCU="ncu --set full --clock-control none --import-source on"
for pad in 0 1; do
  timeout 900 $NCU -k regex:msm_accumulate -s 1 -c 1 -o gpurun_out/${T}_accumulate_pad$pad -f python tools/ncu_targets.py --what msm20 --table-pad $pad --reps 2 > gpurun_out/${T}_ncu_acc_pad$pad.log 2>&1; echo "ncu acc pad$pad rc=$?"
done
timeout 900 $NCU -k regex:msm_accumulate -s 1 -c 1 -o gpurun_out/${T}_accumulate_g2 -f python tools/ncu_targets.py --what g2msm --reps 2 > gpurun_out/${T}_ncu_g2.log 2>&1; echo "ncu g2 rc=$?"
timeout 900 $NCU -k regex:fixed_base_mul -s 2 -c 2 -o gpurun_out/${T}_fixed_base_mul -f python tools/ncu_targets.py --what setup --reps 2 > gpurun_out/${T}_ncu_setup.log 2>&1; echo "ncu setup rc=$?"
timeout 900 $NCU -k regex:poly_mul_trunc -s 20 -c 3 -o gpurun_out/${T}_poly_mul_trunc -f python tools/ncu_targets.py --what quotient --reps 2 > gpurun_out/${T}_ncu_quot.log 2>&1; echo "ncu quotient rc=$?"
timeout 900 $NCU -k regex:msm_tail -s 2 -c 1 -o gpurun_out/${T}_msm_tail_2p12 -f python tools/ncu_targets.py --what msm12 --reps 2 > gpurun_out/${T}_ncu_tail.log 2>&1; echo "ncu tail rc=$?"
ls -la gpurun_out/*.ncu-rep
