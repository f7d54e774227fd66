#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
T=${TAG:-r2f}
timeout 1500 python -m pytest tests/test_gpu_dropin.py tests/test_gpu_parity.py tests/test_gpu_kat_large.py -q --timeout 900 > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/${T}_pytest.log
cp gpurun_out/benchmark-bls12381-gpu.txt gpurun_out/${T}_reference_benchmark_bls12381_on_gpu.txt 2>/dev/null
cp gpurun_out/benchmark-bn254-gpu.txt gpurun_out/${T}_reference_benchmark_bn254_on_gpu.txt 2>/dev/null
cat gpurun_out/${T}_reference_benchmark_bls12381_on_gpu.txt
NCU="ncu --set full --clock-control none --import-source on"
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_fmaheavy,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,launch__grid_size,launch__block_size,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,dram__throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,smsp__average_warp_latency_issue_stalled_long_scoreboard_per_warp_active.pct,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,launch__occupancy_limit_registers,sm__maximum_warps_per_active_cycle_pct"
cap() {  # name  kernel-regex  skip  count  target-args...
  local name=$1 k=$2 s=$3 c=$4; shift 4
  timeout 900 $NCU -k regex:$k -s $s -c $c -o gpurun_out/${T}_$name -f python tools/ncu_targets.py "$@" > gpurun_out/${T}_ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
  ncu -i gpurun_out/${T}_$name.ncu-rep --page raw --csv > gpurun_out/${T}_$name.raw.csv 2>/dev/null
  rm -f gpurun_out/${T}_$name.ncu-rep
}
cap accumulate_pad0 msm_accumulate 1 1 --what msm20 --table-pad 0 --reps 2
cap accumulate_pad1 msm_accumulate 1 1 --what msm20 --table-pad 1 --reps 2
cap accumulate_g2 msm_accumulate 1 1 --what g2msm --reps 2
cap fixed_base_mul fixed_base_mul 2 2 --what setup --reps 2
cap poly_mul_trunc poly_mul_trunc 20 3 --what quotient --reps 2
cap msm_tail_2p12 msm_tail 2 1 --what msm12 --reps 2
cap msm_tail_2p20 msm_tail 3 1 --what msm20 --reps 2
ls -la gpurun_out/ | tail -20; du -sh gpurun_out
