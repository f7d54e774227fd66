set -u
T=${TAG:-r2common2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_dropin.py tests/test_gpu_tail_and_group.py -x -q --timeout 900 > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/${T}_pytest.log
( timeout ${COMMON_TIMEOUT:-240} oracle/_ref/dropin/benchmark-bls12381-gpu --benchmark-common > gpurun_out/${T}_benchmark_common_bls12381.txt 2> gpurun_out/${T}_benchmark_common_bls12381.err ; echo "rc=$?" >> gpurun_out/${T}_benchmark_common_bls12381.txt ) 
tail -16 gpurun_out/${T}_benchmark_common_bls12381.txt
tail -5 gpurun_out/${T}_benchmark_common_bls12381.err
