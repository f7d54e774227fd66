set -u
T=${TAG:-r2common}
mkdir -p gpurun_out
free -g | head -2
nproc
( timeout 540 oracle/_ref/dropin/benchmark-bls12381-gpu --benchmark-common > gpurun_out/${T}_benchmark_common_bls12381.txt 2> gpurun_out/${T}_benchmark_common_bls12381.err ; echo "rc=$?" >> gpurun_out/${T}_benchmark_common_bls12381.txt ) 
tail -25 gpurun_out/${T}_benchmark_common_bls12381.txt
tail -5 gpurun_out/${T}_benchmark_common_bls12381.err
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
